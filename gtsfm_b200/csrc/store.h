// Device-side views handed to the kernels: the descriptor / keypoint store and the per-chunk work buffers.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "types.h"

namespace g2v {

// Descriptor / keypoint store (one row per keypoint; every image starts on a multiple of kRowPad rows).
struct StoreView {
    const __nv_bfloat16* desc_bf16;   // [rows][kpad]    bf16 copy fed to the tensor cores (zero for NaN / padding rows)
    const uint8_t* desc_u8;           // [rows][kpad]    exact integer copy for the resolve kernels (integer path)
    const float* desc_f32;            // [rows][dim]     the caller's float32 descriptors, untouched
    const int* norm_i;                // [rows]          integer squared norm (integer path)
    const float* norm_f;              // [rows]          float32 squared norm (float path)
    const uint8_t* valid;             // [rows]          0 for rows containing NaN (twoway_matcher.py:75-76) and padding
    const double2* kps;               // [rows]          pixel coordinates (Keypoints.coordinates)
    int dim;                          // descriptor dimension as uploaded
    int kpad;                         // dim rounded up to a multiple of 64
};

// Per-chunk buffers of the matcher.
struct MatchWork {
    const PairDesc* pairs;
    int n_pairs;
    // screening output
    const float* rp_v1;
    const float* rp_v2;
    const uint8_t* rp_g;
    const float* colraw;
    // exact per-row / per-column nearest neighbours, indexed by PairDesc::out_base (+row) and cb_base (+column)
    int* row_best;  float* row_d1;  float* row_d2;
    int* col_best;  float* col_d1;  float* col_d2;
    // resolve step (counting sort by bucket): key = bucket | slot in the bucket (rows: half block << 3 | group | slot << 16,
    // columns: lane class | slot << 8); per-pair bucket counters / offsets (rows: 8 max_hb + 1 per pair, columns: 129 per
    // pair); member lists indexed like the per-row / per-column arrays
    unsigned int* row_key;
    unsigned int* col_key;
    int* row_cnt;  int* col_cnt;
    int* row_list; int* col_list;
    const int* row_blk_pair;            // float screening path: pair (chunk-relative) of every block of kRowPad query rows
    const int* col_blk_pair;            //                       ... of every block of kRowPad train columns
    int max_hb;                         // largest number of 128-column half blocks of a pair of the chunk
    // rows / columns whose screening result was ambiguous: (pair << 33 | dir << 32 | index), counted by n_flag
    unsigned long long* col_cells;      // float path: 16 B per train column (0xff-filled): merge cell of exact_tile_kernel
    unsigned long long* flag_scratch;   // 16 B per flagged entry (0xff-filled): merge cell of the segmented exact scan
    unsigned long long* flag_list;
    int* n_flag;
    int flag_cap;
};

}  // namespace g2v
