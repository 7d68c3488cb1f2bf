// Internal device/host structures shared by the kernels of the two-view front-end path.
#pragma once
#include <cstdint>

#include <cuda_runtime.h>

namespace g2v {

constexpr int kRowPad = 256;   // every image occupies a multiple of 256 rows of the descriptor store
constexpr int kTileQ = 128;    // query rows per MMA tile (TMEM lanes)
constexpr int kBlockT = 256;   // train rows held stationary per work item (MMA N)
constexpr int kHalfT = 128;    // train columns scanned by one epilogue warpgroup
constexpr int kGroup = 16;     // train columns per "group" whose minimum the screening kernel keeps
constexpr int kMaxClassRows = 512;   // resolve step: rows of one lane class kept in shared memory (nq <= 65536)
constexpr int kAugK = 16;      // one extra K step (16 bf16) that folds (|q|^2 + |t|^2) / 2 into the accumulator
constexpr float kHuge = 1.2676506e30f;   // 2^100: squared norm given to padding / NaN rows so they never win
// Float path (fp16 screening): rows are eligible when |x|^2 <= kF16MaxNorm2, so every d^2 / 2 of an eligible pair stays
// below 2 kF16MaxNorm2 = 2048 < kF16Never / 2, the half squared norm given to padding / NaN / ineligible rows
constexpr float kF16MaxNorm2 = 1024.0f;
constexpr float kF16Never = 32768.0f;
constexpr int kGroupF = 4;               // float path: train columns per group of the row partials (32 groups per half block)
constexpr int kFlushTiles = 4;           // float path: query tiles per column-minimum flush (a lane class holds <= 4 rows)
constexpr int kImgFlagInts = 4;          // ints per image in the flag array written by the prepare kernel

// One image pair of a chunk.  Row indices are rows of the descriptor store.
struct PairDesc {
    int q_row0, t_row0;      // first store row of image 1 (query side) and image 2 (train side)
    int nq, nt;              // real keypoint counts
    int nq_pad, nt_pad;      // padded counts (multiples of kRowPad)
    int item0;               // first screening work item of this pair (one item per kBlockT train rows)
    int path;                // 0 = integer descriptors (tensor-core screening is exact), 1 = float (exact fp32 scan),
                             // 2 = float, fp16 tensor-core screening + exact recompute of the winning group (ratio test off)
                             // 3 = binary descriptors, HAMMING distance (match_hamming.cu)
    long long rp_base;       // entry offset into the row-partial arrays: [nt_pad / kHalfT][nq_pad]
    long long cr_base;       // float offset into colraw: path 0 [{v1, v2, class, pad}][nt_pad]; path 2: [nt_pad][ceil(nq_pad / 512)][kTileQ]
    long long out_base;      // offset of this pair's slot in the per-query-row arrays and the match slot (capacity nq_pad)
    long long cb_base;       // offset of this pair's slot in the per-train-column arrays (capacity nt_pad)
    int img_q, img_t;        // image slots (for keypoints / calibration)
    int job0;                // first exact-scan job of this pair (float path only: ceil(nq/32) + ceil(nt/32) jobs)
    int fitem0;              // path 2: first work item of the fp16 screening kernel (one item per kBlockT train rows)
    float q_maxn2, t_maxn2;  // path 2: max |x|^2 over the valid rows of image 1 / image 2 (error bound of the screening)
    int pad2_;
};

// Per-image record of the store.
struct ImageRec {
    int row0, n, n_pad, dim;  // n = descriptor rows (0 when the image was uploaded with keypoints only)
    int n_kps;               // keypoints uploaded (0 when none)
    int n_valid;             // rows without NaN
    int path;                // 0 integer, 1 float, 3 binary (uint8 rows uploaded with gtsfm_image_upload_u8, HAMMING distance)
    int f16_ok;              // float image whose rows all qualify for the fp16 screening (|x|^2 <= kF16MaxNorm2)
    float max_norm2;         // max |x|^2 over the valid rows
    double cal[5];           // f, k1, k2, u0, v0
    int has_kps;
};

}  // namespace g2v

namespace g2v {
// "This kernel's attributes are set" flag of the CURRENT device: cudaFuncSetAttribute applies to one device, and a process
// may hold contexts on several devices and call in from several threads (a repeated set is harmless).
struct DeviceOnce {
    int v[64] = {0};
    int& cur() { int dev = 0; cudaGetDevice(&dev); return v[dev & 63]; }
};
}  // namespace g2v
