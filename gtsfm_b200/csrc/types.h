// Internal device/host structures shared by the kernels of the two-view front-end path.
#pragma once
#include <cstdint>

namespace g2v {

constexpr int kRowPad = 256;   // every image occupies a multiple of 256 rows of the descriptor store
constexpr int kTileQ = 128;    // query rows per MMA tile (TMEM lanes)
constexpr int kBlockT = 256;   // train rows held stationary per work item (MMA N)
constexpr int kHalfT = 128;    // train columns scanned by one epilogue warpgroup
constexpr int kGroup = 16;     // train columns per "group" whose minimum the screening kernel keeps
constexpr int kMaxClassRows = 512;   // resolve step: rows of one lane class kept in shared memory (nq <= 65536)
constexpr int kAugK = 16;      // one extra K step (16 bf16) that folds (|q|^2 + |t|^2) / 2 into the accumulator
constexpr float kHuge = 1.2676506e30f;   // 2^100: squared norm given to padding / NaN rows so they never win

// One image pair of a chunk.  Row indices are rows of the descriptor store.
struct PairDesc {
    int q_row0, t_row0;      // first store row of image 1 (query side) and image 2 (train side)
    int nq, nt;              // real keypoint counts
    int nq_pad, nt_pad;      // padded counts (multiples of kRowPad)
    int item0;               // first screening work item of this pair (one item per kBlockT train rows)
    int path;                // 0 = integer descriptors (tensor-core screening is exact), 1 = float (exact fp32 scan)
    long long rp_base;       // entry offset into the row-partial arrays: [nt_pad / kHalfT][nq_pad]
    long long cr_base;       // float offset into colraw: [nt_pad][kTileQ]
    long long out_base;      // offset of this pair's slot in the per-query-row arrays and the match slot (capacity nq_pad)
    long long cb_base;       // offset of this pair's slot in the per-train-column arrays (capacity nt_pad)
    int img_q, img_t;        // image slots (for keypoints / calibration)
    int job0;                // first exact-scan job of this pair (float path only: ceil(nq/32) + ceil(nt/32) jobs)
    int pad_;
};

// Per-image record of the store.
struct ImageRec {
    int row0, n, n_pad, dim;  // n = descriptor rows (0 when the image was uploaded with keypoints only)
    int n_kps;               // keypoints uploaded (0 when none)
    int n_valid;             // rows without NaN
    int path;                // 0 integer, 1 float
    double cal[5];           // f, k1, k2, u0, v0
    int has_kps;
};

}  // namespace g2v
