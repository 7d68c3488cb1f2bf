// Structures and the counter-based 5-point sampler shared by the verifier's host and device code.
#pragma once
#include <cstdint>

namespace g2v {

struct VPair {
    long long m0;            // first match row of the pair in the batch-wide match arrays
    int M;                   // putative matches
    int row1, row2;          // first store rows (keypoints) of image 1 / image 2
    int status;              // GTSFM_V_TOO_FEW is decided on the host (M < 6, opencv_verifier_base.py:69-78)
    double cal1[5], cal2[5]; // f, k1, k2, u0, v0
    double thr2;             // (threshold_px / max(fx1, fx2))^2   (ransac.py:79, opencv_verifier_base.py:86)
    unsigned long long key;  // sampler key
};

// Device buffers of one verifier chunk (all pointers are device memory).
struct VerifyBufs {
    const VPair* pairs;        // [P] (this chunk)
    const void* match;         // batch-wide (i1, i2) rows, uint2
    void* corr_d;              // [total] x1 y1 x2 y2 (normalised), double4
    void* corr_f;              // [total] float4
    float* E32;                // [P][H][10][9] essential matrices, slot r = r-th real root (valid slots only)
    unsigned short* vmask;     // [P][H] bit r set: root r of the hypothesis gave a valid essential matrix (slot r of E32)
    float* cost;               // [P][H][10]
    const int* samples;        // optional explicit [P][H][5] (chunk-relative) or nullptr
    // outputs (batch-wide pointers already offset to this chunk's first pair)
    double* R; double* t; double* E; int* ninl; int* status; unsigned char* mask;
    unsigned long long* stats; // running totals: [0] candidates (real roots), [1] hypotheses solved, [2] Sampson evaluations
                               // (algorithmic: candidates x correspondences), [3] evaluations executed after early exit
    int* topi;                 // [P][8] (this chunk) candidate ids (h * 10 + root) of the lowest fp32 costs, INT_MAX when absent
    double* Ek;                // [P][8][9] (this chunk) their essential matrices re-solved in fp64 (NaN when absent)
    unsigned int* best_cost;   // [P] (this chunk) bit pattern of the lowest complete fp32 MSAC cost seen so far
    // ---- fundamental-matrix mode (verify_fmat.cu); corr_f then holds the float32 PIXEL coordinates cv2 works on
    double* Fd;                // [P][H][3][9] fundamental matrices of the 7-point solves, slot r = r-th real root
    float* Ff;                 // [P][H][3][12] fp32 copy of each matrix (9), its rounding-error band K (1), padding (2)
    unsigned char* nroot;      // [P][H] number of matrices of the hypothesis (0..3)
    int* cnt;                  // [P][H][3] inlier count of every candidate, -1 when it left the race early
    float* cmax;               // [P][4] max |x1|, |y1|, |x2|, |y2| over the pair's correspondences (error bound of the fp32 test)
    int* best_cnt;             // [P] highest complete inlier count seen so far
};

struct VerifyDims {
    int H;                   // hypotheses per pair
    int lo_iters;
    unsigned long long seed;
    double thr_px;           // fundamental-matrix mode: threshold in pixels (ransac.py:103-109)
};

#if defined(__CUDACC__)
#define G2V_HD __host__ __device__
#else
#define G2V_HD
#endif

G2V_HD inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// Hypothesis h of pair `key`: K distinct match rows in [0, M), M >= K (5 for the essential-matrix mode, 7 for the
// fundamental-matrix mode).  Pure integer arithmetic, so the host copy (gtsfm_ransac_samples / _k) and the device copy
// agree bit for bit.
template <int K>
G2V_HD inline void sample_k(uint64_t seed, uint64_t key, uint32_t h, uint32_t M, int* out)
{
    const uint64_t state = splitmix64(splitmix64(seed) ^ splitmix64(key * 0xD1342543DE82EF95ull + h));
    int k = 0;
    uint64_t ctr = 0;
    while (k < K) {
        const uint64_t r = splitmix64(state + ctr++);
        const int idx = static_cast<int>(((r >> 32) * static_cast<uint64_t>(M)) >> 32);
        bool dup = false;
        for (int j = 0; j < k; ++j) dup |= (out[j] == idx);
        if (!dup) out[k++] = idx;
    }
}
G2V_HD inline void sample5(uint64_t seed, uint64_t key, uint32_t h, uint32_t M, int* out) { sample_k<5>(seed, key, h, M, out); }

}  // namespace g2v
