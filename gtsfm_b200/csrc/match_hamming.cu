// HAMMING distance matching of binary descriptors (ORB / BRISK / AKAZE-like uint8 rows):
// TwoWayMatcher(distance_type=HAMMING) -> cv2.BFMatcher(NORM_HAMMING).knnMatch / match
// (gtsfm/frontend/matcher/twoway_matcher.py:17-21,95-98,136,139).  The distance is the number of differing bits, an exact
// integer, reported by cv2 as float32; the two nearest train rows per query row with cv2's strict '<' insertion (lowest train
// index on ties).  Byte/integer work with no tensor-core path on sm_100: XOR + POPC over 32-bit words, the other side staged
// in shared memory tile by tile (broadcast reads), one scan-side row per thread held in registers.  The ratio test, the
// mutual check and the ordering are the same kernels as for the Euclidean case (match_finalize.cu).
#include <cfloat>

#include "store.h"

namespace g2v {

constexpr int kHamTile = 128;      // other-side rows per shared-memory tile

// DIR 0: scan side = query rows of image 1 against the rows of image 2; DIR 1: the reverse direction.
// One CTA per block of kRowPad (256) scan rows: blockIdx -> pair through the host-built block table.
template <int DIR, int W>
__global__ void __launch_bounds__(kRowPad) hamming_kernel(StoreView st, MatchWork w)
{
    __shared__ uint4 s_tile[kHamTile * (W / 4)];
    const long long base = static_cast<long long>(blockIdx.x) * kRowPad;
    const int p = (DIR == 0 ? w.row_blk_pair : w.col_blk_pair)[blockIdx.x];
    const PairDesc pd = w.pairs[p];
    if (pd.path != 3) return;
    const int i0 = static_cast<int>(base - (DIR == 0 ? pd.out_base : pd.cb_base));
    const int n_scan = DIR == 0 ? pd.nq : pd.nt, other_n = DIR == 0 ? pd.nt : pd.nq;
    if (i0 >= n_scan) return;
    const int scan_row0 = (DIR == 0 ? pd.q_row0 : pd.t_row0) + i0;
    const int other_row0 = DIR == 0 ? pd.t_row0 : pd.q_row0;
    const int kw = st.kpad / 4;                           // 32-bit words per stored row (zero padded)
    const int r = threadIdx.x;
    const bool active = i0 + r < n_scan;
    unsigned q[W];
    {
        const unsigned* src = reinterpret_cast<const unsigned*>(st.desc_u8) + static_cast<long long>(scan_row0 + (active ? r : 0)) * kw;
#pragma unroll
        for (int k = 0; k < W; k += 4) {
            const uint4 v = *reinterpret_cast<const uint4*>(src + k);
            q[k] = v.x; q[k + 1] = v.y; q[k + 2] = v.z; q[k + 3] = v.w;
        }
    }
    int b1 = INT_MAX, b2 = INT_MAX, i1 = -1;
    for (int t0 = 0; t0 < other_n; t0 += kHamTile) {
        const int tn = min(kHamTile, other_n - t0);
        __syncthreads();
        for (int k = threadIdx.x; k < tn * (W / 4); k += blockDim.x) {
            const int row = k / (W / 4), c = k - row * (W / 4);
            s_tile[k] = *reinterpret_cast<const uint4*>(reinterpret_cast<const unsigned*>(st.desc_u8) +
                                                        static_cast<long long>(other_row0 + t0 + row) * kw + 4 * c);
        }
        __syncthreads();
        for (int j = 0; j < tn; ++j) {
            int d = 0;
#pragma unroll
            for (int k = 0; k < W / 4; ++k) {
                const uint4 v = s_tile[j * (W / 4) + k];          // same address for the whole warp: broadcast
                d += __popc(q[4 * k] ^ v.x) + __popc(q[4 * k + 1] ^ v.y) + __popc(q[4 * k + 2] ^ v.z) + __popc(q[4 * k + 3] ^ v.w);
            }
            // cv2's top-2 insertion: strict '<', so the lowest train index wins ties
            if (d < b2) {
                if (d < b1) { b2 = b1; b1 = d; i1 = t0 + j; } else b2 = d;
            }
        }
    }
    if (active) {
        const long long o = (DIR == 0 ? pd.out_base : pd.cb_base) + i0 + r;
        int* o_best = DIR == 0 ? w.row_best : w.col_best;
        float* o_d1 = DIR == 0 ? w.row_d1 : w.col_d1;
        float* o_d2 = DIR == 0 ? w.row_d2 : w.col_d2;
        o_best[o] = i1;
        o_d1[o] = i1 >= 0 ? static_cast<float>(b1) : FLT_MAX;
        o_d2[o] = b2 != INT_MAX ? static_cast<float>(b2) : FLT_MAX;
    }
}

// words: 32-bit words that hold a descriptor (8 for <= 32 bytes, 16 for <= 64 bytes)
cudaError_t launch_hamming(const StoreView& st, const MatchWork& w, long long total_rows, long long total_cols, int words, cudaStream_t s)
{
    if (words != 8 && words != 16) return cudaErrorInvalidValue;
    if (total_rows > 0) {
        if (words == 8) hamming_kernel<0, 8><<<static_cast<unsigned>(total_rows / kRowPad), kRowPad, 0, s>>>(st, w);
        else hamming_kernel<0, 16><<<static_cast<unsigned>(total_rows / kRowPad), kRowPad, 0, s>>>(st, w);
    }
    if (total_cols > 0) {
        if (words == 8) hamming_kernel<1, 8><<<static_cast<unsigned>(total_cols / kRowPad), kRowPad, 0, s>>>(st, w);
        else hamming_kernel<1, 16><<<static_cast<unsigned>(total_cols / kRowPad), kRowPad, 0, s>>>(st, w);
    }
    return cudaGetLastError();
}

}  // namespace g2v
