// K2 -- exact nearest / second-nearest neighbours per query row and per train column, from the screening output.
//
// Restates, on the GPU, what cv::BFMatcher::knnMatch(k=2) returns for each direction (reference call sites
// gtsfm/frontend/matcher/twoway_matcher.py:136,139): the best train index with the LOWEST index among equal float32
// distances, the best distance sqrtf(d^2) and the second-best distance.
//
// Integer path: the screening kernel gives, per query row, the best and second-best GROUP minimum (groups of 16 train
// columns) and, per train column, the minimum of each of the 128 lane classes.  Only the winning group / class is
// recomputed here, exactly, with integer dot products (dp4a) on the uint8 copy of the descriptors.  Whenever the
// runner-up group (class) could tie with the winner in float32-distance space the row (column) is flagged and
// re-scanned exhaustively by exact_scan_kernel (match_exact.cu), so the result never depends on the screening order.
#include <cfloat>

#include "store.h"

namespace g2v {

__device__ __forceinline__ int find_by_out_base(const PairDesc* pairs, int n, long long pos)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].out_base <= pos) lo = mid; else hi = mid - 1;
    }
    return lo;
}
__device__ __forceinline__ int find_by_cb_base(const PairDesc* pairs, int n, long long pos)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].cb_base <= pos) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ void flag(const MatchWork& w, int pair, int dir, int idx)
{
    const int k = atomicAdd(w.n_flag, 1);
    if (k < w.flag_cap)
        w.flag_list[k] = (static_cast<unsigned long long>(pair) << 33) | (static_cast<unsigned long long>(dir) << 32) |
                         static_cast<unsigned int>(idx);
}

// exact integer squared distance between two uint8 rows of kpad bytes (kpad % 16 == 0)
__device__ __forceinline__ int dot_u8(const uint4* __restrict__ a, const uint4* __restrict__ b, int n16)
{
    unsigned acc = 0;
    for (int k = 0; k < n16; ++k) {
        const uint4 x = a[k], y = __ldg(b + k);
        acc = __dp4a(x.x, y.x, acc);
        acc = __dp4a(x.y, y.y, acc);
        acc = __dp4a(x.z, y.z, acc);
        acc = __dp4a(x.w, y.w, acc);
    }
    return static_cast<int>(acc);
}

// float32 distance exactly as cv2 produces it for integer-valued descriptors: the squared distance is an exact
// integer below 2^24 in any summation order, then sqrtf (correctly rounded).
__device__ __forceinline__ float dist_from_d2(int d2) { return sqrtf(static_cast<float>(d2)); }
__device__ __forceinline__ float dist_from_half(float h) { return h >= kHuge * 0.25f ? FLT_MAX : sqrtf(h + h); }

// ---------------------------------------------------------------------------------------------------------------
// one thread per query row
__global__ void __launch_bounds__(128) resolve_rows_int_kernel(StoreView st, MatchWork w, long long total_rows)
{
    const long long pos = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= total_rows) return;
    const int p = find_by_out_base(w.pairs, w.n_pairs, static_cast<long long>(blockIdx.x) * blockDim.x);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0) return;
    const int i = static_cast<int>(pos - pd.out_base);
    if (i >= pd.nq) return;
    if (!st.valid[pd.q_row0 + i]) { w.row_best[pos] = -1; w.row_d1[pos] = FLT_MAX; w.row_d2[pos] = FLT_MAX; return; }

    // merge the (best group minimum, second-best group minimum, best group id) partials of all half blocks
    const int n_hb = pd.nt_pad / kHalfT;
    float k1 = kHuge * 8.0f, k2 = kHuge * 8.0f;
    int hb1 = 0, g1 = 0;
    const long long e0 = pd.rp_base + i;
    for (int hb = 0; hb < n_hb; ++hb) {
        const long long e = e0 + static_cast<long long>(hb) * pd.nq_pad;
        const float a = w.rp_v1[e], b = w.rp_v2[e];
        if (a < k1) { k2 = fminf(k1, b); k1 = a; hb1 = hb; g1 = w.rp_g[e]; }
        else k2 = fminf(k2, a);
    }
    // exact recompute of the winning group
    const int n16 = st.kpad / 16;
    uint4 q[16];
    const uint4* qrow = reinterpret_cast<const uint4*>(st.desc_u8 + static_cast<long long>(pd.q_row0 + i) * st.kpad);
#pragma unroll
    for (int k = 0; k < 16; ++k) if (k < n16) q[k] = __ldg(qrow + k);
    const int nq2 = st.norm_i[pd.q_row0 + i];
    const int j0 = hb1 * kHalfT + g1 * kGroup;
    float f1 = FLT_MAX, f2 = FLT_MAX;
    int j1 = -1, dmin = INT_MAX;
    for (int c = 0; c < kGroup; ++c) {
        const int j = j0 + c;
        if (j >= pd.nt || !st.valid[pd.t_row0 + j]) continue;
        const uint4* trow = reinterpret_cast<const uint4*>(st.desc_u8 + static_cast<long long>(pd.t_row0 + j) * st.kpad);
        unsigned acc = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) if (k < n16) {
            const uint4 y = __ldg(trow + k);
            acc = __dp4a(q[k].x, y.x, acc); acc = __dp4a(q[k].y, y.y, acc);
            acc = __dp4a(q[k].z, y.z, acc); acc = __dp4a(q[k].w, y.w, acc);
        }
        const int d2 = nq2 + st.norm_i[pd.t_row0 + j] - 2 * static_cast<int>(acc);
        dmin = min(dmin, d2);
        const float f = dist_from_d2(d2);
        if (f < f2) { if (f < f1) { f2 = f1; f1 = f; j1 = j; } else f2 = f; }
    }
    const float f2o = dist_from_half(k2);
    // Ambiguous when another group may hold an equal (float32) distance at a lower index, or when the recomputed
    // minimum does not reproduce the screened value (never expected; keeps the result independent of screening).
    const bool ambiguous = (j1 < 0) || (f2o <= f1) || (static_cast<float>(dmin) != k1 + k1);
    if (ambiguous) flag(w, p, 0, i);
    w.row_best[pos] = j1;
    w.row_d1[pos] = f1;
    w.row_d2[pos] = fminf(f2, f2o);
}

// ---------------------------------------------------------------------------------------------------------------
// one warp per train column
__global__ void __launch_bounds__(256) resolve_cols_int_kernel(StoreView st, MatchWork w, long long total_cols)
{
    const int lane = threadIdx.x & 31;
    const long long pos = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
    if (pos >= total_cols) return;
    const int p = find_by_cb_base(w.pairs, w.n_pairs, static_cast<long long>(blockIdx.x) * 8);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0) return;
    const int j = static_cast<int>(pos - pd.cb_base);
    if (j >= pd.nt) return;
    if (!st.valid[pd.t_row0 + j]) {
        if (lane == 0) { w.col_best[pos] = -1; w.col_d1[pos] = FLT_MAX; w.col_d2[pos] = FLT_MAX; }
        return;
    }
    // best and second-best lane class
    const float4 c4 = *reinterpret_cast<const float4*>(w.colraw + pd.cr_base + static_cast<long long>(j) * kTileQ + lane * 4);
    const float cv[4] = {c4.x, c4.y, c4.z, c4.w};
    float k1 = kHuge * 8.0f, k2 = kHuge * 8.0f;
    int c1 = 0;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        if (cv[u] < k1) { k2 = k1; k1 = cv[u]; c1 = lane * 4 + u; } else k2 = fminf(k2, cv[u]);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const float o1 = __shfl_xor_sync(0xffffffffu, k1, off), o2 = __shfl_xor_sync(0xffffffffu, k2, off);
        const int oc = __shfl_xor_sync(0xffffffffu, c1, off);
        const bool take = (o1 < k1) || (o1 == k1 && oc < c1);
        const float lose = take ? k1 : o1;
        k2 = fminf(fminf(k2, o2), lose);
        if (take) { k1 = o1; c1 = oc; }
    }
    // exact recompute of the winning class: query rows c1, c1 + 128, ...
    const int n16 = st.kpad / 16;
    const uint4* trow = reinterpret_cast<const uint4*>(st.desc_u8 + static_cast<long long>(pd.t_row0 + j) * st.kpad);
    const int nt2 = st.norm_i[pd.t_row0 + j];
    float f1 = FLT_MAX, f2 = FLT_MAX;
    int i1 = INT_MAX, dmin = INT_MAX;
    for (int m = c1 + lane * kTileQ; m < pd.nq; m += 32 * kTileQ) {
        if (!st.valid[pd.q_row0 + m]) continue;
        const uint4* qrow = reinterpret_cast<const uint4*>(st.desc_u8 + static_cast<long long>(pd.q_row0 + m) * st.kpad);
        const int d2 = nt2 + st.norm_i[pd.q_row0 + m] - 2 * dot_u8(qrow, trow, n16);
        dmin = min(dmin, d2);
        const float f = dist_from_d2(d2);
        if (f < f2) { if (f < f1) { f2 = f1; f1 = f; i1 = m; } else f2 = f; }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const float o1 = __shfl_xor_sync(0xffffffffu, f1, off), o2 = __shfl_xor_sync(0xffffffffu, f2, off);
        const int oi = __shfl_xor_sync(0xffffffffu, i1, off);
        const int od = __shfl_xor_sync(0xffffffffu, dmin, off);
        dmin = min(dmin, od);
        const bool take = (o1 < f1) || (o1 == f1 && oi < i1);
        const float lose = take ? f1 : o1;
        f2 = fminf(fminf(f2, o2), lose);
        if (take) { f1 = o1; i1 = oi; }
    }
    if (lane == 0) {
        const float f2o = dist_from_half(k2);
        const bool ambiguous = (i1 == INT_MAX) || (f2o <= f1) || (static_cast<float>(dmin) != k1 + k1);
        if (ambiguous) flag(w, p, 1, j);
        w.col_best[pos] = (i1 == INT_MAX) ? -1 : i1;
        w.col_d1[pos] = f1;
        w.col_d2[pos] = fminf(f2, f2o);
    }
}

// which: 0 = query rows, 1 = train columns
cudaError_t launch_resolve_int(const StoreView& st, const MatchWork& w, int which, long long total, cudaStream_t s)
{
    if (total <= 0) return cudaSuccess;
    if (which == 0) resolve_rows_int_kernel<<<static_cast<unsigned>((total + 127) / 128), 128, 0, s>>>(st, w, total);
    else resolve_cols_int_kernel<<<static_cast<unsigned>((total + 7) / 8), 256, 0, s>>>(st, w, total);
    return cudaGetLastError();
}

}  // namespace g2v
