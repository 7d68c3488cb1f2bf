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
//
// Two kernels per direction:
//   merge   : reduces the screening partials of a row (column) to (k1, k2, bucket key); coalesced, HBM/L2 streaming.
//   bucket  : one CTA per bucket -- (pair, 128-column half block) for rows, (pair, lane class) for columns.  The CTA
//             stages the <= 128 (class: nq/128) descriptor rows every member of the bucket is compared against in
//             shared memory ONCE, collects the members by scanning the key array, and recomputes the candidates from
//             shared memory.  Per pair this reads each descriptor row about once (2 MB) instead of gathering the same
//             rows once per member through L2 (80 MB for 8192 x 8192).
#include <cfloat>

#include "store.h"

namespace g2v {

constexpr int kListCap = 4096;

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
__device__ __forceinline__ int find_by_item0(const PairDesc* pairs, int n, int item)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].item0 <= item) lo = mid; else hi = mid - 1;
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

// float32 distance exactly as cv2 produces it for integer-valued descriptors: the squared distance is an exact
// integer below 2^24 in any summation order, then sqrtf (correctly rounded).
__device__ __forceinline__ float dist_from_d2(int d2) { return sqrtf(static_cast<float>(d2)); }
__device__ __forceinline__ float dist_from_half(float h) { return h >= kHuge * 0.25f ? FLT_MAX : sqrtf(h + h); }

// ---------------------------------------------------------------------------------------------------------------
// merge, rows: one thread per query row.  row_d1 <- k1, row_d2 <- k2 (halved squared distances), row_key <- bucket.
__global__ void __launch_bounds__(128) merge_rows_kernel(StoreView st, MatchWork w, long long total_rows)
{
    const long long pos = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= total_rows) return;
    const int p = find_by_out_base(w.pairs, w.n_pairs, static_cast<long long>(blockIdx.x) * blockDim.x);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0) return;
    const int i = static_cast<int>(pos - pd.out_base);
    if (i >= pd.nq || !st.valid[pd.q_row0 + i]) {
        w.row_key[pos] = 0xffffu;
        if (i < pd.nq) { w.row_best[pos] = -1; w.row_d1[pos] = FLT_MAX; w.row_d2[pos] = FLT_MAX; }
        return;
    }
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
    w.row_d1[pos] = k1;
    w.row_d2[pos] = k2;
    w.row_key[pos] = static_cast<unsigned short>((hb1 << 3) | g1);
}

// merge, columns: one warp per train column, kColsPerCta columns per CTA (every CTA lies inside one pair because column
// ranges start on multiples of 256, so the pair lookup is done once per CTA).  col_d1 <- k1, col_d2 <- k2, col_key <-
// winning lane class.
constexpr int kColsPerCta = 64;
__global__ void __launch_bounds__(256) merge_cols_kernel(StoreView st, MatchWork w, long long total_cols)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long pos0 = static_cast<long long>(blockIdx.x) * kColsPerCta;
    if (pos0 >= total_cols) return;
    const int p = find_by_cb_base(w.pairs, w.n_pairs, pos0);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0) return;
#pragma unroll 2
    for (int u0 = warp; u0 < kColsPerCta; u0 += 8) {
        const long long pos = pos0 + u0;
        const int j = static_cast<int>(pos - pd.cb_base);
        if (j >= pd.nt || !st.valid[pd.t_row0 + j]) {
            if (lane == 0) {
                w.col_key[pos] = 0xffu;
                if (j < pd.nt) { w.col_best[pos] = -1; w.col_d1[pos] = FLT_MAX; w.col_d2[pos] = FLT_MAX; }
            }
            continue;
        }
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
        if (lane == 0) {
            w.col_d1[pos] = k1;
            w.col_d2[pos] = k2;
            w.col_key[pos] = static_cast<uint8_t>(c1);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
template <int N16> struct RowRegs { uint4 v[N16]; };

template <int N16>
__device__ __forceinline__ unsigned dot_regs_smem(const RowRegs<N16>& a, const uint4* __restrict__ b)
{
    unsigned acc = 0;
#pragma unroll
    for (int k = 0; k < N16; ++k) {
        const uint4 y = b[k];
        acc = __dp4a(a.v[k].x, y.x, acc); acc = __dp4a(a.v[k].y, y.y, acc);
        acc = __dp4a(a.v[k].z, y.z, acc); acc = __dp4a(a.v[k].w, y.w, acc);
    }
    return acc;
}

// bucket, rows: one CTA per (pair, half block of 128 train columns).  Staged rows are padded by 16 B per row and a
// further 16 B per group of 16, so threads working on different groups of the half block hit different banks.
template <int N16>
__global__ void __launch_bounds__(128) bucket_rows_kernel(StoreView st, MatchWork w, int n_half)
{
    constexpr int kStride = N16 + 1;
    __shared__ uint4 s_t[kHalfT * kStride + kHalfT / kGroup];
    __shared__ int s_norm[kHalfT];
    __shared__ uint8_t s_valid[kHalfT];
    __shared__ int s_list[kListCap];
    __shared__ int s_n;

    const int tid = threadIdx.x;
    const int p = find_by_item0(w.pairs, w.n_pairs, blockIdx.x >> 1);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0 || pd.nq == 0) return;
    const int hb = static_cast<int>(blockIdx.x) - 2 * pd.item0;
    if (hb * kHalfT >= pd.nt) return;                              // only padding columns: never a winner
    const long long trow0 = static_cast<long long>(pd.t_row0) + hb * kHalfT;
    const uint4* tsrc = reinterpret_cast<const uint4*>(st.desc_u8) + trow0 * N16;
    for (int idx = tid; idx < kHalfT * N16; idx += 128) {
        const int r = idx / N16, k = idx % N16;
        s_t[r * kStride + (r >> 4) + k] = __ldg(tsrc + idx);
    }
    s_norm[tid] = st.norm_i[trow0 + tid];
    s_valid[tid] = st.valid[trow0 + tid];
    if (tid == 0) s_n = 0;
    __syncthreads();

    const unsigned short* keys = w.row_key + pd.out_base;
    for (int base = 0; base < pd.nq_pad; base += 1024) {
        const int i0 = base + tid * 8;
        if (i0 < pd.nq_pad) {
            const uint4 kk = *reinterpret_cast<const uint4*>(keys + i0);
            const unsigned kw[4] = {kk.x, kk.y, kk.z, kk.w};
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const unsigned key = (kw[u >> 1] >> ((u & 1) * 16)) & 0xffffu;
                if (key != 0xffffu && static_cast<int>(key >> 3) == hb) s_list[atomicAdd(&s_n, 1)] = ((i0 + u) << 3) | (key & 7);
            }
        }
        __syncthreads();
        const int n = s_n;
        const bool last = base + 1024 >= pd.nq_pad;
        if (n > kListCap - 1024 || last) {
            for (int idx = tid; idx < n; idx += 128) {
                const int e = s_list[idx];
                const int i = e >> 3, g = e & 7;
                const long long pos = pd.out_base + i;
                RowRegs<N16> q;
                const uint4* qrow = reinterpret_cast<const uint4*>(st.desc_u8) + static_cast<long long>(pd.q_row0 + i) * N16;
#pragma unroll
                for (int k = 0; k < N16; ++k) q.v[k] = __ldg(qrow + k);
                const int nq2 = st.norm_i[pd.q_row0 + i];
                const float k1 = w.row_d1[pos], k2 = w.row_d2[pos];
                float f1 = FLT_MAX, f2 = FLT_MAX;
                int j1 = -1, dmin = INT_MAX;
#pragma unroll 4
                for (int c = 0; c < kGroup; ++c) {
                    const int r = g * kGroup + c;
                    const int j = hb * kHalfT + r;
                    if (j >= pd.nt || !s_valid[r]) continue;
                    const int d2 = nq2 + s_norm[r] - 2 * static_cast<int>(dot_regs_smem<N16>(q, s_t + r * kStride + g));
                    dmin = min(dmin, d2);
                    const float f = dist_from_d2(d2);
                    if (f < f2) { if (f < f1) { f2 = f1; f1 = f; j1 = j; } else f2 = f; }
                }
                const float f2o = dist_from_half(k2);
                // Ambiguous when another group may hold an equal (float32) distance at a lower index, or when the
                // recomputed minimum does not reproduce the screened value (never expected; keeps the result
                // independent of screening).
                const bool ambiguous = (j1 < 0) || (f2o <= f1) || (static_cast<float>(dmin) != k1 + k1);
                if (ambiguous) flag(w, p, 0, i);
                w.row_best[pos] = j1;
                w.row_d1[pos] = f1;
                w.row_d2[pos] = fminf(f2, f2o);
            }
            __syncthreads();
            if (tid == 0) s_n = 0;
            __syncthreads();
        }
    }
}

// bucket, columns: one CTA per (pair, lane class c).  The class rows c, c + 128, ... of the query image are staged in
// shared memory; two threads share a column (even / odd class rows).
template <int N16>
__global__ void __launch_bounds__(128) bucket_cols_kernel(StoreView st, MatchWork w)
{
    constexpr int kStride = N16 + 1;
    extern __shared__ uint4 s_dyn[];
    __shared__ int s_list[kListCap];
    __shared__ int s_n;

    const int tid = threadIdx.x;
    const int p = blockIdx.x >> 7, c = blockIdx.x & 127;
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0 || pd.nq == 0 || pd.nt == 0 || c >= pd.nq) return;
    const int n_cls = pd.nq_pad / kTileQ;
    uint4* s_q = s_dyn;
    int* s_norm = reinterpret_cast<int*>(s_dyn + n_cls * kStride);
    uint8_t* s_valid = reinterpret_cast<uint8_t*>(s_norm + n_cls);
    const uint4* qsrc = reinterpret_cast<const uint4*>(st.desc_u8) + (static_cast<long long>(pd.q_row0) + c) * N16;
    for (int idx = tid; idx < n_cls * N16; idx += 128) {
        const int t = idx / N16, k = idx % N16;
        s_q[t * kStride + k] = __ldg(qsrc + static_cast<long long>(t) * kTileQ * N16 + k);
    }
    for (int t = tid; t < n_cls; t += 128) {
        const long long row = static_cast<long long>(pd.q_row0) + c + t * kTileQ;
        s_norm[t] = st.norm_i[row];
        s_valid[t] = (c + t * kTileQ < pd.nq) ? st.valid[row] : 0;
    }
    if (tid == 0) s_n = 0;
    __syncthreads();

    const uint8_t* keys = w.col_key + pd.cb_base;
    const int par = tid & 1;
    for (int base = 0; base < pd.nt_pad; base += 2048) {
        const int j0 = base + tid * 16;
        if (j0 < pd.nt_pad) {
            const uint4 kk = *reinterpret_cast<const uint4*>(keys + j0);
            const unsigned kw[4] = {kk.x, kk.y, kk.z, kk.w};
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const unsigned key = (kw[u >> 2] >> ((u & 3) * 8)) & 0xffu;
                if (static_cast<int>(key) == c) s_list[atomicAdd(&s_n, 1)] = j0 + u;
            }
        }
        __syncthreads();
        const int n = s_n;
        const bool last = base + 2048 >= pd.nt_pad;
        if (n > kListCap - 2048 || last) {
            const int n_round = (n + 63) & ~63;
            for (int idx = tid >> 1; idx < n_round; idx += 64) {
                const bool active = idx < n;
                const int j = active ? s_list[idx] : 0;
                const long long pos = pd.cb_base + j;
                RowRegs<N16> tr;
                const uint4* trow = reinterpret_cast<const uint4*>(st.desc_u8) + static_cast<long long>(pd.t_row0 + j) * N16;
#pragma unroll
                for (int k = 0; k < N16; ++k) tr.v[k] = __ldg(trow + k);
                const int nt2 = st.norm_i[pd.t_row0 + j];
                float f1 = FLT_MAX, f2 = FLT_MAX;
                int i1 = INT_MAX, dmin = INT_MAX;
                if (active) {
                    for (int t = par; t < n_cls; t += 2) {
                        if (!s_valid[t]) continue;
                        const int d2 = nt2 + s_norm[t] - 2 * static_cast<int>(dot_regs_smem<N16>(tr, s_q + t * kStride));
                        dmin = min(dmin, d2);
                        const float f = dist_from_d2(d2);
                        if (f < f2) { if (f < f1) { f2 = f1; f1 = f; i1 = c + t * kTileQ; } else f2 = f; }
                    }
                }
                {
                    const float o1 = __shfl_xor_sync(0xffffffffu, f1, 1), o2 = __shfl_xor_sync(0xffffffffu, f2, 1);
                    const int oi = __shfl_xor_sync(0xffffffffu, i1, 1);
                    const int od = __shfl_xor_sync(0xffffffffu, dmin, 1);
                    dmin = min(dmin, od);
                    const bool take = (o1 < f1) || (o1 == f1 && oi < i1);
                    const float lose = take ? f1 : o1;
                    f2 = fminf(fminf(f2, o2), lose);
                    if (take) { f1 = o1; i1 = oi; }
                }
                if (active && par == 0) {
                    const float k1 = w.col_d1[pos], k2 = w.col_d2[pos];
                    const float f2o = dist_from_half(k2);
                    const bool ambiguous = (i1 == INT_MAX) || (f2o <= f1) || (static_cast<float>(dmin) != k1 + k1);
                    if (ambiguous) flag(w, p, 1, j);
                    w.col_best[pos] = (i1 == INT_MAX) ? -1 : i1;
                    w.col_d1[pos] = f1;
                    w.col_d2[pos] = fminf(f2, f2o);
                }
            }
            __syncthreads();
            if (tid == 0) s_n = 0;
            __syncthreads();
        }
    }
}

template <int N16> static cudaError_t launch_bucket_cols(const StoreView& st, const MatchWork& w, int max_class_rows, cudaStream_t s)
{
    static int configured = 0;
    const int smem = max_class_rows * ((N16 + 1) * 16 + 5) + 16;
    if (smem > configured) {
        const int want = kMaxClassRows * ((N16 + 1) * 16 + 5) + 16;
        cudaError_t e = cudaFuncSetAttribute(bucket_cols_kernel<N16>, cudaFuncAttributeMaxDynamicSharedMemorySize, want);
        if (e != cudaSuccess) return e;
        configured = want;
    }
    bucket_cols_kernel<N16><<<static_cast<unsigned>(w.n_pairs) * kTileQ, 128, smem, s>>>(st, w);
    return cudaGetLastError();
}

// which: 0 = query rows, 1 = train columns.  n_items = 256-column screening items of the chunk.
cudaError_t launch_resolve_int(const StoreView& st, const MatchWork& w, int which, long long total, int n_items, int max_class_rows,
                               cudaStream_t s)
{
    if (total <= 0 || n_items <= 0) return cudaSuccess;
    if (st.kpad != 64 && st.kpad != 128) return cudaErrorInvalidValue;
    if (which == 0) {
        merge_rows_kernel<<<static_cast<unsigned>((total + 127) / 128), 128, 0, s>>>(st, w, total);
        if (st.kpad == 64) bucket_rows_kernel<4><<<2 * n_items, 128, 0, s>>>(st, w, 2 * n_items);
        else bucket_rows_kernel<8><<<2 * n_items, 128, 0, s>>>(st, w, 2 * n_items);
    } else {
        merge_cols_kernel<<<static_cast<unsigned>((total + kColsPerCta - 1) / kColsPerCta), 256, 0, s>>>(st, w, total);
        cudaError_t e = st.kpad == 64 ? launch_bucket_cols<4>(st, w, max_class_rows, s) : launch_bucket_cols<8>(st, w, max_class_rows, s);
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

}  // namespace g2v
