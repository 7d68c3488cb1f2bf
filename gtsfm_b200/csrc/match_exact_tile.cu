// Float path of the matcher: exact fp32 distances in OpenCV's own summation order, every distance computed ONCE and
// used for both directions (cv2.BFMatcher.knnMatch(desc1, desc2) and knnMatch(desc2, desc1), twoway_matcher.py:136,139).
//
// cv::batchDistance -> normL2Sqr_(float): blocks of 16 elements feed sixteen partial sums (multiply and add rounded
// separately), combined ((a0+a1)+a2)+a3 lane-wise, then (s0+s2)+(s1+s3), then sqrtf -- see match_exact.cu, where a
// half warp holds the sixteen partial sums of one distance.  Here ONE THREAD holds them (16 registers per distance) and
// computes a 2 x 2 block of distances, so the kernel is plain register-tiled SIMT arithmetic without shuffles:
// 3 FP32 instructions per element pair, operands from shared memory, ~5 x the throughput of the half-warp scan.
//
// CTA = strip of 32 query rows (stationary in shared memory) x all train rows (streamed in tiles of 32).
//   rows   : the CTA sees every train row, so the row top-2 is final when the CTA ends (no atomics);
//   columns: each CTA contributes the top-2 of its 32 query rows; CTAs meet in one 16-byte cell per column
//            (64-bit atomicMin on (distance bits, query index) + atomicMin on the best losing distance), resolved by
//            finish_cols_kernel.  Lowest index wins among equal distances in both directions, as in cv2.
#include <cfloat>

#include "store.h"

namespace g2v {

constexpr int kTQ = 32;
constexpr int kTT = 32;

__device__ __forceinline__ int tile_find_by_job0(const PairDesc* pairs, int n, int job)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].job0 <= job) lo = mid; else hi = mid - 1;
    }
    return lo;
}
__device__ __forceinline__ int tile_find_by_cb(const PairDesc* pairs, int n, long long pos)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].cb_base <= pos) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ void load_rows(float* dst, const float* src, int rows_valid, int dim, int stride)
{
    const int d4 = dim >> 2;
    for (int k = threadIdx.x; k < kTQ * d4; k += blockDim.x) {
        const int r = k / d4, c = (k - r * d4) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < rows_valid) v = *reinterpret_cast<const float4*>(src + static_cast<long long>(r) * dim + c);
        *reinterpret_cast<float4*>(dst + r * stride + c) = v;
    }
}

__global__ void __launch_bounds__(256, 2) exact_tile_kernel(StoreView st, MatchWork w, int n_jobs)
{
    extern __shared__ float sm[];
    __shared__ unsigned long long s_ckey[kTT];
    __shared__ unsigned int s_cm2[kTT];
    __shared__ unsigned char s_tvalid[kTT], s_qvalid[kTQ];

    const int job = blockIdx.x;
    if (job >= n_jobs) return;
    const int p = tile_find_by_job0(w.pairs, w.n_pairs, job);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 1) return;
    const int local = job - pd.job0;
    if (local >= (pd.nq + kTQ - 1) / kTQ) return;         // the remaining jobs of the pair belonged to the column scan
    const int dim = st.dim, stride = dim + 4, nblk = dim / 16;
    float* sq = sm;
    float* stt = sm + kTQ * stride;
    const int q0 = local * kTQ;
    const int nqr = min(kTQ, pd.nq - q0);
    load_rows(sq, st.desc_f32 + static_cast<long long>(pd.q_row0 + q0) * dim, nqr, dim, stride);
    if (threadIdx.x < kTQ) s_qvalid[threadIdx.x] = threadIdx.x < nqr ? st.valid[pd.q_row0 + q0 + threadIdx.x] : 0;

    const int tt = threadIdx.x & 15, tq = threadIdx.x >> 4;
    const float* qa = sq + (2 * tq) * stride;
    const float* qb = qa + stride;
    const float* ta = stt + tt * stride;                  // train rows tt and tt + 16: a warp's sixteen 16-byte loads then
    const float* tb = ta + 16 * stride;                   // fall on distinct banks in two wavefronts (row stride = dim + 4)
    float f1[2] = {FLT_MAX, FLT_MAX}, f2[2] = {FLT_MAX, FLT_MAX};
    int i1[2] = {INT_MAX, INT_MAX};

    for (int t0 = 0; t0 < pd.nt; t0 += kTT) {
        __syncthreads();
        const int ntr = min(kTT, pd.nt - t0);
        load_rows(stt, st.desc_f32 + static_cast<long long>(pd.t_row0 + t0) * dim, ntr, dim, stride);
        if (threadIdx.x < kTT) {
            s_tvalid[threadIdx.x] = threadIdx.x < ntr ? st.valid[pd.t_row0 + t0 + threadIdx.x] : 0;
            s_ckey[threadIdx.x] = ~0ull;
            s_cm2[threadIdx.x] = ~0u;
        }
        __syncthreads();
        // Packed fp32x2 subtract and multiply (FADD2 / FMUL2) on adjacent element pairs, scalar adds: ptxas contracts a
        // packed multiply followed by a packed add into FFMA2 even with explicit .rn, which would break cv2's
        // separately rounded multiply and add.
        float acc[2][2][16];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b)
#pragma unroll
                for (int l = 0; l < 16; ++l) acc[a][b][l] = 0.f;
        for (int blk = 0; blk < nblk; ++blk) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int o = 16 * blk + 8 * h;
                const ulonglong2 x0 = *reinterpret_cast<const ulonglong2*>(qa + o), x1 = *reinterpret_cast<const ulonglong2*>(qa + o + 4);
                const ulonglong2 y0 = *reinterpret_cast<const ulonglong2*>(qb + o), y1 = *reinterpret_cast<const ulonglong2*>(qb + o + 4);
                const ulonglong2 u0 = *reinterpret_cast<const ulonglong2*>(ta + o), u1 = *reinterpret_cast<const ulonglong2*>(ta + o + 4);
                const ulonglong2 v0 = *reinterpret_cast<const ulonglong2*>(tb + o), v1 = *reinterpret_cast<const ulonglong2*>(tb + o + 4);
                const unsigned long long q[2][4] = {{x0.x, x0.y, x1.x, x1.y}, {y0.x, y0.y, y1.x, y1.y}};
                const unsigned long long t[2][4] = {{u0.x, u0.y, u1.x, u1.y}, {v0.x, v0.y, v1.x, v1.y}};
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int b = 0; b < 2; ++b)
#pragma unroll
                        for (int l = 0; l < 4; ++l) {
                            unsigned long long d, m;
                            float m0, m1;
                            asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(q[a][l]), "l"(t[b][l]));
                            asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(m) : "l"(d), "l"(d));
                            asm("mov.b64 {%0, %1}, %2;" : "=f"(m0), "=f"(m1) : "l"(m));
                            acc[a][b][8 * h + 2 * l] = __fadd_rn(acc[a][b][8 * h + 2 * l], m0);
                            acc[a][b][8 * h + 2 * l + 1] = __fadd_rn(acc[a][b][8 * h + 2 * l + 1], m1);
                        }
            }
        }
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const int jr = tt + 16 * b;
            if (jr >= ntr || !s_tvalid[jr]) continue;
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const int ir = 2 * tq + a;
                if (!s_qvalid[ir]) continue;
                float s[4];
#pragma unroll
                for (int l = 0; l < 4; ++l)
                    s[l] = __fadd_rn(__fadd_rn(__fadd_rn(acc[a][b][l], acc[a][b][4 + l]), acc[a][b][8 + l]), acc[a][b][12 + l]);
                const float f = sqrtf(__fadd_rn(__fadd_rn(s[0], s[2]), __fadd_rn(s[1], s[3])));
                // query direction: train rows arrive in ascending order, strict '<' keeps the lowest index on ties
                if (f < f2[a]) {
                    if (f < f1[a]) { f2[a] = f1[a]; f1[a] = f; i1[a] = t0 + jr; } else f2[a] = f;
                }
                // train direction: CTA-level merge cell of column jr
                const unsigned long long key = (static_cast<unsigned long long>(__float_as_uint(f)) << 32) | static_cast<unsigned int>(q0 + ir);
                const unsigned long long old = atomicMin(&s_ckey[jr], key);
                const unsigned long long loser = old > key ? old : key;
                if (loser != ~0ull) atomicMin(&s_cm2[jr], static_cast<unsigned int>(loser >> 32));
            }
        }
        __syncthreads();
        if (threadIdx.x < ntr && s_ckey[threadIdx.x] != ~0ull) {
            const unsigned long long key = s_ckey[threadIdx.x];
            unsigned long long* cell = w.col_cells + 2 * (pd.cb_base + t0 + threadIdx.x);
            const unsigned long long old = atomicMin(cell, key);
            const unsigned long long loser = old > key ? old : key;
            const unsigned int m2 = min(static_cast<unsigned int>(loser >> 32), s_cm2[threadIdx.x]);   // ~0 >> 32 = 0xffffffff
            atomicMin(reinterpret_cast<unsigned int*>(cell + 1), m2);
        }
    }
    // merge the 16 threads that share the two query rows (they sit in one half warp)
#pragma unroll
    for (int a = 0; a < 2; ++a) {
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) {
            const float o1 = __shfl_xor_sync(0xffffffffu, f1[a], off), o2 = __shfl_xor_sync(0xffffffffu, f2[a], off);
            const int oi = __shfl_xor_sync(0xffffffffu, i1[a], off);
            const bool take = (o1 < f1[a]) || (o1 == f1[a] && oi < i1[a]);
            const float lose = take ? f1[a] : o1;
            f2[a] = fminf(fminf(f2[a], o2), lose);
            if (take) { f1[a] = o1; i1[a] = oi; }
        }
        const int ir = 2 * tq + a;
        if (tt == 0 && ir < nqr) {
            const long long pos = pd.out_base + q0 + ir;
            const bool valid = s_qvalid[ir] && i1[a] != INT_MAX;
            w.row_best[pos] = valid ? i1[a] : -1;
            w.row_d1[pos] = valid ? f1[a] : FLT_MAX;
            w.row_d2[pos] = valid ? f2[a] : FLT_MAX;
        }
    }
}

// column merge cells -> (col_best, col_d1, col_d2) for the float-path pairs of the chunk
__global__ void __launch_bounds__(256) finish_cols_kernel(StoreView st, MatchWork w, long long total_cols)
{
    const long long pos = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= total_cols) return;
    const int p = tile_find_by_cb(w.pairs, w.n_pairs, static_cast<long long>(blockIdx.x) * blockDim.x);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 1) return;
    const int j = static_cast<int>(pos - pd.cb_base);
    if (j >= pd.nt) return;
    const unsigned long long key = w.col_cells[2 * pos];
    const unsigned int m2 = static_cast<unsigned int>(w.col_cells[2 * pos + 1] & 0xffffffffull);
    const bool valid = st.valid[pd.t_row0 + j] && key != ~0ull;
    w.col_best[pos] = valid ? static_cast<int>(key & 0xffffffffull) : -1;
    w.col_d1[pos] = valid ? __uint_as_float(static_cast<unsigned int>(key >> 32)) : FLT_MAX;
    w.col_d2[pos] = (valid && m2 < 0x7f800000u) ? __uint_as_float(m2) : FLT_MAX;
}

// true when the tile kernel can take the float-path pairs of this store (otherwise match_exact.cu scans them)
bool exact_tile_supported(const StoreView& st) { return st.dim >= 16 && st.dim % 16 == 0 && st.dim <= 512; }

cudaError_t launch_exact_tile(const StoreView& st, const MatchWork& w, int n_jobs, long long total_cols, cudaStream_t s)
{
    if (n_jobs <= 0) return cudaSuccess;
    const size_t smem = static_cast<size_t>(kTQ + kTT) * (st.dim + 4) * sizeof(float);
    static size_t configured_all[64] = {};             // per device: cudaFuncSetAttribute applies to the current device only
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& configured = configured_all[dev & 63];
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(exact_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    cudaError_t e = cudaMemsetAsync(w.col_cells, 0xff, 16 * static_cast<size_t>(total_cols), s);
    if (e != cudaSuccess) return e;
    exact_tile_kernel<<<n_jobs, 256, smem, s>>>(st, w, n_jobs);
    finish_cols_kernel<<<static_cast<unsigned>((total_cols + 255) / 256), 256, 0, s>>>(st, w, total_cols);
    return cudaGetLastError();
}

}  // namespace g2v
