// Exact fp32 nearest / second-nearest scan in OpenCV's own summation order.
//
// cv::BFMatcher(NORM_L2) -> cv::batchDistance -> normL2Sqr_(float): blocks of 16 elements feed four 4-lane
// accumulators (multiply and add rounded separately), combined ((a0+a1)+a2)+a3 lane-wise, horizontal sum
// (v0+v2)+(v1+v3), scalar tail, then sqrtf (the un-vendored dependency behind twoway_matcher.py:136,139; the same
// order is pinned bit-for-bit against cv2 by the CPU oracle's tests).  Sixteen lanes reproduce the sixteen partial
// sums, so a half warp produces one distance with exactly cv2's rounding.
//
// Used (a) for every row and column of pairs whose descriptors are not integer-valued (float path), and (b) for the
// rows / columns the integer-path resolve kernels flagged as ambiguous.
#include <cfloat>

#include "store.h"

namespace g2v {

constexpr int kScanRows = 32;   // scan-side rows per CTA (4 per warp)
constexpr int kScanTile = 64;   // other-side rows staged per shared-memory tile
constexpr int kScanSeg = 64;    // flagged rows: the other side is scanned by this many CTAs, merged with atomics

__device__ __forceinline__ int find_by_job0(const PairDesc* pairs, int n, int job)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].job0 <= job) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// mode 0: jobs enumerated from PairDesc::job0 (float-path pairs); mode 1: one job per flagged row / column.
// NU = scan rows per warp: 4 in mode 0, 1 in mode 1 (a flagged job is ONE row; computing four cost 4x the arithmetic).
template <int NU>
__global__ void __launch_bounds__(256) exact_scan_kernel(StoreView st, MatchWork w, int mode, int n_range_jobs)
{
    extern __shared__ float sm[];
    const int dim = st.dim;
    const int tstride = dim + 16;                     // half warps read rows j and j+1: offset them by 16 banks
    float* sq = sm;                                   // [kScanRows][dim]
    float* stile = sm + kScanRows * dim;              // [kScanTile][tstride]
    __shared__ unsigned char svalid[kScanTile];

    const int n_jobs = mode == 0 ? n_range_jobs : min(*w.n_flag, w.flag_cap) * kScanSeg;
    for (int item = blockIdx.x; item < n_jobs; item += gridDim.x) {
    __syncthreads();
    const int job = mode == 0 ? item : item / kScanSeg;
    const int seg = mode == 0 ? 0 : item - job * kScanSeg;
    int p, dir, row0, cnt;
    if (mode == 0) {
        p = find_by_job0(w.pairs, w.n_pairs, job);
        const PairDesc& pd = w.pairs[p];
        const int local = job - pd.job0;
        const int nqj = (pd.nq + kScanRows - 1) / kScanRows;
        dir = local < nqj ? 0 : 1;
        row0 = (dir == 0 ? local : local - nqj) * kScanRows;
        cnt = min(kScanRows, (dir == 0 ? pd.nq : pd.nt) - row0);
    } else {
        const unsigned long long e = w.flag_list[job];
        p = static_cast<int>(e >> 33);
        dir = static_cast<int>((e >> 32) & 1);
        row0 = static_cast<int>(e & 0xffffffffu);
        cnt = 1;
    }
    const PairDesc pd = w.pairs[p];
    const int scan_base = dir == 0 ? pd.q_row0 : pd.t_row0;
    const int other_base = dir == 0 ? pd.t_row0 : pd.q_row0;
    const int other_n = dir == 0 ? pd.nt : pd.nq;
    const long long out0 = (dir == 0 ? pd.out_base : pd.cb_base) + row0;
    int* o_best = dir == 0 ? w.row_best : w.col_best;
    float* o_d1 = dir == 0 ? w.row_d1 : w.col_d1;
    float* o_d2 = dir == 0 ? w.row_d2 : w.col_d2;

    for (int k = threadIdx.x; k < cnt * dim; k += blockDim.x) {
        const int r = k / dim, c = k - r * dim;
        sq[r * dim + c] = st.desc_f32[static_cast<long long>(scan_base + row0 + r) * dim + c];
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = lane >> 4, r16 = lane & 15;
    const int nblk = dim / 16;
    float f1[NU], f2[NU];
    int i1[NU];
#pragma unroll
    for (int u = 0; u < NU; ++u) { f1[u] = FLT_MAX; f2[u] = FLT_MAX; i1[u] = INT_MAX; }
    const bool single = mode == 1;                     // one scan row: the 8 warps split each tile's rows
    const bool warp_active = single || warp * 4 < cnt;
    const int qrow0 = single ? 0 : warp * 4;
    const int it0 = single ? warp * (kScanTile / 16) : 0, it1 = single ? (warp + 1) * (kScanTile / 16) : kScanTile / 2;

    // mode 1: this CTA scans segment `seg` of the other side; the segments of a job meet in w.flag_scratch
    const int seg_len = single ? ((other_n + kScanSeg * kScanTile - 1) / (kScanSeg * kScanTile)) * kScanTile : other_n;
    const int t_begin = single ? min(other_n, seg * seg_len) : 0;
    const int t_end = single ? min(other_n, t_begin + seg_len) : other_n;
    if constexpr (NU == 1) __syncthreads();           // the scan row in `sq` is complete
    for (int t0 = t_begin; t0 < t_end; t0 += kScanTile) {
        const int tn = min(kScanTile, t_end - t0);
        // A flagged job uses every element of the other side once: its half warps stream the rows straight from global
        // memory (same row -> half-warp assignment and arithmetic as the staged path, so the results are identical).
        if constexpr (NU != 1) {
        __syncthreads();
        if ((dim & 3) == 0) {
            const int d4 = dim >> 2;
            for (int k = threadIdx.x; k < tn * d4; k += blockDim.x) {
                const int r = k / d4, c = (k - r * d4) * 4;
                *reinterpret_cast<float4*>(stile + r * tstride + c) =
                    *reinterpret_cast<const float4*>(st.desc_f32 + static_cast<long long>(other_base + t0 + r) * dim + c);
            }
        } else {
            for (int k = threadIdx.x; k < tn * dim; k += blockDim.x) {
                const int r = k / dim, c = k - r * dim;
                stile[r * tstride + c] = st.desc_f32[static_cast<long long>(other_base + t0 + r) * dim + c];
            }
        }
        if (threadIdx.x < kScanTile) svalid[threadIdx.x] = threadIdx.x < tn ? st.valid[other_base + t0 + threadIdx.x] : 0;
        __syncthreads();
        }
        if (!warp_active) continue;
        for (int it = it0; it < it1; ++it) {
            const int jr = 2 * it + h;
            const long long grow = static_cast<long long>(other_base) + t0 + (jr < tn ? jr : tn - 1);
            const float* trow = NU == 1 ? st.desc_f32 + grow * dim : stile + jr * tstride;
            float acc[NU];
#pragma unroll
            for (int u = 0; u < NU; ++u) acc[u] = 0.f;
#pragma unroll(NU == 1 ? 8 : 1)
            for (int b = 0; b < nblk; ++b) {
                const float tv = trow[16 * b + r16];
#pragma unroll
                for (int u = 0; u < NU; ++u) {
                    const float d = __fsub_rn(sq[(qrow0 + u) * dim + 16 * b + r16], tv);
                    acc[u] = __fadd_rn(acc[u], __fmul_rn(d, d));
                }
            }
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                const int l = r16 & 3;
                const float a0 = __shfl_sync(0xffffffffu, acc[u], l, 16), a1 = __shfl_sync(0xffffffffu, acc[u], 4 + l, 16);
                const float a2 = __shfl_sync(0xffffffffu, acc[u], 8 + l, 16), a3 = __shfl_sync(0xffffffffu, acc[u], 12 + l, 16);
                const float s = __fadd_rn(__fadd_rn(__fadd_rn(a0, a1), a2), a3);
                const float s0 = __shfl_sync(0xffffffffu, s, 0, 16), s1 = __shfl_sync(0xffffffffu, s, 1, 16);
                const float s2 = __shfl_sync(0xffffffffu, s, 2, 16), s3 = __shfl_sync(0xffffffffu, s, 3, 16);
                float d = __fadd_rn(__fadd_rn(s0, s2), __fadd_rn(s1, s3));
                for (int e = 16 * nblk; e < dim; ++e) {
                    const float t = __fsub_rn(sq[(qrow0 + u) * dim + e], trow[e]);
                    d = __fadd_rn(d, __fmul_rn(t, t));
                }
                const float f = sqrtf(d);
                const bool ok = jr < tn && (NU == 1 ? st.valid[grow] != 0 : svalid[jr] != 0);
                // cv2's top-2 insertion: strict '<', so the lowest train index wins ties
                if (ok && f < f2[u]) {
                    if (f < f1[u]) { f2[u] = f1[u]; f1[u] = f; i1[u] = t0 + jr; } else f2[u] = f;
                }
            }
        }
    }
    if (!warp_active) continue;
#pragma unroll
    for (int u = 0; u < NU; ++u) {
        const float o1 = __shfl_xor_sync(0xffffffffu, f1[u], 16), o2 = __shfl_xor_sync(0xffffffffu, f2[u], 16);
        const int oi = __shfl_xor_sync(0xffffffffu, i1[u], 16);
        const bool take = (o1 < f1[u]) || (o1 == f1[u] && oi < i1[u]);
        const float lose = take ? f1[u] : o1;
        f2[u] = fminf(fminf(f2[u], o2), lose);
        if (take) { f1[u] = o1; i1[u] = oi; }
        if (single) {
            // merge the 8 warps' partial results for the single scan row (u == 0) through shared memory
            __shared__ float m_f1[8], m_f2[8];
            __shared__ int m_i1[8];
            if (u == 0) {
                if (lane == 0) { m_f1[warp] = f1[0]; m_f2[warp] = f2[0]; m_i1[warp] = i1[0]; }
                __syncthreads();
                if (threadIdx.x == 0) {
                    float b1 = m_f1[0], b2 = m_f2[0];
                    int bi = m_i1[0];
                    for (int k = 1; k < 8; ++k) {
                        const bool tk = (m_f1[k] < b1) || (m_f1[k] == b1 && m_i1[k] < bi);
                        const float ls = tk ? b1 : m_f1[k];
                        b2 = fminf(fminf(b2, m_f2[k]), ls);
                        if (tk) { b1 = m_f1[k]; bi = m_i1[k]; }
                    }
                    // Merge with the other segments: lowest (distance, index) by a 64-bit atomicMin; every key that
                    // loses a comparison is a candidate for the second-best distance.  0xff-filled scratch = +infinity.
                    unsigned long long* k1p = w.flag_scratch + 2 * static_cast<long long>(job);
                    unsigned int* m2p = reinterpret_cast<unsigned int*>(k1p + 1);
                    unsigned int* cntp = m2p + 1;
                    if (bi != INT_MAX) {
                        const unsigned long long key = (static_cast<unsigned long long>(__float_as_uint(b1)) << 32) | static_cast<unsigned int>(bi);
                        const unsigned long long old = atomicMin(k1p, key);
                        const unsigned long long loser = old > key ? old : key;
                        atomicMin(m2p, min(static_cast<unsigned int>(loser >> 32), __float_as_uint(b2)));
                    }
                    __threadfence();
                    const unsigned int arrived = atomicAdd(cntp, 1u);              // starts at 0xffffffff
                    if (arrived == static_cast<unsigned int>(kScanSeg - 2)) {        // last segment of the job
                        __threadfence();
                        const unsigned long long kf = *reinterpret_cast<volatile unsigned long long*>(k1p);
                        const unsigned int m2 = *reinterpret_cast<volatile unsigned int*>(m2p);
                        const bool valid = st.valid[scan_base + row0] && kf != ~0ull;
                        o_best[out0] = valid ? static_cast<int>(kf & 0xffffffffu) : -1;
                        o_d1[out0] = valid ? __uint_as_float(static_cast<unsigned int>(kf >> 32)) : FLT_MAX;
                        o_d2[out0] = (valid && m2 < 0x7f800000u) ? __uint_as_float(m2) : FLT_MAX;
                    }
                }
            }
            continue;
        }
        const int r = warp * 4 + u;
        if (lane == 0 && r < cnt) {
            const bool valid = st.valid[scan_base + row0 + r] && i1[u] != INT_MAX;
            o_best[out0 + r] = valid ? i1[u] : -1;
            o_d1[out0 + r] = valid ? f1[u] : FLT_MAX;
            o_d2[out0 + r] = valid ? f2[u] : FLT_MAX;
        }
    }
    }
}

// n_jobs: number of range jobs (mode 0) or an upper bound on the flagged entries (mode 1: a fixed grid strides over them).
cudaError_t launch_exact_scan(const StoreView& st, const MatchWork& w, int mode, int n_jobs, cudaStream_t s)
{
    if (n_jobs <= 0) return cudaSuccess;
    const size_t smem = static_cast<size_t>(kScanRows * st.dim + kScanTile * (st.dim + 16)) * sizeof(float);
    static size_t configured_all[64][2] = {};             // per device (the attribute is per device) and per instantiation
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& configured = configured_all[dev & 63][mode == 1 ? 1 : 0];
    if (smem > configured) {
        cudaError_t e = mode == 1 ? cudaFuncSetAttribute(exact_scan_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem))
                                  : cudaFuncSetAttribute(exact_scan_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    const long long items = mode == 0 ? n_jobs : static_cast<long long>(n_jobs) * kScanSeg;
    const int grid = static_cast<int>(items < 1184 ? items : 1184);
    if (mode == 1) exact_scan_kernel<1><<<grid, 256, smem, s>>>(st, w, mode, n_jobs);
    else exact_scan_kernel<4><<<grid, 256, smem, s>>>(st, w, mode, n_jobs);
    return cudaGetLastError();
}

}  // namespace g2v
