// K2f -- exact resolve of the fp16-screened float path (ratio test off: only the nearest neighbour is needed).
//
// The screening kernel (match_screen_f.cu) leaves, per query row, the minimum of every group of 16 train columns (best /
// second-best group per 128-column half block) and, per train column, the minimum over every lane class
// {q : q mod 128 == lane}.  Its values are d^2 / 2 up to the rounding of the fp16 operands and the fp32 accumulation:
//     |s(q,t) - d^2(q,t) / 2|  <=  eps(q,t) = 1.1e-3 |q||t| + 1e-5 (|q|^2 + |t|^2) + 1e-6 (|q| + |t|)
// (2^-10 |q||t| from the two 11-bit significands, the rest covers fp16 subnormals, the 3-term norm expansion, the float32
// norms and the accumulation over K <= 272).
//
// For a row (column) the winning group (class) is recomputed EXACTLY in cv2's float32 summation order
// (cv::batchDistance -> normL2Sqr_, the arithmetic behind twoway_matcher.py:136,139; see match_exact.cu): the result is
// final when every other group's screened minimum r2 satisfies  r2 - eps_max > (f1^2 / 2)(1 + 1e-5),  i.e. no train row
// outside the group can reach or tie the float32 distance f1 (cv2's own summation error is < 2e-6 relative).  Otherwise
// (about 1 % of the rows of SuperPoint-like inputs) a second stage recomputes every group / class whose screened minimum
// is not above that threshold; only rows with more than kStage2Budget such candidates (degenerate inputs: many equal
// rows) are appended to the flag list for the exhaustive scan (exact_scan_kernel, mode 1).
#include <cfloat>

#include "store.h"

namespace g2v {

constexpr int kResF = 32;          // scan-side rows per CTA
constexpr int kStage2Budget = 1024; // candidates a row / column may recompute in the second stage before it is handed to
                                    // the exhaustive scan

__device__ __forceinline__ float screen_eps(float n2a, float n2b)
{
    const float na = sqrtf(n2a), nb = sqrtf(n2b);
    return 1.1e-3f * na * nb + 1e-5f * (n2a + n2b) + 1e-6f * (na + nb) + 1e-30f;
}

// cv2's float32 L2 distance by a half warp (r16 = lane & 15 holds partial sum r16 of the 16); dim % 16 tail handled.
__device__ __forceinline__ float cv2_l2_halfwarp(const float* a, const float* b, int dim, int r16)
{
    const int nblk = dim >> 4;
    float acc = 0.f;
    for (int k = 0; k < nblk; ++k) {
        const float d = __fsub_rn(a[16 * k + r16], b[16 * k + r16]);
        acc = __fadd_rn(acc, __fmul_rn(d, d));
    }
    const int l = r16 & 3;
    const float a0 = __shfl_sync(0xffffffffu, acc, l, 16), a1 = __shfl_sync(0xffffffffu, acc, 4 + l, 16);
    const float a2 = __shfl_sync(0xffffffffu, acc, 8 + l, 16), a3 = __shfl_sync(0xffffffffu, acc, 12 + l, 16);
    const float s = __fadd_rn(__fadd_rn(__fadd_rn(a0, a1), a2), a3);
    const float s0 = __shfl_sync(0xffffffffu, s, 0, 16), s1 = __shfl_sync(0xffffffffu, s, 1, 16);
    const float s2 = __shfl_sync(0xffffffffu, s, 2, 16), s3 = __shfl_sync(0xffffffffu, s, 3, 16);
    float d = __fadd_rn(__fadd_rn(s0, s2), __fadd_rn(s1, s3));
    for (int e = 16 * nblk; e < dim; ++e) {
        const float t = __fsub_rn(a[e], b[e]);
        d = __fadd_rn(d, __fmul_rn(t, t));
    }
    return sqrtf(d);
}

__device__ __forceinline__ void push_flag_f(const MatchWork& w, int pair, int dir, int idx)
{
    const int k = atomicAdd(w.n_flag, 1);
    if (k < w.flag_cap)
        w.flag_list[k] = (static_cast<unsigned long long>(pair) << 33) | (static_cast<unsigned long long>(dir) << 32) |
                         static_cast<unsigned int>(idx);
}

// DIR 0: one CTA per 32 query rows (candidates: the 16 train columns of the winning group).
// DIR 1: one CTA per 32 train columns (candidates: the nq_pad / 128 query rows of the winning lane class).
template <int DIR>
__global__ void __launch_bounds__(256) resolve_f_kernel(StoreView st, MatchWork w)
{
    extern __shared__ float s_scan[];                 // [kResF][dim] the CTA's scan-side rows
    __shared__ float s_r2[kResF];
    __shared__ int s_c0[kResF];
    const long long base = static_cast<long long>(blockIdx.x) * kResF;
    // pair of this block of rows / columns
    int lo = 0, hi = w.n_pairs - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if ((DIR == 0 ? w.pairs[mid].out_base : w.pairs[mid].cb_base) <= base) lo = mid; else hi = mid - 1;
    }
    const int p = lo;
    const PairDesc pd = w.pairs[p];
    if (pd.path != 2) return;
    const int i0 = static_cast<int>(base - (DIR == 0 ? pd.out_base : pd.cb_base));
    const int n_scan = DIR == 0 ? pd.nq : pd.nt, n_scan_pad = DIR == 0 ? pd.nq_pad : pd.nt_pad;
    if (i0 >= n_scan_pad || i0 >= n_scan) return;
    const int scan_row0 = (DIR == 0 ? pd.q_row0 : pd.t_row0) + i0;
    const int other_row0 = DIR == 0 ? pd.t_row0 : pd.q_row0;
    const int other_n = DIR == 0 ? pd.nt : pd.nq;
    const int dim = st.dim;
    const int cnt = min(kResF, n_scan - i0);
    for (int k = threadIdx.x; k < cnt * dim; k += blockDim.x) s_scan[k] = st.desc_f32[static_cast<long long>(scan_row0) * dim + k];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- phase 1: winning group / class and the smallest screened value outside it
    if (DIR == 0) {
        if (threadIdx.x < kResF) {
            const int q = i0 + threadIdx.x;
            const int nhb = pd.nt_pad / kHalfT;
            float bv = FLT_MAX, r2 = FLT_MAX;
            int bc = 0;
            for (int hb = 0; hb < nhb; ++hb) {
                const long long e = pd.rp_base + static_cast<long long>(hb) * pd.nq_pad + q;
                const float v1 = w.rp_v1[e], v2 = w.rp_v2[e];
                const int g = w.rp_g[e];
                if (v1 < bv) { r2 = fminf(r2, bv); bv = v1; bc = hb * kHalfT + g * kGroup; }
                else r2 = fminf(r2, v1);
                r2 = fminf(r2, v2);
            }
            s_r2[threadIdx.x] = r2;
            s_c0[threadIdx.x] = bc;
        }
    } else {
#pragma unroll 1
        for (int u = 0; u < 4; ++u) {
            const int r = warp * 4 + u;
            const int t = i0 + r;
            const float4 v = *reinterpret_cast<const float4*>(w.colraw + pd.cr_base + static_cast<long long>(t) * kTileQ + 4 * lane);
            float b = v.x, s2 = FLT_MAX;
            int bi = 4 * lane;
            const float vv[3] = {v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (vv[k] < b) { s2 = fminf(s2, b); b = vv[k]; bi = 4 * lane + 1 + k; } else s2 = fminf(s2, vv[k]);
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const float ob = __shfl_xor_sync(0xffffffffu, b, off), os = __shfl_xor_sync(0xffffffffu, s2, off);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
                const bool take = ob < b || (ob == b && oi < bi);
                s2 = fminf(fminf(s2, os), take ? b : ob);
                if (take) { b = ob; bi = oi; }
            }
            if (lane == 0) { s_r2[r] = s2; s_c0[r] = bi; }
        }
    }
    __syncthreads();

    // ---- phase 2: exact recompute of the candidates, a half warp per distance
    const int h = lane >> 4, r16 = lane & 15;
    const int n_cand = DIR == 0 ? kGroup : pd.nq_pad / kTileQ;
    const float other_maxn2 = DIR == 0 ? pd.t_maxn2 : pd.q_maxn2;
    int* o_best = DIR == 0 ? w.row_best : w.col_best;
    float* o_d1 = DIR == 0 ? w.row_d1 : w.col_d1;
    float* o_d2 = DIR == 0 ? w.row_d2 : w.col_d2;
    const long long out0 = (DIR == 0 ? pd.out_base : pd.cb_base) + i0;
#pragma unroll 1
    for (int u = 0; u < 4; ++u) {
        const int r = warp * 4 + u;
        if (r >= cnt) break;
        if (!st.valid[scan_row0 + r]) {
            if (lane == 0) { o_best[out0 + r] = -1; o_d1[out0 + r] = FLT_MAX; o_d2[out0 + r] = FLT_MAX; }
            continue;
        }
        const float* a = s_scan + r * dim;
        float f1 = FLT_MAX;
        int i1 = INT_MAX;
        // exact distances to n candidates c0, c0 + stride, ...: the half warps alternate, both always run the same rounds
        auto scan_cands = [&](int c0, int stride, int n) {
            for (int j = h; j < n; j += 2) {
                const int c = c0 + j * stride;
                const bool ok = c < other_n && st.valid[other_row0 + c];
                const float f = cv2_l2_halfwarp(a, st.desc_f32 + static_cast<long long>(other_row0 + (ok ? c : 0)) * dim, dim, r16);
                if (ok && (f < f1 || (f == f1 && c < i1))) { f1 = f; i1 = c; }       // the lowest index wins ties
            }
        };
        const int c0 = s_c0[r];
        scan_cands(c0, DIR == 0 ? 1 : kTileQ, n_cand);
        {
            const float of = __shfl_xor_sync(0xffffffffu, f1, 16);
            const int oi = __shfl_xor_sync(0xffffffffu, i1, 16);
            if (of < f1 || (of == f1 && oi < i1)) { f1 = of; i1 = oi; }
        }
        const float eps = screen_eps(st.norm_f[scan_row0 + r], other_maxn2);
        // every candidate whose screened value exceeds thr is farther than f1 in cv2's float32 arithmetic
        const double thr_d = 0.5 * static_cast<double>(f1) * static_cast<double>(f1) * (1.0 + 1e-5) + static_cast<double>(eps) + 1e-30;
        bool done = i1 != INT_MAX && static_cast<double>(s_r2[r]) > thr_d;
        if (!done && i1 != INT_MAX) {
            // second stage: recompute every other group / class the screening cannot rule out.  Rows: a half block whose
            // second-best group is a suspect is recomputed whole (the screening keeps only the id of its best group).
            const float thr = __double2float_ru(thr_d);
            int budget = kStage2Budget;                         // candidates this row may still recompute
            if (DIR == 0) {
                const int nhb = pd.nt_pad / kHalfT;
                for (int hb = 0; hb < nhb && budget >= 0; ++hb) {
                    const long long e = pd.rp_base + static_cast<long long>(hb) * pd.nq_pad + (i0 + r);
                    const float v1 = w.rp_v1[e], v2 = w.rp_v2[e];
                    const int g = w.rp_g[e];
                    const int cg = hb * kHalfT + g * kGroup;
                    if (v2 <= thr) {
                        budget -= kHalfT;
                        if (budget >= 0) scan_cands(hb * kHalfT, 1, kHalfT);        // includes the best group again: harmless
                    } else if (v1 <= thr && cg != c0) {
                        budget -= kGroup;
                        if (budget >= 0) scan_cands(cg, 1, kGroup);
                    }
                }
            } else {
                const float* cr = w.colraw + pd.cr_base + static_cast<long long>(i0 + r) * kTileQ;
                for (int cl = 0; cl < kTileQ && budget >= 0; ++cl) {
                    if (cl != c0 && cr[cl] <= thr) {
                        budget -= n_cand;
                        if (budget >= 0) scan_cands(cl, kTileQ, n_cand);
                    }
                }
            }
            const float of = __shfl_xor_sync(0xffffffffu, f1, 16);
            const int oi = __shfl_xor_sync(0xffffffffu, i1, 16);
            if (of < f1 || (of == f1 && oi < i1)) { f1 = of; i1 = oi; }
            done = budget >= 0;
        }
        if (lane == 0) {
            if (done) { o_best[out0 + r] = i1; o_d1[out0 + r] = f1; o_d2[out0 + r] = FLT_MAX; }
            else push_flag_f(w, p, DIR, i0 + r);                // too many suspects (or no valid candidate): exhaustive scan
        }
    }
}

// total_rows / total_cols: rows (columns) of the chunk over all pairs (multiples of 256).
cudaError_t launch_resolve_f(const StoreView& st, const MatchWork& w, long long total_rows, long long total_cols, cudaStream_t s)
{
    const size_t smem = static_cast<size_t>(kResF) * st.dim * sizeof(float);
    static size_t configured = 0;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(resolve_f_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(resolve_f_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    if (total_rows > 0) resolve_f_kernel<0><<<static_cast<unsigned>(total_rows / kResF), 256, smem, s>>>(st, w);
    if (total_cols > 0) resolve_f_kernel<1><<<static_cast<unsigned>(total_cols / kResF), 256, smem, s>>>(st, w);
    return cudaGetLastError();
}

}  // namespace g2v
