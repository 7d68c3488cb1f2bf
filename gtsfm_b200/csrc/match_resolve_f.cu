// K2f -- exact resolve of the fp16-screened float path: the nearest neighbour, and with the ratio test on the decision
// d1 <= ratio * d2 (twoway_matcher.py:137) as well.
//
// The screening kernel (match_screen_f.cu) leaves, per query row, the minimum of every group of 4 train columns (best /
// second-best group per 128-column half block) and, per train column, the minimum over every lane class
// {q = (4 tg + j) 128 + lane, j < 4} (at most four query rows).  Its values are d^2 / 2 up to the rounding of the fp16 operands and the fp32 accumulation:
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
//
// Ratio test on (TwoWayMatcher(ratio_test_threshold=r) on float descriptors: configs/correspondence/kaze.yaml,
// d2net.yaml).  The second-nearest distance d2 is either inside the winning group (exact, it was recomputed) or outside it,
// where the screening brackets it: the smallest screened value outside the group, r2, gives
//     sqrt(2 (r2 - eps)) (1 - 2e-6)  <=  d_out  <=  sqrt(2 (r2 + eps)) (1 + 2e-6)
// for the nearest outside candidate in cv2's float32 arithmetic.  The test is CERTAIN when both ends of the bracket of
// d2 = min(d2_inside, d_out) give the same answer; the row then carries d2 = FLT_MAX (keep) or -1 (drop) for the finalize
// kernel, or the exact d2 when it is the inside one.  Otherwise (a fraction of a percent of the rows) the second stage
// recomputes every group that can hold one of the two nearest neighbours -- screened minimum not above the level of an
// upper bound of d2 -- and writes the exact pair (d1, d2).
#include <cfloat>

#include "store.h"

namespace g2v {

constexpr int kResF = 16;          // scan-side rows per CTA (one per half warp)
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

// NB distances at once (same arithmetic per distance): the scan-side element is loaded once for all of them and their
// loads are independent.
template <int NB>
__device__ __forceinline__ void cv2_l2_halfwarp_n(const float* a, const float* const* b, int dim, int r16, float* out)
{
    const int nblk = dim >> 4;
    float acc[NB];
#pragma unroll
    for (int u = 0; u < NB; ++u) acc[u] = 0.f;
#pragma unroll 4
    for (int k = 0; k < nblk; ++k) {
        const float av = a[16 * k + r16];
#pragma unroll
        for (int u = 0; u < NB; ++u) {
            const float d = __fsub_rn(av, __ldg(b[u] + 16 * k + r16));
            acc[u] = __fadd_rn(acc[u], __fmul_rn(d, d));
        }
    }
    const int l = r16 & 3;
#pragma unroll
    for (int u = 0; u < NB; ++u) {
        const float a0 = __shfl_sync(0xffffffffu, acc[u], l, 16), a1 = __shfl_sync(0xffffffffu, acc[u], 4 + l, 16);
        const float a2 = __shfl_sync(0xffffffffu, acc[u], 8 + l, 16), a3 = __shfl_sync(0xffffffffu, acc[u], 12 + l, 16);
        const float s = __fadd_rn(__fadd_rn(__fadd_rn(a0, a1), a2), a3);
        const float s0 = __shfl_sync(0xffffffffu, s, 0, 16), s1 = __shfl_sync(0xffffffffu, s, 1, 16);
        const float s2 = __shfl_sync(0xffffffffu, s, 2, 16), s3 = __shfl_sync(0xffffffffu, s, 3, 16);
        float d = __fadd_rn(__fadd_rn(s0, s2), __fadd_rn(s1, s3));
        for (int e = 16 * nblk; e < dim; ++e) {
            const float t = __fsub_rn(a[e], b[u][e]);
            d = __fadd_rn(d, __fmul_rn(t, t));
        }
        out[u] = sqrtf(d);
    }
}

__device__ __forceinline__ void push_flag_f(const MatchWork& w, int pair, int dir, int idx)
{
    const int k = atomicAdd(w.n_flag, 1);
    if (k < w.flag_cap)
        w.flag_list[k] = (static_cast<unsigned long long>(pair) << 33) | (static_cast<unsigned long long>(dir) << 32) |
                         static_cast<unsigned int>(idx);
}

// A half warp owns one scan-side row end to end (no shared memory, no block barrier), kResF rows per CTA:
//   DIR 0: a query row; candidates = the 4 train columns of its winning group.
//   DIR 1: a train column; candidates = the <= 4 query rows of its winning lane class.
template <int DIR>
__global__ void __launch_bounds__(256) resolve_f_kernel(StoreView st, MatchWork w, int use_ratio, double ratio)
{
    const long long base = static_cast<long long>(blockIdx.x) * kResF;
    // pair of this block of rows / columns (table built by the host: a binary search over the pair list would put a chain
    // of dependent global loads in front of every CTA)
    const int p = (DIR == 0 ? w.row_blk_pair : w.col_blk_pair)[base / kRowPad];
    const PairDesc pd = w.pairs[p];
    if (pd.path != 2) return;
    const int i0 = static_cast<int>(base - (DIR == 0 ? pd.out_base : pd.cb_base));
    const int n_scan = DIR == 0 ? pd.nq : pd.nt;
    if (i0 >= n_scan) return;
    const int scan_row0 = (DIR == 0 ? pd.q_row0 : pd.t_row0) + i0;
    const int other_row0 = DIR == 0 ? pd.t_row0 : pd.q_row0;
    const int other_n = DIR == 0 ? pd.nt : pd.nq;
    const int dim = st.dim;
    const int cnt = min(kResF, n_scan - i0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = lane >> 4, r16 = lane & 15;
    // DIR 1: lane class cl = tg * 128 + lane holds the query rows (kFlushTiles tg + j) 128 + lane, j < kFlushTiles
    const int n_qtiles = pd.nq_pad / kTileQ;
    const int n_cls = (n_qtiles + kFlushTiles - 1) / kFlushTiles * kTileQ;
    auto class_row0 = [&](int cl) { return (cl / kTileQ) * kFlushTiles * kTileQ + (cl % kTileQ); };
    auto class_rows = [&](int cl) { return min(kFlushTiles, n_qtiles - (cl / kTileQ) * kFlushTiles); };
    const float other_maxn2 = DIR == 0 ? pd.t_maxn2 : pd.q_maxn2;
    int* o_best = DIR == 0 ? w.row_best : w.col_best;
    float* o_d1 = DIR == 0 ? w.row_d1 : w.col_d1;
    float* o_d2 = DIR == 0 ? w.row_d2 : w.col_d2;
    const long long out0 = (DIR == 0 ? pd.out_base : pd.cb_base) + i0;

    const int r = warp * 2 + h;                          // this half warp's row of the block
    const bool in_range = r < cnt;
    const int rs = in_range ? r : 0;
    const bool active = in_range && st.valid[scan_row0 + rs];
    const float* a = st.desc_f32 + static_cast<long long>(scan_row0 + rs) * dim;

    // ---- phase 1: winning group / class (c0) and the smallest screened value outside it (r2)
    float bv = FLT_MAX, r2 = FLT_MAX;
    int c0 = 0;
    if (DIR == 0) {
        const int nhb = pd.nt_pad / kHalfT;
        for (int hb = r16; hb < nhb; hb += 16) {                     // ascending half blocks per lane
            const long long e = pd.rp_base + static_cast<long long>(hb) * pd.nq_pad + (i0 + rs);
            const float v1 = w.rp_v1[e], v2 = w.rp_v2[e];
            const int g = w.rp_g[e];
            if (v1 < bv) { r2 = fminf(r2, bv); bv = v1; c0 = hb * kHalfT + g * kGroupF; }
            else r2 = fminf(r2, v1);
            r2 = fminf(r2, v2);
        }
    } else {
        const float* crp = w.colraw + pd.cr_base + static_cast<long long>(i0 + rs) * n_cls;
        for (int k0 = 4 * r16; k0 < n_cls; k0 += 64) {               // ascending class ids per lane
            const float4 v = __ldg(reinterpret_cast<const float4*>(crp + k0));
            const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (vv[k] < bv) { r2 = fminf(r2, bv); bv = vv[k]; c0 = k0 + k; } else r2 = fminf(r2, vv[k]);
            }
        }
    }
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) {                          // within the half warp
        const float ob = __shfl_xor_sync(0xffffffffu, bv, off), os = __shfl_xor_sync(0xffffffffu, r2, off);
        const int oc = __shfl_xor_sync(0xffffffffu, c0, off);
        const bool take = ob < bv || (ob == bv && oc < c0);
        r2 = fminf(fminf(r2, os), take ? bv : ob);
        if (take) { bv = ob; c0 = oc; }
    }

    // ---- phase 2, fast pass: the (<= 4) candidates in one batch, so all their loads are in flight together
    float f1 = FLT_MAX, f2in = FLT_MAX;                  // nearest and second nearest inside the winning group (exact)
    int i1 = INT_MAX;
    bool done = false;
    {
        const int cbase = DIR == 0 ? c0 : class_row0(c0), cstride = DIR == 0 ? 1 : kTileQ;
        const int cn = DIR == 0 ? kGroupF : class_rows(c0);
        int c[4];
        bool ok[4];
        const float* bp[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            c[k] = cbase + k * cstride;
            ok[k] = active && k < cn && c[k] < other_n && st.valid[other_row0 + (c[k] < other_n ? c[k] : 0)];
            bp[k] = st.desc_f32 + static_cast<long long>(other_row0 + (ok[k] ? c[k] : 0)) * dim;
        }
        float f[4];
        cv2_l2_halfwarp_n<4>(a, bp, dim, r16, f);
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (ok[k]) {                                                         // ascending index: the lowest wins ties
                if (f[k] < f1) { f2in = f1; f1 = f[k]; i1 = c[k]; }
                else f2in = fminf(f2in, f[k]);
            }
    }
    const float eps = screen_eps(st.norm_f[scan_row0 + rs], other_maxn2);
    // every candidate whose screened value exceeds thr is farther than f1 in cv2's float32 arithmetic
    double thr_d = 0.5 * static_cast<double>(f1) * static_cast<double>(f1) * (1.0 + 1e-5) + static_cast<double>(eps) + 1e-30;
    if (active && i1 != INT_MAX) done = static_cast<double>(r2) > thr_d;
    float d2_out = FLT_MAX;                              // what the finalize kernel sees as the second-nearest distance
    if (use_ratio && active && i1 != INT_MAX) {
        // bracket of the nearest candidate OUTSIDE the group in cv2's float32 arithmetic (r2 = FLT_MAX: there is none)
        double lo_out = 1e300, up_out = 1e300;
        if (r2 < 1e37f) {
            const double a0 = static_cast<double>(r2) - static_cast<double>(eps), a1 = static_cast<double>(r2) + static_cast<double>(eps);
            lo_out = a0 > 0 ? sqrt(2.0 * a0) * (1.0 - 2e-6) : 0.0;
            up_out = sqrt(2.0 * a1) * (1.0 + 2e-6);
        }
        const double d1 = f1, d2i = f2in < FLT_MAX ? static_cast<double>(f2in) : 1e300;
        const double d2_lo = fmin(d2i, lo_out), d2_up = fmin(d2i, up_out);
        if (done) {
            if (d2i <= lo_out) d2_out = f2in;                                    // the second nearest is inside the group: exact
            else if (d1 <= ratio * d2_lo && d1 <= ratio * d2_up) d2_out = FLT_MAX;   // certainly kept
            else if (!(d1 <= ratio * d2_lo) && !(d1 <= ratio * d2_up)) d2_out = -1.0f;   // certainly dropped
            else done = false;                                                   // undecided: exact second stage
        }
        if (!done) {
            // the second stage must see every candidate that can be one of the two nearest: an upper bound of d2 is
            // max(f1, min(f2in, up_out)) (two candidates are known to lie at or below it)
            const double U = fmax(d1, d2_up);
            thr_d = fmax(thr_d, 0.5 * U * U * (1.0 + 1e-5) + static_cast<double>(eps) + 1e-30);
        }
    }
    if (r16 == 0 && in_range) {
        if (!active) { o_best[out0 + r] = -1; o_d1[out0 + r] = FLT_MAX; o_d2[out0 + r] = FLT_MAX; }
        else if (done) { o_best[out0 + r] = i1; o_d1[out0 + r] = f1; o_d2[out0 + r] = d2_out; }
    }
    const unsigned need = __ballot_sync(0xffffffffu, active && !done);
    if (need == 0) return;

    // ---- second stage (about 1 % of the rows): the whole warp recomputes every other group / class the screening cannot
    // rule out.  Rows: a half block whose second-best group is a suspect is recomputed whole (the screening keeps only the
    // id of its best group).
#pragma unroll 1
    for (int u = 0; u < 2; ++u) {
        if (!((need >> (16 * u)) & 1u)) continue;
        const int ru = warp * 2 + u;
        const float* au = st.desc_f32 + static_cast<long long>(scan_row0 + ru) * dim;
        const int c0u = __shfl_sync(0xffffffffu, c0, 16 * u);
        const float thr = __double2float_ru(__shfl_sync(0xffffffffu, thr_d, 16 * u));
        // result of the fast pass (the winning group, exact); each half warp collects what IT finds beyond that group in
        // (g1 at j1, g2), the three disjoint sets are merged at the end
        const float s1 = __shfl_sync(0xffffffffu, f1, 16 * u), s2 = __shfl_sync(0xffffffffu, f2in, 16 * u);
        const int si = __shfl_sync(0xffffffffu, i1, 16 * u);
        float g1 = FLT_MAX, g2 = FLT_MAX;
        int j1 = INT_MAX;
        if (si == INT_MAX) { if (lane == 0) push_flag_f(w, p, DIR, i0 + ru); continue; }   // no valid candidate: exhaustive scan
        // index range of the winning group's candidates (rows: 4 consecutive columns; columns: rows 128 apart)
        const int g0lo = DIR == 0 ? c0u : class_row0(c0u), g0step = DIR == 0 ? 1 : kTileQ;
        const int g0hi = g0lo + (DIR == 0 ? kGroupF : class_rows(c0u)) * g0step - (g0step - 1);
        // exact distances to n candidates cb, cb + stride, ...: the half warps alternate, both always run the same rounds
        auto scan_cands = [&](int cb, int stride, int n) {                       // n is even
            constexpr int NB = 2;                                                // candidates per half warp and round
            for (int m0 = 0; m0 < (n >> 1); m0 += NB) {
                int c[NB];
                bool ok[NB];
                const float* bp[NB];
#pragma unroll
                for (int k = 0; k < NB; ++k) {
                    c[k] = cb + (h + 2 * (m0 + k)) * stride;
                    ok[k] = (m0 + k) < (n >> 1) && c[k] < other_n && st.valid[other_row0 + (c[k] < other_n ? c[k] : 0)];
                    bp[k] = st.desc_f32 + static_cast<long long>(other_row0 + (ok[k] ? c[k] : 0)) * dim;
                }
                float f[NB];
                cv2_l2_halfwarp_n<NB>(au, bp, dim, r16, f);
#pragma unroll
                for (int k = 0; k < NB; ++k) {
                    // the winning group was recomputed in the fast pass: its four candidates are in (g1, g2) already
                    if (!ok[k] || (c[k] >= g0lo && c[k] < g0hi && ((c[k] - g0lo) % g0step) == 0)) continue;
                    if (f[k] < g1 || (f[k] == g1 && c[k] < j1)) { g2 = g1; g1 = f[k]; j1 = c[k]; }   // the lowest index wins ties
                    else g2 = fminf(g2, f[k]);
                }
            }
        };
        int budget = kStage2Budget;                             // candidates this row may still recompute
        if (DIR == 0) {
            const int nhb = pd.nt_pad / kHalfT;
            for (int hb = 0; hb < nhb && budget >= 0; ++hb) {
                const long long e = pd.rp_base + static_cast<long long>(hb) * pd.nq_pad + (i0 + ru);
                const float v1 = w.rp_v1[e], v2 = w.rp_v2[e];
                const int cg = hb * kHalfT + w.rp_g[e] * kGroupF;
                if (v2 <= thr) {
                    budget -= kHalfT;
                    if (budget >= 0) scan_cands(hb * kHalfT, 1, kHalfT);        // includes the best group again: harmless
                } else if (v1 <= thr && cg != c0u) {
                    budget -= kGroupF;
                    if (budget >= 0) scan_cands(cg, 1, kGroupF);
                }
            }
        } else {
            const float* cr = w.colraw + pd.cr_base + static_cast<long long>(i0 + ru) * n_cls;
            for (int cl0 = 0; cl0 < n_cls && budget >= 0; cl0 += 32) {          // 32 classes per step, suspects by ballot
                unsigned sus = __ballot_sync(0xffffffffu, (cl0 + lane) < n_cls && (cl0 + lane) != c0u && cr[cl0 + lane] <= thr);
                while (sus && budget >= 0) {
                    const int cl = cl0 + __ffs(sus) - 1;
                    sus &= sus - 1;
                    budget -= kFlushTiles;
                    if (budget >= 0) scan_cands(class_row0(cl), kTileQ, class_rows(cl));
                }
            }
        }
        // merge: winning group (s1 at si, s2), this half (g1 at j1, g2), the other half
        const float of = __shfl_xor_sync(0xffffffffu, g1, 16), of2 = __shfl_xor_sync(0xffffffffu, g2, 16);
        const int oi = __shfl_xor_sync(0xffffffffu, j1, 16);
        {
            float b1 = s1, b2 = s2;
            int bi = si;
            auto add_best = [&](float v, int idx) {            // a candidate with its index: the lowest index wins ties
                if (idx == INT_MAX) return;
                if (v < b1 || (v == b1 && idx < bi)) { b2 = b1; b1 = v; bi = idx; } else b2 = fminf(b2, v);
            };
            add_best(g1, j1); b2 = fminf(b2, g2);
            add_best(of, oi); b2 = fminf(b2, of2);
            g1 = b1; g2 = b2; j1 = bi;
        }
        if (lane == 0) {
            if (budget >= 0) { o_best[out0 + ru] = j1; o_d1[out0 + ru] = g1; o_d2[out0 + ru] = use_ratio ? g2 : FLT_MAX; }
            else push_flag_f(w, p, DIR, i0 + ru);               // too many suspects: exhaustive scan
        }
    }
}

// total_rows / total_cols: rows (columns) of the chunk over all pairs (multiples of 256).
cudaError_t launch_resolve_f(const StoreView& st, const MatchWork& w, long long total_rows, long long total_cols, int use_ratio, double ratio,
                             cudaStream_t s)
{
    if (total_rows > 0) resolve_f_kernel<0><<<static_cast<unsigned>(total_rows / kResF), 256, 0, s>>>(st, w, use_ratio, ratio);
    if (total_cols > 0) resolve_f_kernel<1><<<static_cast<unsigned>(total_cols / kResF), 256, 0, s>>>(st, w, use_ratio, ratio);
    return cudaGetLastError();
}

}  // namespace g2v
