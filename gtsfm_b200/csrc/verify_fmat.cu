// Fundamental-matrix mode of the Ransac verifier, batched over pairs: Ransac(use_intrinsics_in_verification=False)
//   Ransac.estimate_F -> cv2.findFundamentalMat(FM_RANSAC, thr px)      gtsfm/frontend/verifier/ransac.py:84-111
//   E = K2^T F K1, then the same pose recovery as the E mode             opencv_verifier_base.py:90-108, utils/verification.py:97-110
//
//   K3F gather_pixels   match rows -> float32 pixel coordinates (what cv2 works on) + calibrated fp64 copy (pose recovery)
//                       + per-pair coordinate maxima (the error bound of the fp32 inlier test)
//   K4F solve7          one 7-point hypothesis per thread, fp64: null space of the 7x9 system, det(l N0 + (1-l) N1) = 0
//   K5F count           inlier count of every candidate over all correspondences.  cv2's rule is evaluated in double on
//                       float32 points, rounded to float32 and compared with thr^2; here a point is classified by an fp32
//                       evaluation with a RIGOROUS rounding-error band and only the points inside the band (a few per
//                       million) take the double path, so the counts are exact.  Exact early exit: a candidate that can no
//                       longer reach the pair's best complete count is dropped.
//   K6F finalize_F      arg max count (first found among equals, as cv2), mask, E = K2^T F K1, pose recovery
#include <cfloat>

#include "store.h"
#include "verify_block.cuh"
#include "verify_types.h"

namespace g2v {

constexpr int kFMaxCand = 3;

// ---------------------------------------------------------------------------------------------------------------- K3F
__global__ void __launch_bounds__(256) gather_pixels_kernel(StoreView st, VerifyBufs b, int n_pairs)
{
    const int p = blockIdx.y;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    float mx[4] = {0.f, 0.f, 0.f, 0.f};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < vp.M; i += gridDim.x * blockDim.x) {
        const uint2 m = static_cast<const uint2*>(b.match)[vp.m0 + i];
        const double2 k1 = st.kps[vp.row1 + m.x], k2 = st.kps[vp.row2 + m.y];
        double x1, y1, x2, y2;
        vm::calibrate(k1.x, k1.y, vp.cal1, &x1, &y1);
        vm::calibrate(k2.x, k2.y, vp.cal2, &x2, &y2);
        static_cast<double4*>(b.corr_d)[vp.m0 + i] = make_double4(x1, y1, x2, y2);
        // cv2 converts the (M, 2) float64 arrays to float32 before the estimation
        const float4 f = make_float4(static_cast<float>(k1.x), static_cast<float>(k1.y), static_cast<float>(k2.x), static_cast<float>(k2.y));
        static_cast<float4*>(b.corr_f)[vp.m0 + i] = f;
        mx[0] = fmaxf(mx[0], fabsf(f.x)); mx[1] = fmaxf(mx[1], fabsf(f.y)); mx[2] = fmaxf(mx[2], fabsf(f.z)); mx[3] = fmaxf(mx[3], fabsf(f.w));
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float v = mx[k];
        if (!(v <= FLT_MAX)) v = FLT_MAX;            // NaN / inf coordinates: the bound degenerates and every test goes exact
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
        if ((threadIdx.x & 31) == 0 && v > 0.f) atomicMax(reinterpret_cast<unsigned int*>(b.cmax + 4 * p + k), __float_as_uint(v));
    }
}

// ---------------------------------------------------------------------------------------------------------------- K4F
constexpr int kF7Threads = 64;
constexpr int kF7Doubles = 63;                       // the 7x9 system, element i of the calling thread at base[i * kF7Threads]
constexpr int kF7Smem = kF7Doubles * kF7Threads * 8;

struct Scr7 {
    double* base;
    __device__ __forceinline__ double& operator()(int i) const { return base[i * kF7Threads]; }
};

__device__ __forceinline__ double det3c(const double* a, const double* b, const double* c)   // det of the matrix with COLUMNS a, b, c
{
    return a[0] * (b[1] * c[2] - b[2] * c[1]) - b[0] * (a[1] * c[2] - a[2] * c[1]) + c[0] * (a[1] * b[2] - a[2] * b[1]);
}

__device__ __forceinline__ double poly3(const double* p, int deg, double z)
{
    double v = p[3];
    if (deg < 3) v = 0.0;
    v = v * z + p[2];
    v = v * z + p[1];
    return v * z + p[0];
}

// One root of p in (lo, hi) with p(lo) p(hi) < 0: Newton from the midpoint, bisection whenever the step leaves the
// bracket or converges too slowly (the oracle's loop, oracle_verify.c real_roots_sturm).
__device__ inline double root_in_bracket(const double* p, int deg, double lo, double hi, double flo)
{
    double x = 0.5 * (lo + hi), dxold = hi - lo;
    for (int it = 0; it < 200; ++it) {
        const double f = poly3(p, deg, x);
        if (f == 0.0) break;
        if ((f > 0) == (flo > 0)) lo = x; else hi = x;
        const double dp = (3.0 * (deg == 3 ? p[3] : 0.0) * x + 2.0 * p[2]) * x + p[1];
        double xn = (dp != 0.0) ? x - f / dp : x;
        if (!(xn > lo && xn < hi) || fabs(xn - x) > 0.5 * fabs(dxold)) xn = 0.5 * (lo + hi);
        dxold = xn - x;
        if (fabs(dxold) <= 4e-16 * (1.0 + fabs(x))) { x = xn; break; }
        x = xn;
    }
    return x;
}

// Real roots (ascending) of a polynomial of degree <= 3; the normalisation, the degree test and the root bound follow the
// oracle's general routine; the roots are isolated by the critical points instead of a Sturm chain.
__device__ inline int cubic_real_roots(const double* pin, double* roots)
{
    double p[4];
    double mx = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) mx = fmax(mx, fabs(pin[i]));
    if (!(mx > 0) || !isfinite(mx)) return 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) p[i] = pin[i] / mx;
    int deg = 3;
    if (fabs(p[3]) < 1e-13) { deg = 2; if (fabs(p[2]) < 1e-13) { deg = 1; if (fabs(p[1]) < 1e-13) return 0; } }
    if (deg == 1) { roots[0] = -p[0] / p[1]; return 1; }
    const double lead = (deg == 3) ? p[3] : p[2];
    // Fujiwara bound 2 max_k |a_{n-k} / a_n|^(1/k)
    double B;
    if (deg == 3) B = fmax(fabs(p[2] / lead), fmax(sqrt(fabs(p[1] / lead)), cbrt(fabs(p[0] / lead))));
    else B = fmax(fabs(p[1] / lead), sqrt(fabs(p[0] / lead)));
    B = 2.0 * B + 1e-9;
    double brk[4];
    int nb = 0;
    brk[nb++] = -B;
    if (deg == 3) {
        const double disc = p[2] * p[2] - 3.0 * p[3] * p[1];                  // quarter discriminant of p'
        if (disc > 0) {
            const double sq = sqrt(disc), tq = -(p[2] + (p[2] >= 0 ? sq : -sq));
            double ca = tq / (3.0 * p[3]), cb = (tq != 0.0) ? p[1] / tq : ca;
            if (ca > cb) { const double t = ca; ca = cb; cb = t; }
            ca = fmin(fmax(ca, -B), B); cb = fmin(fmax(cb, -B), B);
            brk[nb++] = ca; brk[nb++] = cb;
        }
    } else {
        const double c = fmin(fmax(-p[1] / (2.0 * p[2]), -B), B);
        brk[nb++] = c;
    }
    brk[nb++] = B;
    int nr = 0;
    double flo = poly3(p, deg, brk[0]);
    for (int k = 0; k + 1 < nb && nr < 3; ++k) {
        const double lo = brk[k], hi = brk[k + 1], fhi = poly3(p, deg, hi);
        if (fhi == 0.0) { if (hi > lo) roots[nr++] = hi; }
        else if (flo != 0.0 && (flo > 0) != (fhi > 0)) roots[nr++] = root_in_bracket(p, deg, lo, hi, flo);
        flo = fhi;
    }
    return nr;
}

// cv::FMEstimatorCallback::run7Point restated (oracle_verify.c oracle_sevenpt): returns the number of matrices written to
// `out` (global memory, 9 doubles each, unit Frobenius norm, ascending root).
__device__ inline int sevenpt_solve(const double* x1, const double* y1, const double* x2, const double* y2, const Scr7& A, double* out)
{
#pragma unroll
    for (int r = 0; r < 7; ++r) {
        A(r * 9 + 0) = x2[r] * x1[r]; A(r * 9 + 1) = x2[r] * y1[r]; A(r * 9 + 2) = x2[r];
        A(r * 9 + 3) = y2[r] * x1[r]; A(r * 9 + 4) = y2[r] * y1[r]; A(r * 9 + 5) = y2[r];
        A(r * 9 + 6) = x1[r];         A(r * 9 + 7) = y1[r];         A(r * 9 + 8) = 1.0;
    }
    unsigned long long perm = 0x876543210ull;        // nibble c = original column now at position c
    for (int k = 0; k < 7; ++k) {
        int pr = k, pc = k;
        double best = -1.0;
        for (int r = k; r < 7; ++r)
            for (int c = k; c < 9; ++c) { const double v = fabs(A(r * 9 + c)); if (v > best) { best = v; pr = r; pc = c; } }
        if (!(best > 1e-14)) return 0;
        if (pr != k)
            for (int c = 0; c < 9; ++c) { const double t = A(k * 9 + c); A(k * 9 + c) = A(pr * 9 + c); A(pr * 9 + c) = t; }
        if (pc != k) {
            for (int r = 0; r < 7; ++r) { const double t = A(r * 9 + k); A(r * 9 + k) = A(r * 9 + pc); A(r * 9 + pc) = t; }
            const unsigned long long a = (perm >> (4 * k)) & 15ull, bq = (perm >> (4 * pc)) & 15ull;
            perm &= ~((15ull << (4 * k)) | (15ull << (4 * pc)));
            perm |= (bq << (4 * k)) | (a << (4 * pc));
        }
        const double inv = 1.0 / A(k * 9 + k);
        double rowk[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) {
            rowk[c] = 0.0;
            if (c >= k) { rowk[c] = A(k * 9 + c) * inv; A(k * 9 + c) = rowk[c]; }
        }
        for (int r = 0; r < 7; ++r) {
            if (r == k) continue;
            const double f = A(r * 9 + k);
            if (f != 0.0) {
#pragma unroll
                for (int c = 0; c < 9; ++c) if (c >= k) A(r * 9 + c) -= f * rowk[c];
            }
        }
    }
    double N0[9], N1[9];
    {
        double v0[7], v1[7];
        double n0 = 1.0, n1 = 1.0;
#pragma unroll
        for (int r = 0; r < 7; ++r) { v0[r] = -A(r * 9 + 7); v1[r] = -A(r * 9 + 8); n0 += v0[r] * v0[r]; n1 += v1[r] * v1[r]; }
        n0 = 1.0 / sqrt(n0); n1 = 1.0 / sqrt(n1);
        // un-permute through the scratch (dynamic index), then read back with static indices
#pragma unroll
        for (int c = 0; c < 9; ++c) {
            const int o = static_cast<int>((perm >> (4 * c)) & 15ull);
            A(o) = (c < 7 ? v0[c < 7 ? c : 0] : (c == 7 ? 1.0 : 0.0)) * n0;
            A(9 + o) = (c < 7 ? v1[c < 7 ? c : 0] : (c == 8 ? 1.0 : 0.0)) * n1;
        }
#pragma unroll
        for (int i = 0; i < 9; ++i) { N0[i] = A(i); N1[i] = A(9 + i); }
    }
    double Ad[9], a[3][3], bm[3][3];
#pragma unroll
    for (int i = 0; i < 9; ++i) Ad[i] = N0[i] - N1[i];                        // F(l) = l Ad + N1
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i) { a[j][i] = Ad[3 * i + j]; bm[j][i] = N1[3 * i + j]; }
    double p[4], roots[3];
    p[3] = det3c(a[0], a[1], a[2]);
    p[2] = det3c(bm[0], a[1], a[2]) + det3c(a[0], bm[1], a[2]) + det3c(a[0], a[1], bm[2]);
    p[1] = det3c(a[0], bm[1], bm[2]) + det3c(bm[0], a[1], bm[2]) + det3c(bm[0], bm[1], a[2]);
    p[0] = det3c(bm[0], bm[1], bm[2]);
    const int nr = cubic_real_roots(p, roots);
    int n = 0;
    for (int k = 0; k < nr && n < kFMaxCand; ++k) {
        double F[9], nn = 0;
#pragma unroll
        for (int i = 0; i < 9; ++i) { F[i] = roots[k] * Ad[i] + N1[i]; nn += F[i] * F[i]; }
        if (!(nn > 0) || !isfinite(nn)) continue;
        nn = 1.0 / sqrt(nn);
#pragma unroll
        for (int i = 0; i < 9; ++i) out[9 * n + i] = F[i] * nn;
        ++n;
    }
    return n;
}

__global__ void __launch_bounds__(kF7Threads) solve7_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    extern __shared__ double s_a[];
    const long long total = static_cast<long long>(n_pairs) * d.H;
    const long long g = static_cast<long long>(blockIdx.x) * kF7Threads + threadIdx.x;
    if (g >= total) return;
    const int p = static_cast<int>(g / d.H), h = static_cast<int>(g - static_cast<long long>(p) * d.H);
    const VPair& vp = b.pairs[p];
    int n = 0;
    if (vp.status == 0) {
        int s[7];
        if (b.samples) {
            const int* src = b.samples + (static_cast<long long>(p) * d.H + h) * 7;
#pragma unroll
            for (int k = 0; k < 7; ++k) s[k] = src[k];
        } else {
            sample_k<7>(d.seed, vp.key, static_cast<uint32_t>(h), static_cast<uint32_t>(vp.M), s);
        }
        double x1[7], y1[7], x2[7], y2[7];
        const float4* corr = static_cast<const float4*>(b.corr_f) + vp.m0;
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            const float4 c = corr[min(max(s[k], 0), vp.M - 1)];
            x1[k] = c.x; y1[k] = c.y; x2[k] = c.z; y2[k] = c.w;
        }
        n = sevenpt_solve(x1, y1, x2, y2, Scr7{s_a + threadIdx.x}, b.Fd + g * (9 * kFMaxCand));
    }
    b.nroot[g] = static_cast<unsigned char>(n);
    if (b.stats) {
        const unsigned act = __activemask();
        const int nv = __reduce_add_sync(act, n);
        if ((threadIdx.x & 31) == __ffs(act) - 1 && nv) atomicAdd(b.stats, static_cast<unsigned long long>(nv));
        if (threadIdx.x == 0) {
            const long long left = total - static_cast<long long>(blockIdx.x) * kF7Threads;
            atomicAdd(b.stats + 1, static_cast<unsigned long long>(left < kF7Threads ? left : kF7Threads));
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------- K5F
// cv::FMEstimatorCallback::computeError: double arithmetic on float32 points, separately rounded operations, the result
// rounded to float32 (oracle_verify.c oracle_fm_error).
__device__ __forceinline__ float fm_error_exact(const double* F, float fx1, float fy1, float fx2, float fy2)
{
    const double x1 = fx1, y1 = fy1, x2 = fx2, y2 = fy2;
    double a = __dadd_rn(__dadd_rn(__dmul_rn(F[0], x1), __dmul_rn(F[1], y1)), F[2]);
    double bb = __dadd_rn(__dadd_rn(__dmul_rn(F[3], x1), __dmul_rn(F[4], y1)), F[5]);
    double c = __dadd_rn(__dadd_rn(__dmul_rn(F[6], x1), __dmul_rn(F[7], y1)), F[8]);
    const double s2 = __ddiv_rn(1.0, __dadd_rn(__dmul_rn(a, a), __dmul_rn(bb, bb)));
    const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(x2, a), __dmul_rn(y2, bb)), c);
    a = __dadd_rn(__dadd_rn(__dmul_rn(F[0], x2), __dmul_rn(F[3], y2)), F[6]);
    bb = __dadd_rn(__dadd_rn(__dmul_rn(F[1], x2), __dmul_rn(F[4], y2)), F[7]);
    c = __dadd_rn(__dadd_rn(__dmul_rn(F[2], x2), __dmul_rn(F[5], y2)), F[8]);
    const double s1 = __ddiv_rn(1.0, __dadd_rn(__dmul_rn(a, a), __dmul_rn(bb, bb)));
    const double d1 = __dadd_rn(__dadd_rn(__dmul_rn(x1, a), __dmul_rn(y1, bb)), c);
    const double e1 = __dmul_rn(__dmul_rn(d1, d1), s1), e2 = __dmul_rn(__dmul_rn(d2, d2), s2);
    return static_cast<float>(e1 > e2 ? e1 : e2);
}

constexpr int kCntHyp = 128;                         // hypotheses per CTA (the staged points are shared by their candidates)
constexpr int kCntMaxCand = kCntHyp * kFMaxCand;
constexpr int kCntTile = 4096;                       // correspondences per shared-memory tile (64 KB)
constexpr int kCntCheck = 512;                       // correspondences between two early-exit checks

// fp32 classification of one correspondence against one candidate.  With d = x2^T F x1, n1 = |(F x1)_{0,1}|,
// n2 = |(F^T x2)_{0,1}| the rule "max(d^2 / n1^2, d^2 / n2^2) <= thr^2" is |d| <= thr min(n1, n2).  All fp32 rounding
// (the conversion of F, every FMA, the squares, the square root) is covered by the candidate's constant band K
// (count_kernel) and the 2e-6 relative slack of thr_lo / thr_hi, so "in" and "out" are certain; anything else is decided
// by fm_error_exact.
__global__ void __launch_bounds__(256) count_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    extern __shared__ float4 s_pts[];                  // [kCntTile] x1 y1 x2 y2 (float32 pixels)
    __shared__ int s_cnt[kCntMaxCand];
    __shared__ unsigned short s_cand[kCntMaxCand];     // (local hypothesis << 2) | root
    __shared__ unsigned char s_dead[kCntMaxCand];
    __shared__ unsigned long long s_exec;
    __shared__ int s_n, s_next;
    __shared__ int s_wcnt[kCntHyp / 32];
    const int p = blockIdx.y;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    if (vp.status != 0) return;
    const int h0 = blockIdx.x * kCntHyp;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int nr = 0, incl = 0;
    if (threadIdx.x < kCntHyp) {
        nr = (h0 + threadIdx.x < d.H) ? b.nroot[static_cast<long long>(p) * d.H + h0 + threadIdx.x] : 0;
        incl = nr;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += v; }
        if (lane == 31) s_wcnt[warp] = incl;
    }
    for (int k = threadIdx.x; k < kCntMaxCand; k += blockDim.x) { s_cnt[k] = 0; s_dead[k] = 0; }
    if (threadIdx.x == 0) { s_exec = 0ull; s_next = 0; }
    __syncthreads();
    if (threadIdx.x < kCntHyp) {
        int pos = incl - nr;
        for (int w2 = 0; w2 < warp; ++w2) pos += s_wcnt[w2];
        for (int c = 0; c < nr; ++c) s_cand[pos++] = static_cast<unsigned short>((threadIdx.x << 2) | c);
        if (threadIdx.x == kCntHyp - 1) s_n = pos;
    }
    __syncthreads();
    const int n = s_n;
    if (n == 0) return;
    if (threadIdx.x == 0 && b.stats) atomicAdd(b.stats + 2, static_cast<unsigned long long>(n) * static_cast<unsigned long long>(vp.M));
    const double thr = d.thr_px, thr2 = thr * thr;
    const float thr_lo = __double2float_rd(thr * (1.0 - 2e-6)), thr_hi = __double2float_ru(thr * (1.0 + 2e-6));
    const double X1 = b.cmax[4 * p], Y1 = b.cmax[4 * p + 1], X2 = b.cmax[4 * p + 2], Y2 = b.cmax[4 * p + 3];
    const double* Fbase = b.Fd + (static_cast<long long>(p) * d.H + h0) * (9 * kFMaxCand);
    const volatile int* best = reinterpret_cast<const volatile int*>(b.best_cnt + p);
    for (int t0 = 0; t0 < vp.M; t0 += kCntTile) {
        const int tn = min(kCntTile, vp.M - t0);
        const bool last_tile = t0 + kCntTile >= vp.M;
        __syncthreads();
        for (int q = threadIdx.x; q < tn; q += blockDim.x) s_pts[q] = static_cast<const float4*>(b.corr_f)[vp.m0 + t0 + q];
        if (threadIdx.x == 0) s_next = 0;
        __syncthreads();
        for (;;) {
            int ci = 0;
            if (lane == 0) ci = atomicAdd(&s_next, 1);
            ci = __shfl_sync(0xffffffffu, ci, 0);
            if (ci >= n) break;
            if (s_dead[ci]) continue;
            const int code = s_cand[ci];
            const double* Fp = Fbase + ((code >> 2) * kFMaxCand + (code & 3)) * 9;
            float f[9], Kb;
            {
                double Fd9[9];
#pragma unroll
                for (int e = 0; e < 9; ++e) { Fd9[e] = __ldg(Fp + e); f[e] = static_cast<float>(Fd9[e]); }
                // rounding-error band of the fp32 test (u = 2^-24): |delta d| <= 6 u S, |delta n| <= 3.2 u max(A1 + B1, A2 + B2)
                const double A1 = fabs(Fd9[0]) * X1 + fabs(Fd9[1]) * Y1 + fabs(Fd9[2]);
                const double B1 = fabs(Fd9[3]) * X1 + fabs(Fd9[4]) * Y1 + fabs(Fd9[5]);
                const double C1 = fabs(Fd9[6]) * X1 + fabs(Fd9[7]) * Y1 + fabs(Fd9[8]);
                const double A2 = fabs(Fd9[0]) * X2 + fabs(Fd9[3]) * Y2 + fabs(Fd9[6]);
                const double B2 = fabs(Fd9[1]) * X2 + fabs(Fd9[4]) * Y2 + fabs(Fd9[7]);
                const double S = X2 * A1 + Y2 * B1 + C1;
                const double u = 5.9604644775390625e-8;
                const double kb = (6.0 * u * S + thr * 3.2 * u * fmax(A1 + B1, A2 + B2)) * 1.01 + 1e-37;
                Kb = __double2float_ru(kb);
                if (!(Kb <= FLT_MAX)) Kb = FLT_MAX;
            }
            const int before = s_cnt[ci];
            int mine = 0, done = 0, total_w = 0;
            bool dead = false;
            for (int c0 = 0; c0 < tn; c0 += kCntCheck) {
                const int c1 = min(tn, c0 + kCntCheck);
                const int bestv = *best;                   // issued here, consumed after the segment
#pragma unroll 2
                for (int k = c0 + lane; k < c1; k += 32) {
                    const float4 c = s_pts[k];
                    const float a1 = fmaf(f[0], c.x, fmaf(f[1], c.y, f[2])), b1 = fmaf(f[3], c.x, fmaf(f[4], c.y, f[5])),
                                c1v = fmaf(f[6], c.x, fmaf(f[7], c.y, f[8]));
                    const float dd = fabsf(fmaf(c.z, a1, fmaf(c.w, b1, c1v)));
                    const float a2 = fmaf(f[0], c.z, fmaf(f[3], c.w, f[6])), b2 = fmaf(f[1], c.z, fmaf(f[4], c.w, f[7]));
                    const float n1 = fmaf(a1, a1, b1 * b1), n2 = fmaf(a2, a2, b2 * b2);
                    const float sn = sqrtf(fminf(n1, n2));
                    const bool in = dd <= fmaf(thr_lo, sn, -Kb);
                    const bool out = dd > fmaf(thr_hi, sn, Kb);
                    if (in) ++mine;
                    else if (!out) {                       // inside the band (or NaN): cv2's double-precision rule decides
                        double Fd9[9];
#pragma unroll
                        for (int e = 0; e < 9; ++e) Fd9[e] = __ldg(Fp + e);
                        if (fm_error_exact(Fd9, c.x, c.y, c.z, c.w) <= thr2) ++mine;
                    }
                }
                done = c1;
                // exact early exit: even if every remaining correspondence were an inlier the candidate would stay below
                // the best complete count of the pair (equal counts are kept: the first found wins)
                total_w = __reduce_add_sync(0xffffffffu, mine);
                if (before + total_w + (vp.M - t0 - c1) < bestv) { dead = true; break; }
            }
            if (lane == 0) {
                s_cnt[ci] = before + total_w;
                if (dead) s_dead[ci] = 1;
                else if (last_tile) atomicMax(b.best_cnt + p, before + total_w);
                atomicAdd(&s_exec, static_cast<unsigned long long>(done));
            }
        }
    }
    __syncthreads();
    for (int ci = threadIdx.x; ci < n; ci += blockDim.x) {
        const int code = s_cand[ci];
        b.cnt[(static_cast<long long>(p) * d.H + h0 + (code >> 2)) * kFMaxCand + (code & 3)] = s_dead[ci] ? -1 : s_cnt[ci];
    }
    if (threadIdx.x == 0 && b.stats) atomicAdd(b.stats + 3, s_exec);
}

// ---------------------------------------------------------------------------------------------------------------- K6F
__global__ void __launch_bounds__(kFinThreads, 4) finalize_f_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    __shared__ double s_red[kFinThreads / 32][24];
    __shared__ double s_out[24];
    __shared__ double s_pose[21];
    __shared__ double s_E[9];
    __shared__ long long s_key[kFinThreads / 32];
    const int p = blockIdx.x;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < vp.M; i += kFinThreads) b.mask[vp.m0 + i] = 0;
    if (tid < 9) { b.R[p * 9 + tid] = 0; b.E[p * 9 + tid] = 0; }
    if (tid < 3) b.t[p * 3 + tid] = 0;
    if (tid == 0) { b.ninl[p] = 0; b.status[p] = vp.status; }
    if (vp.status != 0) return;

    // ---- the model with the most inliers, the first found among equal counts; cv2 accepts a model only when it has more
    // inliers than max(previous best, modelPoints - 1 = 6)
    long long key = -1;
    const int* cnt = b.cnt + static_cast<long long>(p) * d.H * kFMaxCand;
    const unsigned char* nroot = b.nroot + static_cast<long long>(p) * d.H;
    for (int h = tid; h < d.H; h += kFinThreads) {
        const int nr = nroot[h];
        for (int c = 0; c < nr; ++c) {
            const int v = cnt[h * kFMaxCand + c];
            if (v > 6) {
                const long long k2 = (static_cast<long long>(v) << 32) | static_cast<long long>(0x7fffffff - (h * kFMaxCand + c));
                key = k2 > key ? k2 : key;
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) { const long long o = __shfl_xor_sync(0xffffffffu, key, off); key = o > key ? o : key; }
    if (lane == 0) s_key[warp] = key;
    __syncthreads();
    for (int w = 0; w < kFinThreads / 32; ++w) key = s_key[w] > key ? s_key[w] : key;
    if (key < 0) { if (tid == 0) b.status[p] = 2; return; }
    const int best_id = 0x7fffffff - static_cast<int>(key & 0x7fffffffll);
    double F[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) F[e] = b.Fd[(static_cast<long long>(p) * d.H * kFMaxCand + best_id) * 9 + e];

    // ---- mask: cv2's error of the winning model <= thr^2
    const double thr2 = d.thr_px * d.thr_px;
    {
        double c1[1] = {0};
        for (int i = tid; i < vp.M; i += kFinThreads) {
            const float4 c = static_cast<const float4*>(b.corr_f)[vp.m0 + i];
            const bool in = fm_error_exact(F, c.x, c.y, c.z, c.w) <= thr2;
            b.mask[vp.m0 + i] = in ? 1 : 0;
            c1[0] += in ? 1.0 : 0.0;
        }
        block_sum<1>(c1, s_red, s_out);
    }
    const int n_in = static_cast<int>(s_out[0]);
    // ---- E = K2^T F K1 (utils/verification.py:110), K = [f 0 u0; 0 f v0; 0 0 1]
    if (tid == 0) {
        const double K1[9] = {vp.cal1[0], 0, vp.cal1[3], 0, vp.cal1[0], vp.cal1[4], 0, 0, 1};
        const double K2t[9] = {vp.cal2[0], 0, 0, 0, vp.cal2[0], 0, vp.cal2[3], vp.cal2[4], 1};
        double T[9];
        vm::mat3_mul(K2t, F, T);
        vm::mat3_mul(T, K1, s_E);
        b.ninl[p] = n_in;
    }
    __syncthreads();
    if (tid < 9) b.E[p * 9 + tid] = s_E[tid];
    if (n_in == 0) { if (tid == 0) b.status[p] = 2; return; }
    double E[9];
#pragma unroll
    for (int e = 0; e < 9; ++e) E[e] = s_E[e];
    recover_pose_block(b, vp, p, E, s_pose, s_red, s_out);
}

// ---------------------------------------------------------------------------------------------------------------- launch
// step: 0 gather, 1 solve7, 2 count, 3 finalize
cudaError_t launch_verify_step_f(int step, const StoreView& st, const VerifyBufs& b, const VerifyDims& d, int n_pairs, int max_m, cudaStream_t s)
{
    if (n_pairs <= 0) return cudaSuccess;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kCntTile * 16);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    if (step == 0) {
        const int gx = max_m > 0 ? (max_m + 255) / 256 : 1;
        gather_pixels_kernel<<<dim3(gx, n_pairs), 256, 0, s>>>(st, b, n_pairs);
    } else if (step == 1) {
        const long long nb = (static_cast<long long>(n_pairs) * d.H + kF7Threads - 1) / kF7Threads;
        solve7_kernel<<<static_cast<unsigned>(nb), kF7Threads, kF7Smem, s>>>(b, d, n_pairs);
    } else if (step == 2) {
        count_kernel<<<dim3((d.H + kCntHyp - 1) / kCntHyp, n_pairs), 256, kCntTile * 16, s>>>(b, d, n_pairs);
    } else {
        finalize_f_kernel<<<n_pairs, kFinThreads, 0, s>>>(b, d, n_pairs);
    }
    return cudaGetLastError();
}

}  // namespace g2v
