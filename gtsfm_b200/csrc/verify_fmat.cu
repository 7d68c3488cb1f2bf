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
        // fp32 copy of every matrix + the rounding-error band of its fp32 inlier test (count_kernel), u = 2^-24:
        // |delta d| <= 6 u S, |delta n| <= 3.2 u max(A1 + B1, A2 + B2); thr * 1.2e-19 covers an argument of the approximate
        // square root flushed to zero (< 2^-126).  S bounds every fp32 intermediate: when it is not far below the fp32
        // overflow threshold of a square the band is infinite and the candidate is counted entirely in double.
        const double X1 = b.cmax[4 * p], Y1 = b.cmax[4 * p + 1], X2 = b.cmax[4 * p + 2], Y2 = b.cmax[4 * p + 3];
        for (int c = 0; c < n; ++c) {
            const double* F = b.Fd + (g * kFMaxCand + c) * 9;
            const double A1 = fabs(F[0]) * X1 + fabs(F[1]) * Y1 + fabs(F[2]);
            const double B1 = fabs(F[3]) * X1 + fabs(F[4]) * Y1 + fabs(F[5]);
            const double C1 = fabs(F[6]) * X1 + fabs(F[7]) * Y1 + fabs(F[8]);
            const double A2 = fabs(F[0]) * X2 + fabs(F[3]) * Y2 + fabs(F[6]);
            const double B2 = fabs(F[1]) * X2 + fabs(F[4]) * Y2 + fabs(F[7]);
            const double S = X2 * A1 + Y2 * B1 + C1;
            const double u = 5.9604644775390625e-8;
            const double kb = (6.0 * u * S + d.thr_px * 3.2 * u * fmax(A1 + B1, A2 + B2)) * 1.01 + d.thr_px * 1.2e-19 + 1e-37;
            float Kb = __double2float_ru(kb);
            if (!(S < 1e15) || !(fmax(A1 + B1, A2 + B2) < 1e15) || !(Kb <= FLT_MAX)) Kb = __int_as_float(0x7f800000);
            // half width of an interval [c - Wh, c + Wh], c = thr_lo sn - K + Wh, that contains the band
            // [thr_lo sn - K, thr_hi sn + K] for every sn the kernel can compute (sn <= (A1 + B1) (1 + 1e-5))
            const float Wh = __double2float_ru(static_cast<double>(Kb) + 2.1e-6 * d.thr_px * (A1 + B1) * (1.0 + 1e-5));
            float4* o = reinterpret_cast<float4*>(b.Ff) + (g * kFMaxCand + c) * 3;
            o[0] = make_float4(static_cast<float>(F[0]), static_cast<float>(F[1]), static_cast<float>(F[2]), static_cast<float>(F[3]));
            o[1] = make_float4(static_cast<float>(F[4]), static_cast<float>(F[5]), static_cast<float>(F[6]), static_cast<float>(F[7]));
            o[2] = make_float4(static_cast<float>(F[8]), Kb, Wh, 0.f);
        }
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
constexpr int kCntTile = 4096;                       // correspondences per shared-memory tile (64 KB; three CTAs per SM)
constexpr int kCntCheck = 512;                       // correspondences between two early-exit checks

// packed fp32x2 arithmetic (FFMA2 / FMUL2 on sm_100a): two correspondences per lane and instruction
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk2(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float sqrt_fast(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// -1 when a <= b, else 0 (also for NaN): one FSET, so that two tests are counted by a single three-input add
__device__ __forceinline__ int set_le(float a, float b) { int r; asm("set.le.s32.f32 %0, %1, %2;" : "=r"(r) : "f"(a), "f"(b)); return r; }

struct CandF32 {
    f32x2 p[9];          // the candidate's matrix, each entry duplicated in both halves
    f32x2 tl, th;        // thr (1 - 2e-6) rounded down, thr (1 + 2e-6) rounded up
    f32x2 nk, pk;        // -K, +K: the candidate's absolute rounding-error band
    f32x2 kc;            // Wh - K: centre offset of the fast test
    float w_in, w_band;  // -Wh (1 + 1e-6): certain inlier below; Wh (1 + 2e-6): band suspected at or below (absolute value)
};

// fp32 classification of two correspondences (the packed pair at index k of the tile) against one candidate.  With
// d = x2^T F x1, n1 = |(F x1)_{0,1}|, n2 = |(F^T x2)_{0,1}| the rule "max(d^2 / n1^2, d^2 / n2^2) <= thr^2" is
// |d| <= thr min(n1, n2).  All fp32 rounding (the conversion of F, every FMA, the squares, the approximate square root
// with relative error 2^-23, its flush of subnormal arguments) is covered by the candidate's constant band K
// (solve7_kernel) and the 2e-6 relative slack of tl / th: |d| <= lo is a certain inlier, |d| > hi a certain outlier,
// anything in between is decided by fm_error_exact.
__device__ __forceinline__ void classify2(const CandF32& c, const float4* sA, const float4* sB, int k,
                                          float& da, float& db, float& loa, float& lob, float& hia, float& hib)
{
    const ulonglong2 A = *reinterpret_cast<const ulonglong2*>(sA + k);   // x1 pair, y1 pair
    const ulonglong2 B = *reinterpret_cast<const ulonglong2*>(sB + k);   // x2 pair, y2 pair
    const f32x2 a1 = fma2(c.p[0], A.x, fma2(c.p[1], A.y, c.p[2])), b1 = fma2(c.p[3], A.x, fma2(c.p[4], A.y, c.p[5])),
                c1 = fma2(c.p[6], A.x, fma2(c.p[7], A.y, c.p[8]));
    const f32x2 dd = fma2(B.x, a1, fma2(B.y, b1, c1));
    const f32x2 a2 = fma2(c.p[0], B.x, fma2(c.p[3], B.y, c.p[6])), b2 = fma2(c.p[1], B.x, fma2(c.p[4], B.y, c.p[7]));
    const f32x2 n1 = fma2(a1, a1, mul2(b1, b1)), n2 = fma2(a2, a2, mul2(b2, b2));
    float n1a, n1b, n2a, n2b;
    upk2(n1, n1a, n1b);
    upk2(n2, n2a, n2b);
    const f32x2 sn = pk2(sqrt_fast(fminf(n1a, n2a)), sqrt_fast(fminf(n1b, n2b)));
    upk2(fma2(c.tl, sn, c.nk), loa, lob);
    upk2(fma2(c.th, sn, c.pk), hia, hib);
    upk2(dd, da, db);
    da = fabsf(da); db = fabsf(db);
}

// Fast path of classify2: z = |d| - c with c the centre of an interval that contains the whole band.  z <= -Wh is a
// certain inlier (|d| <= lo); |z| > Wh a certain inlier or outlier; the smallest |z| of a lane's share tells whether
// any of its correspondences might be in the band, in which case the share is recounted with classify2.
__device__ __forceinline__ void classify2_fast(const CandF32& c, const float4* sA, const float4* sB, int k, int& cnt, float& zmin)
{
    const ulonglong2 A = *reinterpret_cast<const ulonglong2*>(sA + k);   // x1 pair, y1 pair
    const ulonglong2 B = *reinterpret_cast<const ulonglong2*>(sB + k);   // x2 pair, y2 pair
    const f32x2 a1 = fma2(c.p[0], A.x, fma2(c.p[1], A.y, c.p[2])), b1 = fma2(c.p[3], A.x, fma2(c.p[4], A.y, c.p[5])),
                c1 = fma2(c.p[6], A.x, fma2(c.p[7], A.y, c.p[8]));
    const f32x2 dd = fma2(B.x, a1, fma2(B.y, b1, c1));
    const f32x2 a2 = fma2(c.p[0], B.x, fma2(c.p[3], B.y, c.p[6])), b2 = fma2(c.p[1], B.x, fma2(c.p[4], B.y, c.p[7]));
    const f32x2 n1 = fma2(a1, a1, mul2(b1, b1)), n2 = fma2(a2, a2, mul2(b2, b2));
    float n1a, n1b, n2a, n2b, da, db, ca, cb;
    upk2(n1, n1a, n1b);
    upk2(n2, n2a, n2b);
    const f32x2 sn = pk2(sqrt_fast(fminf(n1a, n2a)), sqrt_fast(fminf(n1b, n2b)));
    upk2(fma2(c.tl, sn, c.kc), ca, cb);
    upk2(dd, da, db);
    const float za = fabsf(da) - ca, zb = fabsf(db) - cb;
    cnt = cnt - set_le(za, c.w_in) - set_le(zb, c.w_in);
    zmin = fminf(zmin, fminf(fabsf(za), fabsf(zb)));
}

constexpr int kCntPairs = kCntCheck / 2;             // packed pairs between two early-exit checks

// Shared-memory tile: correspondences 2q and 2q + 1 are stored as A[q] = (x1a, x1b, y1a, y1b), B[q] = (x2a, x2b, y2a, y2b)
// so that one LDS.128 yields two packed operands.
__global__ void __launch_bounds__(256, 3) count_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    extern __shared__ float4 s_pts[];                  // [kCntTile / 2] A, then [kCntTile / 2] B
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
    const double* Fbase = b.Fd + (static_cast<long long>(p) * d.H + h0) * (9 * kFMaxCand);
    const float4* Ffbase = reinterpret_cast<const float4*>(b.Ff) + (static_cast<long long>(p) * d.H + h0) * (3 * kFMaxCand);
    const volatile int* best = reinterpret_cast<const volatile int*>(b.best_cnt + p);
    float4* sA = s_pts;
    float4* sB = s_pts + kCntTile / 2;
    unsigned long long exec_w = 0;                     // correspondences this warp evaluated (lane 0)
    for (int t0 = 0; t0 < vp.M; t0 += kCntTile) {
        const int tn = min(kCntTile, vp.M - t0);
        const int nq = tn >> 1;                        // complete pairs of correspondences
        const bool last_tile = t0 + kCntTile >= vp.M;
        __syncthreads();
        for (int q = threadIdx.x; q < (tn + 1) >> 1; q += blockDim.x) {
            const float4 ca = static_cast<const float4*>(b.corr_f)[vp.m0 + t0 + 2 * q];
            const float4 cb = (2 * q + 1 < tn) ? static_cast<const float4*>(b.corr_f)[vp.m0 + t0 + 2 * q + 1] : ca;
            sA[q] = make_float4(ca.x, cb.x, ca.y, cb.y);
            sB[q] = make_float4(ca.z, cb.z, ca.w, cb.w);
        }
        if (threadIdx.x == 0) s_next = 0;
        __syncthreads();
        // candidates are handed out dynamically (early exits make their cost very uneven); the matrix of the next one is
        // fetched while the current one is counted
        auto grab = [&]() { int ci = 0; if (lane == 0) ci = atomicAdd(&s_next, 1); return __shfl_sync(0xffffffffu, ci, 0); };
        auto fetch = [&](int ci, float4& q0, float4& q1, float4& q2) {
            if (ci < n) {
                const int code = s_cand[ci];
                const float4* src = Ffbase + ((code >> 2) * kFMaxCand + (code & 3)) * 3;
                q0 = __ldg(src); q1 = __ldg(src + 1); q2 = __ldg(src + 2);
            }
        };
        int ci_next = grab();
        float4 q0 = make_float4(0, 0, 0, 0), q1 = q0, q2 = q0;
        fetch(ci_next, q0, q1, q2);
        for (;;) {
            const int ci = ci_next;
            if (ci >= n) break;
            const float4 g0 = q0, g1 = q1, g2 = q2;
            ci_next = grab();
            fetch(ci_next, q0, q1, q2);
            if (s_dead[ci]) continue;
            const int code = s_cand[ci];
            const double* Fp = Fbase + ((code >> 2) * kFMaxCand + (code & 3)) * 9;
            CandF32 c;
            c.p[0] = pk2(g0.x, g0.x); c.p[1] = pk2(g0.y, g0.y); c.p[2] = pk2(g0.z, g0.z); c.p[3] = pk2(g0.w, g0.w);
            c.p[4] = pk2(g1.x, g1.x); c.p[5] = pk2(g1.y, g1.y); c.p[6] = pk2(g1.z, g1.z); c.p[7] = pk2(g1.w, g1.w);
            c.p[8] = pk2(g2.x, g2.x);
            const float Kb = g2.y;
            const bool all_exact = !(Kb <= FLT_MAX);       // fp32 could overflow on this candidate: count it in double
            c.tl = pk2(thr_lo, thr_lo); c.th = pk2(thr_hi, thr_hi); c.nk = pk2(-Kb, -Kb); c.pk = pk2(Kb, Kb);
            {
                const float Wh = g2.z, kc = __fsub_rn(Wh, Kb);
                c.kc = pk2(kc, kc);
                c.w_in = -__fmul_ru(Wh, 1.000001f);
                c.w_band = __fmul_ru(Wh, 1.000002f);
            }
            // cv2's double-precision rule for the halves `bits` of the packed pair k
            auto exact = [&](int k, unsigned bits) {
                double Fd9[9];
#pragma unroll
                for (int e = 0; e < 9; ++e) Fd9[e] = __ldg(Fp + e);
                const float4 a4 = sA[k], b4 = sB[k];
                int add = 0;
                if ((bits & 1u) && fm_error_exact(Fd9, a4.x, a4.z, b4.x, b4.z) <= thr2) ++add;
                if ((bits & 2u) && fm_error_exact(Fd9, a4.y, a4.w, b4.y, b4.w) <= thr2) ++add;
                return add;
            };
            const int before = s_cnt[ci];
            int mine = 0, done = 0, total_w = 0;
            bool dead = false;
            for (int c0 = 0; c0 < nq; c0 += kCntPairs) {
                const int c1 = min(nq, c0 + kCntPairs);
                const int bestv = *best;                   // issued here, consumed after the segment
                // certain inliers of the lane's share of the segment and its closest approach to the band; a correspondence
                // that may be inside the band (a few per million) makes the lane recount its share, with the
                // double-precision rule for the correspondences that are
                int lo = 0;
                float zmin = FLT_MAX;
                if (!all_exact) {
                    int k = c0 + lane;
#pragma unroll 4
                    for (; k < c1; k += 32) classify2_fast(c, sA, sB, k, lo, zmin);
                }
                if (zmin <= c.w_band || all_exact) {
                    lo = 0;
#pragma unroll 1
                    for (int k = c0 + lane; k < c1; k += 32) {
                        float da, db, loa, lob, hia, hib;
                        unsigned bits = 3u;
                        if (!all_exact) {
                            classify2(c, sA, sB, k, da, db, loa, lob, hia, hib);
                            const bool ina = da <= loa, inb = db <= lob;
                            lo += (ina ? 1 : 0) + (inb ? 1 : 0);
                            bits = ((ina || da > hia) ? 0u : 1u) | ((inb || db > hib) ? 0u : 2u);
                        }
                        if (bits) lo += exact(k, bits);
                    }
                }
                mine += lo;
                done = c1;
                // exact early exit: even if every remaining correspondence were an inlier the candidate would stay below
                // the best complete count of the pair (equal counts are kept: the first found wins)
                total_w = __reduce_add_sync(0xffffffffu, mine);
                if (before + total_w + (vp.M - t0 - 2 * c1) < bestv) { dead = true; break; }
            }
            if (!dead && (tn & 1) && lane == 0) {      // odd tail: first half of the last (duplicated) pair, decided exactly
                total_w += exact(nq, 1u);
            }
            if (lane == 0) {
                s_cnt[ci] = before + total_w;
                if (dead) s_dead[ci] = 1;
                else if (last_tile) atomicMax(b.best_cnt + p, before + total_w);
                exec_w += static_cast<unsigned long long>(2 * done + (dead ? 0 : (tn & 1)));
            }
        }
    }
    if (lane == 0) atomicAdd(&s_exec, exec_w);
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
    __shared__ double4 s_q[kFinThreads / 32][64];      // recover_pose_block: per-warp inlier queue
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
    recover_pose_block(b, vp, p, E, s_pose, s_red, s_out, s_q);
}

// ---------------------------------------------------------------------------------------------------------------- launch
// step: 0 gather, 1 solve7, 2 count, 3 finalize
cudaError_t launch_verify_step_f(int step, const StoreView& st, const VerifyBufs& b, const VerifyDims& d, int n_pairs, int max_m, cudaStream_t s)
{
    if (n_pairs <= 0) return cudaSuccess;
    static g2v::DeviceOnce once;
    int& configured = once.cur();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kCntTile * 16);
        if (e != cudaSuccess) return e;
        configured = 1;
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
