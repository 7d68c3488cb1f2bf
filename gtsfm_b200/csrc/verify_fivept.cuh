// Nister's five-point minimal solver (Nister, TPAMI 2004 -- the algorithm cited at gtsfm/frontend/verifier/ransac.py:6,
// executed in the reference by cv2.findEssentialMat at ransac.py:74), one hypothesis per thread, fp64.
//
// Pipeline: null space of the 5x9 epipolar system (Gauss-Jordan, complete pivoting) -> ten cubic constraints in
// (x, y, z) on E = x N0 + y N1 + z N2 + N3 -> Gauss-Jordan on the ten highest monomials -> degree-10 polynomial in z
// -> real roots by Sturm-sequence isolation + safeguarded Newton -> (x, y) from the 3x3 null vector -> Gauss-Newton
// polish on the ten constraints.  Output: up to 10 unit-Frobenius-norm essential matrices in ascending z.
// The CPU oracle holds an independent C restatement that is pinned against cv2's own 5-point solver.
#pragma once
#include <cstddef>

#include "verify_math.cuh"

namespace vm {

constexpr int kMaxCand = 10;

// Per-thread scratch of kWorkDoubles doubles: element i of the calling thread sits at base[i * STRIDE].  The solve kernel
// keeps it in shared memory ([element][thread], STRIDE = threads per CTA: conflict-free and out of the L1/L2 path that
// thread-local arrays take); the finalize kernel does the same for its few re-solving threads.
constexpr int kWorkDoubles = 200;
template <int STRIDE> struct Work {
    double* base;
    __device__ __forceinline__ double& operator()(int i) const { return base[i * STRIDE]; }
};

// Monomial index tables.  They are constexpr functions so that, after full unrolling, every accumulator index is a
// compile-time constant and the accumulators live in registers (a table in constant memory makes every "+=" a
// dependent local-memory load / store pair).
__host__ __device__ constexpr int idx11(int i, int j)
{
    constexpr int t[4][4] = {{0, 3, 4, 6}, {3, 1, 5, 7}, {4, 5, 2, 8}, {6, 7, 8, 9}};
    return t[i][j];
}
__host__ __device__ constexpr int idx21(int i, int j)
{
    constexpr int t[10][4] = {{0, 2, 4, 5},   {3, 1, 6, 7},   {10, 13, 16, 17}, {2, 3, 8, 9},     {4, 8, 10, 11},
                              {8, 6, 13, 14}, {5, 9, 11, 12}, {9, 7, 14, 15},   {11, 14, 17, 18}, {12, 15, 18, 19}};
    return t[i][j];
}

// c (degree 2, 10 coefficients) += +-a (degree 1) * b (degree 1);  monomials x^2 y^2 z^2 xy xz yz x y z 1
template <bool NEG> __device__ __forceinline__ void p11_acc(const double* a, const double* b, double* c)
{
    double av[4], bv[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { av[i] = NEG ? -a[i] : a[i]; bv[i] = b[i]; }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[idx11(i, j)] = fma(av[i], bv[j], c[idx11(i, j)]);
}
// c (degree 3, 20 coefficients) += +-a (degree 2) * b (degree 1)
template <bool NEG> __device__ __forceinline__ void p21_acc(const double* a, const double* b, double* c)
{
    double bv[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) bv[j] = NEG ? -b[j] : b[j];
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const double ai = a[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) c[idx21(i, j)] = fma(ai, bv[j], c[idx21(i, j)]);
    }
}

// Null space of the 5x9 epipolar system by Gauss-Jordan with complete pivoting.  The matrix lives in the caller's
// scratch (element i at wk(i)): with the [element][thread] shared-memory layout a per-thread row / column index costs
// nothing (every thread stays in its own bank pair), whereas a thread-local array indexed that way scatters a warp's
// access over up to 64 cache lines.
template <class W>
__device__ inline bool nullspace_5x9(const double* x1, const double* y1, const double* x2, const double* y2, double N[4][9], const W& wk)
{
    // A[r][c] = wk(r * 9 + c); the null vectors are assembled in wk(45 + j * 9 + c)
    int perm[9];
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        wk(r * 9 + 0) = x2[r] * x1[r]; wk(r * 9 + 1) = x2[r] * y1[r]; wk(r * 9 + 2) = x2[r];
        wk(r * 9 + 3) = y2[r] * x1[r]; wk(r * 9 + 4) = y2[r] * y1[r]; wk(r * 9 + 5) = y2[r];
        wk(r * 9 + 6) = x1[r];         wk(r * 9 + 7) = y1[r];         wk(r * 9 + 8) = 1.0;
    }
#pragma unroll
    for (int c = 0; c < 9; ++c) perm[c] = c;
#pragma unroll 1
    for (int k = 0; k < 5; ++k) {
        int pr = k, pc = k;
        double best = -1.0;
        for (int r = k; r < 5; ++r)
            for (int c = k; c < 9; ++c) { const double v = fabs(wk(r * 9 + c)); if (v > best) { best = v; pr = r; pc = c; } }
        if (!(best > 1e-14)) return false;
        if (pr != k) for (int c = 0; c < 9; ++c) { const double t = wk(k * 9 + c); wk(k * 9 + c) = wk(pr * 9 + c); wk(pr * 9 + c) = t; }
        if (pc != k) {
            for (int r = 0; r < 5; ++r) { const double t = wk(r * 9 + k); wk(r * 9 + k) = wk(r * 9 + pc); wk(r * 9 + pc) = t; }
            const int t = perm[k]; perm[k] = perm[pc]; perm[pc] = t;
        }
        const double inv = 1.0 / wk(k * 9 + k);
        double prow[9];
#pragma unroll
        for (int c = 0; c < 9; ++c) { prow[c] = wk(k * 9 + c) * inv; wk(k * 9 + c) = prow[c]; }
#pragma unroll
        for (int r = 0; r < 5; ++r) {
            if (r == k) continue;
            const double f = -wk(r * 9 + k);
#pragma unroll
            for (int c = 0; c < 9; ++c) wk(r * 9 + c) = fma(f, prow[c], wk(r * 9 + c));
        }
    }
#pragma unroll 1
    for (int j = 0; j < 4; ++j) {
        double v[9];
#pragma unroll
        for (int r = 0; r < 5; ++r) v[r] = -wk(r * 9 + 5 + j);
#pragma unroll
        for (int c = 5; c < 9; ++c) v[c] = (c == 5 + j) ? 1.0 : 0.0;
        double nn = 0;
#pragma unroll
        for (int c = 0; c < 9; ++c) nn += v[c] * v[c];
        nn = 1.0 / sqrt(nn);
#pragma unroll
        for (int c = 0; c < 9; ++c) wk(45 + j * 9 + perm[c]) = v[c] * nn;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int c = 0; c < 9; ++c) N[j][c] = wk(45 + j * 9 + c);
    return true;
}

__device__ __forceinline__ double horner(const double* p, int deg, double z)
{
    double v = p[deg];
    for (int i = deg - 1; i >= 0; --i) v = v * z + p[i];
    return v;
}
__device__ __forceinline__ void pmul(const double* a, int da, const double* b, int db, double* c)
{
    for (int i = 0; i <= da + db; ++i) c[i] = 0;
    for (int i = 0; i <= da; ++i) for (int j = 0; j <= db; ++j) c[i + j] += a[i] * b[j];
}

// Sturm chain of a polynomial of degree <= 10, packed triangularly: chain polynomial s (degree <= 10 - s) starts at
// soff(s) and has 11 - s coefficients, zero-filled above its true degree, so the sign-change count is one fixed,
// fully unrolled sequence of 55 FMAs for every thread of a warp (no data-dependent control flow).
constexpr int kSturmLen = 66;
__host__ __device__ constexpr int soff(int s) { return s * 11 - s * (s - 1) / 2; }

// number of sign changes of the chain at t; *sg0 = sign of the polynomial itself
template <class W> __device__ __forceinline__ int sturm_changes(const W& S, double t, int* sg0)
{
    int cnt = 0, last = 0;
#pragma unroll
    for (int s = 0; s < 11; ++s) {
        double v = S(soff(s) + 10 - s);
#pragma unroll
        for (int i = 9 - s; i >= 0; --i) v = v * t + S(soff(s) + i);
        const int sg = (v > 0) - (v < 0);
        if (s == 0) *sg0 = sg;
        cnt += (sg != 0 && last != 0 && sg != last) ? 1 : 0;
        last = (sg != 0) ? sg : last;
    }
    return cnt;
}

// Normalise pin (degree <= 10) into p, build its Sturm chain and isolate every real root: on return lo[r] < hi[r]
// (ascending, disjoint) brackets exactly one sign change of p when bit r of *mid is clear, and the root is taken as the
// midpoint of [lo, hi] when the bit is set (multiple / numerically unresolved root, or a root hit exactly).
// Returns the number of roots; *deg_out = degree after dropping negligible leading coefficients.
template <class W>
__device__ inline int isolate_roots(const double* pin, double* p, int* deg_out, double* lo_out, double* hi_out, unsigned* mid, const W& S)
{
    int deg = 10;
    *mid = 0;
    *deg_out = 0;
    double mx = 0;
    for (int i = 0; i <= deg; ++i) if (fabs(pin[i]) > mx) mx = fabs(pin[i]);
    if (!(mx > 0) || !isfinite(mx)) return 0;
    const double imx = 1.0 / mx;
    for (int i = 0; i <= 10; ++i) p[i] = pin[i] * imx;
    while (deg > 0 && fabs(p[deg]) < 1e-13) --deg;
    for (int i = deg + 1; i <= 10; ++i) p[i] = 0;
    *deg_out = deg;
    if (deg == 0) return 0;

    for (int i = 0; i < kSturmLen; ++i) S(i) = 0;
    int sd[12];
    for (int i = 0; i <= deg; ++i) S(i) = p[i];
    sd[0] = deg;
    for (int i = 1; i <= deg; ++i) S(soff(1) + i - 1) = i * p[i];
    sd[1] = deg - 1;
    int ns = 2;
    while (sd[ns - 1] > 0 && ns < 11) {
        double r[11], d[11];
        const int dr = sd[ns - 2], dd = sd[ns - 1];
        const int o2 = soff(ns - 2), o1 = soff(ns - 1);
        for (int i = 0; i <= dr; ++i) r[i] = S(o2 + i);
        for (int i = 0; i <= dd; ++i) d[i] = S(o1 + i);
        const double idl = 1.0 / d[dd];
        for (int k = dr; k >= dd; --k) {
            const double f = r[k] * idl;
            for (int j = 0; j <= dd; ++j) r[k - dd + j] -= f * d[j];
            r[k] = 0;
        }
        int nd = dd - 1;
        double rm = 0;
        for (int i = 0; i <= nd; ++i) if (fabs(r[i]) > rm) rm = fabs(r[i]);
        if (!(rm > 1e-300)) break;
        while (nd > 0 && fabs(r[nd]) < 1e-14 * rm) --nd;
        const double irm = -1.0 / rm;
        const int o0 = soff(ns);
        for (int i = 0; i <= nd; ++i) S(o0 + i) = r[i] * irm;
        sd[ns] = nd;
        ++ns;
    }
    // Fujiwara bound 2 max_k |a_{n-k} / a_n|^(1/k), evaluated in float and inflated
    float fb = 0.f;
    const float ilead = 1.0f / static_cast<float>(fabs(p[deg]));
    for (int k = 1; k <= deg; ++k) {
        const float a = static_cast<float>(fabs(p[deg - k])) * ilead;
        const float v = a > 0.f ? exp2f(__log2f(a) / static_cast<float>(k)) : 0.f;
        fb = fmaxf(fb, v);
    }
    const double B = 2.004 * static_cast<double>(fb) + 1e-9;

    // interval stack: entries are disjoint and each holds >= 1 root, so at most 10 are alive
    double slo[12], shi[12];
    int sa[12], sb[12];            // sign changes at lo / hi, | (sign of p + 1) << 8
    int sp = 1, nr = 0, sg;
    slo[0] = -B; shi[0] = B;
    sa[0] = sturm_changes(S, -B, &sg); sa[0] |= (sg + 1) << 8;
    sb[0] = sturm_changes(S, B, &sg);  sb[0] |= (sg + 1) << 8;
    while (sp > 0 && nr < 10) {
        --sp;
        const double lo = slo[sp], hi = shi[sp];
        const int A = sa[sp], Bq = sb[sp];
        const int a = A & 0xff, b = Bq & 0xff, sgl = (A >> 8) - 1, sgh = (Bq >> 8) - 1;
        const int cnt = a - b;
        if (cnt <= 0) continue;
        const bool tiny = !((hi - lo) > 1e-13 * (1.0 + fabs(lo) + fabs(hi)));
        if (cnt == 1 && sgh == 0) { lo_out[nr] = hi; hi_out[nr] = hi; *mid |= 1u << nr; ++nr; continue; }
        if (cnt == 1 && sgl * sgh < 0) { lo_out[nr] = lo; hi_out[nr] = hi; ++nr; continue; }
        if (tiny) { lo_out[nr] = lo; hi_out[nr] = hi; *mid |= 1u << nr; ++nr; continue; }
        const double m = 0.5 * (lo + hi);
        int sgm;
        const int cm = sturm_changes(S, m, &sgm);
        const int M = cm | ((sgm + 1) << 8);
        if (cm - b > 0) { slo[sp] = m; shi[sp] = hi; sa[sp] = M; sb[sp] = Bq; ++sp; }
        if (a - cm > 0) { slo[sp] = lo; shi[sp] = m; sa[sp] = A; sb[sp] = M; ++sp; }
    }
    return nr;
}

// Offsets of the HypState fields inside the hand-off block of kHypDoubles doubles.
constexpr int kOffN = 0, kOffBx = 36, kOffBy = 48, kOffBc = 60, kOffP = 75, kOffLo = 86, kOffHi = 96, kOffS = 106;

// Root of p (degree <= 10, zero-filled above its degree) inside the isolating bracket [lo, hi] (opposite signs at the
// ends): safeguarded Newton.  p and p' come from one unrolled Horner recurrence.  The root only has to be good to
// ~1e-10: the Gauss-Newton polish on (x, y, z) that follows converges quadratically from there.
template <class HV> __device__ inline double refine_root(const HV& h, double lo, double hi)
{
    double pc[11];
#pragma unroll
    for (int i = 0; i <= 10; ++i) pc[i] = h(kOffP + i);
    double flo = pc[10];
#pragma unroll
    for (int i = 9; i >= 0; --i) flo = flo * lo + pc[i];
    double x = 0.5 * (lo + hi);
    double dxold = hi - lo;
    for (int it = 0; it < 200; ++it) {
        double f = pc[10], dp = 0.0;
#pragma unroll
        for (int i = 9; i >= 0; --i) { dp = dp * x + f; f = f * x + pc[i]; }
        if (f == 0.0) break;
        if ((f > 0) == (flo > 0)) lo = x; else hi = x;
        double xn = (dp != 0.0) ? x - f / dp : x;
        if (!(xn > lo && xn < hi) || fabs(xn - x) > 0.5 * fabs(dxold)) xn = 0.5 * (lo + hi);
        const double dx = xn - x;
        const bool done = fabs(dx) <= 1e-10 * (1.0 + fabs(x));
        dxold = dx;
        x = xn;
        if (done) break;
    }
    return x;
}

// r[0] = det E, r[1..9] = 2 E E^T E - tr(E E^T) E
__device__ inline void ess_residual(const double* E, double* r)
{
    double EEt[9], T[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
    const double tr = EEt[0] + EEt[4] + EEt[8];
    mat3_mul(EEt, E, T);
    r[0] = E[0] * (E[4] * E[8] - E[5] * E[7]) - E[1] * (E[3] * E[8] - E[5] * E[6]) + E[2] * (E[3] * E[7] - E[4] * E[6]);
    for (int k = 0; k < 9; ++k) r[1 + k] = 2.0 * T[k] - tr * E[k];
}

// Gauss-Newton polish of a root (x, y, z) of the ten cubic constraints; Jacobian columns are the analytic directional
// derivatives of (det, 2 E E^T E - tr(E E^T) E) along N0, N1, N2.
template <class HV> __device__ inline void polish_xyz(const HV& h, double* xyz)
{
    for (int it = 0; it < 3; ++it) {
        double E[9], r[10], J[10][3];
#pragma unroll
        for (int e = 0; e < 9; ++e) E[e] = xyz[0] * h(kOffN + e) + xyz[1] * h(kOffN + 9 + e) + xyz[2] * h(kOffN + 18 + e) + h(kOffN + 27 + e);
        ess_residual(E, r);
        double EEt[9], EtE[9], cof[9];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) {
                EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
                EtE[3 * i + j] = E[i] * E[j] + E[3 + i] * E[3 + j] + E[6 + i] * E[6 + j];
            }
        const double tr = EEt[0] + EEt[4] + EEt[8];
        cof[0] = E[4] * E[8] - E[5] * E[7]; cof[1] = E[5] * E[6] - E[3] * E[8]; cof[2] = E[3] * E[7] - E[4] * E[6];
        cof[3] = E[2] * E[7] - E[1] * E[8]; cof[4] = E[0] * E[8] - E[2] * E[6]; cof[5] = E[1] * E[6] - E[0] * E[7];
        cof[6] = E[1] * E[5] - E[2] * E[4]; cof[7] = E[2] * E[3] - E[0] * E[5]; cof[8] = E[0] * E[4] - E[1] * E[3];
        for (int d = 0; d < 3; ++d) {
            double D[9];
#pragma unroll
            for (int k = 0; k < 9; ++k) D[k] = h(kOffN + 9 * d + k);
            double T1[9], T2[9], T3[9], Dt[9], EDt[9], ip = 0, dd = 0;
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Dt[3 * i + j] = D[3 * j + i];
            mat3_mul(D, EtE, T1); mat3_mul(E, Dt, EDt); mat3_mul(EDt, E, T2); mat3_mul(EEt, D, T3);
            for (int k = 0; k < 9; ++k) { ip += E[k] * D[k]; dd += cof[k] * D[k]; }
            J[0][d] = dd;
            for (int k = 0; k < 9; ++k) J[1 + k][d] = 2.0 * (T1[k] + T2[k] + T3[k]) - 2.0 * ip * E[k] - tr * D[k];
        }
        double A[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, b[3] = {0, 0, 0};
        for (int k = 0; k < 10; ++k)
            for (int i = 0; i < 3; ++i) { b[i] += J[k][i] * r[k]; for (int j = 0; j < 3; ++j) A[i][j] += J[k][i] * J[k][j]; }
        const double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                           A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
        if (!(fabs(det) > 0) || !isfinite(det)) return;
        double dx[3];
        dx[0] = (b[0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (b[1] * A[2][2] - A[1][2] * b[2]) + A[0][2] * (b[1] * A[2][1] - A[1][1] * b[2])) / det;
        dx[1] = (A[0][0] * (b[1] * A[2][2] - A[1][2] * b[2]) - b[0] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) + A[0][2] * (A[1][0] * b[2] - b[1] * A[2][0])) / det;
        dx[2] = (A[0][0] * (A[1][1] * b[2] - b[1] * A[2][1]) - A[0][1] * (A[1][0] * b[2] - b[1] * A[2][0]) + b[0] * (A[1][0] * A[2][1] - A[1][1] * A[2][0])) / det;
        if (!isfinite(dx[0]) || !isfinite(dx[1]) || !isfinite(dx[2])) return;
        for (int i = 0; i < 3; ++i) xyz[i] -= dx[i];
    }
}

// Everything the per-root phase needs about one hypothesis.  The first kHypDoubles doubles are plain data so the solve
// kernel can hand the state from the per-hypothesis phase to the per-root phase through shared memory.
struct HypState {
    double N[4][9];                          // null-space basis: E = x N0 + y N1 + z N2 + N3
    double Bx[3][4], By[3][4], Bc[3][5];     // B(z) [x y 1]^T = 0
    double p[11];                            // normalised degree-10 polynomial det B(z)
    double lo[10], hi[10];                   // isolating brackets of its real roots, ascending
    double sx1[5], sy1[5], sx2[5], sy2[5];   // the five sample correspondences (model validity test)
    int deg, nr;
    unsigned mid;                            // bit r: root r = midpoint of its bracket
};
constexpr int kHypDoubles = 36 + 12 + 12 + 15 + 11 + 20 + 20;
static_assert(offsetof(HypState, deg) == kHypDoubles * 8, "HypState layout");
static_assert(offsetof(HypState, Bx) == 36 * 8 && offsetof(HypState, p) == 75 * 8 && offsetof(HypState, lo) == 86 * 8 && offsetof(HypState, hi) == 96 * 8 &&
              offsetof(HypState, sx1) == 106 * 8, "HypState layout");

// Phase A (once per hypothesis): five normalised correspondences -> null space, the ten cubic constraints, their
// Gauss-Jordan reduction, det B(z) and the isolating brackets of its real roots.  Returns the number of roots.
template <class W>
__device__ inline int fivept_isolate(const double* x1, const double* y1, const double* x2, const double* y2, HypState& hs, const W& wk)
{
    hs.nr = 0; hs.deg = 0; hs.mid = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) { hs.sx1[k] = x1[k]; hs.sy1[k] = y1[k]; hs.sx2[k] = x2[k]; hs.sy2[k] = y2[k]; }
    if (!nullspace_5x9(x1, y1, x2, y2, hs.N, wk)) return 0;
    double Ep[9][4];
    for (int e = 0; e < 9; ++e) for (int k = 0; k < 4; ++k) Ep[e][k] = hs.N[k][e];
    // C[r][c] = wk(r * 20 + c)
    {   // det(E) = 0
        double m0[10], m1[10], m2[10], row[20];
#pragma unroll
        for (int k = 0; k < 10; ++k) { m0[k] = 0; m1[k] = 0; m2[k] = 0; }
#pragma unroll
        for (int k = 0; k < 20; ++k) row[k] = 0;
        p11_acc<false>(Ep[4], Ep[8], m0); p11_acc<true>(Ep[5], Ep[7], m0);
        p11_acc<false>(Ep[3], Ep[8], m1); p11_acc<true>(Ep[5], Ep[6], m1);
        p11_acc<false>(Ep[3], Ep[7], m2); p11_acc<true>(Ep[4], Ep[6], m2);
        p21_acc<false>(m0, Ep[0], row); p21_acc<true>(m1, Ep[1], row); p21_acc<false>(m2, Ep[2], row);
#pragma unroll
        for (int k = 0; k < 20; ++k) wk(k) = row[k];
    }
    {   // (E E^T - 1/2 tr(E E^T) I) E = 0
        double G[3][3][10];
#pragma unroll 1
        for (int i = 0; i < 3; ++i)
#pragma unroll 1
            for (int j = i; j < 3; ++j) {
                double g[10];
#pragma unroll
                for (int k = 0; k < 10; ++k) g[k] = 0;
#pragma unroll
                for (int k = 0; k < 3; ++k) p11_acc<false>(Ep[3 * i + k], Ep[3 * j + k], g);
#pragma unroll
                for (int k = 0; k < 10; ++k) { G[i][j][k] = g[k]; G[j][i][k] = g[k]; }
            }
#pragma unroll
        for (int k = 0; k < 10; ++k) {
            const double tr = 0.5 * (G[0][0][k] + G[1][1][k] + G[2][2][k]);
            G[0][0][k] -= tr; G[1][1][k] -= tr; G[2][2][k] -= tr;
        }
#pragma unroll 1
        for (int i = 0; i < 3; ++i)
#pragma unroll 1
            for (int j = 0; j < 3; ++j) {
                double row[20];
#pragma unroll
                for (int k = 0; k < 20; ++k) row[k] = 0;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    double g[10];
#pragma unroll
                    for (int u = 0; u < 10; ++u) g[u] = G[i][k][u];
                    p21_acc<false>(g, Ep[3 * k + j], row);
                }
#pragma unroll
                for (int k = 0; k < 20; ++k) wk((1 + 3 * i + j) * 20 + k) = row[k];
            }
    }
    // Gaussian elimination with row pivoting on the first ten columns.  Rows are not swapped: the pivot row of column k
    // is read out with a statically indexed, predicated copy (a per-thread row index would scatter the warp's local
    // memory accesses over 64 cache lines), scaled to unit pivot and kept as P[k]; only rows 4..9 are needed later.
    double P[6][15];                 // pivot rows 4..9: columns 5..9 (index c - 5) and 10..19 (index c - 5)
    unsigned used = 0;
#pragma unroll 1
    for (int k = 0; k < 10; ++k) {
        int pr = -1;
        double best = -1.0, pv = 0.0;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            const double v = wk(r * 20 + k), av = fabs(v);
            if (!((used >> r) & 1u) && av > best) { best = av; pr = r; pv = v; }
        }
        if (!(best > 1e-300)) return 0;
        used |= 1u << pr;
        double prow[20];
#pragma unroll
        for (int c = 0; c < 20; ++c) prow[c] = 0;
#pragma unroll
        for (int r = 0; r < 10; ++r)
            if (r == pr) {
#pragma unroll
                for (int c = 0; c < 20; ++c) if (c >= 10 || c > k) prow[c] = wk(r * 20 + c);
            }
        const double inv = 1.0 / pv;
#pragma unroll
        for (int c = 0; c < 20; ++c) prow[c] *= inv;
        if (k >= 4) {
#pragma unroll
            for (int c = 5; c < 20; ++c) P[k - 4][c - 5] = prow[c];
        }
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            if ((used >> r) & 1u) continue;
            const double f = -wk(r * 20 + k);
#pragma unroll
            for (int c = 0; c < 20; ++c) if (c >= 10 || c > k) wk(r * 20 + c) = fma(f, prow[c], wk(r * 20 + c));
        }
    }
    // back substitution of the right half, pivot rows 8..4
#pragma unroll
    for (int i = 8; i >= 4; --i)
#pragma unroll
        for (int k = i + 1; k < 10; ++k) {
            const double f = -P[i - 4][k - 5];
#pragma unroll
            for (int c = 10; c < 20; ++c) P[i - 4][c - 5] = fma(f, P[k - 4][c - 5], P[i - 4][c - 5]);
        }
    // B(z) [x y 1]^T = 0; columns 10 xz^2, 11 xz, 12 x, 13 yz^2, 14 yz, 15 y, 16 z^3, 17 z^2, 18 z, 19 1
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const double* a = P[2 * q];          // column c sits at index c - 5
        const double* b = P[2 * q + 1];
        hs.Bx[q][0] = a[7];  hs.Bx[q][1] = a[6] - b[7];   hs.Bx[q][2] = a[5] - b[6];  hs.Bx[q][3] = -b[5];
        hs.By[q][0] = a[10]; hs.By[q][1] = a[9] - b[10];  hs.By[q][2] = a[8] - b[9];  hs.By[q][3] = -b[8];
        hs.Bc[q][0] = a[14]; hs.Bc[q][1] = a[13] - b[14]; hs.Bc[q][2] = a[12] - b[13]; hs.Bc[q][3] = a[11] - b[12]; hs.Bc[q][4] = -b[11];
    }
    double poly[11], t1[8], t2[8], t3[11];
    for (int i = 0; i < 11; ++i) poly[i] = 0;
    pmul(hs.By[1], 3, hs.Bc[2], 4, t1); pmul(hs.Bc[1], 4, hs.By[2], 3, t2); for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    pmul(hs.Bx[0], 3, t1, 7, t3); for (int i = 0; i < 11; ++i) poly[i] += t3[i];
    pmul(hs.Bx[1], 3, hs.Bc[2], 4, t1); pmul(hs.Bc[1], 4, hs.Bx[2], 3, t2); for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    pmul(hs.By[0], 3, t1, 7, t3); for (int i = 0; i < 11; ++i) poly[i] -= t3[i];
    {
        double u1[7], u2[7];
        pmul(hs.Bx[1], 3, hs.By[2], 3, u1); pmul(hs.By[1], 3, hs.Bx[2], 3, u2);
        for (int i = 0; i < 7; ++i) u1[i] -= u2[i];
        pmul(hs.Bc[0], 4, u1, 6, t3);
        for (int i = 0; i < 11; ++i) poly[i] += t3[i];
    }
    hs.nr = isolate_roots(poly, hs.p, &hs.deg, hs.lo, hs.hi, &hs.mid, wk);
    return hs.nr;
}

// Model validity test of USAC-style estimators (the oriented epipolar constraint of Chum, Werner and Matas, 2004; cv2's
// USAC applies it to every minimal-sample model before scoring): with e2 the epipole in image 2 (E^T e2 = 0), the
// vectors e2 x x2 and E x1 must point the same way for every sample correspondence, or the opposite way for every one.
// A model that mixes signs cannot have all five points in front of both cameras.
__device__ __forceinline__ bool oriented_sample_ok(const double* E, const double* x1, const double* y1, const double* x2, const double* y2)
{
    const double c0[3] = {E[0], E[3], E[6]}, c1[3] = {E[1], E[4], E[7]}, c2[3] = {E[2], E[5], E[8]};
    double v01[3], v02[3], v12[3];
    cross3(c0, c1, v01); cross3(c0, c2, v02); cross3(c1, c2, v12);
    const double n01 = v01[0] * v01[0] + v01[1] * v01[1] + v01[2] * v01[2];
    const double n02 = v02[0] * v02[0] + v02[1] * v02[1] + v02[2] * v02[2];
    const double n12 = v12[0] * v12[0] + v12[1] * v12[1] + v12[2] * v12[2];
    double e[3] = {v01[0], v01[1], v01[2]};
    double nb = n01;
    if (n02 > nb) { nb = n02; e[0] = v02[0]; e[1] = v02[1]; e[2] = v02[2]; }
    if (n12 > nb) { nb = n12; e[0] = v12[0]; e[1] = v12[1]; e[2] = v12[2]; }
    int pos = 0, neg = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const double l0 = E[0] * x1[k] + E[1] * y1[k] + E[2], l1 = E[3] * x1[k] + E[4] * y1[k] + E[5], l2 = E[6] * x1[k] + E[7] * y1[k] + E[8];
        const double w0 = e[1] - e[2] * y2[k], w1 = e[2] * x2[k] - e[0], w2 = e[0] * y2[k] - e[1] * x2[k];   // e x (x2, y2, 1)
        const double sgn = w0 * l0 + w1 * l1 + w2 * l2;
        pos += sgn > 0.0;
        neg += sgn < 0.0;
    }
    return pos == 5 || neg == 5;
}

// Read-only views of a hand-off block: element f of the block at base[f * STRIDE].
template <int STRIDE> struct HypView {
    const double* base;
    __device__ __forceinline__ double operator()(int f) const { return base[f * STRIDE]; }
};

// Phase B1 (once per real root): root r -> z (Newton), (x, y) from the null vector of B(z), and the validity of the
// unpolished model.  false when the root yields no finite, valid essential matrix: most wrong roots leave here.
template <class HV> __device__ inline bool fivept_root_xyz(const HV& h, unsigned mid, int r, double* xyz)
{
    const double lo = h(kOffLo + r), hi = h(kOffHi + r);
    const double z = ((mid >> r) & 1u) ? 0.5 * (lo + hi) : refine_root(h, lo, hi);
    double bm[3][3];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        double vx = h(kOffBx + 4 * q + 3), vy = h(kOffBy + 4 * q + 3), vc = h(kOffBc + 5 * q + 4);
#pragma unroll
        for (int i = 2; i >= 0; --i) { vx = vx * z + h(kOffBx + 4 * q + i); vy = vy * z + h(kOffBy + 4 * q + i); }
#pragma unroll
        for (int i = 3; i >= 0; --i) vc = vc * z + h(kOffBc + 5 * q + i);
        bm[q][0] = vx; bm[q][1] = vy; bm[q][2] = vc;
    }
    double best = -1, bv[3] = {0, 0, 0};
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = a + 1; b < 3; ++b) {
            double v[3];
            cross3(bm[a], bm[b], v);
            if (fabs(v[2]) > best) { best = fabs(v[2]); bv[0] = v[0]; bv[1] = v[1]; bv[2] = v[2]; }
        }
    if (!(best > 0)) return false;
    xyz[0] = bv[0] / bv[2]; xyz[1] = bv[1] / bv[2]; xyz[2] = z;
    if (!isfinite(xyz[0]) || !isfinite(xyz[1])) return false;
    double sx1[5], sy1[5], sx2[5], sy2[5], E0[9];
#pragma unroll
    for (int k = 0; k < 5; ++k) { sx1[k] = h(kOffS + k); sy1[k] = h(kOffS + 5 + k); sx2[k] = h(kOffS + 10 + k); sy2[k] = h(kOffS + 15 + k); }
#pragma unroll
    for (int e = 0; e < 9; ++e) E0[e] = xyz[0] * h(kOffN + e) + xyz[1] * h(kOffN + 9 + e) + xyz[2] * h(kOffN + 18 + e) + h(kOffN + 27 + e);
    return oriented_sample_ok(E0, sx1, sy1, sx2, sy2);
}

// Phase B2 (once per root that passed B1): Gauss-Newton polish, unit-norm E, validity of the final model.
template <class HV> __device__ inline bool fivept_root_polish(const HV& h, double* xyz, double* Eout)
{
    polish_xyz(h, xyz);
    double E[9], nn = 0;
#pragma unroll
    for (int e = 0; e < 9; ++e) {
        E[e] = xyz[0] * h(kOffN + e) + xyz[1] * h(kOffN + 9 + e) + xyz[2] * h(kOffN + 18 + e) + h(kOffN + 27 + e);
        nn += E[e] * E[e];
    }
    if (!(nn > 0) || !isfinite(nn)) return false;
    double sx1[5], sy1[5], sx2[5], sy2[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) { sx1[k] = h(kOffS + k); sy1[k] = h(kOffS + 5 + k); sx2[k] = h(kOffS + 10 + k); sy2[k] = h(kOffS + 15 + k); }
    if (!oriented_sample_ok(E, sx1, sy1, sx2, sy2)) return false;
    nn = 1.0 / sqrt(nn);
#pragma unroll
    for (int e = 0; e < 9; ++e) Eout[e] = E[e] * nn;
    return true;
}

template <class HV> __device__ inline bool fivept_root(const HV& h, unsigned mid, int r, double* Eout)
{
    double xyz[3];
    return fivept_root_xyz(h, mid, r, xyz) && fivept_root_polish(h, xyz, Eout);
}

}  // namespace vm
