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
// keeps it in shared memory ([element][thread], STRIDE = threads per CTA: conflict-free, a per-thread dynamic index costs
// nothing, and it stays out of the L1/L2 path that thread-local arrays take); the finalize kernel does the same for its
// few re-solving threads.  Everything the solver indexes dynamically lives here, never in a thread-local array.
//
// Use of the 200 slots over the life of a hypothesis:
//   null space     0..80    the 5 x 9 system and the four null vectors
//   constraints    0..199   the 10 x 20 coefficient matrix C[r][c] at r * 20 + c, eliminated in place
//   hand-off       kOffN (36)  null-space basis      kOffBx/By/Bc (39)  B(z)        kOffLo/kOffHi (10 + 10) root brackets
//                  kOffS (20)  the five sample correspondences (written last)       kOffP (11) normalised polynomial
//   isolation      kOffSturm (66) Sturm chain, its first polynomial IS the hand-off polynomial; kOffS.. doubles as the
//                  remainder buffer while the chain is built and as the interval stack afterwards, kOffTag as its tags
//   per-root phase kOffXyz.. (x, y, z) of the roots that passed the first validity test (solve kernel only)
constexpr int kWorkDoubles = 200;
constexpr int kOffN = 0, kOffBx = 36, kOffBy = 48, kOffBc = 60, kOffTag = 75, kOffLo = 86, kOffHi = 96, kOffS = 106;
constexpr int kOffSturm = 134, kOffP = kOffSturm, kOffXyz = 145;
template <int STRIDE> struct Work {
    double* base;
    __device__ __forceinline__ double& operator()(int i) const { return base[i * STRIDE]; }
};

// 1 / a to within an ulp or two: hardware seed (about 20 bits) and two Newton steps.  Used where the quotient only scales
// a row or steps an iteration, so the last bit does not matter; far shorter than the IEEE division sequence.
__device__ __forceinline__ double rcp_nr(double a)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
    double e = fma(-a, r, 1.0);
    r = fma(r, e, r);
    e = fma(-a, r, 1.0);
    return fma(r, e, r);
}

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
    // A[r][c] = wk(r * 9 + c); the null vectors are assembled in wk(45 + j * 9 + c); column permutation in nibbles
    unsigned long long perm = 0x876543210ull;
#pragma unroll
    for (int r = 0; r < 5; ++r) {
        wk(r * 9 + 0) = x2[r] * x1[r]; wk(r * 9 + 1) = x2[r] * y1[r]; wk(r * 9 + 2) = x2[r];
        wk(r * 9 + 3) = y2[r] * x1[r]; wk(r * 9 + 4) = y2[r] * y1[r]; wk(r * 9 + 5) = y2[r];
        wk(r * 9 + 6) = x1[r];         wk(r * 9 + 7) = y1[r];         wk(r * 9 + 8) = 1.0;
    }
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
            const unsigned long long x = ((perm >> (4 * k)) ^ (perm >> (4 * pc))) & 15ull;
            perm ^= (x << (4 * k)) | (x << (4 * pc));
        }
        const double inv = rcp_nr(wk(k * 9 + k));
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
        for (int c = 0; c < 9; ++c) wk(45 + j * 9 + static_cast<int>((perm >> (4 * c)) & 15ull)) = v[c] * nn;
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
template <class W> __device__ __forceinline__ int sturm_changes(const W& wk, double t, int* sg0)
{
    int cnt = 0, last = 0;
#pragma unroll
    for (int s = 0; s < 11; ++s) {
        double v = wk(kOffSturm + soff(s) + 10 - s);
#pragma unroll
        for (int i = 9 - s; i >= 0; --i) v = v * t + wk(kOffSturm + soff(s) + i);
        const int sg = (v > 0) - (v < 0);
        if (s == 0) *sg0 = sg;
        cnt += (sg != 0 && last != 0 && sg != last) ? 1 : 0;
        last = (sg != 0) ? sg : last;
    }
    return cnt;
}

// Normalise pin (degree <= 10) into the hand-off polynomial, build its Sturm chain and isolate every real root: on return
// the brackets lo[r] < hi[r] (ascending, disjoint; hand-off slots kOffLo + r, kOffHi + r) hold exactly one sign change of
// p when bit r of *mid is clear, and the root is taken as the midpoint of [lo, hi] when the bit is set (multiple /
// numerically unresolved root, or a root hit exactly).  Returns the number of roots.
template <class W>
__device__ inline int isolate_roots(const double* pin, unsigned* mid, const W& wk)
{
    *mid = 0;
    double mx = 0;
#pragma unroll
    for (int i = 0; i <= 10; ++i) mx = fmax(mx, fabs(pin[i]));
    if (!(mx > 0) || !isfinite(mx)) return 0;
    const double imx = 1.0 / mx;
    int deg = 0;                                  // degree after dropping negligible leading coefficients
#pragma unroll
    for (int i = 1; i <= 10; ++i) if (fabs(pin[i] * imx) >= 1e-13) deg = i;
    if (deg == 0) return 0;
    const auto S = [&wk](int i) -> double& { return wk(kOffSturm + i); };
    const auto R = [&wk](int i) -> double& { return wk(kOffS + i); };       // remainder buffer (11)
#pragma unroll
    for (int i = 0; i <= 10; ++i) S(i) = (i <= deg) ? pin[i] * imx : 0.0;
#pragma unroll
    for (int i = 11; i < kSturmLen; ++i) S(i) = 0.0;
#pragma unroll
    for (int i = 1; i <= 10; ++i) if (i <= deg) S(soff(1) + i - 1) = i * (pin[i] * imx);
    unsigned long long sd = static_cast<unsigned long long>(deg) | (static_cast<unsigned long long>(deg - 1) << 4);   // chain degrees, one nibble each
    int ns = 2;
    while (static_cast<int>((sd >> (4 * (ns - 1))) & 15ull) > 0 && ns < 11) {
        const int dr = static_cast<int>((sd >> (4 * (ns - 2))) & 15ull), dd = static_cast<int>((sd >> (4 * (ns - 1))) & 15ull);
        const int o2 = soff(ns - 2), o1 = soff(ns - 1);
        for (int i = 0; i <= dr; ++i) R(i) = S(o2 + i);
        const double idl = rcp_nr(S(o1 + dd));
        for (int k = dr; k >= dd; --k) {
            const double f = R(k) * idl;
            for (int j = 0; j < dd; ++j) R(k - dd + j) = fma(-f, S(o1 + j), R(k - dd + j));     // R(k) itself is never read again
        }
        int nd = dd - 1;
        double rm = 0;
        for (int i = 0; i <= nd; ++i) rm = fmax(rm, fabs(R(i)));
        if (!(rm > 1e-300)) break;
        while (nd > 0 && fabs(R(nd)) < 1e-14 * rm) --nd;
        const double irm = -rcp_nr(rm);
        const int o0 = soff(ns);
        for (int i = 0; i <= nd; ++i) S(o0 + i) = R(i) * irm;
        sd |= static_cast<unsigned long long>(nd) << (4 * ns);
        ++ns;
    }
    // Fujiwara bound 2 max_k |a_{n-k} / a_n|^(1/k), evaluated in float and inflated
    float fb = 0.f;
    const float ilead = 1.0f / static_cast<float>(fabs(S(deg)));
    for (int k = 1; k <= deg; ++k) {
        const float a = static_cast<float>(fabs(S(deg - k))) * ilead;
        const float v = a > 0.f ? exp2f(__log2f(a) / static_cast<float>(k)) : 0.f;
        fb = fmaxf(fb, v);
    }
    const double B = 2.004 * static_cast<double>(fb) + 1e-9;

    // interval stack: entries are disjoint and each holds >= 1 root, so at most 10 are alive.  lo at kOffS + i, hi at
    // kOffS + 10 + i, tag at kOffTag + i = (sign changes | (sign of p + 1) << 8) at lo (high word) and at hi (low word)
    int sp = 1, nr = 0, sg;
    {
        int A = sturm_changes(wk, -B, &sg); A |= (sg + 1) << 8;
        int Bq = sturm_changes(wk, B, &sg); Bq |= (sg + 1) << 8;
        wk(kOffS) = -B; wk(kOffS + 10) = B; wk(kOffTag) = __hiloint2double(A, Bq);
    }
    while (sp > 0 && nr < 10) {
        --sp;
        const double lo = wk(kOffS + sp), hi = wk(kOffS + 10 + sp), tag = wk(kOffTag + sp);
        const int A = __double2hiint(tag), Bq = __double2loint(tag);
        const int a = A & 0xff, b = Bq & 0xff, sgl = (A >> 8) - 1, sgh = (Bq >> 8) - 1;
        const int cnt = a - b;
        if (cnt <= 0) continue;
        const bool tiny = !((hi - lo) > 1e-13 * (1.0 + fabs(lo) + fabs(hi)));
        if (cnt == 1 && sgh == 0) { wk(kOffLo + nr) = hi; wk(kOffHi + nr) = hi; *mid |= 1u << nr; ++nr; continue; }
        if (cnt == 1 && sgl * sgh < 0) { wk(kOffLo + nr) = lo; wk(kOffHi + nr) = hi; ++nr; continue; }
        if (tiny) { wk(kOffLo + nr) = lo; wk(kOffHi + nr) = hi; *mid |= 1u << nr; ++nr; continue; }
        const double m = 0.5 * (lo + hi);
        int sgm;
        const int cm = sturm_changes(wk, m, &sgm);
        const int M = cm | ((sgm + 1) << 8);
        if (cm - b > 0 && sp < 10) { wk(kOffS + sp) = m; wk(kOffS + 10 + sp) = hi; wk(kOffTag + sp) = __hiloint2double(M, Bq); ++sp; }
        if (a - cm > 0 && sp < 10) { wk(kOffS + sp) = lo; wk(kOffS + 10 + sp) = m; wk(kOffTag + sp) = __hiloint2double(A, M); ++sp; }
    }
    return nr;
}

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
        double xn = fma(-f, rcp_nr(dp), x);                    // dp = 0 gives a non-finite step: bisection below
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

// Phase A (once per hypothesis): five normalised correspondences -> null space, the ten cubic constraints, their
// Gauss-Jordan reduction, det B(z) and the isolating brackets of its real roots.  Returns the number of roots and leaves
// the hand-off block (all but the sample correspondences at kOffS, which the caller writes) in the scratch.
template <class W>
__device__ inline int fivept_isolate(const double* x1, const double* y1, const double* x2, const double* y2, const W& wk, unsigned* mid)
{
    *mid = 0;
    double N[4][9];
    if (!nullspace_5x9(x1, y1, x2, y2, N, wk)) return 0;
    double Ep[9][4];
    for (int e = 0; e < 9; ++e) for (int k = 0; k < 4; ++k) Ep[e][k] = N[k][e];
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
    // Gauss-Jordan on the first ten columns with partial pivoting.  The pivot row is swapped into place: in the
    // [element][thread] scratch a per-thread row index is free, so row k is the pivot row of column k afterwards and
    // the update loop only touches the rows below it (their range is the same for every thread of the warp).  Only
    // pivot rows 4..9 are needed later, so the rows above the pivot are never updated and rows 0..3 not written back.
#pragma unroll 1
    for (int k = 0; k < 10; ++k) {
        int pr = k;
        double best = -1.0, pv = 0.0;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            if (r >= k) {
                const double v = wk(r * 20 + k), av = fabs(v);
                if (av > best) { best = av; pr = r; pv = v; }
            }
        }
        if (!(best > 1e-300)) return 0;
        const double inv = rcp_nr(pv);
        wk(pr * 20 + k) = wk(k * 20 + k);
        double prow[20];
#pragma unroll
        for (int c = 0; c < 20; ++c) {
            prow[c] = 0;
            if (c >= 10 || c > k) {
                const double a = wk(pr * 20 + c);
                wk(pr * 20 + c) = wk(k * 20 + c);
                prow[c] = a * inv;
                if (k >= 4) wk(k * 20 + c) = prow[c];
            }
        }
#pragma unroll
        for (int r = 1; r < 10; ++r) {
            if (r > k) {
                const double f = -wk(r * 20 + k);
#pragma unroll
                for (int c = 0; c < 20; ++c) if (c >= 10 || c > k) wk(r * 20 + c) = fma(f, prow[c], wk(r * 20 + c));
            }
        }
    }
    // back substitution of the right half, pivot rows 8..4, in place
#pragma unroll
    for (int i = 8; i >= 4; --i) {
        double acc[10];
#pragma unroll
        for (int c = 0; c < 10; ++c) acc[c] = wk(i * 20 + 10 + c);
#pragma unroll
        for (int k = i + 1; k < 10; ++k) {
            const double f = -wk(i * 20 + k);
#pragma unroll
            for (int c = 0; c < 10; ++c) acc[c] = fma(f, wk(k * 20 + 10 + c), acc[c]);
        }
#pragma unroll
        for (int c = 0; c < 10; ++c) wk(i * 20 + 10 + c) = acc[c];
    }
    // hand-off: null-space basis (rows 0..3 of C are dead by now)
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 9; ++e) wk(kOffN + 9 * k + e) = N[k][e];
    // B(z) [x y 1]^T = 0; columns 10 xz^2, 11 xz, 12 x, 13 yz^2, 14 yz, 15 y, 16 z^3, 17 z^2, 18 z, 19 1
    double Bx[3][4], By[3][4], Bc[3][5];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        double a[10], b[10];                       // pivot rows 4 + 2q and 5 + 2q, columns 10..19
#pragma unroll
        for (int c = 0; c < 10; ++c) { a[c] = wk((4 + 2 * q) * 20 + 10 + c); b[c] = wk((5 + 2 * q) * 20 + 10 + c); }
        Bx[q][0] = a[2]; Bx[q][1] = a[1] - b[2]; Bx[q][2] = a[0] - b[1]; Bx[q][3] = -b[0];
        By[q][0] = a[5]; By[q][1] = a[4] - b[5]; By[q][2] = a[3] - b[4]; By[q][3] = -b[3];
        Bc[q][0] = a[9]; Bc[q][1] = a[8] - b[9]; Bc[q][2] = a[7] - b[8]; Bc[q][3] = a[6] - b[7]; Bc[q][4] = -b[6];
    }
#pragma unroll
    for (int q = 0; q < 3; ++q) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { wk(kOffBx + 4 * q + i) = Bx[q][i]; wk(kOffBy + 4 * q + i) = By[q][i]; }
#pragma unroll
        for (int i = 0; i < 5; ++i) wk(kOffBc + 5 * q + i) = Bc[q][i];
    }
    double poly[11], t1[8], t2[8], t3[11];
    for (int i = 0; i < 11; ++i) poly[i] = 0;
    pmul(By[1], 3, Bc[2], 4, t1); pmul(Bc[1], 4, By[2], 3, t2); for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    pmul(Bx[0], 3, t1, 7, t3); for (int i = 0; i < 11; ++i) poly[i] += t3[i];
    pmul(Bx[1], 3, Bc[2], 4, t1); pmul(Bc[1], 4, Bx[2], 3, t2); for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    pmul(By[0], 3, t1, 7, t3); for (int i = 0; i < 11; ++i) poly[i] -= t3[i];
    {
        double u1[7], u2[7];
        pmul(Bx[1], 3, By[2], 3, u1); pmul(By[1], 3, Bx[2], 3, u2);
        for (int i = 0; i < 7; ++i) u1[i] -= u2[i];
        pmul(Bc[0], 4, u1, 6, t3);
        for (int i = 0; i < 11; ++i) poly[i] += t3[i];
    }
    return isolate_roots(poly, mid, wk);
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
