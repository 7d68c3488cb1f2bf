// Nister's five-point minimal solver (Nister, TPAMI 2004 -- the algorithm cited at gtsfm/frontend/verifier/ransac.py:6,
// executed in the reference by cv2.findEssentialMat at ransac.py:74), one hypothesis per thread, fp64.
//
// Pipeline: null space of the 5x9 epipolar system (Gauss-Jordan, complete pivoting) -> ten cubic constraints in
// (x, y, z) on E = x N0 + y N1 + z N2 + N3 -> Gauss-Jordan on the ten highest monomials -> degree-10 polynomial in z
// -> real roots by Sturm-sequence isolation + safeguarded Newton -> (x, y) from the 3x3 null vector -> Gauss-Newton
// polish on the ten constraints.  Output: up to 10 unit-Frobenius-norm essential matrices in ascending z.
// The CPU oracle holds an independent C restatement that is pinned against cv2's own 5-point solver.
#pragma once
#include "verify_math.cuh"

namespace vm {

constexpr int kMaxCand = 10;

__device__ __constant__ int c_idx11[4][4] = {{0, 3, 4, 6}, {3, 1, 5, 7}, {4, 5, 2, 8}, {6, 7, 8, 9}};
__device__ __constant__ int c_idx21[10][4] = {{0, 2, 4, 5},   {3, 1, 6, 7},   {10, 13, 16, 17}, {2, 3, 8, 9},     {4, 8, 10, 11},
                                              {8, 6, 13, 14}, {5, 9, 11, 12}, {9, 7, 14, 15},   {11, 14, 17, 18}, {12, 15, 18, 19}};

// c (degree 2, 10 coefficients) += s * a (degree 1) * b (degree 1);  monomials x^2 y^2 z^2 xy xz yz x y z 1
__device__ __forceinline__ void p11_acc(const double* a, const double* b, double s, double* c)
{
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[c_idx11[i][j]] += s * a[i] * b[j];
}
// c (degree 3, 20 coefficients) += s * a (degree 2) * b (degree 1)
__device__ __forceinline__ void p21_acc(const double* a, const double* b, double s, double* c)
{
    for (int i = 0; i < 10; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) c[c_idx21[i][j]] += s * a[i] * b[j];
}

__device__ inline bool nullspace_5x9(const double* x1, const double* y1, const double* x2, const double* y2, double N[4][9])
{
    double A[5][9];
    int perm[9];
    for (int r = 0; r < 5; ++r) {
        A[r][0] = x2[r] * x1[r]; A[r][1] = x2[r] * y1[r]; A[r][2] = x2[r];
        A[r][3] = y2[r] * x1[r]; A[r][4] = y2[r] * y1[r]; A[r][5] = y2[r];
        A[r][6] = x1[r];         A[r][7] = y1[r];         A[r][8] = 1.0;
    }
    for (int c = 0; c < 9; ++c) perm[c] = c;
    for (int k = 0; k < 5; ++k) {
        int pr = k, pc = k;
        double best = -1.0;
        for (int r = k; r < 5; ++r)
            for (int c = k; c < 9; ++c) { const double v = fabs(A[r][c]); if (v > best) { best = v; pr = r; pc = c; } }
        if (!(best > 1e-14)) return false;
        if (pr != k) for (int c = 0; c < 9; ++c) { const double t = A[k][c]; A[k][c] = A[pr][c]; A[pr][c] = t; }
        if (pc != k) {
            for (int r = 0; r < 5; ++r) { const double t = A[r][k]; A[r][k] = A[r][pc]; A[r][pc] = t; }
            const int t = perm[k]; perm[k] = perm[pc]; perm[pc] = t;
        }
        const double inv = 1.0 / A[k][k];
        for (int c = k; c < 9; ++c) A[k][c] *= inv;
        for (int r = 0; r < 5; ++r)
            if (r != k) { const double f = A[r][k]; if (f != 0.0) for (int c = k; c < 9; ++c) A[r][c] -= f * A[k][c]; }
    }
    for (int j = 0; j < 4; ++j) {
        double v[9];
        for (int r = 0; r < 5; ++r) v[r] = -A[r][5 + j];
        for (int c = 5; c < 9; ++c) v[c] = (c == 5 + j) ? 1.0 : 0.0;
        double nn = 0;
        for (int c = 0; c < 9; ++c) nn += v[c] * v[c];
        nn = 1.0 / sqrt(nn);
        for (int c = 0; c < 9; ++c) N[j][perm[c]] = v[c] * nn;
    }
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

struct Sturm {
    double S[12][11];
    int sd[12];
    int ns;
    __device__ int changes(double t) const
    {
        int cnt = 0, last = 0;
        for (int s = 0; s < ns; ++s) {
            const double v = horner(S[s], sd[s], t);
            const int sg = (v > 0) - (v < 0);
            if (sg != 0) { if (last != 0 && sg != last) ++cnt; last = sg; }
        }
        return cnt;
    }
};

// Real roots (ascending) of a polynomial of degree <= 10; coefficient i multiplies z^i.
__device__ inline int real_roots_sturm(const double* pin, int deg, double* roots)
{
    double p[11];
    double mx = 0;
    for (int i = 0; i <= deg; ++i) if (fabs(pin[i]) > mx) mx = fabs(pin[i]);
    if (!(mx > 0) || !isfinite(mx)) return 0;
    for (int i = 0; i <= deg; ++i) p[i] = pin[i] / mx;
    while (deg > 0 && fabs(p[deg]) < 1e-13) --deg;
    if (deg == 0) return 0;
    Sturm st;
    for (int i = 0; i <= deg; ++i) st.S[0][i] = p[i];
    st.sd[0] = deg;
    for (int i = 1; i <= deg; ++i) st.S[1][i - 1] = i * p[i];
    st.sd[1] = deg - 1;
    st.ns = 2;
    while (st.sd[st.ns - 1] > 0 && st.ns < 12) {
        double r[11];
        const int dr = st.sd[st.ns - 2];
        for (int i = 0; i <= dr; ++i) r[i] = st.S[st.ns - 2][i];
        const int dd = st.sd[st.ns - 1];
        const double* d = st.S[st.ns - 1];
        for (int k = dr; k >= dd; --k) {
            const double f = r[k] / d[dd];
            for (int j = 0; j <= dd; ++j) r[k - dd + j] -= f * d[j];
            r[k] = 0;
        }
        int nd = dd - 1;
        double rm = 0;
        for (int i = 0; i <= nd; ++i) if (fabs(r[i]) > rm) rm = fabs(r[i]);
        if (!(rm > 1e-300)) break;
        while (nd > 0 && fabs(r[nd]) < 1e-14 * rm) --nd;
        for (int i = 0; i <= nd; ++i) st.S[st.ns][i] = -r[i] / rm;
        st.sd[st.ns] = nd;
        ++st.ns;
    }
    double B = 0;
    for (int k = 1; k <= deg; ++k) { const double v = pow(fabs(p[deg - k] / p[deg]), 1.0 / k); if (v > B) B = v; }
    B = 2.0 * B + 1e-9;
    double lo_s[64], hi_s[64];
    int nlo_s[64], nhi_s[64];
    int sp = 0, nr = 0;
    lo_s[0] = -B; hi_s[0] = B; nlo_s[0] = st.changes(-B); nhi_s[0] = st.changes(B); sp = 1;
    while (sp > 0 && nr < 10) {
        --sp;
        double lo = lo_s[sp], hi = hi_s[sp];
        int a = nlo_s[sp];
        const int b = nhi_s[sp];
        const int cnt = a - b;
        if (cnt <= 0) continue;
        if (cnt > 1 && (hi - lo) > 1e-13 * (1.0 + fabs(lo) + fabs(hi))) {
            const double mid = 0.5 * (lo + hi);
            const int m = st.changes(mid);
            if (sp + 2 > 64) continue;
            lo_s[sp] = mid; hi_s[sp] = hi; nlo_s[sp] = m; nhi_s[sp] = b; ++sp;
            lo_s[sp] = lo; hi_s[sp] = mid; nlo_s[sp] = a; nhi_s[sp] = m; ++sp;
            continue;
        }
        if (cnt > 1) { roots[nr++] = 0.5 * (lo + hi); continue; }
        const double flo = horner(p, deg, lo), fhi = horner(p, deg, hi);
        if (fhi == 0.0) { roots[nr++] = hi; continue; }
        double x = 0.5 * (lo + hi);
        if ((flo > 0) == (fhi > 0)) {
            for (int it = 0; it < 80 && (hi - lo) > 1e-15 * (1.0 + fabs(x)); ++it) {
                x = 0.5 * (lo + hi);
                const int m = st.changes(x);
                if (a - m >= 1) hi = x; else { lo = x; a = m; }
            }
            roots[nr++] = 0.5 * (lo + hi);
            continue;
        }
        double dxold = hi - lo;
        for (int it = 0; it < 200; ++it) {
            const double f = horner(p, deg, x);
            if (f == 0.0) break;
            if ((f > 0) == (flo > 0)) lo = x; else hi = x;
            double dp = 0;
            for (int i = deg; i >= 1; --i) dp = dp * x + i * p[i];
            double xn = (dp != 0.0) ? x - f / dp : x;
            if (!(xn > lo && xn < hi) || fabs(xn - x) > 0.5 * fabs(dxold)) xn = 0.5 * (lo + hi);
            dxold = xn - x;
            if (fabs(dxold) <= 4e-16 * (1.0 + fabs(x))) { x = xn; break; }
            x = xn;
        }
        roots[nr++] = x;
    }
    return nr;
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
__device__ inline void polish_xyz(const double N[4][9], double* xyz)
{
    for (int it = 0; it < 3; ++it) {
        double E[9], r[10], J[10][3];
        for (int e = 0; e < 9; ++e) E[e] = xyz[0] * N[0][e] + xyz[1] * N[1][e] + xyz[2] * N[2][e] + N[3][e];
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
            const double* D = N[d];
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

// Five normalised correspondences -> up to kMaxCand essential matrices (row-major, unit Frobenius norm), ascending z.
__device__ inline int fivept(const double* x1, const double* y1, const double* x2, const double* y2, double* Eout)
{
    double N[4][9];
    if (!nullspace_5x9(x1, y1, x2, y2, N)) return 0;
    double Ep[9][4];
    for (int e = 0; e < 9; ++e) for (int k = 0; k < 4; ++k) Ep[e][k] = N[k][e];
    double C[10][20];
    for (int i = 0; i < 10; ++i) for (int j = 0; j < 20; ++j) C[i][j] = 0;
    {   // det(E) = 0
        double m[3][10];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 10; ++j) m[i][j] = 0;
        p11_acc(Ep[4], Ep[8], 1.0, m[0]); p11_acc(Ep[5], Ep[7], -1.0, m[0]);
        p11_acc(Ep[3], Ep[8], 1.0, m[1]); p11_acc(Ep[5], Ep[6], -1.0, m[1]);
        p11_acc(Ep[3], Ep[7], 1.0, m[2]); p11_acc(Ep[4], Ep[6], -1.0, m[2]);
        p21_acc(m[0], Ep[0], 1.0, C[0]); p21_acc(m[1], Ep[1], -1.0, C[0]); p21_acc(m[2], Ep[2], 1.0, C[0]);
    }
    {   // (E E^T - 1/2 tr(E E^T) I) E = 0
        double G[3][3][10];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) for (int k = 0; k < 10; ++k) G[i][j][k] = 0;
        for (int i = 0; i < 3; ++i)
            for (int j = i; j < 3; ++j) {
                for (int k = 0; k < 3; ++k) p11_acc(Ep[3 * i + k], Ep[3 * j + k], 1.0, G[i][j]);
                if (j != i) for (int k = 0; k < 10; ++k) G[j][i][k] = G[i][j][k];
            }
        double tr[10];
        for (int k = 0; k < 10; ++k) tr[k] = 0.5 * (G[0][0][k] + G[1][1][k] + G[2][2][k]);
        for (int i = 0; i < 3; ++i) for (int k = 0; k < 10; ++k) G[i][i][k] -= tr[k];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 3; ++k) p21_acc(G[i][k], Ep[3 * k + j], 1.0, C[1 + 3 * i + j]);
    }
    for (int k = 0; k < 10; ++k) {   // Gauss-Jordan with row pivoting on the first ten columns
        int pr = k;
        double best = fabs(C[k][k]);
        for (int r = k + 1; r < 10; ++r) if (fabs(C[r][k]) > best) { best = fabs(C[r][k]); pr = r; }
        if (!(best > 1e-300)) return 0;
        if (pr != k) for (int c = 0; c < 20; ++c) { const double t = C[k][c]; C[k][c] = C[pr][c]; C[pr][c] = t; }
        const double inv = 1.0 / C[k][k];
        for (int c = k; c < 20; ++c) C[k][c] *= inv;
        for (int r = 0; r < 10; ++r)
            if (r != k) { const double f = C[r][k]; if (f != 0.0) for (int c = k; c < 20; ++c) C[r][c] -= f * C[k][c]; }
    }
    // B(z) [x y 1]^T = 0; columns 10 xz^2, 11 xz, 12 x, 13 yz^2, 14 yz, 15 y, 16 z^3, 17 z^2, 18 z, 19 1
    double Bx[3][4], By[3][4], Bc[3][5];
    for (int q = 0; q < 3; ++q) {
        const double* a = C[4 + 2 * q];
        const double* b = C[5 + 2 * q];
        Bx[q][0] = a[12]; Bx[q][1] = a[11] - b[12]; Bx[q][2] = a[10] - b[11]; Bx[q][3] = -b[10];
        By[q][0] = a[15]; By[q][1] = a[14] - b[15]; By[q][2] = a[13] - b[14]; By[q][3] = -b[13];
        Bc[q][0] = a[19]; Bc[q][1] = a[18] - b[19]; Bc[q][2] = a[17] - b[18]; Bc[q][3] = a[16] - b[17]; Bc[q][4] = -b[16];
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
    double roots[10];
    const int nr = real_roots_sturm(poly, 10, roots);
    int ne = 0;
    for (int r = 0; r < nr && ne < kMaxCand; ++r) {
        const double z = roots[r];
        double bm[3][3];
        for (int q = 0; q < 3; ++q) { bm[q][0] = horner(Bx[q], 3, z); bm[q][1] = horner(By[q], 3, z); bm[q][2] = horner(Bc[q], 4, z); }
        double best = -1, bv[3] = {0, 0, 0};
        for (int a = 0; a < 3; ++a)
            for (int b = a + 1; b < 3; ++b) {
                double v[3];
                cross3(bm[a], bm[b], v);
                if (fabs(v[2]) > best) { best = fabs(v[2]); bv[0] = v[0]; bv[1] = v[1]; bv[2] = v[2]; }
            }
        if (!(best > 0)) continue;
        double xyz[3] = {bv[0] / bv[2], bv[1] / bv[2], z};
        if (!isfinite(xyz[0]) || !isfinite(xyz[1])) continue;
        polish_xyz(N, xyz);
        double E[9], nn = 0;
        for (int e = 0; e < 9; ++e) { E[e] = xyz[0] * N[0][e] + xyz[1] * N[1][e] + xyz[2] * N[2][e] + N[3][e]; nn += E[e] * E[e]; }
        if (!(nn > 0) || !isfinite(nn)) continue;
        nn = 1.0 / sqrt(nn);
        for (int e = 0; e < 9; ++e) Eout[9 * ne + e] = E[e] * nn;
        ++ne;
    }
    return ne;
}

}  // namespace vm
