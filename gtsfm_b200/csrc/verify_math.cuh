// fp64 device math of the verifier: small dense helpers, symmetric Jacobi eigen-solver, essential-matrix
// decomposition, DLT triangulation, squared Sampson distance, calibration.
//
//   squared Sampson distance      gtsfm/utils/verification.py:170-214
//   normalize_coordinates         gtsfm/utils/features.py:40-50  (gtsam Cal3Bundler.calibrate, un-vendored gtsam 4.2)
//   cv2.recoverPose               call site gtsfm/utils/verification.py:81 (un-vendored opencv-python): SVD of E ->
//                                 {R1,R2} x {+t,-t}, DLT triangulation of every inlier, cheirality vote (distance 50)
#pragma once
#include <cstdint>

namespace vm {

__device__ __forceinline__ void mat3_mul(const double* A, const double* B, double* C)
{
    double T[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) T[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
#pragma unroll
    for (int i = 0; i < 9; ++i) C[i] = T[i];
}
__device__ __forceinline__ void cross3(const double* a, const double* b, double* c)
{
    const double t0 = a[1] * b[2] - a[2] * b[1], t1 = a[2] * b[0] - a[0] * b[2], t2 = a[0] * b[1] - a[1] * b[0];
    c[0] = t0; c[1] = t1; c[2] = t2;
}
__device__ __forceinline__ double norm3(const double* a) { return sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }

// Cyclic Jacobi eigen-decomposition of a symmetric N x N matrix (N = 3 or 4).  A is destroyed; w = eigenvalues,
// V columns = eigenvectors (row-major).
template <int N> __device__ void jacobi_eig(double* A, double* w, double* V)
{
    for (int i = 0; i < N; ++i) for (int j = 0; j < N; ++j) V[i * N + j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0, diag = 0;
        for (int i = 0; i < N; ++i) {
            diag += A[i * N + i] * A[i * N + i];
            for (int j = i + 1; j < N; ++j) off += A[i * N + j] * A[i * N + j];
        }
        if (off <= 1e-32 * diag || off == 0.0) break;
        for (int p = 0; p < N - 1; ++p)
            for (int q = p + 1; q < N; ++q) {
                const double apq = A[p * N + q];
                if (apq == 0.0) continue;
                const double theta = (A[q * N + q] - A[p * N + p]) / (2.0 * apq);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < N; ++k) {
                    const double akp = A[k * N + p], akq = A[k * N + q];
                    A[k * N + p] = c * akp - s * akq; A[k * N + q] = s * akp + c * akq;
                }
                for (int k = 0; k < N; ++k) {
                    const double apk = A[p * N + k], aqk = A[q * N + k];
                    A[p * N + k] = c * apk - s * aqk; A[q * N + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < N; ++k) {
                    const double vkp = V[k * N + p], vkq = V[k * N + q];
                    V[k * N + p] = c * vkp - s * vkq; V[k * N + q] = s * vkp + c * vkq;
                }
            }
    }
    for (int i = 0; i < N; ++i) w[i] = A[i * N + i];
}

// SVD of a (near-)essential matrix with proper rotations U, V: E ~ U diag(s1, s2, ~0) V^T.
__device__ inline void essential_svd(const double* E, double* U, double* V)
{
    double A[9], w[3], Q[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) A[3 * i + j] = E[i] * E[j] + E[3 + i] * E[3 + j] + E[6 + i] * E[6 + j];
    jacobi_eig<3>(A, w, Q);
    int i0 = 0, i1 = 1, i2 = 2, t;
    if (w[i0] < w[i1]) { t = i0; i0 = i1; i1 = t; }
    if (w[i0] < w[i2]) { t = i0; i0 = i2; i2 = t; }
    if (w[i1] < w[i2]) { t = i1; i1 = i2; i2 = t; }
    double v1[3] = {Q[i0], Q[3 + i0], Q[6 + i0]}, v2[3] = {Q[i1], Q[3 + i1], Q[6 + i1]}, v3[3];
    cross3(v1, v2, v3);
    double u1[3], u2[3], u3[3];
    for (int i = 0; i < 3; ++i) {
        u1[i] = E[3 * i] * v1[0] + E[3 * i + 1] * v1[1] + E[3 * i + 2] * v1[2];
        u2[i] = E[3 * i] * v2[0] + E[3 * i + 1] * v2[1] + E[3 * i + 2] * v2[2];
    }
    const double n1 = norm3(u1);
    for (int i = 0; i < 3; ++i) u1[i] /= n1;
    const double d = u1[0] * u2[0] + u1[1] * u2[1] + u1[2] * u2[2];
    for (int i = 0; i < 3; ++i) u2[i] -= d * u1[i];
    const double n2 = norm3(u2);
    for (int i = 0; i < 3; ++i) u2[i] /= n2;
    cross3(u1, u2, u3);
    for (int i = 0; i < 3; ++i) {
        U[3 * i] = u1[i]; U[3 * i + 1] = u2[i]; U[3 * i + 2] = u3[i];
        V[3 * i] = v1[i]; V[3 * i + 1] = v2[i]; V[3 * i + 2] = v3[i];
    }
}

// cv::decomposeEssentialMat: R1 = U W V^T, R2 = U W^T V^T, t = U[:,2]
__device__ inline void decompose_essential(const double* E, double* R1, double* R2, double* t)
{
    double U[9], V[9], Vt[9], T[9];
    const double W[9] = {0, 1, 0, -1, 0, 0, 0, 0, 1}, Wt[9] = {0, -1, 0, 1, 0, 0, 0, 0, 1};
    essential_svd(E, U, V);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Vt[3 * i + j] = V[3 * j + i];
    mat3_mul(U, W, T); mat3_mul(T, Vt, R1);
    mat3_mul(U, Wt, T); mat3_mul(T, Vt, R2);
    t[0] = U[2]; t[1] = U[5]; t[2] = U[8];
}

// (x2^T F x1)^2 / ((F x1)_0^2 + (F x1)_1^2 + (F^T x2)_0^2 + (F^T x2)_1^2)
__device__ __forceinline__ double sampson_sq(const double* F, double x1, double y1, double x2, double y2)
{
    const double l0 = F[0] * x1 + F[1] * y1 + F[2], l1 = F[3] * x1 + F[4] * y1 + F[5], l2 = F[6] * x1 + F[7] * y1 + F[8];
    const double m0 = F[0] * x2 + F[3] * y2 + F[6], m1 = F[1] * x2 + F[4] * y2 + F[7];
    const double num = x2 * l0 + y2 * l1 + l2, den = l0 * l0 + l1 * l1 + m0 * m0 + m1 * m1;
    return num * num / den;
}

// gtsam::Cal3Bundler::calibrate: exact ((u-u0)/f, (v-v0)/f) when k1 = k2 = 0, fixed-point undistortion otherwise.
__device__ __forceinline__ void calibrate(double u, double v, const double* cal, double* xn, double* yn)
{
    const double f = cal[0], k1 = cal[1], k2 = cal[2], u0 = cal[3], v0 = cal[4];
    const double px = (u - u0) / f, py = (v - v0) / f;
    const double rr = px * px + py * py;
    const double g = 1.0 + (k1 + k2 * rr) * rr;
    double x = px / g, y = py / g;
    if (k1 != 0.0 || k2 != 0.0) {
        for (int it = 0; it < 10; ++it) {
            const double r2 = x * x + y * y, gg = 1.0 + (k1 + k2 * r2) * r2;
            const double uu = f * gg * x + u0, vv = f * gg * y + v0;
            const double du = uu - u, dv = vv - v;
            if (du * du + dv * dv <= 1e-5) break;
            x = px / gg; y = py / gg;
        }
    }
    *xn = x; *yn = y;
}

// Right singular vector of the smallest singular value of the 4x4 DLT matrix A (rows r0..r3), i.e. the eigenvector of
// the smallest eigenvalue of M = A^T A, by LDL^T + inverse iteration: with camera 0 = [I | 0] the leading pivots are
// >= 1, the last pivot carries the (tiny) smallest eigenvalue, L^-T e3 is already the null vector up to
// O(lambda_min / lambda_2), and each further solve with M gains that factor again.  ~200 flops, no iteration on the
// matrix itself (a cyclic Jacobi sweep per inlier and pose candidate dominated the finalize kernel).
__device__ __forceinline__ void dlt_null4(const double* A, double* Q)
{
    double m[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) m[i][j] = A[i] * A[j] + A[4 + i] * A[4 + j] + A[8 + i] * A[8 + j] + A[12 + i] * A[12 + j];
    const double eps = 1e-15 * (m[0][0] + m[1][1] + m[2][2] + m[3][3]) + 1e-300;
    const double d0 = m[0][0] + eps, i0 = 1.0 / d0;
    const double l10 = m[1][0] * i0, l20 = m[2][0] * i0, l30 = m[3][0] * i0;
    const double d1 = m[1][1] + eps - l10 * m[1][0], i1 = 1.0 / d1;
    const double t21 = m[2][1] - l20 * m[1][0], t31 = m[3][1] - l30 * m[1][0];
    const double l21 = t21 * i1, l31 = t31 * i1;
    const double d2 = m[2][2] + eps - l20 * m[2][0] - l21 * t21, i2 = 1.0 / d2;
    const double t32 = m[3][2] - l30 * m[2][0] - l31 * t21;
    const double l32 = t32 * i2;
    double d3 = m[3][3] + eps - l30 * m[3][0] - l31 * t31 - l32 * t32;
    if (!(d3 > 0)) d3 = eps;
    const double i3 = 1.0 / d3;
    // x = L^-T e3
    double x3 = 1.0, x2 = -l32, x1 = -l21 * x2 - l31, x0 = -l10 * x1 - l20 * x2 - l30;
#pragma unroll
    for (int it = 0; it < 3; ++it) {
        const double sc = 1.0 / fmax(fmax(fabs(x0), fabs(x1)), fmax(fabs(x2), fabs(x3)));
        x0 *= sc; x1 *= sc; x2 *= sc; x3 *= sc;
        // L z = x
        const double z0 = x0, z1 = x1 - l10 * z0, z2 = x2 - l20 * z0 - l21 * z1, z3 = x3 - l30 * z0 - l31 * z1 - l32 * z2;
        // L^T y = D^-1 z
        x3 = z3 * i3;
        x2 = z2 * i2 - l32 * x3;
        x1 = z1 * i1 - l21 * x2 - l31 * x3;
        x0 = z0 * i0 - l10 * x1 - l20 * x2 - l30 * x3;
    }
    const double n = 1.0 / sqrt(x0 * x0 + x1 * x1 + x2 * x2 + x3 * x3);
    Q[0] = x0 * n; Q[1] = x1 * n; Q[2] = x2 * n; Q[3] = x3 * n;
}

// cv::triangulatePoints for one point + the cheirality test of cv::recoverPose (distance threshold 50).
// P1 = [R | t] row-major 3x4; camera 0 is [I | 0].
__device__ inline bool cheirality_ok(const double* P1, double ax, double ay, double bx, double by)
{
    double A[16], Q[4];
    A[0] = -1.0; A[1] = 0.0; A[2] = ax; A[3] = 0.0;
    A[4] = 0.0; A[5] = -1.0; A[6] = ay; A[7] = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        A[8 + k] = bx * P1[8 + k] - P1[k];
        A[12 + k] = by * P1[8 + k] - P1[4 + k];
    }
    dlt_null4(A, Q);
    const double Q0 = Q[0], Q1 = Q[1], Q2 = Q[2], Q3 = Q[3];
    bool ok = (Q2 * Q3) > 0;
    const double X = Q0 / Q3, Y = Q1 / Q3, Z = Q2 / Q3;
    ok = ok && (Z < 50.0);
    const double z2 = P1[8] * X + P1[9] * Y + P1[10] * Z + P1[11];
    return ok && (z2 > 0) && (z2 < 50.0);
}

}  // namespace vm
