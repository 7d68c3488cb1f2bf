/*
 * ORACLE (test infrastructure only -- never linked, imported or executed by the product path).
 *
 * CPU restatement (fp64) of the reference's Ransac verifier in essential-matrix mode, driven by an
 * explicit (H,5) hypothesis-sample index array so that a GPU run with the same samples can be
 * compared mask-for-mask (BASELINE.json north_star).
 *
 * Follows:
 *   gtsfm/frontend/verifier/opencv_verifier_base.py:45-109  gates (M<5, M<6), fx=max(fx1,fx2),
 *                                                           mask -> v_corr_idxs, inlier ratio, pose
 *   gtsfm/frontend/verifier/ransac.py:52-82                 identity K, threshold = px / fx
 *   gtsfm/utils/features.py:40-50                           normalize_coordinates (Cal3Bundler.calibrate)
 *   gtsfm/utils/verification.py:170-214                     squared Sampson distance
 *   gtsfm/utils/verification.py:52-94                       recoverPose (cv2) restated in recover_pose()
 * The robust estimator itself lives in the un-vendored dependency opencv-python (cv2.findEssentialMat,
 * USAC_ACCURATE) whose sample sets are not observable; what is restated here is its published
 * structure: Nister's 5-point minimal solver (Nister, TPAMI 2004 -- cited at ransac.py:6), MSAC
 * scoring sum(min(sampson^2, thr^2)), best model = lowest cost (ties: lowest hypothesis/candidate
 * index) among the models that pass the oriented epipolar constraint on their own sample, a local optimisation (Levenberg-Marquardt on the Sampson residuals of the current inliers)
 * and the final mask sampson^2 < thr^2 (strict).  PARITY NOTE: the minimal solver is pinned against
 * cv2's own 5-point solver per sample (tests/test_oracle_verify.py), recover_pose against
 * cv2.recoverPose, the Sampson arithmetic against the reference's numpy functions and the reference
 * tests' known answers; the sample-set-driven RANSAC loop as a whole is "parity unpinned" versus
 * black-box USAC_ACCURATE (only statistical agreement is reported).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MAXC 10 /* max essential matrices per 5-point sample */

/* ---------------------------------------------------------------- small dense helpers */

static void mat3_mul(const double* A, const double* B, double* C)
{
    double T[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            T[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
    memcpy(C, T, sizeof(T));
}

static void cross3(const double* a, const double* b, double* c)
{
    double t0 = a[1] * b[2] - a[2] * b[1];
    double t1 = a[2] * b[0] - a[0] * b[2];
    double t2 = a[0] * b[1] - a[1] * b[0];
    c[0] = t0; c[1] = t1; c[2] = t2;
}

static double norm3(const double* a) { return sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }

/* Cyclic Jacobi eigen-decomposition of a symmetric n x n matrix (n <= 4).  A is destroyed; on exit
 * w = eigenvalues, V columns = eigenvectors (V row-major n x n). */
static void jacobi_eig(double* A, int n, double* w, double* V)
{
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) V[i * n + j] = (i == j);
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0, diag = 0;
        for (int i = 0; i < n; ++i) { diag += A[i * n + i] * A[i * n + i]; for (int j = i + 1; j < n; ++j) off += A[i * n + j] * A[i * n + j]; }
        if (off <= 1e-32 * diag || off == 0.0) break;
        for (int p = 0; p < n - 1; ++p)
            for (int q = p + 1; q < n; ++q) {
                double apq = A[p * n + q];
                if (apq == 0.0) continue;
                double theta = (A[q * n + q] - A[p * n + p]) / (2.0 * apq);
                double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < n; ++k) {           /* A <- A J */
                    double akp = A[k * n + p], akq = A[k * n + q];
                    A[k * n + p] = c * akp - s * akq; A[k * n + q] = s * akp + c * akq;
                }
                for (int k = 0; k < n; ++k) {           /* A <- J^T A */
                    double apk = A[p * n + k], aqk = A[q * n + k];
                    A[p * n + k] = c * apk - s * aqk; A[q * n + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < n; ++k) {
                    double vkp = V[k * n + p], vkq = V[k * n + q];
                    V[k * n + p] = c * vkp - s * vkq; V[k * n + q] = s * vkp + c * vkq;
                }
            }
    }
    for (int i = 0; i < n; ++i) w[i] = A[i * n + i];
}

/* SVD of a (near-)essential 3x3 matrix, returning proper rotations U, V (det = +1) such that
 * E ~ U diag(s1,s2,~0) V^T.  This is what cv::decomposeEssentialMat needs (it flips U, Vt to
 * det>0).  Columns: v1,v2 = top eigenvectors of E^T E, v3 = v1 x v2; u1 = E v1/|.|, u2 =
 * Gram-Schmidt(E v2), u3 = u1 x u2. */
static void essential_svd(const double* E, double* U, double* V)
{
    double A[9], w[3], Q[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            A[3 * i + j] = E[i] * E[j] + E[3 + i] * E[3 + j] + E[6 + i] * E[6 + j];
    jacobi_eig(A, 3, w, Q);
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
    double n1 = norm3(u1);
    for (int i = 0; i < 3; ++i) u1[i] /= n1;
    double d = u1[0] * u2[0] + u1[1] * u2[1] + u1[2] * u2[2];
    for (int i = 0; i < 3; ++i) u2[i] -= d * u1[i];
    double n2 = norm3(u2);
    for (int i = 0; i < 3; ++i) u2[i] /= n2;
    cross3(u1, u2, u3);
    for (int i = 0; i < 3; ++i) {
        U[3 * i] = u1[i]; U[3 * i + 1] = u2[i]; U[3 * i + 2] = u3[i];
        V[3 * i] = v1[i]; V[3 * i + 1] = v2[i]; V[3 * i + 2] = v3[i];
    }
}

/* cv::decomposeEssentialMat: R1 = U W V^T, R2 = U W^T V^T, t = U[:,2], W = [0 1 0; -1 0 0; 0 0 1]. */
void oracle_decompose_essential(const double* E, double* R1, double* R2, double* t)
{
    double U[9], V[9], Vt[9], UW[9], UWt[9];
    const double W[9] = {0, 1, 0, -1, 0, 0, 0, 0, 1}, Wt[9] = {0, -1, 0, 1, 0, 0, 0, 0, 1};
    essential_svd(E, U, V);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Vt[3 * i + j] = V[3 * j + i];
    mat3_mul(U, W, UW); mat3_mul(UW, Vt, R1);
    mat3_mul(U, Wt, UWt); mat3_mul(UWt, Vt, R2);
    t[0] = U[2]; t[1] = U[5]; t[2] = U[8];
}

/* ---------------------------------------------------------------- normalisation + Sampson */

/* gtsam::Cal3Bundler::calibrate (gtsam 4.2, un-vendored; call site gtsfm/utils/features.py:50):
 * pn = ((u-u0)/f, (v-v0)/f) / g(pn), g = 1 + k1 r^2 + k2 r^4, fixed-point iteration (<= 10 rounds,
 * tolerance 1e-5 on the re-projected pixel).  With k1 = k2 = 0 (every pipeline run,
 * gtsfm/loader/loader_base.py:218-224) this is exactly ((u-u0)/f, (v-v0)/f). */
void oracle_calibrate(double u, double v, double f, double k1, double k2, double u0, double v0,
                      double* xn, double* yn)
{
    double px = (u - u0) / f, py = (v - v0) / f;
    double rr = px * px + py * py;
    double g = 1.0 + (k1 + k2 * rr) * rr;
    double x = px / g, y = py / g;
    if (k1 != 0.0 || k2 != 0.0) {
        for (int it = 0; it < 10; ++it) {
            double r2 = x * x + y * y, gg = 1.0 + (k1 + k2 * r2) * r2;
            double uu = f * gg * x + u0, vv = f * gg * y + v0;
            double du = uu - u, dv = vv - v;
            if (du * du + dv * dv <= 1e-5) break;
            x = px / gg; y = py / gg;
        }
    }
    *xn = x; *yn = y;
}

/* Squared Sampson distance exactly as gtsfm/utils/verification.py:206-214:
 * (x2^T F x1)^2 / ((F x1)_0^2 + (F x1)_1^2 + (F^T x2)_0^2 + (F^T x2)_1^2). */
double oracle_sampson_sq(const double* F, double x1, double y1, double x2, double y2)
{
    double l0 = F[0] * x1 + F[1] * y1 + F[2];
    double l1 = F[3] * x1 + F[4] * y1 + F[5];
    double l2 = F[6] * x1 + F[7] * y1 + F[8];
    double m0 = F[0] * x2 + F[3] * y2 + F[6];
    double m1 = F[1] * x2 + F[4] * y2 + F[7];
    double num = x2 * l0 + y2 * l1 + l2;
    double den = l0 * l0 + l1 * l1 + m0 * m0 + m1 * m1;
    return num * num / den;
}

void oracle_sampson_sq_batch(const double* F, const double* x1, const double* y1, const double* x2,
                             const double* y2, int n, double* out)
{
    for (int i = 0; i < n; ++i) out[i] = oracle_sampson_sq(F, x1[i], y1[i], x2[i], y2[i]);
}

static double msac_cost(const double* E, const double* x1, const double* y1, const double* x2,
                        const double* y2, int n, double thr2, int* ninl)
{
    double c = 0; int k = 0;
    for (int i = 0; i < n; ++i) {
        double s = oracle_sampson_sq(E, x1[i], y1[i], x2[i], y2[i]);
        if (s < thr2) { c += s; ++k; } else c += thr2;    /* NaN counts as outlier */
    }
    if (ninl) *ninl = k;
    return c;
}

/* ---------------------------------------------------------------- cv::recoverPose restatement */

/* cv::triangulatePoints for one point: X = right singular vector of the 4x4 DLT matrix with the
 * smallest singular value (= eigenvector of A^T A with the smallest eigenvalue). */
static void triangulate_dlt(const double* P0, const double* P1, double ax, double ay, double bx, double by, double* X)
{
    double A[16], AtA[16], w[4], V[16];
    for (int k = 0; k < 4; ++k) {
        A[0 * 4 + k] = ax * P0[8 + k] - P0[k];
        A[1 * 4 + k] = ay * P0[8 + k] - P0[4 + k];
        A[2 * 4 + k] = bx * P1[8 + k] - P1[k];
        A[3 * 4 + k] = by * P1[8 + k] - P1[4 + k];
    }
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0;
            for (int r = 0; r < 4; ++r) s += A[r * 4 + i] * A[r * 4 + j];
            AtA[i * 4 + j] = s;
        }
    jacobi_eig(AtA, 4, w, V);
    int m = 0;
    for (int i = 1; i < 4; ++i) if (w[i] < w[m]) m = i;
    for (int i = 0; i < 4; ++i) X[i] = V[i * 4 + m];
}

/* Cheirality vote of cv::recoverPose (distanceThresh = 50): for each of (R1,t),(R2,t),(R1,-t),
 * (R2,-t) count points with z*w>0, z/w<50 in camera 1 and 0<z<50 in camera 2; the first candidate
 * that is >= all others wins.  x1,y1,x2,y2 are normalised coordinates of the inliers.
 * Outputs R (row-major 3x3), t (3) and the four counts. */
void oracle_recover_pose(const double* E, const double* x1, const double* y1, const double* x2,
                         const double* y2, int n, double* R, double* t, int* good)
{
    double R1[9], R2[9], tt[3];
    oracle_decompose_essential(E, R1, R2, tt);
    const double P0[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    const double dist = 50.0;
    for (int c = 0; c < 4; ++c) {
        const double* Rc = (c & 1) ? R2 : R1;
        double sg = (c & 2) ? -1.0 : 1.0;
        double P1[12];
        for (int i = 0; i < 3; ++i) { P1[4 * i] = Rc[3 * i]; P1[4 * i + 1] = Rc[3 * i + 1]; P1[4 * i + 2] = Rc[3 * i + 2]; P1[4 * i + 3] = sg * tt[i]; }
        int g = 0;
        for (int k = 0; k < n; ++k) {
            double Q[4];
            triangulate_dlt(P0, P1, x1[k], y1[k], x2[k], y2[k], Q);
            int ok = (Q[2] * Q[3]) > 0;
            double X = Q[0] / Q[3], Y = Q[1] / Q[3], Z = Q[2] / Q[3];
            ok = ok && (Z < dist);
            double z2 = P1[8] * X + P1[9] * Y + P1[10] * Z + P1[11];
            ok = ok && (z2 > 0) && (z2 < dist);
            g += ok;
        }
        good[c] = g;
    }
    int best;
    if (good[0] >= good[1] && good[0] >= good[2] && good[0] >= good[3]) best = 0;
    else if (good[1] >= good[0] && good[1] >= good[2] && good[1] >= good[3]) best = 1;
    else if (good[2] >= good[0] && good[2] >= good[1] && good[2] >= good[3]) best = 2;
    else best = 3;
    memcpy(R, (best & 1) ? R2 : R1, sizeof(double) * 9);
    double sg = (best & 2) ? -1.0 : 1.0;
    for (int i = 0; i < 3; ++i) t[i] = sg * tt[i];
}

/* ---------------------------------------------------------------- 5-point minimal solver (Nister 2004) */

/* Monomials of degree <= 3 in (x,y,z), ordered so that Gauss-Jordan on the first ten columns leaves
 * the rows for x^2 z, x^2, y^2 z, y^2, xyz, xy expressed in [x z^2, x z, x, y z^2, y z, y, z^3, z^2, z, 1]. */
static const int M3[20][3] = {
    {3,0,0},{0,3,0},{2,1,0},{1,2,0},{2,0,1},{2,0,0},{0,2,1},{0,2,0},{1,1,1},{1,1,0},
    {1,0,2},{1,0,1},{1,0,0},{0,1,2},{0,1,1},{0,1,0},{0,0,3},{0,0,2},{0,0,1},{0,0,0}};
/* degree <= 2: x^2 y^2 z^2 xy xz yz x y z 1 */
static const int M2[10][3] = {{2,0,0},{0,2,0},{0,0,2},{1,1,0},{1,0,1},{0,1,1},{1,0,0},{0,1,0},{0,0,1},{0,0,0}};
/* degree <= 1: x y z 1 */
static const int M1[4][3] = {{1,0,0},{0,1,0},{0,0,1},{0,0,0}};

static int idx11[4][4], idx21[10][4], tables_ready = 0;

static void build_tables(void)
{
    if (tables_ready) return;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            for (int k = 0; k < 10; ++k)
                if (M1[i][0] + M1[j][0] == M2[k][0] && M1[i][1] + M1[j][1] == M2[k][1] && M1[i][2] + M1[j][2] == M2[k][2]) idx11[i][j] = k;
    for (int i = 0; i < 10; ++i)
        for (int j = 0; j < 4; ++j)
            for (int k = 0; k < 20; ++k)
                if (M2[i][0] + M1[j][0] == M3[k][0] && M2[i][1] + M1[j][1] == M3[k][1] && M2[i][2] + M1[j][2] == M3[k][2]) idx21[i][j] = k;
    tables_ready = 1;
}

/* c (deg 2) += s * a (deg 1) * b (deg 1) */
static void p11_acc(const double* a, const double* b, double s, double* c)
{
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) c[idx11[i][j]] += s * a[i] * b[j];
}
/* c (deg 3) += s * a (deg 2) * b (deg 1) */
static void p21_acc(const double* a, const double* b, double s, double* c)
{
    for (int i = 0; i < 10; ++i) for (int j = 0; j < 4; ++j) c[idx21[i][j]] += s * a[i] * b[j];
}

/* Null space of the 5x9 epipolar constraint matrix by Gauss-Jordan with complete pivoting.
 * Row for a correspondence: x2^T E x1 = 0 with E row-major ->
 * [x2x1, x2y1, x2, y2x1, y2y1, y2, x1, y1, 1].  Returns 4 basis vectors N[4][9]; 0 on rank loss. */
static int nullspace_5x9(const double* x1, const double* y1, const double* x2, const double* y2, double N[4][9])
{
    double A[5][9]; int perm[9];
    for (int r = 0; r < 5; ++r) {
        A[r][0] = x2[r] * x1[r]; A[r][1] = x2[r] * y1[r]; A[r][2] = x2[r];
        A[r][3] = y2[r] * x1[r]; A[r][4] = y2[r] * y1[r]; A[r][5] = y2[r];
        A[r][6] = x1[r];         A[r][7] = y1[r];         A[r][8] = 1.0;
    }
    for (int c = 0; c < 9; ++c) perm[c] = c;
    for (int k = 0; k < 5; ++k) {
        int pr = k, pc = k; double best = -1.0;
        for (int r = k; r < 5; ++r) for (int c = k; c < 9; ++c) { double v = fabs(A[r][c]); if (v > best) { best = v; pr = r; pc = c; } }
        if (!(best > 1e-14)) return 0;
        if (pr != k) for (int c = 0; c < 9; ++c) { double t = A[k][c]; A[k][c] = A[pr][c]; A[pr][c] = t; }
        if (pc != k) { for (int r = 0; r < 5; ++r) { double t = A[r][k]; A[r][k] = A[r][pc]; A[r][pc] = t; } int t = perm[k]; perm[k] = perm[pc]; perm[pc] = t; }
        double inv = 1.0 / A[k][k];
        for (int c = k; c < 9; ++c) A[k][c] *= inv;
        for (int r = 0; r < 5; ++r) if (r != k) { double f = A[r][k]; if (f != 0.0) for (int c = k; c < 9; ++c) A[r][c] -= f * A[k][c]; }
    }
    /* A = [I5 | B] in permuted columns; null vectors: free column 5+j set to 1, pivots = -B[:,j] */
    for (int j = 0; j < 4; ++j) {
        double v[9];
        for (int r = 0; r < 5; ++r) v[r] = -A[r][5 + j];
        for (int c = 5; c < 9; ++c) v[c] = (c == 5 + j) ? 1.0 : 0.0;
        double nn = 0; for (int c = 0; c < 9; ++c) nn += v[c] * v[c];
        nn = 1.0 / sqrt(nn);
        for (int c = 0; c < 9; ++c) N[j][perm[c]] = v[c] * nn;
    }
    return 1;
}

/* ---- polynomial helpers in one variable (coefficient i multiplies z^i) */
static double horner(const double* p, int deg, double z) { double v = p[deg]; for (int i = deg - 1; i >= 0; --i) v = v * z + p[i]; return v; }

static void pmul(const double* a, int da, const double* b, int db, double* c) /* c = a*b, deg da+db */
{
    for (int i = 0; i <= da + db; ++i) c[i] = 0;
    for (int i = 0; i <= da; ++i) for (int j = 0; j <= db; ++j) c[i + j] += a[i] * b[j];
}

/* Real roots of a polynomial of degree <= 10 by Sturm-sequence isolation + safeguarded Newton.
 * Returns the number of roots written to `roots` (ascending). */
static int real_roots_sturm(const double* pin, int deg, double* roots)
{
    double p[11];
    double mx = 0; for (int i = 0; i <= deg; ++i) if (fabs(pin[i]) > mx) mx = fabs(pin[i]);
    if (!(mx > 0) || !isfinite(mx)) return 0;
    for (int i = 0; i <= deg; ++i) p[i] = pin[i] / mx;
    while (deg > 0 && fabs(p[deg]) < 1e-13) --deg;
    if (deg == 0) return 0;
    /* Sturm chain */
    double S[12][11]; int sd[12]; int ns = 2;
    for (int i = 0; i <= deg; ++i) S[0][i] = p[i];
    sd[0] = deg;
    for (int i = 1; i <= deg; ++i) S[1][i - 1] = i * p[i];
    sd[1] = deg - 1;
    while (sd[ns - 1] > 0 && ns < 12) {
        /* r = -(S[ns-2] mod S[ns-1]) */
        double r[11]; int dr = sd[ns - 2];
        for (int i = 0; i <= dr; ++i) r[i] = S[ns - 2][i];
        int dd = sd[ns - 1]; const double* d = S[ns - 1];
        for (int k = dr; k >= dd; --k) {
            double f = r[k] / d[dd];
            for (int j = 0; j <= dd; ++j) r[k - dd + j] -= f * d[j];
            r[k] = 0;
        }
        int nd = dd - 1; double rm = 0;
        for (int i = 0; i <= nd; ++i) if (fabs(r[i]) > rm) rm = fabs(r[i]);
        if (!(rm > 1e-300)) break;                         /* exact gcd reached (multiple root) */
        while (nd > 0 && fabs(r[nd]) < 1e-14 * rm) --nd;
        for (int i = 0; i <= nd; ++i) S[ns][i] = -r[i] / rm;
        sd[ns] = nd; ++ns;
    }
    /* Fujiwara root bound: 2 max_k |a_{n-k}/a_n|^(1/k) */
    double B = 0;
    for (int k = 1; k <= deg; ++k) { double v = pow(fabs(p[deg - k] / p[deg]), 1.0 / k); if (v > B) B = v; }
    B = 2.0 * B + 1e-9;
#define SIGNCHG(t, out) do { int cnt = 0, last = 0; for (int s_ = 0; s_ < ns; ++s_) { double v_ = horner(S[s_], sd[s_], (t)); int sg_ = (v_ > 0) - (v_ < 0); if (sg_ != 0) { if (last != 0 && sg_ != last) ++cnt; last = sg_; } } (out) = cnt; } while (0)
    double lo_s[64], hi_s[64]; int nlo_s[64], nhi_s[64]; int sp = 0, nr = 0;
    int nl, nh; SIGNCHG(-B, nl); SIGNCHG(B, nh);
    lo_s[0] = -B; hi_s[0] = B; nlo_s[0] = nl; nhi_s[0] = nh; sp = 1;
    while (sp > 0 && nr < 10) {
        --sp;
        double lo = lo_s[sp], hi = hi_s[sp]; int a = nlo_s[sp], b = nhi_s[sp];
        int cnt = a - b;
        if (cnt <= 0) continue;
        if (cnt > 1 && (hi - lo) > 1e-13 * (1.0 + fabs(lo) + fabs(hi))) {
            double mid = 0.5 * (lo + hi); int m; SIGNCHG(mid, m);
            if (sp + 2 > 64) continue;
            /* push right first so the left half is processed first (ascending output) */
            lo_s[sp] = mid; hi_s[sp] = hi; nlo_s[sp] = m; nhi_s[sp] = b; ++sp;
            lo_s[sp] = lo; hi_s[sp] = mid; nlo_s[sp] = a; nhi_s[sp] = m; ++sp;
            continue;
        }
        if (cnt > 1) { roots[nr++] = 0.5 * (lo + hi); continue; }   /* unresolved cluster: one representative */
        /* exactly one root in (lo, hi]: bisection on the sign of p, then Newton */
        double flo = horner(p, deg, lo), fhi = horner(p, deg, hi);
        if (fhi == 0.0) { roots[nr++] = hi; continue; }
        double x = 0.5 * (lo + hi);
        if ((flo > 0) == (fhi > 0)) {
            /* sign pattern unusable (root of even multiplicity in the chain's eyes): bisect on Sturm counts */
            for (int it = 0; it < 80 && (hi - lo) > 1e-15 * (1.0 + fabs(x)); ++it) { x = 0.5 * (lo + hi); int m; SIGNCHG(x, m); if (a - m >= 1) hi = x; else { lo = x; a = m; } }
            roots[nr++] = 0.5 * (lo + hi); continue;
        }
        double dxold = hi - lo;
        for (int it = 0; it < 200; ++it) {
            double f = horner(p, deg, x);
            if (f == 0.0) break;
            if ((f > 0) == (flo > 0)) lo = x; else hi = x;
            /* Newton step from x; bisect instead when it leaves the bracket or converges too slowly (rtsafe) */
            double dp = 0; for (int i = deg; i >= 1; --i) dp = dp * x + i * p[i];
            double xn = (dp != 0.0) ? x - f / dp : x;
            if (!(xn > lo && xn < hi) || fabs(xn - x) > 0.5 * fabs(dxold)) xn = 0.5 * (lo + hi);
            dxold = xn - x;
            if (fabs(dxold) <= 4e-16 * (1.0 + fabs(x))) { x = xn; break; }
            x = xn;
        }
        roots[nr++] = x;
    }
#undef SIGNCHG
    return nr;
}


/* Residual of the ten essential-matrix constraints at E: r[0] = det E, r[1..9] = 2 E E^T E - tr(E E^T) E. */
static void ess_residual(const double* E, double* r)
{
    double EEt[9], T[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
    double tr = EEt[0] + EEt[4] + EEt[8];
    mat3_mul(EEt, E, T);
    r[0] = E[0] * (E[4] * E[8] - E[5] * E[7]) - E[1] * (E[3] * E[8] - E[5] * E[6]) + E[2] * (E[3] * E[7] - E[4] * E[6]);
    for (int k = 0; k < 9; ++k) r[1 + k] = 2.0 * T[k] - tr * E[k];
}

/* Gauss-Newton polish of a root (x,y,z) of the ten cubic constraints with E = x N0 + y N1 + z N2 + N3.
 * Jacobian columns are the analytic directional derivatives along N0, N1, N2. */
static void polish_xyz(const double N[4][9], double* xyz)
{
    for (int it = 0; it < 3; ++it) {
        double E[9], r[10], J[10][3];
        for (int e = 0; e < 9; ++e) E[e] = xyz[0] * N[0][e] + xyz[1] * N[1][e] + xyz[2] * N[2][e] + N[3][e];
        ess_residual(E, r);
        double EEt[9], EtE[9], cof[9];
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) {
            EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
            EtE[3 * i + j] = E[i] * E[j] + E[3 + i] * E[3 + j] + E[6 + i] * E[6 + j];
        }
        double tr = EEt[0] + EEt[4] + EEt[8];
        cof[0] = E[4] * E[8] - E[5] * E[7]; cof[1] = E[5] * E[6] - E[3] * E[8]; cof[2] = E[3] * E[7] - E[4] * E[6];
        cof[3] = E[2] * E[7] - E[1] * E[8]; cof[4] = E[0] * E[8] - E[2] * E[6]; cof[5] = E[1] * E[6] - E[0] * E[7];
        cof[6] = E[1] * E[5] - E[2] * E[4]; cof[7] = E[2] * E[3] - E[0] * E[5]; cof[8] = E[0] * E[4] - E[1] * E[3];
        for (int d = 0; d < 3; ++d) {
            /* D det[D] = <cof, D>;  D g[D] = 2 (D E^T E + E D^T E + E E^T D) - 2 <E, D> E - tr(E E^T) D */
            const double* D = N[d];
            double T1[9], T2[9], T3[9], Dt[9], EDt[9], ip = 0, dd = 0;
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Dt[3 * i + j] = D[3 * j + i];
            mat3_mul(D, EtE, T1); mat3_mul(E, Dt, EDt); mat3_mul(EDt, E, T2); mat3_mul(EEt, D, T3);
            for (int k = 0; k < 9; ++k) { ip += E[k] * D[k]; dd += cof[k] * D[k]; }
            J[0][d] = dd;
            for (int k = 0; k < 9; ++k) J[1 + k][d] = 2.0 * (T1[k] + T2[k] + T3[k]) - 2.0 * ip * E[k] - tr * D[k];
        }
        double A[3][3] = {{0}}, b[3] = {0};
        for (int k = 0; k < 10; ++k) for (int i = 0; i < 3; ++i) { b[i] += J[k][i] * r[k]; for (int j = 0; j < 3; ++j) A[i][j] += J[k][i] * J[k][j]; }
        double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) + A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
        if (!(fabs(det) > 0) || !isfinite(det)) return;
        double dx[3];
        dx[0] = (b[0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (b[1] * A[2][2] - A[1][2] * b[2]) + A[0][2] * (b[1] * A[2][1] - A[1][1] * b[2])) / det;
        dx[1] = (A[0][0] * (b[1] * A[2][2] - A[1][2] * b[2]) - b[0] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) + A[0][2] * (A[1][0] * b[2] - b[1] * A[2][0])) / det;
        dx[2] = (A[0][0] * (A[1][1] * b[2] - b[1] * A[2][1]) - A[0][1] * (A[1][0] * b[2] - b[1] * A[2][0]) + b[0] * (A[1][0] * A[2][1] - A[1][1] * A[2][0])) / det;
        if (!isfinite(dx[0]) || !isfinite(dx[1]) || !isfinite(dx[2])) return;
        for (int i = 0; i < 3; ++i) xyz[i] -= dx[i];
    }
}

/* Nister's five-point solver.  Inputs: five normalised correspondences.  Output: up to 10 essential
 * matrices (row-major, unit Frobenius norm) in ascending order of the root z.  Returns the count. */
static int fivept_impl(const double* x1, const double* y1, const double* x2, const double* y2, double* Eout, double* dbg)
{
    build_tables();
    double N[4][9];
    if (!nullspace_5x9(x1, y1, x2, y2, N)) return 0;
    /* E = x N0 + y N1 + z N2 + N3 : entry (i,j) as a degree-1 polynomial [x,y,z,1] */
    double Ep[9][4];
    for (int e = 0; e < 9; ++e) for (int k = 0; k < 4; ++k) Ep[e][k] = N[k][e];
    /* 10 cubic constraints, 20 monomials each */
    double C[10][20];
    memset(C, 0, sizeof(C));
    /* det(E) = 0 */
    {
        double m[3][10]; memset(m, 0, sizeof(m));
        p11_acc(Ep[4], Ep[8], 1.0, m[0]); p11_acc(Ep[5], Ep[7], -1.0, m[0]);   /* E11E22 - E12E21 */
        p11_acc(Ep[3], Ep[8], 1.0, m[1]); p11_acc(Ep[5], Ep[6], -1.0, m[1]);   /* E10E22 - E12E20 */
        p11_acc(Ep[3], Ep[7], 1.0, m[2]); p11_acc(Ep[4], Ep[6], -1.0, m[2]);   /* E10E21 - E11E20 */
        p21_acc(m[0], Ep[0], 1.0, C[0]); p21_acc(m[1], Ep[1], -1.0, C[0]); p21_acc(m[2], Ep[2], 1.0, C[0]);
    }
    /* 2 E E^T E - tr(E E^T) E = 0  <=>  (E E^T - 1/2 tr(E E^T) I) E = 0 */
    {
        double G[3][3][10]; memset(G, 0, sizeof(G));
        for (int i = 0; i < 3; ++i)
            for (int j = i; j < 3; ++j) {
                for (int k = 0; k < 3; ++k) p11_acc(Ep[3 * i + k], Ep[3 * j + k], 1.0, G[i][j]);
                if (j != i) memcpy(G[j][i], G[i][j], sizeof(double) * 10);
            }
        double tr[10];
        for (int k = 0; k < 10; ++k) tr[k] = 0.5 * (G[0][0][k] + G[1][1][k] + G[2][2][k]);
        for (int i = 0; i < 3; ++i) for (int k = 0; k < 10; ++k) G[i][i][k] -= tr[k];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 3; ++k) p21_acc(G[i][k], Ep[3 * k + j], 1.0, C[1 + 3 * i + j]);
    }
    /* Gauss-Jordan with partial (row) pivoting on the first ten columns */
    for (int k = 0; k < 10; ++k) {
        int pr = k; double best = fabs(C[k][k]);
        for (int r = k + 1; r < 10; ++r) if (fabs(C[r][k]) > best) { best = fabs(C[r][k]); pr = r; }
        if (!(best > 1e-300)) return 0;
        if (pr != k) for (int c = 0; c < 20; ++c) { double t = C[k][c]; C[k][c] = C[pr][c]; C[pr][c] = t; }
        double inv = 1.0 / C[k][k];
        for (int c = k; c < 20; ++c) C[k][c] *= inv;
        for (int r = 0; r < 10; ++r) if (r != k) { double f = C[r][k]; if (f != 0.0) for (int c = k; c < 20; ++c) C[r][c] -= f * C[k][c]; }
    }
    /* B(z) [x y 1]^T = 0 from <k>=row4 - z row5, <l>=row6 - z row7, <m>=row8 - z row9.
     * right-hand block columns: 10 xz^2, 11 xz, 12 x, 13 yz^2, 14 yz, 15 y, 16 z^3, 17 z^2, 18 z, 19 1 */
    double Bx[3][4], By[3][4], Bc[3][5];
    for (int q = 0; q < 3; ++q) {
        const double* a = C[4 + 2 * q]; const double* b = C[5 + 2 * q];
        Bx[q][0] = a[12];         Bx[q][1] = a[11] - b[12]; Bx[q][2] = a[10] - b[11]; Bx[q][3] = -b[10];
        By[q][0] = a[15];         By[q][1] = a[14] - b[15]; By[q][2] = a[13] - b[14]; By[q][3] = -b[13];
        Bc[q][0] = a[19];         Bc[q][1] = a[18] - b[19]; Bc[q][2] = a[17] - b[18]; Bc[q][3] = a[16] - b[17]; Bc[q][4] = -b[16];
    }
    /* det B(z): degree 10 */
    double poly[11], t1[8], t2[8], t3[11];
    for (int i = 0; i < 11; ++i) poly[i] = 0;
    /* kx (ly mc - lc my) */
    pmul(By[1], 3, Bc[2], 4, t1); pmul(Bc[1], 4, By[2], 3, t2); for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    pmul(Bx[0], 3, t1, 7, t3); for (int i = 0; i < 11; ++i) poly[i] += t3[i];
    /* - ky (lx mc - lc mx) */
    pmul(Bx[1], 3, Bc[2], 4, t1); pmul(Bc[1], 4, Bx[2], 3, t2); for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    pmul(By[0], 3, t1, 7, t3); for (int i = 0; i < 11; ++i) poly[i] -= t3[i];
    /* + kc (lx my - ly mx) */
    { double u1[7], u2[7]; pmul(Bx[1], 3, By[2], 3, u1); pmul(By[1], 3, Bx[2], 3, u2); for (int i = 0; i < 7; ++i) u1[i] -= u2[i];
      pmul(Bc[0], 4, u1, 6, t3); for (int i = 0; i < 11; ++i) poly[i] += t3[i]; }
    double roots[10];
    int nr = real_roots_sturm(poly, 10, roots), ne = 0;
    if (dbg) { memcpy(dbg, N, sizeof(double) * 36); memcpy(dbg + 36, poly, sizeof(double) * 11); dbg[47] = nr; memcpy(dbg + 48, roots, sizeof(double) * (size_t)nr); }
    for (int r = 0; r < nr && ne < MAXC; ++r) {
        double z = roots[r];
        double bm[3][3];
        for (int q = 0; q < 3; ++q) { bm[q][0] = horner(Bx[q], 3, z); bm[q][1] = horner(By[q], 3, z); bm[q][2] = horner(Bc[q], 4, z); }
        /* null vector of the 3x3 B(z): cross product of the row pair with the largest |w| */
        double best = -1, bv[3] = {0, 0, 0};
        for (int a = 0; a < 3; ++a) for (int b = a + 1; b < 3; ++b) { double v[3]; cross3(bm[a], bm[b], v); if (fabs(v[2]) > best) { best = fabs(v[2]); bv[0] = v[0]; bv[1] = v[1]; bv[2] = v[2]; } }
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

int oracle_fivept(const double* x1, const double* y1, const double* x2, const double* y2, double* Eout)
{
    return fivept_impl(x1, y1, x2, y2, Eout, 0);
}
/* debug view: dbg[0:36] null-space basis, [36:47] degree-10 polynomial, [47] #roots, [48:58] roots */
int oracle_fivept_debug(const double* x1, const double* y1, const double* x2, const double* y2, double* Eout, double* dbg)
{
    return fivept_impl(x1, y1, x2, y2, Eout, dbg);
}
/* exported for tests: real roots of a polynomial (coefficient i multiplies z^i), degree <= 10 */
int oracle_real_roots(const double* p, int deg, double* roots) { return real_roots_sturm(p, deg, roots); }

/* ---------------------------------------------------------------- local optimisation (LM on Sampson residuals) */

static void skew_times(const double* t, const double* R, double* E)   /* E = [t]x R */
{
    for (int j = 0; j < 3; ++j) {
        E[j]     = -t[2] * R[3 + j] + t[1] * R[6 + j];
        E[3 + j] =  t[2] * R[j]     - t[0] * R[6 + j];
        E[6 + j] = -t[1] * R[j]     + t[0] * R[3 + j];
    }
}

static void rodrigues(const double* w, double* R)
{
    double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
    double a, b;                       /* R = I + a [w]x + b [w]x^2 */
    if (th < 1e-8) { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; }
    else { a = sin(th) / th; b = (1.0 - cos(th)) / th2; }
    double K[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0}, K2[9];
    mat3_mul(K, K, K2);
    for (int i = 0; i < 9; ++i) R[i] = ((i % 4) == 0) + a * K[i] + b * K2[i];
}

/* tangent basis of the unit sphere at t: b1 = normalize(t x e_k) with k the index of the smallest |t_k|
 * (lowest k on ties), b2 = t x b1 */
static void tangent_basis(const double* t, double* b1, double* b2)
{
    int k = 0;
    if (fabs(t[1]) < fabs(t[k])) k = 1;
    if (fabs(t[2]) < fabs(t[k])) k = 2;
    double e[3] = {0, 0, 0}; e[k] = 1.0;
    cross3(t, e, b1);
    double n = norm3(b1);
    for (int i = 0; i < 3; ++i) b1[i] /= n;
    cross3(t, b1, b2);
}

/* One pass over all correspondences at model (R,t): MSAC cost, inlier count, and the Gauss-Newton
 * normal equations JtJ (5x5, row-major) / Jtr (5) of the signed Sampson residuals of the inliers.
 * Parameters: R <- R Exp(w) (3), t <- normalize(t + b1 a + b2 b) (2). */
static double lo_pass(const double* R, const double* t, const double* x1, const double* y1, const double* x2,
                      const double* y2, int n, double thr2, int* ninl, double* JtJ, double* Jtr)
{
    double E[9], b1[3], b2[3], dE[5][9];
    skew_times(t, R, E);
    tangent_basis(t, b1, b2);
    /* dE/dw_k = E [e_k]x ; dE/da = [b1]x R ; dE/db = [b2]x R */
    for (int i = 0; i < 3; ++i) {
        dE[0][3 * i] = 0;             dE[0][3 * i + 1] = E[3 * i + 2];  dE[0][3 * i + 2] = -E[3 * i + 1];
        dE[1][3 * i] = -E[3 * i + 2]; dE[1][3 * i + 1] = 0;             dE[1][3 * i + 2] = E[3 * i];
        dE[2][3 * i] = E[3 * i + 1];  dE[2][3 * i + 1] = -E[3 * i];     dE[2][3 * i + 2] = 0;
    }
    skew_times(b1, R, dE[3]); skew_times(b2, R, dE[4]);
    for (int i = 0; i < 25; ++i) JtJ[i] = 0;
    for (int i = 0; i < 5; ++i) Jtr[i] = 0;
    double cost = 0; int k = 0;
    for (int i = 0; i < n; ++i) {
        double a1 = x1[i], c1 = y1[i], a2 = x2[i], c2 = y2[i];
        double l0 = E[0] * a1 + E[1] * c1 + E[2], l1 = E[3] * a1 + E[4] * c1 + E[5], l2 = E[6] * a1 + E[7] * c1 + E[8];
        double m0 = E[0] * a2 + E[3] * c2 + E[6], m1 = E[1] * a2 + E[4] * c2 + E[7];
        double num = a2 * l0 + c2 * l1 + l2, den = l0 * l0 + l1 * l1 + m0 * m0 + m1 * m1;
        double s = num * num / den;
        if (!(s < thr2)) { cost += thr2; continue; }
        cost += s; ++k;
        double isd = 1.0 / sqrt(den), r = num * isd, q = r / den;   /* dr = dnum*isd - q * (l0 dl0 + l1 dl1 + m0 dm0 + m1 dm1) */
        double J[5];
        for (int p = 0; p < 5; ++p) {
            const double* D = dE[p];
            double dl0 = D[0] * a1 + D[1] * c1 + D[2], dl1 = D[3] * a1 + D[4] * c1 + D[5], dl2 = D[6] * a1 + D[7] * c1 + D[8];
            double dm0 = D[0] * a2 + D[3] * c2 + D[6], dm1 = D[1] * a2 + D[4] * c2 + D[7];
            double dnum = a2 * dl0 + c2 * dl1 + dl2;
            J[p] = dnum * isd - q * (l0 * dl0 + l1 * dl1 + m0 * dm0 + m1 * dm1);
        }
        for (int p = 0; p < 5; ++p) { Jtr[p] += J[p] * r; for (int u = 0; u < 5; ++u) JtJ[5 * p + u] += J[p] * J[u]; }
    }
    *ninl = k;
    return cost;
}

/* Solve the damped 5x5 system (JtJ + lambda diag(JtJ)) d = -Jtr by Cholesky.  Returns 0 on failure. */
static int solve_lm5(const double* JtJ, const double* Jtr, double lambda, double* d)
{
    double L[25];
    for (int i = 0; i < 25; ++i) L[i] = JtJ[i];
    for (int i = 0; i < 5; ++i) L[6 * i] += lambda * JtJ[6 * i] + 1e-300;
    for (int j = 0; j < 5; ++j) {
        double s = L[6 * j];
        for (int k = 0; k < j; ++k) s -= L[5 * j + k] * L[5 * j + k];
        if (!(s > 0)) return 0;
        double ljj = sqrt(s);
        L[6 * j] = ljj;
        for (int i = j + 1; i < 5; ++i) {
            double v = L[5 * i + j];
            for (int k = 0; k < j; ++k) v -= L[5 * i + k] * L[5 * j + k];
            L[5 * i + j] = v / ljj;
        }
    }
    double y[5];
    for (int i = 0; i < 5; ++i) { double v = -Jtr[i]; for (int k = 0; k < i; ++k) v -= L[5 * i + k] * y[k]; y[i] = v / L[6 * i]; }
    for (int i = 4; i >= 0; --i) { double v = y[i]; for (int k = i + 1; k < 5; ++k) v -= L[5 * k + i] * d[k]; d[i] = v / L[6 * i]; }
    return 1;
}

#define LO_MAX_ITERS 20

/* Starting point of the local optimisation.  E_in = +-[t]x R has two equivalent decompositions, (Ra, ta) and its twisted
 * pair (Rb, -ta), and the minimal solver fixes neither the sign of E_in nor which of them an SVD returns first.  The
 * retraction R Exp(w) differs between the two (to second order in the step), so the start must not depend on those
 * accidents: take the rotation with the larger trace (the smaller rotation angle; the other one is turned by 180 degrees
 * about the baseline) and the sign of t that reproduces E_in.  Starting from -E_in then mirrors the whole trajectory
 * (same R, opposite t), so the result is independent of the sign of the minimal model. */
static void lo_canonical_start(const double* E_in, double* R, double* t)
{
    double R1[9], R2[9], Et[9], dot = 0;
    oracle_decompose_essential(E_in, R1, R2, t);
    const double tr1 = R1[0] + R1[4] + R1[8], tr2 = R2[0] + R2[4] + R2[8];
    memcpy(R, tr1 >= tr2 ? R1 : R2, sizeof(double) * 9);
    skew_times(t, R, Et);
    for (int i = 0; i < 9; ++i) dot += Et[i] * E_in[i];
    if (dot < 0) { t[0] = -t[0]; t[1] = -t[1]; t[2] = -t[2]; }
}

/* diagnostics: Levenberg-Marquardt trial steps taken by the last oracle_lo_refine call */
int oracle_lo_trials_last = 0;

/* Local optimisation: Levenberg-Marquardt on (R,t) starting from the decomposition (R1, t) of E_in,
 * minimising the MSAC cost; each accepted step re-selects the inliers.  E_out = [t]x R (norm sqrt 2). */
double oracle_lo_refine(const double* E_in, const double* x1, const double* y1, const double* x2, const double* y2,
                        int n, double thr2, int max_iters, double* E_out, int* ninl_out)
{
    double R[9], R2[9], t[3], JtJ[25], Jtr[5];
    int ninl;
    lo_canonical_start(E_in, R, t);
    (void)R2;
    double cost = lo_pass(R, t, x1, y1, x2, y2, n, thr2, &ninl, JtJ, Jtr);
    double lambda = 1e-4;
    oracle_lo_trials_last = 0;
    for (int it = 0; it < max_iters; ++it) {
        if (ninl < 5) break;
        ++oracle_lo_trials_last;
        double d[5];
        if (!solve_lm5(JtJ, Jtr, lambda, d)) { lambda *= 10.0; if (lambda > 1e4) break; continue; }
        double dR[9], Rn[9], tn[3], b1[3], b2[3], JtJn[25], Jtrn[5];
        int ninl_n;
        rodrigues(d, dR);
        mat3_mul(R, dR, Rn);
        tangent_basis(t, b1, b2);
        for (int i = 0; i < 3; ++i) tn[i] = t[i] + b1[i] * d[3] + b2[i] * d[4];
        double nn = norm3(tn);
        for (int i = 0; i < 3; ++i) tn[i] /= nn;
        double cn = lo_pass(Rn, tn, x1, y1, x2, y2, n, thr2, &ninl_n, JtJn, Jtrn);
        if (cn < cost) {
            double gain = cost - cn;
            memcpy(R, Rn, sizeof(R)); memcpy(t, tn, sizeof(t)); memcpy(JtJ, JtJn, sizeof(JtJ)); memcpy(Jtr, Jtrn, sizeof(Jtr));
            int stop = !(gain > 1e-10 * cost);
            cost = cn; ninl = ninl_n;
            lambda *= 0.1; if (lambda < 1e-9) lambda = 1e-9;
            if (stop) break;
        } else {
            lambda *= 10.0;
            if (lambda > 1e4) break;
        }
    }
    skew_times(t, R, E_out);
    if (ninl_out) *ninl_out = ninl;
    return cost;
}

/* ---------------------------------------------------------------- the verifier, one pair */

/* Status codes (shared with include/gtsfm_b200.h) */
#define VSTAT_OK 0
#define VSTAT_TOO_FEW 1      /* M < 5 or M < 6 -> failure tuple (opencv_verifier_base.py:69-78) */
#define VSTAT_NO_MODEL 2     /* no essential matrix from any sample / no inliers */

/* Model validity test of USAC-style estimators: the oriented epipolar constraint (Chum, Werner, Matas, "Epipolar
 * geometry estimation via RANSAC benefits from the oriented epipolar constraint", ICPR 2004), which cv2's USAC framework
 * (the estimator behind ransac.py:74) applies to every model of a minimal sample before scoring it.  With e2 the
 * epipole in image 2 (E^T e2 = 0, taken as the largest cross product of two columns of E), e2 x x2 and E x1 must have
 * the same orientation for all five sample correspondences, or the opposite one for all five. */
int oracle_oriented_ok(const double* E, const double* x1, const double* y1, const double* x2, const double* y2)
{
    const double c0[3] = {E[0], E[3], E[6]}, c1[3] = {E[1], E[4], E[7]}, c2[3] = {E[2], E[5], E[8]};
    double v01[3], v02[3], v12[3], e[3];
    cross3(c0, c1, v01); cross3(c0, c2, v02); cross3(c1, c2, v12);
    double n01 = v01[0] * v01[0] + v01[1] * v01[1] + v01[2] * v01[2];
    double n02 = v02[0] * v02[0] + v02[1] * v02[1] + v02[2] * v02[2];
    double n12 = v12[0] * v12[0] + v12[1] * v12[1] + v12[2] * v12[2];
    double nb = n01;
    memcpy(e, v01, sizeof(e));
    if (n02 > nb) { nb = n02; memcpy(e, v02, sizeof(e)); }
    if (n12 > nb) { nb = n12; memcpy(e, v12, sizeof(e)); }
    int pos = 0, neg = 0;
    for (int k = 0; k < 5; ++k) {
        double l0 = E[0] * x1[k] + E[1] * y1[k] + E[2], l1 = E[3] * x1[k] + E[4] * y1[k] + E[5], l2 = E[6] * x1[k] + E[7] * y1[k] + E[8];
        double w0 = e[1] - e[2] * y2[k], w1 = e[2] * x2[k] - e[0], w2 = e[0] * y2[k] - e[1] * x2[k];
        double sgn = w0 * l0 + w1 * l1 + w2 * l2;
        pos += sgn > 0.0;
        neg += sgn < 0.0;
    }
    return pos == 5 || neg == 5;
}

/* OpencvVerifierBase.verify + Ransac.estimate_E, essential-matrix mode, with explicit samples.
 *   kp1 (n1,2), kp2 (n2,2): pixel coordinates, float64 row-major
 *   mi1, mi2 (M): match indices into kp1 / kp2
 *   cal1, cal2: {f, k1, k2, u0, v0}
 *   samples (H,5) int32: rows of indices into [0, M)
 * Outputs: E_min (best minimal model), E (after local optimisation), mask (M) 0/1, R (9), t (3),
 *          info[0] = #inliers, info[1] = best candidate id (h*10 + c), info[2] = #candidates scored;
 *          cost[0] = best minimal MSAC cost, cost[1] = final MSAC cost.
 * Returns a VSTAT_* code. */
int oracle_verify_pair(const double* kp1, const double* kp2, const int32_t* mi1, const int32_t* mi2, int M,
                       const double* cal1, const double* cal2, double thr_px,
                       const int32_t* samples, int H, int lo_iters,
                       double* E_min, double* E, uint8_t* mask, double* R, double* t, int32_t* info, double* cost)
{
    info[0] = info[1] = info[2] = 0;
    if (M < 5) return VSTAT_TOO_FEW;
    if (M < 6) return VSTAT_TOO_FEW;
    double* x1 = (double*)malloc(sizeof(double) * 4 * (size_t)M);
    double *y1 = x1 + M, *x2 = y1 + M, *y2 = x2 + M;
    for (int i = 0; i < M; ++i) {
        oracle_calibrate(kp1[2 * mi1[i]], kp1[2 * mi1[i] + 1], cal1[0], cal1[1], cal1[2], cal1[3], cal1[4], &x1[i], &y1[i]);
        oracle_calibrate(kp2[2 * mi2[i]], kp2[2 * mi2[i] + 1], cal2[0], cal2[1], cal2[2], cal2[3], cal2[4], &x2[i], &y2[i]);
    }
    double fx = cal1[0] > cal2[0] ? cal1[0] : cal2[0];          /* opencv_verifier_base.py:86 */
    double thr = thr_px / fx, thr2 = thr * thr;                 /* ransac.py:79 */
    double best_cost = INFINITY; int best_id = -1, ncand = 0;
    for (int h = 0; h < H; ++h) {
        double sx1[5], sy1[5], sx2[5], sy2[5], Es[9 * MAXC];
        for (int k = 0; k < 5; ++k) { int s = samples[5 * h + k]; sx1[k] = x1[s]; sy1[k] = y1[s]; sx2[k] = x2[s]; sy2[k] = y2[s]; }
        int ne = oracle_fivept(sx1, sy1, sx2, sy2, Es);
        for (int c = 0; c < ne; ++c) {
            int ni;
            if (!oracle_oriented_ok(Es + 9 * c, sx1, sy1, sx2, sy2)) continue;      /* invalid models are never scored */
            double cst = msac_cost(Es + 9 * c, x1, y1, x2, y2, M, thr2, &ni);
            ++ncand;
            if (cst < best_cost) { best_cost = cst; best_id = h * MAXC + c; memcpy(E_min, Es + 9 * c, sizeof(double) * 9); }
        }
    }
    info[1] = best_id; info[2] = ncand;
    if (best_id < 0) { free(x1); return VSTAT_NO_MODEL; }
    cost[0] = best_cost;
    int ninl;
    if (lo_iters > 0) cost[1] = oracle_lo_refine(E_min, x1, y1, x2, y2, M, thr2, lo_iters, E, &ninl);
    else { memcpy(E, E_min, sizeof(double) * 9); cost[1] = best_cost; }
    ninl = 0;
    for (int i = 0; i < M; ++i) { mask[i] = oracle_sampson_sq(E, x1[i], y1[i], x2[i], y2[i]) < thr2; ninl += mask[i]; }
    info[0] = ninl;
    if (ninl == 0) { free(x1); return VSTAT_NO_MODEL; }
    /* recover pose from the inliers (opencv_verifier_base.py:102-108 -> verification.py:52-94) */
    double* ix1 = (double*)malloc(sizeof(double) * 4 * (size_t)ninl);
    double *iy1 = ix1 + ninl, *ix2 = iy1 + ninl, *iy2 = ix2 + ninl;
    int k = 0;
    for (int i = 0; i < M; ++i) if (mask[i]) { ix1[k] = x1[i]; iy1[k] = y1[i]; ix2[k] = x2[i]; iy2[k] = y2[i]; ++k; }
    int good[4];
    oracle_recover_pose(E, ix1, iy1, ix2, iy2, ninl, R, t, good);
    free(ix1); free(x1);
    return VSTAT_OK;
}

/* ================================================================ fundamental-matrix mode
 * Ransac(use_intrinsics_in_verification=False): Ransac.estimate_F -> cv2.findFundamentalMat(FM_RANSAC, thr px)
 * (gtsfm/frontend/verifier/ransac.py:84-111), then E = K2^T F K1 (utils/verification.py:97-110,
 * opencv_verifier_base.py:90-96) and the same pose recovery as the essential-matrix mode.
 *
 * OpenCV (un-vendored dependency of the reference, opencv-python 4.13) runs its classic RANSAC here: 7-point minimal
 * solver on the float32 pixel coordinates, error = max of the two squared point-to-epipolar-line distances, inlier
 * <=> error <= thr^2, model with the most inliers wins (the first one found among equal counts), no refit.  Restated
 * below with explicit (H,7) sample sets.  PARITY UNPINNED for the loop itself (cv2 neither takes nor reports its samples
 * and stops adaptively); pinned: the 7-point solver against cv2.findFundamentalMat(FM_7POINT) and the error / mask rule
 * against the mask cv2 returns for its own F (tests/golden/fmat_*.npz). */

/* cv::FMEstimatorCallback::computeError: double arithmetic on float32 points, result rounded to float32. */
float oracle_fm_error(const double* F, float fx1, float fy1, float fx2, float fy2)
{
    const double x1 = fx1, y1 = fy1, x2 = fx2, y2 = fy2;
    double a = F[0] * x1 + F[1] * y1 + F[2];
    double b = F[3] * x1 + F[4] * y1 + F[5];
    double c = F[6] * x1 + F[7] * y1 + F[8];
    const double s2 = 1. / (a * a + b * b);
    const double d2 = x2 * a + y2 * b + c;
    a = F[0] * x2 + F[3] * y2 + F[6];
    b = F[1] * x2 + F[4] * y2 + F[7];
    c = F[2] * x2 + F[5] * y2 + F[8];
    const double s1 = 1. / (a * a + b * b);
    const double d1 = x1 * a + y1 * b + c;
    const double e1 = d1 * d1 * s1, e2 = d2 * d2 * s2;
    return (float)(e1 > e2 ? e1 : e2);
}

void oracle_fm_error_batch(const double* F, const float* x1, const float* y1, const float* x2, const float* y2, int n, float* out)
{
    for (int i = 0; i < n; ++i) out[i] = oracle_fm_error(F, x1[i], y1[i], x2[i], y2[i]);
}

/* Null space (2 vectors) of the 7x9 epipolar system, Gauss-Jordan with complete pivoting (as nullspace_5x9). */
static int nullspace_7x9(const double* x1, const double* y1, const double* x2, const double* y2, double N[2][9])
{
    double A[7][9]; int perm[9];
    for (int r = 0; r < 7; ++r) {
        A[r][0] = x2[r] * x1[r]; A[r][1] = x2[r] * y1[r]; A[r][2] = x2[r];
        A[r][3] = y2[r] * x1[r]; A[r][4] = y2[r] * y1[r]; A[r][5] = y2[r];
        A[r][6] = x1[r];         A[r][7] = y1[r];         A[r][8] = 1.0;
    }
    for (int c = 0; c < 9; ++c) perm[c] = c;
    for (int k = 0; k < 7; ++k) {
        int pr = k, pc = k; double best = -1.0;
        for (int r = k; r < 7; ++r) for (int c = k; c < 9; ++c) { double v = fabs(A[r][c]); if (v > best) { best = v; pr = r; pc = c; } }
        if (!(best > 1e-14)) return 0;
        if (pr != k) for (int c = 0; c < 9; ++c) { double t = A[k][c]; A[k][c] = A[pr][c]; A[pr][c] = t; }
        if (pc != k) { for (int r = 0; r < 7; ++r) { double t = A[r][k]; A[r][k] = A[r][pc]; A[r][pc] = t; } int t = perm[k]; perm[k] = perm[pc]; perm[pc] = t; }
        double inv = 1.0 / A[k][k];
        for (int c = k; c < 9; ++c) A[k][c] *= inv;
        for (int r = 0; r < 7; ++r) if (r != k) { double f = A[r][k]; if (f != 0.0) for (int c = k; c < 9; ++c) A[r][c] -= f * A[k][c]; }
    }
    for (int j = 0; j < 2; ++j) {
        double v[9];
        for (int r = 0; r < 7; ++r) v[r] = -A[r][7 + j];
        for (int c = 7; c < 9; ++c) v[c] = (c == 7 + j) ? 1.0 : 0.0;
        double nn = 0; for (int c = 0; c < 9; ++c) nn += v[c] * v[c];
        nn = 1.0 / sqrt(nn);
        for (int c = 0; c < 9; ++c) N[j][perm[c]] = v[c] * nn;
    }
    return 1;
}

static double det3c(const double* a, const double* b, const double* c)   /* det of the matrix with COLUMNS a, b, c */
{
    return a[0] * (b[1] * c[2] - b[2] * c[1]) - b[0] * (a[1] * c[2] - a[2] * c[1]) + c[0] * (a[1] * b[2] - a[2] * b[1]);
}

/* cv::FMEstimatorCallback::run7Point restated: F(l) = l N0 + (1 - l) N1 over the two null vectors, det F(l) = 0 is a
 * cubic in l; one fundamental matrix per real root (ascending l), unit Frobenius norm.  Returns their number (<= 3). */
int oracle_sevenpt(const double* x1, const double* y1, const double* x2, const double* y2, double* Fout)
{
    double N[2][9], A[9], a[3][3], b[3][3];
    if (!nullspace_7x9(x1, y1, x2, y2, N)) return 0;
    for (int i = 0; i < 9; ++i) A[i] = N[0][i] - N[1][i];        /* F(l) = l A + B, B = N1 */
    for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) { a[j][i] = A[3 * i + j]; b[j][i] = N[1][3 * i + j]; }   /* columns */
    double p[4], roots[10];
    p[3] = det3c(a[0], a[1], a[2]);
    p[2] = det3c(b[0], a[1], a[2]) + det3c(a[0], b[1], a[2]) + det3c(a[0], a[1], b[2]);
    p[1] = det3c(a[0], b[1], b[2]) + det3c(b[0], a[1], b[2]) + det3c(b[0], b[1], a[2]);
    p[0] = det3c(b[0], b[1], b[2]);
    int nr = real_roots_sturm(p, 3, roots);
    int n = 0;
    for (int k = 0; k < nr && n < 3; ++k) {
        double F[9], nn = 0;
        for (int i = 0; i < 9; ++i) { F[i] = roots[k] * A[i] + N[1][i]; nn += F[i] * F[i]; }
        if (!(nn > 0) || !isfinite(nn)) continue;
        nn = 1.0 / sqrt(nn);
        for (int i = 0; i < 9; ++i) Fout[9 * n + i] = F[i] * nn;
        ++n;
    }
    return n;
}

/* OpencvVerifierBase.verify in fundamental-matrix mode with explicit samples.
 *   samples (H,7) int32.  Outputs: F (best model), E = K2^T F K1, mask (M), R, t,
 *   info[0] = #inliers, info[1] = best candidate id (h*3 + root), info[2] = #candidates scored. */
int oracle_verify_pair_F(const double* kp1, const double* kp2, const int32_t* mi1, const int32_t* mi2, int M,
                         const double* cal1, const double* cal2, double thr_px, const int32_t* samples, int H,
                         double* F, double* E, uint8_t* mask, double* R, double* t, int32_t* info)
{
    info[0] = info[1] = info[2] = 0;
    if (M < 8) return VSTAT_TOO_FEW;                             /* NUM_MATCHES_REQ_F_MATRIX, opencv_verifier_base.py:38-40,69 */
    float* px = (float*)malloc(sizeof(float) * 4 * (size_t)M);   /* cv2 converts the points to float32 */
    float *fx1 = px, *fy1 = px + M, *fx2 = fy1 + M, *fy2 = fx2 + M;
    for (int i = 0; i < M; ++i) {
        fx1[i] = (float)kp1[2 * mi1[i]]; fy1[i] = (float)kp1[2 * mi1[i] + 1];
        fx2[i] = (float)kp2[2 * mi2[i]]; fy2[i] = (float)kp2[2 * mi2[i] + 1];
    }
    const double thr2 = thr_px * thr_px;
    int best_cnt = 6, best_id = -1, ncand = 0;                   /* cv2: goodCount > max(maxGoodCount, modelPoints - 1) */
    for (int h = 0; h < H; ++h) {
        double sx1[7], sy1[7], sx2[7], sy2[7], Fs[27];
        for (int k = 0; k < 7; ++k) { int s = samples[7 * h + k]; sx1[k] = fx1[s]; sy1[k] = fy1[s]; sx2[k] = fx2[s]; sy2[k] = fy2[s]; }
        int nf = oracle_sevenpt(sx1, sy1, sx2, sy2, Fs);
        for (int c = 0; c < nf; ++c) {
            int cnt = 0;
            for (int i = 0; i < M; ++i) cnt += oracle_fm_error(Fs + 9 * c, fx1[i], fy1[i], fx2[i], fy2[i]) <= thr2;
            ++ncand;
            if (cnt > best_cnt) { best_cnt = cnt; best_id = h * 3 + c; memcpy(F, Fs + 9 * c, sizeof(double) * 9); }
        }
    }
    info[1] = best_id; info[2] = ncand;
    if (best_id < 0) { free(px); return VSTAT_NO_MODEL; }
    int ninl = 0;
    for (int i = 0; i < M; ++i) { mask[i] = oracle_fm_error(F, fx1[i], fy1[i], fx2[i], fy2[i]) <= thr2; ninl += mask[i]; }
    info[0] = ninl;
    /* E = K2^T F K1 (utils/verification.py:110), K = [f 0 u0; 0 f v0; 0 0 1] */
    {
        const double K1[9] = {cal1[0], 0, cal1[3], 0, cal1[0], cal1[4], 0, 0, 1};
        const double K2t[9] = {cal2[0], 0, 0, 0, cal2[0], 0, cal2[3], cal2[4], 1};
        double T[9];
        mat3_mul(K2t, F, T);
        mat3_mul(T, K1, E);
    }
    double* ix1 = (double*)malloc(sizeof(double) * 4 * (size_t)ninl);
    double *iy1 = ix1 + ninl, *ix2 = iy1 + ninl, *iy2 = ix2 + ninl;
    int k = 0;
    for (int i = 0; i < M; ++i) if (mask[i]) {
        oracle_calibrate(kp1[2 * mi1[i]], kp1[2 * mi1[i] + 1], cal1[0], cal1[1], cal1[2], cal1[3], cal1[4], &ix1[k], &iy1[k]);
        oracle_calibrate(kp2[2 * mi2[i]], kp2[2 * mi2[i] + 1], cal2[0], cal2[1], cal2[2], cal2[3], cal2[4], &ix2[k], &iy2[k]);
        ++k;
    }
    int good[4];
    oracle_recover_pose(E, ix1, iy1, ix2, iy2, ninl, R, t, good);
    free(ix1); free(px);
    return VSTAT_OK;
}
