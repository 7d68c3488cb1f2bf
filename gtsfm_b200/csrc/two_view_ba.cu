// Two-view triangulation + bundle adjustment of every pair, batched on the device: the step
// TwoViewEstimator.run_2view performs between verify() and the InlierSupportProcessor when bundle_adjust_2view is True
// (gtsfm/two_view_estimator.py:101-208,316-338 -- the default of configs/sift_front_end.yaml:40).
//
// Reference, per pair (skipped below processor._min_num_inliers_est_model verified correspondences):
//   cameras 0 = (identity, K1), 1 = (i2Ti1^-1, K2) with i2Ti1 = (R, unit t) from the verifier     two_view_estimator.py:160-170
//   Point3dInitializer(NO_RANSAC).triangulate per correspondence: gtsam.triangulatePoint3(DLT, optimize=True), cheirality,
//     every reprojection error < triangulation_options.reproj_error_threshold                      data_association/point3d_initializer.py:225-295
//   TwoViewBundleAdjustment.run_ba: factor graph of GeneralSFMFactor2 reprojection factors (sigma 1 px, Huber 1.345),
//     priors on X(0), K(i) (sigma 1e-5: fixed in effect) and P(0), gtsam LevenbergMarquardtOptimizer (lambda 1e-5, factor 10,
//     relative / absolute error tolerance 1e-5, max_iterations), then filter_landmarks: a track stays when all its
//     reprojection errors are < ba_reproj_error_thresholds[-1]                                    bundle/bundle_adjustment.py:106-357
//   i2Ti1 = wTi2.between(wTi1) -> (Rot3, Unit3(translation)), verified[triangulated][valid]         two_view_estimator.py:196-208
// The arithmetic lives in the un-vendored gtsam==4.2 (absent here): PARITY UNPINNED.  What is restated: the same robust
// reprojection cost and the same Levenberg-Marquardt schedule, on the gauge-fixed minimal parametrisation -- camera 0 at the
// origin, R <- R Exp(w), t on the unit sphere (5 pose parameters), intrinsics constant.  The reference's soft gauge priors
// (X(0) and P(0), sigma 0.1) only select a member of the similarity class the reprojection cost is invariant under, and its
// outputs (relative rotation, translation DIRECTION, track validity) are invariant too, so both formulations share their optima.
//
// One CTA per pair.  A Levenberg-Marquardt trial is two passes over the tracks: (1) linearise -- per track the two residuals,
// their Jacobians, the 3 x 3 point block V, the 5 x 3 coupling W -- and reduce the Schur complement of the points onto the 5
// pose parameters (block sum), (2) after thread 0 solved the 5 x 5 system, back-substitute every point, and evaluate the
// robust cost at the trial state.  The triangulated tracks are compacted into a dense per-pair list first (row order), so the
// passes run with full warps; their points live in two global scratch buffers that swap roles when a step is accepted.
#include <cfloat>
#include <cmath>

#include "ctx.h"
#include "verify_block.cuh"
#include "verify_types.h"

namespace g2v {

struct BaBufs {
    const VPair* pairs; const void* match; const unsigned char* mask; const double* R; const double* t; const int* ninl; const int* status;
    double* X; double* Xn;          // [total][3] point buffers (current / trial, swapped on acceptance), DENSE per pair: entry m0 + j
    int* idx;                       // [total] match row (pair-relative) of dense track j at m0 + j: the triangulated tracks in row order
    unsigned char* ba_mask;         // [total] 1 = track valid after BA (verified, triangulated, reprojection errors below the threshold)
    double* out_R; double* out_t;   // [P][9], [P][3]
    int* out_n; int* out_status;    // [P] valid tracks; 0 = BA ran, 1 = skipped (too few verified / failed pair), 2 = nothing triangulated,
                                    // 3 = BA ran and no track passed the filter (the verifier's pose)
    int* out_iters;                 // [P] accepted LM iterations
    int min_inliers; int max_iters; double tri_thr; double ba_thr; double huber_k; double min_angle_deg;
};

// Cal3Bundler projection of a camera-frame point and its 2 x 3 Jacobian; false when the point is not in front
__device__ __forceinline__ bool ba_project(const double* cal, double x, double y, double z, double* u, double* v, double* J)
{
    if (!(z > 1e-12)) return false;
    const double iz = 1.0 / z, a = x * iz, b = y * iz;
    const double f = cal[0], k1 = cal[1], k2 = cal[2];
    const double r2 = a * a + b * b, g = 1.0 + (k1 + k2 * r2) * r2, dg = 2.0 * (k1 + 2.0 * k2 * r2);   // dg/da = dg * a
    *u = f * g * a + cal[3];
    *v = f * g * b + cal[4];
    if (J) {
        // d(u,v)/d(a,b) = f (g I + (a,b)^T dg (a,b)); d(a,b)/d(x,y,z) = [iz 0 -a iz; 0 iz -b iz]
        const double m00 = f * (g + dg * a * a), m01 = f * dg * a * b, m11 = f * (g + dg * b * b);
        J[0] = m00 * iz; J[1] = m01 * iz; J[2] = -(m00 * a + m01 * b) * iz;
        J[3] = m01 * iz; J[4] = m11 * iz; J[5] = -(m01 * a + m11 * b) * iz;
    }
    return true;
}

__device__ __forceinline__ double huber_weight(double e2, double k) { const double e = sqrt(e2); return e <= k ? 1.0 : k / e; }
__device__ __forceinline__ double huber_loss(double e2, double k) { const double e = sqrt(e2); return e <= k ? 0.5 * e2 : k * (e - 0.5 * k); }

// inverse of a symmetric positive definite 3 x 3 (a00 a01 a02 a11 a12 a22) -> same packing; false when singular
__device__ __forceinline__ bool inv_sym3(const double* a, double* o)
{
    const double c00 = a[3] * a[5] - a[4] * a[4], c01 = a[2] * a[4] - a[1] * a[5], c02 = a[1] * a[4] - a[2] * a[3];
    const double det = a[0] * c00 + a[1] * c01 + a[2] * c02;
    if (!(fabs(det) > 1e-300)) return false;
    const double id = 1.0 / det;
    o[0] = c00 * id; o[1] = c01 * id; o[2] = c02 * id;
    o[3] = (a[0] * a[5] - a[2] * a[2]) * id; o[4] = (a[1] * a[2] - a[0] * a[4]) * id; o[5] = (a[0] * a[3] - a[1] * a[1]) * id;
    return true;
}

struct BaPose { double R[9], t[3], b1[3], b2[3]; };

__device__ inline void ba_tangent(const double* t, double* b1, double* b2)
{
    int k = 0;
    if (fabs(t[1]) < fabs(t[k])) k = 1;
    if (fabs(t[2]) < fabs(t[k])) k = 2;
    double e[3] = {0, 0, 0};
    e[k] = 1.0;
    vm::cross3(t, e, b1);
    const double n = vm::norm3(b1);
    for (int i = 0; i < 3; ++i) b1[i] /= n;
    vm::cross3(t, b1, b2);
}

__device__ inline void ba_rodrigues(const double* w, double* R)
{
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
    double a, b;
    if (th < 1e-8) { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; } else { a = sin(th) / th; b = (1.0 - cos(th)) / th2; }
    const double K[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
    double K2[9];
    vm::mat3_mul(K, K, K2);
    for (int i = 0; i < 9; ++i) R[i] = ((i % 4) == 0 ? 1.0 : 0.0) + a * K[i] + b * K2[i];
}

// Linearisation of one track at (pose, X): residuals e0 / e1 (pixels), Huber weights, A0 = de0/dX, A1 = de1/dX (2 x 3),
// B = de1/d(w, a, b) (2 x 5).  A camera the point is not in front of contributes nothing (gtsam's GeneralSFMFactor2 returns a
// zero error with zero Jacobians on a cheirality exception).
struct BaLin { double e0[2], e1[2], w0, w1, A0[6], A1[6], B[10]; };

__device__ __forceinline__ void ba_linearise(const BaPose& ps, const double* cal1, const double* cal2, const double* X, double2 m1, double2 m2,
                                             double hk, bool jac, BaLin& L)
{
    double u, v, J[6];
    L.w0 = L.w1 = 0.0;
    L.e0[0] = L.e0[1] = L.e1[0] = L.e1[1] = 0.0;
    if (ba_project(cal1, X[0], X[1], X[2], &u, &v, jac ? J : nullptr)) {
        L.e0[0] = u - m1.x; L.e0[1] = v - m1.y;
        L.w0 = huber_weight(L.e0[0] * L.e0[0] + L.e0[1] * L.e0[1], hk);
        if (jac) for (int i = 0; i < 6; ++i) L.A0[i] = J[i];
    } else if (jac) {
        for (int i = 0; i < 6; ++i) L.A0[i] = 0.0;
    }
    const double* R = ps.R;
    const double rx = R[0] * X[0] + R[1] * X[1] + R[2] * X[2], ry = R[3] * X[0] + R[4] * X[1] + R[5] * X[2], rz = R[6] * X[0] + R[7] * X[1] + R[8] * X[2];
    if (ba_project(cal2, rx + ps.t[0], ry + ps.t[1], rz + ps.t[2], &u, &v, jac ? J : nullptr)) {
        L.e1[0] = u - m2.x; L.e1[1] = v - m2.y;
        L.w1 = huber_weight(L.e1[0] * L.e1[0] + L.e1[1] * L.e1[1], hk);
        if (jac) {
            // dXc/dX = R; dXc/dw = -R [X]x (R <- R Exp(w)); dXc/da = b1, dXc/db = b2
            for (int r = 0; r < 2; ++r) {
                const double j0 = J[3 * r], j1 = J[3 * r + 1], j2 = J[3 * r + 2];
                const double q0 = j0 * R[0] + j1 * R[3] + j2 * R[6], q1 = j0 * R[1] + j1 * R[4] + j2 * R[7], q2 = j0 * R[2] + j1 * R[5] + j2 * R[8];   // J R
                L.A1[3 * r] = q0; L.A1[3 * r + 1] = q1; L.A1[3 * r + 2] = q2;
                L.B[5 * r + 3] = j0 * ps.b1[0] + j1 * ps.b1[1] + j2 * ps.b1[2];
                L.B[5 * r + 4] = j0 * ps.b2[0] + j1 * ps.b2[1] + j2 * ps.b2[2];
                // q^T (-[X]x) = -(q0 q1 q2) [0 -X2 X1; X2 0 -X0; -X1 X0 0] = -( q1 X2 - q2 X1, -q0 X2 + q2 X0, q0 X1 - q1 X0 )
                L.B[5 * r] = -(q1 * X[2] - q2 * X[1]);
                L.B[5 * r + 1] = -(-q0 * X[2] + q2 * X[0]);
                L.B[5 * r + 2] = -(q0 * X[1] - q1 * X[0]);
            }
        }
    } else if (jac) {
        for (int i = 0; i < 6; ++i) L.A1[i] = 0.0;
        for (int i = 0; i < 10; ++i) L.B[i] = 0.0;
    }
}

constexpr int kBaThreads = 128;
static_assert(kBaThreads == kFinThreads, "block_sum is written for kFinThreads threads");

__global__ void __launch_bounds__(kBaThreads, 3) two_view_ba_kernel(StoreView st, BaBufs b, int n_pairs)
{
    __shared__ double s_red[kBaThreads / 32][24];
    __shared__ double s_out[24];
    __shared__ BaPose s_pose, s_trial;
    __shared__ double s_dc[5];
    __shared__ int s_ctrl;
    __shared__ int s_wcnt[kBaThreads / 32];
    const int p = blockIdx.x;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    const int tid = threadIdx.x;
    const bool ok = b.status[p] == 0;
    const int n_ver = ok ? b.ninl[p] : 0;
    // ---- outputs default to the verifier's (two_view_estimator.py:334-338: without BA post_ba = pre_ba)
    for (int i = tid; i < vp.M; i += kBaThreads) b.ba_mask[vp.m0 + i] = ok ? b.mask[vp.m0 + i] : 0;
    if (tid < 9) b.out_R[9 * p + tid] = ok ? b.R[9 * p + tid] : 0.0;
    if (tid < 3) b.out_t[3 * p + tid] = ok ? b.t[3 * p + tid] : 0.0;
    if (tid == 0) { b.out_n[p] = n_ver; b.out_status[p] = 1; b.out_iters[p] = 0; }
    if (!ok || n_ver < b.min_inliers) return;          // two_view_estimator.py:316: BA only with enough verified correspondences
    const uint2* match = static_cast<const uint2*>(b.match) + vp.m0;
    const double2* kp1 = st.kps + vp.row1;
    const double2* kp2 = st.kps + vp.row2;
    if (tid == 0) {
        for (int i = 0; i < 9; ++i) s_pose.R[i] = b.R[9 * p + i];
        for (int i = 0; i < 3; ++i) s_pose.t[i] = b.t[3 * p + i];
        ba_tangent(s_pose.t, s_pose.b1, s_pose.b2);
    }
    __syncthreads();

    // ---- triangulation of every verified correspondence: DLT on the calibrated rays, three Gauss-Newton steps on the pixel
    // reprojection error (gtsam.triangulatePoint3(optimize=True)), cheirality in both cameras, reprojection errors below
    // triangulation_options.reproj_error_threshold (point3d_initializer.py:262-281)
    // The tracks that pass are appended to the pair's DENSE list in row order (ballot + prefix over the CTA, deterministic), so
    // the LM passes below run with full warps whatever the inlier ratio is.
    double acc[20];
    int n_tri = 0;                                       // same value in every thread
    int* didx = b.idx + vp.m0;
    double* Xcur = b.X + 3 * vp.m0;
    double* Xtry = b.Xn + 3 * vp.m0;
    const int warp = tid >> 5, lane = tid & 31;
    for (int base = 0; base < vp.M; base += kBaThreads) {
        const int i = base + tid;
        bool good = false;
        double X[3] = {0, 0, 1};
        if (i < vp.M && b.mask[vp.m0 + i]) {
            const uint2 m = match[i];
            const double2 m1 = kp1[m.x], m2 = kp2[m.y];
            double x1, y1, x2, y2;
            vm::calibrate(m1.x, m1.y, vp.cal1, &x1, &y1);
            vm::calibrate(m2.x, m2.y, vp.cal2, &x2, &y2);
            const double* R = s_pose.R;
            const double* t = s_pose.t;
            // rows of the DLT matrix: x P_3 - P_1, y P_3 - P_2 for P = [I | 0] and [R | t]
            double A[16] = {-1, 0, x1, 0, 0, -1, y1, 0,
                            x2 * R[6] - R[0], x2 * R[7] - R[1], x2 * R[8] - R[2], x2 * t[2] - t[0],
                            y2 * R[6] - R[3], y2 * R[7] - R[4], y2 * R[8] - R[5], y2 * t[2] - t[1]};
            double Q[4];
            vm::dlt_null4(A, Q);
            good = fabs(Q[3]) > 1e-300;
            if (good) {
                X[0] = Q[0] / Q[3]; X[1] = Q[1] / Q[3]; X[2] = Q[2] / Q[3];
                for (int it = 0; it < 3; ++it) {
                    BaLin L;
                    ba_linearise(s_pose, vp.cal1, vp.cal2, X, m1, m2, 1e300, true, L);        // no robust weight in the triangulation
                    double V[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
                    for (int r = 0; r < 2; ++r) {
                        const double* a0 = L.A0 + 3 * r;
                        const double* a1 = L.A1 + 3 * r;
                        const double w0 = L.w0 > 0 ? 1.0 : 0.0, w1 = L.w1 > 0 ? 1.0 : 0.0;
                        V[0] += w0 * a0[0] * a0[0] + w1 * a1[0] * a1[0]; V[1] += w0 * a0[0] * a0[1] + w1 * a1[0] * a1[1];
                        V[2] += w0 * a0[0] * a0[2] + w1 * a1[0] * a1[2]; V[3] += w0 * a0[1] * a0[1] + w1 * a1[1] * a1[1];
                        V[4] += w0 * a0[1] * a0[2] + w1 * a1[1] * a1[2]; V[5] += w0 * a0[2] * a0[2] + w1 * a1[2] * a1[2];
                        for (int c = 0; c < 3; ++c) g[c] += w0 * a0[c] * L.e0[r] + w1 * a1[c] * L.e1[r];
                    }
                    double Vi[6];
                    if (!inv_sym3(V, Vi)) break;
                    X[0] -= Vi[0] * g[0] + Vi[1] * g[1] + Vi[2] * g[2];
                    X[1] -= Vi[1] * g[0] + Vi[3] * g[1] + Vi[4] * g[2];
                    X[2] -= Vi[2] * g[0] + Vi[4] * g[1] + Vi[5] * g[2];
                }
                double u, v;
                const double rx = R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + t[0], ry = R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + t[1],
                             rz = R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + t[2];
                good = X[2] > 0 && rz > 0 && isfinite(X[0]) && isfinite(X[1]) && isfinite(X[2]);     // TriangulationCheiralityException
                if (good) {
                    good = ba_project(vp.cal1, X[0], X[1], X[2], &u, &v, nullptr) && hypot(u - m1.x, v - m1.y) < b.tri_thr;
                    good = good && ba_project(vp.cal2, rx, ry, rz, &u, &v, nullptr) && hypot(u - m2.x, v - m2.y) < b.tri_thr;
                }
                if (good && b.min_angle_deg > 0) {
                    // angle subtended at the landmark by the camera centres c1 = 0 and c2 = -R^T t (utils/tracks.py:82-102):
                    // X - c1 = X and X - c2 = R^T (R X + t)
                    const double ax = X[0], ay = X[1], az = X[2];
                    const double bx = R[0] * rx + R[3] * ry + R[6] * rz, by = R[1] * rx + R[4] * ry + R[7] * rz, bz = R[2] * rx + R[5] * ry + R[8] * rz;
                    const double cs = (ax * bx + ay * by + az * bz) / (sqrt(ax * ax + ay * ay + az * az) * sqrt(bx * bx + by * by + bz * bz));
                    good = acos(fmin(1.0, fmax(-1.0, cs))) * 57.29577951308232 >= b.min_angle_deg;
                }
            }
        }
        if (i < vp.M) b.ba_mask[vp.m0 + i] = 0;          // rewritten by the filter at the end for the tracks that stay
        const unsigned bal = __ballot_sync(0xffffffffu, good);
        if (lane == 0) s_wcnt[warp] = __popc(bal);
        __syncthreads();
        int pos = n_tri + __popc(bal & ((1u << lane) - 1u));
        for (int w = 0; w < warp; ++w) pos += s_wcnt[w];
        if (good) {
            didx[pos] = i;
            Xcur[3 * pos] = X[0]; Xcur[3 * pos + 1] = X[1]; Xcur[3 * pos + 2] = X[2];
        }
        for (int w = 0; w < kBaThreads / 32; ++w) n_tri += s_wcnt[w];
        __syncthreads();
    }
    __syncthreads();                                     // the dense list is read by other threads from here on
    if (n_tri == 0) {                                    // two_view_estimator.py:179-180: the initial pose, no correspondences
        if (tid == 0) { b.out_n[p] = 0; b.out_status[p] = 2; }
        return;
    }

    // ---- robust cost at the start: 0.5 sum rho (gtsam's graph.error)
    auto cost_pass = [&](const BaPose& ps, const double* Xs) {
        double c[1] = {0.0};
        for (int j = tid; j < n_tri; j += kBaThreads) {
            const uint2 m = match[didx[j]];
            BaLin L;
            ba_linearise(ps, vp.cal1, vp.cal2, Xs + 3 * j, kp1[m.x], kp2[m.y], b.huber_k, false, L);
            c[0] += huber_loss(L.e0[0] * L.e0[0] + L.e0[1] * L.e0[1], b.huber_k) + huber_loss(L.e1[0] * L.e1[0] + L.e1[1] * L.e1[1], b.huber_k);
        }
        block_sum<1>(c, s_red, s_out);
        const double r = s_out[0];
        __syncthreads();
        return r;
    };
    double cost = cost_pass(s_pose, Xcur);
    double lambda = 1e-5;                                 // gtsam LevenbergMarquardtParams: lambdaInitial 1e-5, lambdaFactor 10,
    int iters = 0;                                        // lambdaUpperBound 1e5, relative / absolute error tolerance 1e-5
    for (int trial = 0; trial < 4 * b.max_iters && iters < b.max_iters; ++trial) {
        // pass 1: Schur complement of the points onto the pose: S = U - sum W (V + lambda I)^-1 W^T, rhs = -gc + sum W (V + lambda I)^-1 gp
        for (int k = 0; k < 20; ++k) acc[k] = 0.0;
        for (int j = tid; j < n_tri; j += kBaThreads) {
            const uint2 m = match[didx[j]];
            BaLin L;
            ba_linearise(s_pose, vp.cal1, vp.cal2, Xcur + 3 * j, kp1[m.x], kp2[m.y], b.huber_k, true, L);
            double V[6] = {lambda, 0, 0, lambda, 0, lambda}, gp[3] = {0, 0, 0}, W[15], gc[5] = {0, 0, 0, 0, 0};
            for (int k = 0; k < 15; ++k) W[k] = 0.0;
            for (int r = 0; r < 2; ++r) {
                const double* a0 = L.A0 + 3 * r;
                const double* a1 = L.A1 + 3 * r;
                const double* bb = L.B + 5 * r;
                V[0] += L.w0 * a0[0] * a0[0] + L.w1 * a1[0] * a1[0]; V[1] += L.w0 * a0[0] * a0[1] + L.w1 * a1[0] * a1[1];
                V[2] += L.w0 * a0[0] * a0[2] + L.w1 * a1[0] * a1[2]; V[3] += L.w0 * a0[1] * a0[1] + L.w1 * a1[1] * a1[1];
                V[4] += L.w0 * a0[1] * a0[2] + L.w1 * a1[1] * a1[2]; V[5] += L.w0 * a0[2] * a0[2] + L.w1 * a1[2] * a1[2];
                for (int c = 0; c < 3; ++c) gp[c] += L.w0 * a0[c] * L.e0[r] + L.w1 * a1[c] * L.e1[r];
                int k = 0;
                for (int u = 0; u < 5; ++u) {
                    gc[u] += L.w1 * bb[u] * L.e1[r];
                    for (int c = 0; c < 3; ++c) W[3 * u + c] += L.w1 * bb[u] * a1[c];
                    for (int v = u; v < 5; ++v) acc[k++] += L.w1 * bb[u] * bb[v];       // U, upper triangle
                }
            }
            double Vi[6];
            if (!inv_sym3(V, Vi)) { for (int u = 0; u < 5; ++u) acc[15 + u] -= gc[u]; continue; }
            // T = W Vi (5 x 3)
            double T[15];
            for (int u = 0; u < 5; ++u) {
                const double w0 = W[3 * u], w1 = W[3 * u + 1], w2 = W[3 * u + 2];
                T[3 * u] = w0 * Vi[0] + w1 * Vi[1] + w2 * Vi[2];
                T[3 * u + 1] = w0 * Vi[1] + w1 * Vi[3] + w2 * Vi[4];
                T[3 * u + 2] = w0 * Vi[2] + w1 * Vi[4] + w2 * Vi[5];
            }
            int k = 0;
            for (int u = 0; u < 5; ++u) {
                acc[15 + u] += -gc[u] + T[3 * u] * gp[0] + T[3 * u + 1] * gp[1] + T[3 * u + 2] * gp[2];
                for (int v = u; v < 5; ++v) acc[k++] -= T[3 * u] * W[3 * v] + T[3 * u + 1] * W[3 * v + 1] + T[3 * u + 2] * W[3 * v + 2];
            }
        }
        block_sum<20>(acc, s_red, s_out);
        if (tid == 0) {
            double S[25], rhs[5];
            int k = 0;
            for (int u = 0; u < 5; ++u) { rhs[u] = s_out[15 + u]; for (int v = u; v < 5; ++v) { S[5 * u + v] = s_out[k]; S[5 * v + u] = s_out[k]; ++k; } }
            for (int u = 0; u < 5; ++u) S[6 * u] += lambda;
            // Cholesky solve of the 5 x 5 reduced system
            bool pd = true;
            for (int j = 0; j < 5 && pd; ++j) {
                double d = S[6 * j];
                for (int q = 0; q < j; ++q) d -= S[5 * j + q] * S[5 * j + q];
                if (!(d > 0)) { pd = false; break; }
                const double l = sqrt(d);
                S[6 * j] = l;
                for (int i = j + 1; i < 5; ++i) {
                    double v = S[5 * i + j];
                    for (int q = 0; q < j; ++q) v -= S[5 * i + q] * S[5 * j + q];
                    S[5 * i + j] = v / l;
                }
            }
            if (pd) {
                double y[5];
                for (int i = 0; i < 5; ++i) { double v = rhs[i]; for (int q = 0; q < i; ++q) v -= S[5 * i + q] * y[q]; y[i] = v / S[6 * i]; }
                for (int i = 4; i >= 0; --i) { double v = y[i]; for (int q = i + 1; q < 5; ++q) v -= S[5 * q + i] * s_dc[q]; s_dc[i] = v / S[6 * i]; }
                double dR[9], tn[3];
                ba_rodrigues(s_dc, dR);
                vm::mat3_mul(s_pose.R, dR, s_trial.R);
                for (int i = 0; i < 3; ++i) tn[i] = s_pose.t[i] + s_pose.b1[i] * s_dc[3] + s_pose.b2[i] * s_dc[4];
                const double nn = vm::norm3(tn);
                for (int i = 0; i < 3; ++i) s_trial.t[i] = tn[i] / nn;
                ba_tangent(s_trial.t, s_trial.b1, s_trial.b2);
            }
            s_ctrl = pd ? 1 : 0;
        }
        __syncthreads();
        if (!s_ctrl) { lambda *= 10.0; if (lambda > 1e5) break; continue; }
        // pass 2: dX = -(V + lambda I)^-1 (gp + W^T dc) per point (same linearisation), trial points, robust cost at the trial state
        double c1[1] = {0.0};
        for (int j = tid; j < n_tri; j += kBaThreads) {
            const uint2 m = match[didx[j]];
            const double2 m1 = kp1[m.x], m2 = kp2[m.y];
            const double* Xc = Xcur + 3 * j;
            BaLin L;
            ba_linearise(s_pose, vp.cal1, vp.cal2, Xc, m1, m2, b.huber_k, true, L);
            double V[6] = {lambda, 0, 0, lambda, 0, lambda}, g[3] = {0, 0, 0};
            for (int r = 0; r < 2; ++r) {
                const double* a0 = L.A0 + 3 * r;
                const double* a1 = L.A1 + 3 * r;
                const double* bb = L.B + 5 * r;
                V[0] += L.w0 * a0[0] * a0[0] + L.w1 * a1[0] * a1[0]; V[1] += L.w0 * a0[0] * a0[1] + L.w1 * a1[0] * a1[1];
                V[2] += L.w0 * a0[0] * a0[2] + L.w1 * a1[0] * a1[2]; V[3] += L.w0 * a0[1] * a0[1] + L.w1 * a1[1] * a1[1];
                V[4] += L.w0 * a0[1] * a0[2] + L.w1 * a1[1] * a1[2]; V[5] += L.w0 * a0[2] * a0[2] + L.w1 * a1[2] * a1[2];
                double bd = 0;                                   // (B dc)_r: the pose step seen by this residual
                for (int u = 0; u < 5; ++u) bd += bb[u] * s_dc[u];
                for (int c = 0; c < 3; ++c) g[c] += L.w0 * a0[c] * L.e0[r] + L.w1 * a1[c] * (L.e1[r] + bd);
            }
            double Vi[6], Xt[3] = {Xc[0], Xc[1], Xc[2]};
            if (inv_sym3(V, Vi)) {
                Xt[0] -= Vi[0] * g[0] + Vi[1] * g[1] + Vi[2] * g[2];
                Xt[1] -= Vi[1] * g[0] + Vi[3] * g[1] + Vi[4] * g[2];
                Xt[2] -= Vi[2] * g[0] + Vi[4] * g[1] + Vi[5] * g[2];
            }
            double* Xo = Xtry + 3 * j;
            Xo[0] = Xt[0]; Xo[1] = Xt[1]; Xo[2] = Xt[2];
            BaLin T;
            ba_linearise(s_trial, vp.cal1, vp.cal2, Xt, m1, m2, b.huber_k, false, T);
            c1[0] += huber_loss(T.e0[0] * T.e0[0] + T.e0[1] * T.e0[1], b.huber_k) + huber_loss(T.e1[0] * T.e1[0] + T.e1[1] * T.e1[1], b.huber_k);
        }
        block_sum<1>(c1, s_red, s_out);
        const double new_cost = s_out[0];
        __syncthreads();
        if (new_cost < cost && isfinite(new_cost)) {
            // accept: the trial state becomes current (the point buffers swap roles)
            double* sw = Xcur; Xcur = Xtry; Xtry = sw;
            if (tid == 0) s_pose = s_trial;
            __syncthreads();
            const double dec = cost - new_cost;
            const bool converged = dec < 1e-5 || dec < 1e-5 * cost;       // absoluteErrorTol, relativeErrorTol
            cost = new_cost;
            lambda *= 0.1;
            ++iters;
            if (converged) break;
        } else {
            lambda *= 10.0;
            if (lambda > 1e5) break;
        }
    }

    // ---- filter_landmarks(ba_reproj_error_thresholds[-1]): a track stays when both reprojection errors are below the threshold
    double nv[1] = {0.0};
    for (int j = tid; j < n_tri; j += kBaThreads) {
        const int i = didx[j];
        const uint2 m = match[i];
        const double2 m1 = kp1[m.x], m2 = kp2[m.y];
        const double* X = Xcur + 3 * j;
        const double* R = s_pose.R;
        double u, v;
        bool good = ba_project(vp.cal1, X[0], X[1], X[2], &u, &v, nullptr) && hypot(u - m1.x, v - m1.y) < b.ba_thr;
        good = good && ba_project(vp.cal2, R[0] * X[0] + R[1] * X[1] + R[2] * X[2] + s_pose.t[0], R[3] * X[0] + R[4] * X[1] + R[5] * X[2] + s_pose.t[1],
                                  R[6] * X[0] + R[7] * X[1] + R[8] * X[2] + s_pose.t[2], &u, &v, nullptr) && hypot(u - m2.x, v - m2.y) < b.ba_thr;
        b.ba_mask[vp.m0 + i] = good ? 1 : 0;
        nv[0] += good ? 1.0 : 0.0;
    }
    block_sum<1>(nv, s_red, s_out);
    const int n_valid = static_cast<int>(s_out[0]);
    // no track left: filter_landmarks drops both cameras, get_camera_poses() yields None and bundle_adjust returns the INITIAL
    // pose with the (empty) valid correspondences (two_view_estimator.py:199-202) -- the defaults written above stay
    if (n_valid > 0) {
        if (tid < 9) b.out_R[9 * p + tid] = s_pose.R[tid];
        if (tid < 3) b.out_t[3 * p + tid] = s_pose.t[tid];
    }
    if (tid == 0) { b.out_n[p] = n_valid; b.out_status[p] = n_valid > 0 ? 0 : 3; b.out_iters[p] = iters; }
}

}  // namespace g2v

using namespace g2v;

// v_bufs slots of this stage: 27 = points (two buffers), 28 = dense track list (int) | valid-after-BA flags per row,
// 29 = per-pair outputs (R 9, t 3 doubles; valid tracks, status, iterations as ints)
extern "C" int gtsfm_two_view_ba_batch(gtsfm_ctx* c, int n_pairs, const gtsfm_two_view_ba_params* prm, double* i2Ri1, double* i2Ui1,
                                       uint8_t* valid_mask, int32_t* num_valid, int32_t* ba_status, int32_t* iterations)
{
    if (!c || !prm || n_pairs < 0) return fail(c, GTSFM_E_ARG, "bad arguments");
    if (prm->max_iterations < 0 || !(prm->ba_reproj_error_thresh > 0) || !(prm->triangulation_reproj_error_thresh > 0) || !(prm->huber_k > 0))
        return fail(c, GTSFM_E_ARG, "two_view_ba_batch: thresholds and the Huber constant have to be positive, max_iterations >= 0");
    cudaSetDevice(c->device);
    if (n_pairs == 0) { c->ba_n = 0; return GTSFM_OK; }
    if (c->verify_n != n_pairs) return fail(c, GTSFM_E_ARG, "two_view_ba_batch needs the pairs of the last verify_batch");
    cudaStream_t s = c->stream;
    DevBuf* vb = c->v_bufs;
    const size_t total = static_cast<size_t>(c->verify_total), P = static_cast<size_t>(n_pairs);
    G2V_CUDA(c, vb[27].reserve(48 * (total + 1), false, s));
    G2V_CUDA(c, vb[28].reserve(5 * (total + 1), false, s));
    G2V_CUDA(c, vb[29].reserve(112 * P, false, s));
    BaBufs b{};
    b.pairs = vb[0].as<VPair>(); b.match = c->match_out.p; b.mask = vb[12].as<unsigned char>();
    b.R = vb[7].as<double>(); b.t = vb[8].as<double>(); b.ninl = vb[10].as<int>(); b.status = vb[11].as<int>();
    b.X = vb[27].as<double>(); b.Xn = b.X + 3 * (total + 1);
    b.idx = vb[28].as<int>(); b.ba_mask = vb[28].as<unsigned char>() + 4 * (total + 1);
    b.out_R = vb[29].as<double>(); b.out_t = b.out_R + 9 * P;
    b.out_n = reinterpret_cast<int*>(b.out_t + 3 * P); b.out_status = b.out_n + P; b.out_iters = b.out_status + P;
    b.min_inliers = prm->min_num_inliers; b.max_iters = prm->max_iterations;
    b.tri_thr = prm->triangulation_reproj_error_thresh; b.ba_thr = prm->ba_reproj_error_thresh; b.huber_k = prm->huber_k;
    b.min_angle_deg = prm->min_triangulation_angle_deg;
    two_view_ba_kernel<<<n_pairs, kBaThreads, 0, s>>>(c->view(), b, n_pairs);
    G2V_CUDA(c, cudaGetLastError());
    c->launches += 1;
    c->ba_n = n_pairs;
    if (i2Ri1) G2V_CUDA(c, cudaMemcpyAsync(i2Ri1, b.out_R, 72 * P, cudaMemcpyDeviceToHost, s));
    if (i2Ui1) G2V_CUDA(c, cudaMemcpyAsync(i2Ui1, b.out_t, 24 * P, cudaMemcpyDeviceToHost, s));
    if (valid_mask && total) G2V_CUDA(c, cudaMemcpyAsync(valid_mask, b.ba_mask, total, cudaMemcpyDeviceToHost, s));
    if (num_valid) G2V_CUDA(c, cudaMemcpyAsync(num_valid, b.out_n, 4 * P, cudaMemcpyDeviceToHost, s));
    if (ba_status) G2V_CUDA(c, cudaMemcpyAsync(ba_status, b.out_status, 4 * P, cudaMemcpyDeviceToHost, s));
    if (iterations) G2V_CUDA(c, cudaMemcpyAsync(iterations, b.out_iters, 4 * P, cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaStreamSynchronize(s));
    return GTSFM_OK;
}
