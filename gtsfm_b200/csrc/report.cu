// Post-verification report of every pair, batched on the device.  Replaces, for a whole batch,
// the per-pair Python stage that follows verify() in TwoViewEstimator.run_2view (gtsfm/two_view_estimator.py:298-349):
//   compute_correspondence_metrics -> epipolar_inlier_correspondences        gtsfm/utils/metrics.py:38-128
//     F_gt = K2^-T [t]x R K1^-1  (essential_to_fundamental_matrix)           gtsfm/utils/verification.py:113-126
//     squared Sampson distance of every VERIFIED correspondence in pixels    gtsfm/utils/verification.py:170-214
//     correct <=> distance^2 < dist_threshold^2
//   compute_relative_pose_metrics (R_error_deg, U_error_deg)                 gtsfm/utils/geometry_comparisons.py:266-309
//   generate_two_view_report (counts, ratios, mean errors)                   gtsfm/two_view_estimator.py:354-393
//   InlierSupportProcessor.run_inlier_support (keep / drop the pair)         gtsfm/frontend/inlier_support_processor.py:39-95
// It reads the device-resident state of the last gtsfm_verify_batch (match rows, inlier mask, poses), so nothing travels to
// the host between verification and report.
#include <cfloat>
#include <cmath>

#include "ctx.h"
#include "verify_block.cuh"
#include "verify_types.h"

namespace g2v {

struct ReportBufs {
    const VPair* pairs; const void* match; const unsigned char* mask; const double* R; const double* t; const int* ninl; const int* status;
    const int* ninl_ratio;                          // verified count the inlier ratio is quoted on (= ninl before BA)
    const double* gt_R; const double* gt_t;         // [P][9], [P][3] ground-truth i2Ri1 / i2ti1 or nullptr
    const double* gt_cal;                           // [P][10] calibrations of the ground-truth cameras or nullptr (-> VPair's)
    double* dist_sq; unsigned char* gt_mask;        // per match row (NaN / 0 for rows that were not verified)
    int* n_gt; double* in_avg; double* out_avg; double* R_err; double* U_err; int* keep;
    double thr2; int min_inl; double min_ratio;
};

__global__ void __launch_bounds__(kFinThreads) two_view_report_kernel(StoreView st, ReportBufs b, int n_pairs)
{
    __shared__ double s_red[kFinThreads / 32][24];
    __shared__ double s_out[24];
    __shared__ double s_F[9];
    const int p = blockIdx.x;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    const int tid = threadIdx.x;
    const int n_inl = b.ninl[p];
    const bool ok = b.status[p] == 0;
    const bool have_gt = b.gt_R != nullptr && b.gt_t != nullptr && b.gt_R[9 * p] == b.gt_R[9 * p];   // NaN row: no GT for this pair
    if (tid == 0) {
        // InlierSupportProcessor.run_inlier_support: the ratio test first, then the count test (only for valid models)
        // after BA the ratio stays the verifier's (two_view_estimator.py:328-329), the count is the tracks BA kept
        const double ratio = (ok && vp.M > 0) ? static_cast<double>(b.ninl_ratio[p]) / vp.M : 0.0;
        const int n_est = ok ? n_inl : 0;
        bool keep = !(ratio < b.min_ratio);
        if (keep && n_est > 0 && n_est < b.min_inl) keep = false;
        b.keep[p] = (keep && ok) ? 1 : 0;
        double re = NAN, ue = NAN;
        if (have_gt && ok) {
            // angle of R_est^T R_gt (norm of the rotation vector), angle between the unit translations
            const double* R1 = b.R + 9 * p;
            const double* R2 = b.gt_R + 9 * p;
            double D[9];
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) D[3 * i + j] = R1[i] * R2[j] + R1[3 + i] * R2[3 + j] + R1[6 + i] * R2[6 + j];
            const double ax = 0.5 * (D[7] - D[5]), ay = 0.5 * (D[2] - D[6]), az = 0.5 * (D[3] - D[1]);
            const double sn = sqrt(ax * ax + ay * ay + az * az), cs = 0.5 * (D[0] + D[4] + D[8] - 1.0);
            re = atan2(sn, cs) * 57.29577951308232;
            const double* t1 = b.t + 3 * p;
            const double* t2 = b.gt_t + 3 * p;
            const double n2 = sqrt(t2[0] * t2[0] + t2[1] * t2[1] + t2[2] * t2[2]);
            double dot = (t1[0] * t2[0] + t1[1] * t2[1] + t1[2] * t2[2]) / n2;
            dot = fmin(1.0, fmax(-1.0, dot));
            ue = acos(dot) * 57.29577951308232;
        }
        b.R_err[p] = re; b.U_err[p] = ue;
        if (have_gt) {
            // F = K2^-T ([t]x R) K1^-1 with t normalised (EssentialMatrix(R, Unit3(t)))
            const double* R = b.gt_R + 9 * p;
            const double* tg = b.gt_t + 3 * p;
            const double nn = 1.0 / sqrt(tg[0] * tg[0] + tg[1] * tg[1] + tg[2] * tg[2]);
            const double tx = tg[0] * nn, ty = tg[1] * nn, tz = tg[2] * nn;
            double E[9];
            for (int j = 0; j < 3; ++j) {
                E[j] = -tz * R[3 + j] + ty * R[6 + j];
                E[3 + j] = tz * R[j] - tx * R[6 + j];
                E[6 + j] = -ty * R[j] + tx * R[3 + j];
            }
            const double* c1 = b.gt_cal ? b.gt_cal + 10 * p : vp.cal1;
            const double* c2 = b.gt_cal ? b.gt_cal + 10 * p + 5 : vp.cal2;
            const double f1 = c1[0], u1 = c1[3], v1 = c1[4], f2 = c2[0], u2 = c2[3], v2 = c2[4];
            const double K1i[9] = {1.0 / f1, 0, -u1 / f1, 0, 1.0 / f1, -v1 / f1, 0, 0, 1};
            const double K2it[9] = {1.0 / f2, 0, 0, 0, 1.0 / f2, 0, -u2 / f2, -v2 / f2, 1};
            double T[9];
            vm::mat3_mul(K2it, E, T);
            vm::mat3_mul(T, K1i, s_F);
        }
    }
    __syncthreads();
    double acc[4] = {0, 0, 0, 0};      // correct count, sum of d^2 over the correct ones, sum over the others (NaN skipped), their count
    for (int i = tid; i < vp.M; i += kFinThreads) {
        double d2 = NAN;
        unsigned char gm = 0;
        if (have_gt && ok && b.mask[vp.m0 + i]) {
            const uint2 m = static_cast<const uint2*>(b.match)[vp.m0 + i];
            const double2 k1 = st.kps[vp.row1 + m.x], k2 = st.kps[vp.row2 + m.y];
            const double* F = s_F;
            // l2 = F x1, l1 = F^T x2, numerator (x1 . l1)^2, denominator |l1[:2]|^2 + |l2[:2]|^2
            const double l2a = F[0] * k1.x + F[1] * k1.y + F[2], l2b = F[3] * k1.x + F[4] * k1.y + F[5];
            const double l1a = F[0] * k2.x + F[3] * k2.y + F[6], l1b = F[1] * k2.x + F[4] * k2.y + F[7], l1c = F[2] * k2.x + F[5] * k2.y + F[8];
            const double num = k1.x * l1a + k1.y * l1b + l1c;
            d2 = num * num / (l1a * l1a + l1b * l1b + l2a * l2a + l2b * l2b);
            gm = d2 < b.thr2 ? 1 : 0;
            if (gm) { acc[0] += 1.0; acc[1] += d2; }
            else if (d2 == d2) { acc[2] += d2; acc[3] += 1.0; }
        }
        b.dist_sq[vp.m0 + i] = d2;
        b.gt_mask[vp.m0 + i] = gm;
    }
    block_sum<4>(acc, s_red, s_out);
    if (tid == 0) {
        b.n_gt[p] = static_cast<int>(s_out[0]);
        b.in_avg[p] = s_out[0] > 0 ? s_out[1] / s_out[0] : NAN;        // np.mean of an empty selection is NaN
        b.out_avg[p] = s_out[3] > 0 ? s_out[2] / s_out[3] : NAN;       // np.nanmean
    }
}

}  // namespace g2v

using namespace g2v;

extern "C" int gtsfm_two_view_report_batch(gtsfm_ctx* c, int n_pairs, const double* gt_i2Ri1, const double* gt_i2ti1,
                                           const gtsfm_report_params* prm, double* dist_sq, uint8_t* gt_mask, int32_t* num_inliers_gt,
                                           double* inlier_avg_err, double* outlier_avg_err, double* R_error_deg, double* U_error_deg,
                                           int32_t* keep)
{
    if (!c || !prm || n_pairs < 0) return fail(c, GTSFM_E_ARG, "bad arguments");
    cudaSetDevice(c->device);
    if (n_pairs == 0) return GTSFM_OK;
    if (c->verify_n != n_pairs) return fail(c, GTSFM_E_ARG, "two_view_report_batch needs the pairs of the last verify_batch");
    if ((gt_i2Ri1 == nullptr) != (gt_i2ti1 == nullptr)) return fail(c, GTSFM_E_ARG, "ground truth needs both the rotations and the translations");
    cudaStream_t s = c->stream;
    DevBuf* vb = c->v_bufs;
    const size_t total = static_cast<size_t>(c->verify_total);
    G2V_CUDA(c, vb[19].reserve(176 * static_cast<size_t>(n_pairs), false, s));         // gt R, t, calibrations
    G2V_CUDA(c, vb[20].reserve(8 * (total + 1), false, s));                              // dist_sq
    G2V_CUDA(c, vb[21].reserve(total + 1, false, s));                                    // gt mask
    G2V_CUDA(c, vb[22].reserve(48 * static_cast<size_t>(n_pairs), false, s));            // n_gt, keep (int) + 4 doubles per pair
    ReportBufs b{};
    b.pairs = vb[0].as<VPair>(); b.match = c->match_out.p; b.mask = vb[12].as<unsigned char>();
    b.R = vb[7].as<double>(); b.t = vb[8].as<double>(); b.ninl = vb[10].as<int>(); b.status = vb[11].as<int>();
    b.ninl_ratio = b.ninl;
    if (prm->after_ba) {
        // the post-BA report (two_view_estimator.py:319-333): poses, valid tracks and their count from gtsfm_two_view_ba_batch
        if (c->ba_n != n_pairs) return fail(c, GTSFM_E_ARG, "two_view_report_batch(after_ba) needs a preceding two_view_ba_batch");
        const size_t P = static_cast<size_t>(n_pairs);
        b.mask = vb[28].as<unsigned char>() + 4 * (total + 1);
        b.R = vb[29].as<double>(); b.t = b.R + 9 * P;
        b.ninl = reinterpret_cast<const int*>(b.t + 3 * P);
    }
    if (gt_i2Ri1) {
        double* g = vb[19].as<double>();
        G2V_CUDA(c, cudaMemcpyAsync(g, gt_i2Ri1, 72 * static_cast<size_t>(n_pairs), cudaMemcpyHostToDevice, s));
        G2V_CUDA(c, cudaMemcpyAsync(g + 9ll * n_pairs, gt_i2ti1, 24 * static_cast<size_t>(n_pairs), cudaMemcpyHostToDevice, s));
        b.gt_R = g; b.gt_t = g + 9ll * n_pairs;
        if (prm->gt_cal) {
            G2V_CUDA(c, cudaMemcpyAsync(g + 12ll * n_pairs, prm->gt_cal, 80 * static_cast<size_t>(n_pairs), cudaMemcpyHostToDevice, s));
            b.gt_cal = g + 12ll * n_pairs;
        }
    }
    b.dist_sq = vb[20].as<double>(); b.gt_mask = vb[21].as<unsigned char>();
    double* pd = vb[22].as<double>();
    b.in_avg = pd; b.out_avg = pd + n_pairs; b.R_err = pd + 2ll * n_pairs; b.U_err = pd + 3ll * n_pairs;
    b.n_gt = reinterpret_cast<int*>(pd + 4ll * n_pairs); b.keep = b.n_gt + n_pairs;
    b.thr2 = prm->dist_threshold_px * prm->dist_threshold_px;
    b.min_inl = prm->min_num_inliers_est_model; b.min_ratio = prm->min_inlier_ratio_est_model;
    two_view_report_kernel<<<n_pairs, kFinThreads, 0, s>>>(c->view(), b, n_pairs);
    G2V_CUDA(c, cudaGetLastError());
    c->launches += 1;
    const size_t np8 = 8 * static_cast<size_t>(n_pairs), np4 = 4 * static_cast<size_t>(n_pairs);
    if (dist_sq && total) G2V_CUDA(c, cudaMemcpyAsync(dist_sq, b.dist_sq, 8 * total, cudaMemcpyDeviceToHost, s));
    if (gt_mask && total) G2V_CUDA(c, cudaMemcpyAsync(gt_mask, b.gt_mask, total, cudaMemcpyDeviceToHost, s));
    if (num_inliers_gt) G2V_CUDA(c, cudaMemcpyAsync(num_inliers_gt, b.n_gt, np4, cudaMemcpyDeviceToHost, s));
    if (keep) G2V_CUDA(c, cudaMemcpyAsync(keep, b.keep, np4, cudaMemcpyDeviceToHost, s));
    if (inlier_avg_err) G2V_CUDA(c, cudaMemcpyAsync(inlier_avg_err, b.in_avg, np8, cudaMemcpyDeviceToHost, s));
    if (outlier_avg_err) G2V_CUDA(c, cudaMemcpyAsync(outlier_avg_err, b.out_avg, np8, cudaMemcpyDeviceToHost, s));
    if (R_error_deg) G2V_CUDA(c, cudaMemcpyAsync(R_error_deg, b.R_err, np8, cudaMemcpyDeviceToHost, s));
    if (U_error_deg) G2V_CUDA(c, cudaMemcpyAsync(U_error_deg, b.U_err, np8, cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaStreamSynchronize(s));
    return GTSFM_OK;
}
