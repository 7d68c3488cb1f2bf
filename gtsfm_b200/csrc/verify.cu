// Host orchestration of the batched Ransac verifier (C ABI: gtsfm_ransac_samples / gtsfm_verify_batch /
// gtsfm_verify_pair).  Replaces the one-Dask-task-per-pair loop of gtsfm/two_view_estimator.py:569-584 for the
// verifier part of run_2view (two_view_estimator.py:290-296).
#include <algorithm>
#include <cmath>
#include <cstring>

#include "ctx.h"
#include "verify_types.h"

namespace g2v {
cudaError_t launch_verify_step(int step, const StoreView& st, const VerifyBufs& b, const VerifyDims& d, int n_pairs, int max_m, cudaStream_t s);
cudaError_t launch_verify_step_f(int step, const StoreView& st, const VerifyBufs& b, const VerifyDims& d, int n_pairs, int max_m, cudaStream_t s);
}  // namespace g2v


namespace g2v {

struct VOut { double* R; double* t; double* E; int* ninl; int* status; };

// Runs the four verifier steps over `n_pairs` pair records (device array d_vps, host copy vps) in chunks sized by the
// workspace budget; per-pair outputs go to `out` (indexed like d_vps), the mask to the batch-wide array (indexed by VPair::m0).
static int run_verify_chunks(gtsfm_ctx* c, const VPair* vps, const VPair* d_vps, int n_pairs, int H, bool fmode,
                             const gtsfm_verify_params* prm, const int* d_samples, const VOut& out)
{
    cudaStream_t s = c->stream;
    DevBuf* vb = c->v_bufs;
    const int ssz = fmode ? 7 : 5;
    const size_t per_pair = static_cast<size_t>(H) * (9 * 10 * 4 + 2 + 10 * 4);   // F mode needs less (216 + 1 + 12 per hypothesis)
    int chunk = static_cast<int>(std::max<long long>(1, std::min<long long>(c->workspace_bytes / static_cast<long long>(per_pair), 32768)));
    chunk = std::min(chunk, n_pairs);
    G2V_CUDA(c, vb[3].reserve(static_cast<size_t>(chunk) * H * 360, false, s));
    G2V_CUDA(c, vb[4].reserve(static_cast<size_t>(chunk) * H * 2, false, s));
    G2V_CUDA(c, vb[5].reserve(static_cast<size_t>(chunk) * H * 40, false, s));
    G2V_CUDA(c, vb[14].reserve(4 * static_cast<size_t>(chunk), false, s));
    G2V_CUDA(c, vb[15].reserve(32 * static_cast<size_t>(chunk), false, s));
    G2V_CUDA(c, vb[16].reserve(576 * static_cast<size_t>(chunk), false, s));
    if (fmode) {
        G2V_CUDA(c, vb[17].reserve(16 * static_cast<size_t>(chunk), false, s));
        G2V_CUDA(c, vb[18].reserve(static_cast<size_t>(chunk) * H * 144, false, s));
    }
    const StoreView st = c->view();
    VerifyDims d{H, prm->lo_iters, prm->seed, prm->threshold_px};
    for (int p0 = 0; p0 < n_pairs; p0 += chunk) {
        const int np = std::min(chunk, n_pairs - p0);
        VerifyBufs b{};
        b.pairs = d_vps + p0;
        b.match = c->match_out.p;
        b.corr_d = vb[1].p; b.corr_f = vb[2].p;
        b.E32 = vb[3].as<float>(); b.vmask = vb[4].as<unsigned short>(); b.cost = vb[5].as<float>();
        b.samples = d_samples ? d_samples + static_cast<long long>(p0) * H * ssz : nullptr;
        b.R = out.R + 9ll * p0; b.t = out.t + 3ll * p0; b.E = out.E + 9ll * p0;
        b.stats = vb[13].as<unsigned long long>();
        b.best_cost = vb[14].as<unsigned int>();
        b.topi = vb[15].as<int>(); b.Ek = vb[16].as<double>();
        if (fmode) {
            b.Fd = vb[3].as<double>(); b.nroot = vb[4].as<unsigned char>(); b.cnt = vb[5].as<int>();
            b.cmax = vb[17].as<float>(); b.best_cnt = vb[14].as<int>(); b.Ff = vb[18].as<float>();
            G2V_CUDA(c, cudaMemsetAsync(vb[14].p, 0, 4 * static_cast<size_t>(np), s));
            G2V_CUDA(c, cudaMemsetAsync(vb[17].p, 0, 16 * static_cast<size_t>(np), s));
        } else {
            G2V_CUDA(c, cudaMemsetAsync(vb[14].p, 0x7f, 4 * static_cast<size_t>(np), s));   // 0x7f7f7f7f = 3.4e38f
        }
        b.ninl = out.ninl + p0; b.status = out.status + p0; b.mask = vb[12].as<unsigned char>();
        int mm = 0;
        for (int p = p0; p < p0 + np; ++p) mm = std::max(mm, vps[p].M);
        for (int step = 0; step < 4; ++step) {
            c->prof_begin(P_GATHER + step);
            G2V_CUDA(c, fmode ? launch_verify_step_f(step, st, b, d, np, mm, s) : launch_verify_step(step, st, b, d, np, mm, s));
            c->prof_end();
        }
        c->launches += fmode ? 4 : 6;
    }
    return GTSFM_OK;
}

// Adaptive rounds: the per-pair outputs of a re-run (sub-list order) replace those of the pairs they belong to.
__global__ void scatter_pair_outputs_kernel(const int* __restrict__ idx, int n, VOut src, VOut dst)
{
    const int k = blockIdx.x;
    if (k >= n) return;
    const int p = idx[k], t = threadIdx.x;
    if (t < 9) { dst.R[9ll * p + t] = src.R[9ll * k + t]; dst.E[9ll * p + t] = src.E[9ll * k + t]; }
    if (t < 3) dst.t[3ll * p + t] = src.t[3ll * k + t];
    if (t == 0) { dst.ninl[p] = src.ninl[k]; dst.status[p] = src.status[k]; }
}

// cv2's RANSACUpdateNumIters (the rule behind findFundamentalMat's confidence / maxIters, ransac.py:22-23,103-110): iterations
// after which a sample of `model_points` inliers has been drawn with probability `conf` at inlier ratio w.
static long long ransac_required_iters(double conf, double w, int model_points, long long max_iters)
{
    if (!(w > 0)) return max_iters;
    const double num = std::log(1.0 - conf);
    const double denom = std::log(1.0 - std::pow(w, model_points));
    if (!(denom < 0) || -num >= static_cast<double>(max_iters) * (-denom)) return max_iters;
    return std::llround(num / denom);
}

}  // namespace g2v

using namespace g2v;

extern "C" {

int gtsfm_ransac_samples(uint64_t seed, uint64_t pair_key, int n_matches, int n_hyp, int32_t* out)
{
    if (!out || n_matches < 5 || n_hyp < 0) return GTSFM_E_ARG;
    for (int h = 0; h < n_hyp; ++h) sample5(seed, pair_key, static_cast<uint32_t>(h), static_cast<uint32_t>(n_matches), out + 5 * h);
    return GTSFM_OK;
}

int gtsfm_ransac_samples_k(uint64_t seed, uint64_t pair_key, int n_matches, int n_hyp, int sample_size, int32_t* out)
{
    if (sample_size == 5) return gtsfm_ransac_samples(seed, pair_key, n_matches, n_hyp, out);
    if (sample_size != 7 || !out || n_matches < 7 || n_hyp < 0) return GTSFM_E_ARG;
    for (int h = 0; h < n_hyp; ++h) sample_k<7>(seed, pair_key, static_cast<uint32_t>(h), static_cast<uint32_t>(n_matches), out + 7 * h);
    return GTSFM_OK;
}

int gtsfm_verify_batch(gtsfm_ctx* c, const int32_t* pairs, int n_pairs, const uint32_t* match_idx, const int64_t* match_offsets,
                       const gtsfm_verify_params* prm, double* R, double* t, int32_t* n_inliers, int32_t* status,
                       uint8_t* inlier_mask, double* E)
{
    if (!c || !prm || n_pairs < 0 || (n_pairs > 0 && !pairs)) return fail(c, GTSFM_E_ARG, "bad arguments");
    cudaSetDevice(c->device);
    c->verify_n = 0; c->ba_n = 0;
    if (n_pairs == 0) return GTSFM_OK;
    if (prm->n_hypotheses <= 0 || prm->n_hypotheses > (1 << 20)) return fail(c, GTSFM_E_ARG, "n_hypotheses out of range");
    if (!(prm->threshold_px > 0)) return fail(c, GTSFM_E_ARG, "threshold_px must be positive");
    cudaStream_t s = c->stream;
    const int n_img = static_cast<int>(c->images.size());
    const int H = prm->n_hypotheses;
    const bool fmode = prm->estimate_fundamental != 0;     // Ransac(use_intrinsics_in_verification=False), ransac.py:84-111
    const int min_m = fmode ? 8 : 6;                       // opencv_verifier_base.py:38-40,69-78
    const int ssz = fmode ? 7 : 5;
    int r;

    // ---- match rows: the device-resident result of the last match batch, or the caller's arrays
    std::vector<long long> off(n_pairs + 1, 0);
    if (match_idx == nullptr && match_offsets == nullptr) {
        if (c->batch_n != n_pairs) return fail(c, GTSFM_E_ARG, "verify_batch without matches needs the pairs of the last match_batch");
        for (int p = 0; p < n_pairs; ++p) {
            if (c->batch_img[2 * p] != pairs[2 * p] || c->batch_img[2 * p + 1] != pairs[2 * p + 1])
                return fail(c, GTSFM_E_ARG, "pair list differs from the last match_batch");
            off[p + 1] = off[p] + c->batch_counts[p];
        }
    } else {
        if (!match_offsets) return fail(c, GTSFM_E_ARG, "match_offsets is NULL");
        if ((r = c->wait_fetch())) return r;                 // match_out is about to be overwritten
        for (int p = 0; p <= n_pairs; ++p) off[p] = match_offsets[p];
        for (int p = 0; p < n_pairs; ++p) if (off[p + 1] < off[p]) return fail(c, GTSFM_E_ARG, "match_offsets must be non-decreasing");
        if (off[n_pairs] > 0 && !match_idx) return fail(c, GTSFM_E_ARG, "match_idx is NULL");
        G2V_CUDA(c, c->match_out.reserve(8 * static_cast<size_t>(off[n_pairs] + 1), false, s));
        if (off[n_pairs] > 0)
            G2V_CUDA(c, cudaMemcpyAsync(c->match_out.p, match_idx, 8 * static_cast<size_t>(off[n_pairs]), cudaMemcpyHostToDevice, s));
        c->batch_n = 0;   // the device match buffer no longer belongs to a match batch
    }
    const long long total = off[n_pairs];

    // ---- per-pair records
    std::vector<VPair> vps(n_pairs);
    int max_m = 0;
    {
        int max_image = -1;
        for (int p = 0; p < n_pairs; ++p) max_image = std::max(max_image, std::max(pairs[2 * p], pairs[2 * p + 1]));
        if ((r = c->wait_images(max_image, false))) return r;        // keypoint uploads run on the copy stream
    }
    for (int p = 0; p < n_pairs; ++p) {
        const int a = pairs[2 * p], b = pairs[2 * p + 1];
        if (a < 0 || b < 0 || a >= n_img || b >= n_img) return fail(c, GTSFM_E_ARG, "pair names an image that was not uploaded");
        const ImageRec& i1 = c->images[a];
        const ImageRec& i2 = c->images[b];
        VPair v{};
        v.m0 = off[p];
        v.M = static_cast<int>(off[p + 1] - off[p]);
        // opencv_verifier_base.py:69-78: fewer than 5 (then fewer than 6) correspondences -> failure tuple
        v.status = v.M < min_m ? GTSFM_V_TOO_FEW : GTSFM_V_OK;
        if (v.status == GTSFM_V_OK && (!i1.has_kps || !i2.has_kps)) return fail(c, GTSFM_E_ARG, "verification needs keypoints for every image of the pair list");
        v.row1 = i1.row0; v.row2 = i2.row0;
        std::memcpy(v.cal1, i1.cal, sizeof(v.cal1));
        std::memcpy(v.cal2, i2.cal, sizeof(v.cal2));
        const double fx = std::max(i1.cal[0], i2.cal[0]);                 // opencv_verifier_base.py:86
        const double thr = prm->threshold_px / fx;                         // ransac.py:79
        v.thr2 = fmode ? prm->threshold_px * prm->threshold_px : thr * thr;   // F mode works in pixels (ransac.py:103-109)
        v.key = prm->pair_keys ? prm->pair_keys[p] : static_cast<unsigned long long>(p);
        max_m = std::max(max_m, v.M);
        vps[p] = v;
        if (match_idx) {
            // caller-supplied rows index the keypoints of the two images (the reference would raise IndexError on a bad one)
            const uint32_t n1 = static_cast<uint32_t>(i1.n_kps), n2 = static_cast<uint32_t>(i2.n_kps);
            const uint32_t* row = match_idx + 2 * off[p];
            for (int k = 0; k < v.M; ++k)
                if (row[2 * k] >= n1 || row[2 * k + 1] >= n2)
                    return fail(c, GTSFM_E_ARG, "pair " + std::to_string(p) + ": match row " + std::to_string(k) + " indexes a keypoint that image " +
                                                    std::to_string(row[2 * k] >= n1 ? a : b) + " does not have");
        }
    }

    // ---- buffers
    DevBuf* vb = c->v_bufs;
    G2V_CUDA(c, vb[0].reserve(sizeof(VPair) * n_pairs, false, s));
    G2V_CUDA(c, vb[1].reserve(32 * static_cast<size_t>(total + 1), false, s));
    G2V_CUDA(c, vb[2].reserve(16 * static_cast<size_t>(total + 1), false, s));
    G2V_CUDA(c, vb[7].reserve(72 * static_cast<size_t>(n_pairs), false, s));
    G2V_CUDA(c, vb[8].reserve(24 * static_cast<size_t>(n_pairs), false, s));
    G2V_CUDA(c, vb[9].reserve(72 * static_cast<size_t>(n_pairs), false, s));
    G2V_CUDA(c, vb[10].reserve(4 * static_cast<size_t>(n_pairs), false, s));
    G2V_CUDA(c, vb[11].reserve(4 * static_cast<size_t>(n_pairs), false, s));
    G2V_CUDA(c, vb[12].reserve(static_cast<size_t>(total + 1), false, s));
    if (!vb[13].p) { G2V_CUDA(c, vb[13].reserve(32, false, s)); G2V_CUDA(c, cudaMemsetAsync(vb[13].p, 0, 32, s)); }
    G2V_CUDA(c, c->h_pairs.reserve(sizeof(VPair) * n_pairs));
    std::memcpy(c->h_pairs.p, vps.data(), sizeof(VPair) * n_pairs);
    G2V_CUDA(c, cudaMemcpyAsync(vb[0].p, c->h_pairs.p, sizeof(VPair) * n_pairs, cudaMemcpyHostToDevice, s));
    if (prm->samples) {
        const size_t nb = static_cast<size_t>(n_pairs) * H * ssz * 4;
        G2V_CUDA(c, vb[6].reserve(nb, false, s));
        G2V_CUDA(c, cudaMemcpyAsync(vb[6].p, prm->samples, nb, cudaMemcpyHostToDevice, s));
    }
    const VOut full{vb[7].as<double>(), vb[8].as<double>(), vb[9].as<double>(), vb[10].as<int>(), vb[11].as<int>()};
    if ((r = run_verify_chunks(c, vps.data(), vb[0].as<VPair>(), n_pairs, H, fmode, prm, prm->samples ? vb[6].as<int>() : nullptr, full))) return r;
    c->verify_hyp.assign(n_pairs, H);

    // ---- fundamental-matrix mode, adaptive budget: cv2.findFundamentalMat(FM_RANSAC, confidence, maxIters) stops after
    // log(1 - conf) / log(1 - w^7) iterations at inlier ratio w (ransac.py:22-23,103-110).  Pairs whose best model does not
    // justify stopping after H hypotheses are run again with twice the budget (up to the cap); hypothesis h
    // draws the same sample whatever the budget, so a re-run can only find a model with at least as many inliers.
    const long long cap = std::min<long long>(prm->adaptive_max_hypotheses, 1ll << 20);
    if (fmode && !prm->samples && cap > H) {
        const double conf = prm->confidence > 0 && prm->confidence < 1 ? prm->confidence : 0.999999;
        std::vector<int> h_ninl(n_pairs);
        while (true) {
            G2V_CUDA(c, cudaMemcpyAsync(h_ninl.data(), vb[10].p, 4 * static_cast<size_t>(n_pairs), cudaMemcpyDeviceToHost, s));
            G2V_CUDA(c, cudaStreamSynchronize(s));
            std::vector<std::pair<int, int>> redo;                  // (budget, pair)
            for (int p = 0; p < n_pairs; ++p) {
                if (vps[p].status != GTSFM_V_OK || c->verify_hyp[p] >= cap) continue;
                const long long need = ransac_required_iters(conf, static_cast<double>(h_ninl[p]) / vps[p].M, 7, cap);
                if (need <= c->verify_hyp[p]) continue;
                // double rather than jump to `need`: the first rounds underestimate w (their best model is contaminated), and
                // cv2 likewise lowers its iteration count every time a better model turns up; total work <= 2x the last round
                redo.emplace_back(static_cast<int>(std::min(2ll * c->verify_hyp[p], cap)), p);
            }
            if (redo.empty()) break;
            std::sort(redo.begin(), redo.end());
            for (size_t a0 = 0; a0 < redo.size();) {
                size_t a1 = a0;
                while (a1 < redo.size() && redo[a1].first == redo[a0].first) ++a1;
                const int n_sub = static_cast<int>(a1 - a0), H2 = redo[a0].first;
                std::vector<VPair> sub(n_sub);
                std::vector<int> idx(n_sub);
                for (int k = 0; k < n_sub; ++k) { idx[k] = redo[a0 + k].second; sub[k] = vps[idx[k]]; c->verify_hyp[idx[k]] = H2; }
                G2V_CUDA(c, vb[19].reserve(sizeof(VPair) * n_sub, false, s));
                G2V_CUDA(c, vb[20].reserve(72 * static_cast<size_t>(n_sub), false, s));
                G2V_CUDA(c, vb[21].reserve(24 * static_cast<size_t>(n_sub), false, s));
                G2V_CUDA(c, vb[22].reserve(72 * static_cast<size_t>(n_sub), false, s));
                G2V_CUDA(c, vb[23].reserve(12 * static_cast<size_t>(n_sub), false, s));
                G2V_CUDA(c, cudaMemcpyAsync(vb[19].p, sub.data(), sizeof(VPair) * n_sub, cudaMemcpyHostToDevice, s));
                int* d_idx = vb[23].as<int>() + 2 * n_sub;
                G2V_CUDA(c, cudaMemcpyAsync(d_idx, idx.data(), 4 * static_cast<size_t>(n_sub), cudaMemcpyHostToDevice, s));
                const VOut part{vb[20].as<double>(), vb[21].as<double>(), vb[22].as<double>(), vb[23].as<int>(), vb[23].as<int>() + n_sub};
                if ((r = run_verify_chunks(c, sub.data(), vb[19].as<VPair>(), n_sub, H2, true, prm, nullptr, part))) return r;
                scatter_pair_outputs_kernel<<<n_sub, 32, 0, s>>>(d_idx, n_sub, part, full);
                G2V_CUDA(c, cudaGetLastError());
                c->launches += 1;
                G2V_CUDA(c, cudaStreamSynchronize(s));        // `sub` / `idx` are released at the end of the scope
                a0 = a1;
            }
        }
    }
    // ---- results back to the host
    if (R) G2V_CUDA(c, cudaMemcpyAsync(R, vb[7].p, 72 * static_cast<size_t>(n_pairs), cudaMemcpyDeviceToHost, s));
    if (t) G2V_CUDA(c, cudaMemcpyAsync(t, vb[8].p, 24 * static_cast<size_t>(n_pairs), cudaMemcpyDeviceToHost, s));
    if (E) G2V_CUDA(c, cudaMemcpyAsync(E, vb[9].p, 72 * static_cast<size_t>(n_pairs), cudaMemcpyDeviceToHost, s));
    if (n_inliers) G2V_CUDA(c, cudaMemcpyAsync(n_inliers, vb[10].p, 4 * static_cast<size_t>(n_pairs), cudaMemcpyDeviceToHost, s));
    if (status) G2V_CUDA(c, cudaMemcpyAsync(status, vb[11].p, 4 * static_cast<size_t>(n_pairs), cudaMemcpyDeviceToHost, s));
    if (inlier_mask && total > 0) G2V_CUDA(c, cudaMemcpyAsync(inlier_mask, vb[12].p, static_cast<size_t>(total), cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaStreamSynchronize(s));
    c->verify_n = n_pairs;
    c->verify_total = total;
    return GTSFM_OK;
}

int gtsfm_verify_hypotheses_used(gtsfm_ctx* c, int32_t* out, int n_pairs)
{
    if (!c || !out || n_pairs < 0 || static_cast<size_t>(n_pairs) > c->verify_hyp.size()) return fail(c, GTSFM_E_ARG, "no verify batch of that size");
    for (int p = 0; p < n_pairs; ++p) out[p] = c->verify_hyp[p];
    return GTSFM_OK;
}

int gtsfm_verify_stats(gtsfm_ctx* c, int64_t* out, int reset)
{
    if (!c || !out) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    out[0] = out[1] = out[2] = out[3] = 0;
    DevBuf& sb = c->v_bufs[13];
    if (!sb.p) return GTSFM_OK;
    unsigned long long h[4] = {0, 0, 0, 0};
    G2V_CUDA(c, cudaMemcpyAsync(h, sb.p, 32, cudaMemcpyDeviceToHost, c->stream));
    if (reset) G2V_CUDA(c, cudaMemsetAsync(sb.p, 0, 32, c->stream));
    G2V_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int k = 0; k < 4; ++k) out[k] = static_cast<int64_t>(h[k]);
    return GTSFM_OK;
}

int gtsfm_verify_pair(gtsfm_ctx* c, const double* kps1, int n1, const double* kps2, int n2, const uint32_t* match_idx, int n_matches,
                      const double* cal1, const double* cal2, const gtsfm_verify_params* prm, double* R, double* t,
                      int32_t* n_inliers, int32_t* status, uint8_t* inlier_mask, double* E)
{
    if (!c || !prm || n_matches < 0 || !cal1 || !cal2) return fail(c, GTSFM_E_ARG, "bad arguments");
    for (int k = 0; k < n_matches; ++k)
        if (match_idx[2 * k] >= static_cast<uint32_t>(n1) || match_idx[2 * k + 1] >= static_cast<uint32_t>(n2))
            return fail(c, GTSFM_E_ARG, "match index out of range");
    int r;
    if ((r = gtsfm_store_clear(c))) return r;
    if ((r = gtsfm_image_upload(c, 0, nullptr, n1, 0, kps1, cal1))) return r;
    if ((r = gtsfm_image_upload(c, 1, nullptr, n2, 0, kps2, cal2))) return r;
    const int32_t pr[2] = {0, 1};
    const int64_t off[2] = {0, n_matches};
    return gtsfm_verify_batch(c, pr, 1, match_idx, off, prm, R, t, n_inliers, status, inlier_mask, E);
}

}  // extern "C"
