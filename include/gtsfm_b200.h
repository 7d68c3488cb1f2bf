/*
 * gtsfm_b200 -- C ABI of the B200-native two-view front-end (descriptor matching + essential-matrix RANSAC).
 *
 * Drop-in boundary for the reference's plugin API.  Each entry point names the reference interface it replaces
 * (paths relative to the GTSfM tree):
 *   MatcherBase.match(...)            gtsfm/frontend/matcher/matcher_base.py:31-63
 *   TwoWayMatcher.match(...)          gtsfm/frontend/matcher/twoway_matcher.py:42-144
 *   VerifierBase.verify(...)          gtsfm/frontend/verifier/verifier_base.py:58-82
 *   OpencvVerifierBase.verify(...)    gtsfm/frontend/verifier/opencv_verifier_base.py:45-109
 *   Ransac.estimate_E(...)            gtsfm/frontend/verifier/ransac.py:52-82
 *   recover_relative_pose_from_essential_matrix   gtsfm/utils/verification.py:52-94
 *   per-pair submission loops         gtsfm/frontend/correspondence_generator/det_desc_correspondence_generator.py:66-85,
 *                                     gtsfm/two_view_estimator.py:567-587   (one call here per BATCH of pairs)
 *
 * Conventions: plain pointers and sizes, host memory unless a name says otherwise, no exceptions; every function
 * returns 0 on success or a GTSFM_E_* code, and gtsfm_last_error() gives the text.  All buffers are borrowed for the
 * duration of the call.  A context is bound to one CUDA device and is not thread-safe: use one per thread / process
 * (the reference runs one matcher per Dask worker process, gtsfm/runner/gtsfm_runner_base.py:288-296).
 * There is no CPU fallback: without a CUDA device gtsfm_ctx_create fails.
 */
#ifndef GTSFM_B200_H
#define GTSFM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gtsfm_ctx gtsfm_ctx;

enum {
    GTSFM_OK = 0,
    GTSFM_E_CUDA = 1,        /* a CUDA call failed (text in gtsfm_last_error) */
    GTSFM_E_ARG = 2,         /* bad argument */
    GTSFM_E_NO_DEVICE = 3,   /* no usable sm_100 device */
    GTSFM_E_KNN_UNPACK = 4,  /* ratio test requested with < 2 valid train descriptors: the reference raises
                                ValueError at twoway_matcher.py:137 ("not enough values to unpack") */
    GTSFM_E_UNSUPPORTED = 5
};

/* per-pair verifier status (matches the oracle's VSTAT_* codes) */
enum {
    GTSFM_V_OK = 0,
    GTSFM_V_TOO_FEW = 1,     /* fewer than 6 (F mode: 8) putative matches -> failure tuple (opencv_verifier_base.py:69-78) */
    GTSFM_V_NO_MODEL = 2     /* no essential matrix / no inliers */
};

/* ---- context ------------------------------------------------------------------------------------------------ */
int gtsfm_ctx_create(int device, gtsfm_ctx** out);
void gtsfm_ctx_destroy(gtsfm_ctx* ctx);
const char* gtsfm_last_error(const gtsfm_ctx* ctx);   /* ctx may be NULL: error of the last failed create */
/* Work-buffer budget in bytes for one chunk of pairs (default 12 GiB). */
int gtsfm_ctx_set_workspace(gtsfm_ctx* ctx, int64_t bytes);
/* Float (non-integer) descriptors, ratio test off: on (default) = fp16 tensor-core screening + exact recompute of the
 * winning candidates + exhaustive re-scan of ambiguous rows; off = every distance by the exhaustive fp32 kernel.  Both give
 * the same matches (cv2's float32 arithmetic); the switch exists for tests and measurements. */
int gtsfm_ctx_set_float_screening(gtsfm_ctx* ctx, int on);
/* Number of hand-written kernels launched by this context since creation (bench.py's gpu_launches). */
int64_t gtsfm_kernel_launches(const gtsfm_ctx* ctx);

/* Per-kernel device timing with CUDA events on the context's stream.  Slots: 0 prepare, 1 screen (tcgen05 GEMM),
 * 2 resolve rows, 3 resolve columns, 4 exact scan, 5 match finalize, 6 gather/normalise, 7 5-point solve,
 * 8 MSAC scoring, 9 verifier finalize.  enable(…, 1) resets the totals; get() synchronises and returns the
 * accumulated milliseconds and launch counts of the first n_slots slots. */
int gtsfm_profile_enable(gtsfm_ctx* ctx, int on);
int gtsfm_profile_get(gtsfm_ctx* ctx, double* ms, int64_t* calls, int n_slots);
/* User events on the context's stream (which in 0..7): record, then elapsed milliseconds between two of them. */
int gtsfm_event_record(gtsfm_ctx* ctx, int which);
int gtsfm_event_elapsed_ms(gtsfm_ctx* ctx, int a, int b, double* ms);

/* ---- descriptor / keypoint store (replaces pickling Keypoints + descriptors to every Dask task) ---------------- */
/* Drop all images. */
int gtsfm_store_clear(gtsfm_ctx* ctx);
/* Upload image `image_id` (ids are dense, 0..n-1, in upload order).
 *   desc  (n, dim) float32 row-major  -- the array MatcherBase.match receives as descriptors_i*
 *   kps   (n, 2) float64 pixel coordinates (Keypoints.coordinates) or NULL when only matching is wanted
 *   cal   {f, k1, k2, u0, v0} of the image's Cal3Bundler or NULL */
int gtsfm_image_upload(gtsfm_ctx* ctx, int image_id, const float* desc, int n, int dim, const double* kps, const double* cal);

/* A list of images in one call (ids first_image_id, first_image_id + 1, ...): desc[i] (n[i], dim) float32, kps[i] (n[i], 2)
 * float64 (kps NULL, or kps[i] NULL: no keypoints), cal (n_images, 5) or NULL.  Same as n_images calls of gtsfm_image_upload
 * without their per-call overhead; the copies are only ENQUEUED (on the context's copy stream), so the host arrays must stay
 * valid until the next synchronising call (gtsfm_sync, or the batch calls for the images they name). */
int gtsfm_images_upload(gtsfm_ctx* ctx, int first_image_id, int n_images, const float* const* desc, const int32_t* n, int dim,
                        const double* const* kps, const double* cal);

/* Binary descriptors for TwoWayMatcher(distance_type=HAMMING) (twoway_matcher.py:17-21,95-98): desc (n, nbytes) uint8 rows,
 * nbytes <= 64.  A store holds either float32 or binary descriptors; gtsfm_match_batch picks the distance accordingly. */
int gtsfm_image_upload_u8(gtsfm_ctx* ctx, int image_id, const uint8_t* desc, int n, int nbytes, const double* kps, const double* cal);

/* ---- TwoWayMatcher over a batch of pairs ---------------------------------------------------------------------- */
/* pairs: (P, 2) int32 image ids (i1, i2).  ratio: Lowe ratio threshold, NaN = ratio test off
 * (TwoWayMatcher(ratio_test_threshold=None)).  On return counts[p] = number of matches of pair p and *total their
 * sum; the matches stay on the device until gtsfm_match_fetch / gtsfm_verify_batch. */
int gtsfm_match_batch(gtsfm_ctx* ctx, const int32_t* pairs, int n_pairs, double ratio, int32_t* counts, int64_t* total);
/* Copy the matches of the last gtsfm_match_batch to the host: out is (total, 2) uint32, pair after pair in input
 * order, rows of a pair ordered exactly like TwoWayMatcher.match's return value. */
int gtsfm_match_fetch(gtsfm_ctx* ctx, uint32_t* out);

/* ---- Ransac verifier (essential-matrix mode) over a batch of pairs --------------------------------------------- */
/* Hypothesis h of pair-key k draws 5 distinct match rows in [0, M): the counter-based sampler used on the device,
 * exported so that a CPU oracle can consume the identical (H, 5) index set.  out: (n_hyp, 5) int32. */
int gtsfm_ransac_samples(uint64_t seed, uint64_t pair_key, int n_matches, int n_hyp, int32_t* out);
/* Same sampler with sample_size 5 or 7 (fundamental-matrix mode): out is (n_hyp, sample_size) int32. */
int gtsfm_ransac_samples_k(uint64_t seed, uint64_t pair_key, int n_matches, int n_hyp, int sample_size, int32_t* out);

typedef struct gtsfm_verify_params {
    double threshold_px;    /* Ransac(estimation_threshold_px) */
    int n_hypotheses;       /* 5-point samples per pair (the reference's cv2 default is maxIters = 1000) */
    int lo_iters;           /* Levenberg-Marquardt local-optimisation iterations on the best model (0 = off) */
    uint64_t seed;          /* sampler seed; pair key = index of the pair in the call unless pair_keys is given */
    const uint64_t* pair_keys;   /* optional (P) sampler keys */
    const int32_t* samples;      /* optional explicit samples (P, n_hypotheses, 5) int32 -- (P, n_hypotheses, 7) in the
                                    fundamental-matrix mode -- overrides the sampler */
    int estimate_fundamental;    /* 0: Ransac(use_intrinsics_in_verification=True), essential matrix from 5 points
                                    (ransac.py:52-82).  1: use_intrinsics_in_verification=False, Ransac.estimate_F =
                                    cv2.findFundamentalMat(FM_RANSAC) on pixel coordinates (ransac.py:84-111): 7-point
                                    samples, inlier <=> max squared epipolar-line distance <= threshold_px^2, the model with
                                    the most inliers, fewer than 8 matches -> GTSFM_V_TOO_FEW, E = K2^T F K1
                                    (utils/verification.py:97-110), then the same pose recovery.  lo_iters is ignored. */
    int adaptive_max_hypotheses; /* fundamental-matrix mode only, ignored with explicit samples.  > n_hypotheses: cv2's stopping
                                    rule -- findFundamentalMat(confidence = 0.999999, maxIters = 1000000), ransac.py:22-23,
                                    103-110 -- in rounds: a pair whose best model has inlier ratio w is finished once
                                    log(1 - confidence) / log(1 - w^7) hypotheses were evaluated; otherwise it is run again with
                                    twice the budget, up to min(this, 2^20).  Hypothesis h
                                    draws the same sample whatever the budget.  0: exactly n_hypotheses per pair. */
    double confidence;           /* of that rule; <= 0 or >= 1 selects the reference's 0.999999 (RANSAC_SUCCESS_PROB) */
} gtsfm_verify_params;

/* Verify the matches produced by the last gtsfm_match_batch (same pairs, same order; nothing is copied back to the
 * host in between), or, when match_idx != NULL, caller-supplied putative matches:
 *   match_idx (total, 2) uint32, match_offsets (P + 1) int64 row offsets.
 * Outputs (any may be NULL): R (P, 9) row-major i2Ri1, t (P, 3) unit i2Ui1, n_inliers (P), status (P) GTSFM_V_*,
 * inlier_mask (total) uint8 aligned with the match rows (v_corr_idxs = match rows with mask 1,
 * inlier_ratio = n_inliers / M, opencv_verifier_base.py:98-101), E (P, 9) the final essential matrix (K2^T F K1 in the
 * fundamental-matrix mode). */
int gtsfm_verify_batch(gtsfm_ctx* ctx, const int32_t* pairs, int n_pairs, const uint32_t* match_idx,
                       const int64_t* match_offsets, const gtsfm_verify_params* params, double* R, double* t,
                       int32_t* n_inliers, int32_t* status, uint8_t* inlier_mask, double* E);

/* Hypotheses evaluated per pair by the last gtsfm_verify_batch (n_hypotheses unless the adaptive fundamental-matrix budget
 * raised it): out (n_pairs) int32.  With it a CPU run can draw the identical sample set (gtsfm_ransac_samples_k). */
int gtsfm_verify_hypotheses_used(gtsfm_ctx* ctx, int32_t* out, int n_pairs);

/* ---- post-verification report over a batch of pairs ----------------------------------------- */
typedef struct gtsfm_report_params {
    double dist_threshold_px;           /* TwoViewEstimator(corr_metric_dist_threshold): correct <=> sampson^2 < threshold^2 */
    int min_num_inliers_est_model;      /* InlierSupportProcessor(min_num_inliers_est_model) */
    double min_inlier_ratio_est_model;  /* InlierSupportProcessor(min_inlier_ratio_est_model) */
    const double* gt_cal;               /* optional (P, 10): {f, k1, k2, u0, v0} of the GROUND-TRUTH cameras of image 1 and image 2
                                           (gt_camera.calibration(), utils/metrics.py:90-97) for F_gt; NULL = the calibrations the
                                           images were uploaded with */
    int after_ba;                       /* 0 = the verifier's results (pre-BA report); 1 = the results of the last
                                           gtsfm_two_view_ba_batch (post-BA report: its poses and valid tracks, the inlier RATIO
                                           stays the verifier's, two_view_estimator.py:325-329) */
} gtsfm_report_params;

/* The stage that follows verify() in TwoViewEstimator.run_2view (gtsfm/two_view_estimator.py:298-349), for every pair of
 * the last gtsfm_verify_batch (same order; its match rows, inlier mask and poses are still on the device):
 *   - compute_correspondence_metrics / epipolar_inlier_correspondences (gtsfm/utils/metrics.py:38-128): squared Sampson
 *     distance in pixels of every VERIFIED correspondence against F_gt = K2^-T [t]x R K1^-1 (utils/verification.py:113-126,
 *     170-214), correct <=> distance^2 < dist_threshold_px^2;
 *   - compute_relative_pose_metrics (utils/geometry_comparisons.py:266-309): rotation / translation-direction error in degrees;
 *   - the aggregates of generate_two_view_report (two_view_estimator.py:354-393);
 *   - InlierSupportProcessor.run_inlier_support (frontend/inlier_support_processor.py:39-95): keep[p] = 0 when the inlier
 *     ratio or the inlier count is insufficient (or verification failed).
 * gt_i2Ri1 (P, 9) row-major and gt_i2ti1 (P, 3) are the ground-truth relative poses, both NULL when there is none (then only
 * keep[] is meaningful, the metrics are NaN / 0); a pair whose gt_i2Ri1 row starts with NaN has no ground truth either (one of
 * its gt cameras is None, two_view_estimator.py:131).  Outputs (any may be NULL): dist_sq (total) float64 aligned with the match
 * rows (NaN for rows that were not verified), gt_mask (total) uint8, num_inliers_gt (P), inlier_avg_err / outlier_avg_err (P)
 * mean squared distance of the correct / incorrect verified correspondences (NaN when there are none), R_error_deg /
 * U_error_deg (P), keep (P). */
int gtsfm_two_view_report_batch(gtsfm_ctx* ctx, int n_pairs, const double* gt_i2Ri1, const double* gt_i2ti1,
                                const gtsfm_report_params* params, double* dist_sq, uint8_t* gt_mask, int32_t* num_inliers_gt,
                                double* inlier_avg_err, double* outlier_avg_err, double* R_error_deg, double* U_error_deg,
                                int32_t* keep);

/* ---- two-view triangulation + bundle adjustment over a batch of pairs ---------------------------
 * TwoViewEstimator.bundle_adjust (gtsfm/two_view_estimator.py:137-208), run when bundle_adjust_2view is True -- the default of
 * configs/sift_front_end.yaml:40 -- for every pair of the last gtsfm_verify_batch whose verified correspondences number at least
 * min_num_inliers (two_view_estimator.py:316).  The arithmetic of the reference lives in the un-vendored gtsam==4.2
 * (triangulatePoint3, GeneralSFMFactor2, LevenbergMarquardtOptimizer): PARITY UNPINNED, see DESIGN.md. */
typedef struct gtsfm_two_view_ba_params {
    int min_num_inliers;                       /* processor._min_num_inliers_est_model: fewer verified rows -> the pair is skipped */
    int max_iterations;                        /* TwoViewEstimator(bundle_adjust_2view_maxiters), default 100 */
    double triangulation_reproj_error_thresh;  /* TriangulationOptions.reproj_error_threshold (point3d_initializer.py:54), pixels */
    double ba_reproj_error_thresh;             /* ba_reproj_error_thresholds[-1] (bundle_adjustment.py:335-357), pixels */
    double huber_k;                            /* gtsam.noiseModel.mEstimator.Huber(1.345) (bundle_adjustment.py:137-141) */
    double min_triangulation_angle_deg;        /* TriangulationOptions.min_triangulation_angle (point3d_initializer.py:287-291), 0 = off */
} gtsfm_two_view_ba_params;

/* Per pair: cameras (I, K1) and (i2Ti1 from the verifier, K2); every verified correspondence is triangulated (DLT + three
 * Gauss-Newton steps on the pixel error, cheirality, both reprojection errors < triangulation_reproj_error_thresh); the pose and
 * the points are refined by Levenberg-Marquardt on the Huber-robust reprojection cost (lambda 1e-5, factor 10, upper bound 1e5,
 * relative / absolute error tolerance 1e-5, at most max_iterations accepted steps; camera 1 fixed, translation on the unit
 * sphere, intrinsics fixed); a track stays valid when both reprojection errors are < ba_reproj_error_thresh.
 * Outputs (any may be NULL), aligned with the last gtsfm_verify_batch: i2Ri1 (P, 9) row-major, i2Ui1 (P, 3) unit vectors,
 * valid_mask (total match rows) uint8 = post_ba_v_corr_idxs as a mask, num_valid (P), ba_status (P): 0 = BA ran, 1 = skipped
 * (verification failed or too few verified rows: pose / mask / count are the verifier's, two_view_estimator.py:334-338), 2 =
 * nothing could be triangulated (the verifier's pose, no correspondences, two_view_estimator.py:179-180), 3 = BA ran and no
 * track passed the filter (the verifier's pose, no correspondences, two_view_estimator.py:199-202); iterations (P) accepted LM
 * steps.  The results stay on the device for gtsfm_two_view_report_batch(after_ba = 1). */
int gtsfm_two_view_ba_batch(gtsfm_ctx* ctx, int n_pairs, const gtsfm_two_view_ba_params* params, double* i2Ri1, double* i2Ui1,
                            uint8_t* valid_mask, int32_t* num_valid, int32_t* ba_status, int32_t* iterations);

/* ---- device-resident result records for the multi-GPU exchange ------------------------------------------
 * What unzip_two_view_results hands to the back end (gtsfm/runner/gtsfm_runner_base.py:415-440) for every pair of the last
 * gtsfm_verify_batch, packed ON THE DEVICE as one block [headers | payload] that a rank contributes to an NCCL all-gather:
 *   header (128 bytes per pair, little endian): int64 pair_id; int32 n_matches, n_inliers, status, pad; double R[9] (i2Ri1,
 *     row-major); double t[3] (i2Ui1); double inlier_ratio (opencv_verifier_base.py:101).  Failed pairs: n_inliers 0, R = t = 0.
 *   payload: the verified correspondences v_corr_idxs (match rows with inlier mask 1, input order, opencv_verifier_base.py:
 *     98-100) of pair 0, pair 1, ... back to back as (i1, i2) uint32 rows; pair p's rows start at the exclusive prefix sum of
 *     n_inliers (the offset table is rebuilt from the headers).
 * gtsfm_result_pack: builds the headers and the offset table; pair_ids (n_pairs) are the caller's global pair ids or NULL
 *   (= 0..n_pairs-1); *n_payload_rows = total number of payload rows (the one number a rank has to announce before the
 *   exchange).
 * gtsfm_result_pack_copy: writes the headers at d_dst and the payload at d_dst + payload_offset_bytes (>= 128 * n_pairs, a
 *   multiple of 8) into a DEVICE buffer of dst_bytes owned by the caller (e.g. a torch tensor); enqueued on the context's
 *   stream, returns without synchronising.
 * gtsfm_stream_wait: orders the context's stream against a caller's CUDA stream without a host synchronisation: direction 0 =
 *   the caller's stream waits for everything the context has enqueued, 1 = the context's stream waits for the caller's. */
int gtsfm_result_pack(gtsfm_ctx* ctx, int n_pairs, const int64_t* pair_ids, int64_t* n_payload_rows);
int gtsfm_result_pack_copy(gtsfm_ctx* ctx, void* d_dst, int64_t dst_bytes, int64_t payload_offset_bytes);
int gtsfm_stream_wait(gtsfm_ctx* ctx, void* cuda_stream, int direction);

/* ---- the retriever step in front of the matcher ---------------------------------------------------
 * Replaces NetVLADRetriever.compute_similarity_matrix (gtsfm/retriever/netvlad_retriever.py:77-149: blocked
 * torch.einsum("id,jd->ij") of the global descriptors) and pairs_from_score_matrix (:196-228: mask + min_score + torch.topk).
 *
 * gtsfm_retriever_similarity: global_desc (n_images, dim) float32 row-major on the host -> sim = G G^T, fp32 accumulation in
 * ascending k (two levels: spans of 128, then the spans).  Like the reference only the blocks with block(j) >= block(i)
 * (block = index / blocksize) are computed, all other entries are 0.  The matrix stays on the device for
 * gtsfm_retriever_topk; sim_out (n_images^2 floats, may be NULL) receives a copy.  More than 10,000 images -> GTSFM_E_ARG
 * (MAX_NUM_IMAGES, netvlad_retriever.py:21,91-92).
 * gtsfm_retriever_topk: for every image i the num_select = min(num_select, n_images) best j > i (upper triangle, no
 * self-matches, netvlad_retriever.py:166-168) with !(score < (float)min_score) and finite score, best first, ties by lower j.
 * min_score NaN = no minimum.  sim_in NULL = the matrix of the last gtsfm_retriever_similarity call, else an (n_images,
 * n_images) float32 host matrix.  out_idx / out_score: (n_images, num_select) row-major, unused slots -1 / -inf. */
int gtsfm_retriever_similarity(gtsfm_ctx* ctx, const float* global_desc, int n_images, int dim, int blocksize, float* sim_out);
int gtsfm_retriever_topk(gtsfm_ctx* ctx, const float* sim_in, int n_images, int num_select, double min_score, int32_t* out_idx,
                         float* out_score);

/* ---- allocation-free retriever step ---------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * gtsfm_retriever_similarity + gtsfm_retriever_topk in one call that allocates nothing: the (n_images, n_images) similarity
 * matrix lives in a fixed __device__ array inside the library (GTSFM_RETRIEVER_NOALLOC_MAX_IMAGES^2 floats = 4 MiB, held by
 * every GPU that loads the library until the process exits).  d_global_desc (n_images, dim) float32, d_out_idx / d_out_score
 * (n_images, num_select) are DEVICE pointers of the current device, everything is enqueued on cuda_stream (a cudaStream_t,
 * NULL = the legacy default stream) and the call returns without synchronising; results (same values as the two regular
 * calls: unused slots -1 / -inf, min_score NaN = no minimum) are ready once the stream reaches that point.  n_images above the
 * maximum, or num_select > n_images, -> GTSFM_E_ARG with the text in gtsfm_last_error(NULL).  No context is involved.  All
 * calls on one device share the array, which keeps the previous call's contents (the part a call uses is cleared first):
 * calls that overlap on the GPU -- different streams without an ordering between them -- race.  The first call on each
 * device looks up the array's address (cudaGetSymbolAddress): make that call outside stream capture. */
#define GTSFM_RETRIEVER_NOALLOC_MAX_IMAGES 1024
int gtsfm_retriever_pairs_noalloc(const float* d_global_desc, int n_images, int dim, int blocksize, int num_select, double min_score,
                                  int32_t* d_out_idx, float* d_out_score, void* cuda_stream);

/* ---- single-pair conveniences used by the MatcherBase / VerifierBase plugin classes ---------------------------- */
/* TwoWayMatcher.match for one pair from host arrays.  out capacity: min(n1, n2) rows of 2 uint32. */
int gtsfm_match_pair(gtsfm_ctx* ctx, const float* desc1, int n1, const float* desc2, int n2, int dim, double ratio,
                     uint32_t* out, int32_t* count);
/* TwoWayMatcher(distance_type=HAMMING).match for one pair of uint8 descriptor arrays. */
int gtsfm_match_pair_u8(gtsfm_ctx* ctx, const uint8_t* desc1, int n1, const uint8_t* desc2, int n2, int nbytes, double ratio,
                        uint32_t* out, int32_t* count);
/* OpencvVerifierBase.verify for one pair from host arrays (kps in pixels, cal = {f,k1,k2,u0,v0}). */
int gtsfm_verify_pair(gtsfm_ctx* ctx, const double* kps1, int n1, const double* kps2, int n2, const uint32_t* match_idx,
                      int n_matches, const double* cal1, const double* cal2, const gtsfm_verify_params* params, double* R,
                      double* t, int32_t* n_inliers, int32_t* status, uint8_t* inlier_mask, double* E);

/* ---- diagnostics (tests and tools only) --------------------------------------------------------------------------- */
/* Pinned (page-locked) host memory for result buffers: device -> host copies into it run at full PCIe speed and
 * asynchronously.  Any host pointer is accepted by the fetch / verify calls; pageable ones just copy slower. */
int gtsfm_host_alloc(void** ptr, int64_t bytes);
void gtsfm_host_free(void* ptr);
/* Like gtsfm_match_fetch, but the copy is only enqueued (on a second stream, ordered after the match kernels) so that it
 * overlaps gtsfm_verify_batch; `out` must stay valid and must not be read until gtsfm_sync() returns. */
int gtsfm_match_fetch_async(gtsfm_ctx* ctx, uint32_t* out);
/* Wait for everything this context has enqueued (kernels and asynchronous copies). */
int gtsfm_sync(gtsfm_ctx* ctx);

/* Running totals of the verifier since the last reset, out[4]: [0] essential-matrix candidates scored (real roots of
 * all 5-point solves, the S_h of SURVEY.md 8(d)), [1] hypotheses solved, [2] algorithmic Sampson evaluations (sum over
 * pairs of candidates x correspondences), [3] evaluations actually executed (the scoring kernel drops a candidate once
 * its partial cost exceeds the pair's best complete cost).  reset != 0 zeroes them afterwards. */
int gtsfm_verify_stats(gtsfm_ctx* ctx, int64_t* out, int reset);

/* Copy an internal device buffer of the last chunk to the host.  what: 0 rp_v1 (f32), 1 rp_v2 (f32), 2 rp_g (u8),
 * 3 colraw (f32), 4 row_best (i32), 5 row_d1 (f32), 6 row_d2 (f32), 7 col_best (i32), 8 col_d1 (f32), 9 col_d2 (f32),
 * 10 number of rows / columns flagged for the exact re-scan (i32). */
int gtsfm_debug_fetch(gtsfm_ctx* ctx, int what, void* dst, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* GTSFM_B200_H */
