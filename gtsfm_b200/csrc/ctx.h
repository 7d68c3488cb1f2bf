// Host-side context: descriptor / keypoint store, work buffers, TMA tensor maps.  Internal to the library.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/gtsfm_b200.h"
#include "store.h"

namespace g2v {

// A device buffer that only ever grows (contents are NOT preserved on growth unless keep = true).
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes, bool keep = false, cudaStream_t s = nullptr);
    void release();
    template <typename T> T* as() const { return static_cast<T*>(p); }
};

struct PinnedBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes);
    void release();
    template <typename T> T* as() const { return static_cast<T*>(p); }
};

}  // namespace g2v

struct gtsfm_ctx {
    int device = 0;
    int n_sm = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;     // device -> host copies that overlap the next kernels (gtsfm_match_fetch_async)
    cudaEvent_t match_done = nullptr;       // recorded on `stream` when the compacted matches of a batch are complete
    cudaEvent_t upload_ev = nullptr;        // store_clear: copy_stream waits for the kernels that may still read the old rows
    // Image uploads (H2D copy + prepare kernel) run on copy_stream; img_ev[i] is recorded behind image i.  copy_stream is in
    // order, so a batch waits for the event of the LAST image it names and starts while later images are still arriving.
    std::vector<cudaEvent_t> img_ev;
    int wait_images(int max_image, bool host);   // `stream` (and the host) wait for the uploads of images 0..max_image
    std::string err;
    int64_t launches = 0;
    int64_t workspace_bytes = 12ll << 30;

    // ---- store
    int dim = 0, kpad = 0;
    int binary = -1;                    // -1 no descriptors yet, 0 float32 descriptors (EUCLIDEAN), 1 uint8 rows (HAMMING)
    long long rows_used = 0, rows_cap = 0;
    std::vector<g2v::ImageRec> images;
    bool flags_dirty = false;
    int flags_synced = 0;               // images whose prepare-kernel flags have been read back
    g2v::DevBuf desc_bf16, desc_u8, desc_f32, augq, augt, norm_i, valid, kps, img_flags, stage;
    g2v::DevBuf desc_h16, augq_h, augt_h, norm_f;     // float path: fp16 operands of the screening GEMM, float32 squared norms
    CUtensorMap tm_desc{}, tm_augq{}, tm_augt{};
    CUtensorMap tm_desc_h{}, tm_augq_h{}, tm_augt_h{};
    bool maps_valid = false;
    bool force_exact_float = false;     // diagnostics: float descriptors always take the exhaustive fp32 kernels

    // ---- matcher work buffers (one chunk of pairs) and batch outputs
    g2v::DevBuf d_pairs, rp_v1, rp_v2, rp_g, colraw, row_best, row_d1, row_d2, col_best, col_d1, col_d2, flag_list, n_flag, row_key, col_key, row_cnt, col_cnt, row_list, col_list, flag_scratch, col_cells;
    g2v::DevBuf keys_scratch, match_slot, match_count, match_offsets, match_out, d_total;
    g2v::PinnedBuf h_pairs, h_small, h_blk_tab;
    g2v::DevBuf blk_tab;                        // float screening path: (row block -> pair) and (column block -> pair) tables
    std::vector<g2v::PairDesc> batch_pairs;     // PairDescs of the last match batch (chunk-relative bases)
    std::vector<int> batch_img;                 // (P, 2) image ids of the last match batch
    std::vector<int> batch_counts;              // matches per pair of the last match batch
    int64_t batch_total = 0;
    int batch_n = 0;

    // ---- per-kernel device timing (CUDA events on this context's stream) and user events
    bool prof_on = false;
    std::vector<cudaEvent_t> prof_pool;          // recycled events
    size_t prof_used = 0;
    struct ProfSpan { int slot; cudaEvent_t a, b; };
    std::vector<ProfSpan> prof_spans;
    double prof_ms[16] = {0};
    int64_t prof_calls[16] = {0};
    cudaEvent_t user_ev[8] = {nullptr};
    cudaEvent_t prof_event();
    void prof_begin(int slot);
    void prof_end();
    int prof_collect();

    // ---- verifier buffers; state of the last gtsfm_verify_batch (read by gtsfm_two_view_report_batch)
    g2v::DevBuf v_bufs[32];
    int verify_n = 0;
    int ba_n = 0;                     // pairs of the last two_view_ba_batch (0 = none for the current verify batch)
    int pack_n = 0;                          // pairs / payload rows of the last gtsfm_result_pack
    long long pack_rows = 0;
    cudaEvent_t fetch_ev = nullptr;          // recorded behind the copy of gtsfm_match_fetch_async (which reads match_out)
    bool fetch_pending = false;
    int wait_fetch();                        // host waits for that copy only (not for the uploads queued behind it)
    cudaEvent_t xs_ev = nullptr;             // gtsfm_stream_wait: ordering against a caller's stream
    std::vector<int> verify_hyp;            // hypotheses evaluated per pair of the last verify batch (adaptive F mode)
    long long verify_total = 0;

    // ---- retriever (retrieve.cu): global descriptors, similarity matrix of the last gtsfm_retriever_similarity, top-K outputs
    g2v::DevBuf r_desc, r_sim, r_tiles, r_out;
    int r_n = 0;

    g2v::StoreView view() const;
};

namespace g2v {
int fail(gtsfm_ctx* c, int code, const std::string& msg);
int cuda_fail(gtsfm_ctx* c, cudaError_t e, const char* what);
#define G2V_CUDA(ctx, call)                                                 \
    do {                                                                    \
        cudaError_t e__ = (call);                                           \
        if (e__ != cudaSuccess) return g2v::cuda_fail((ctx), e__, #call);   \
    } while (0)

int store_reserve_rows(gtsfm_ctx* c, long long rows);
int store_sync_flags(gtsfm_ctx* c, int max_image = -1);
enum ProfSlot { P_PREPARE = 0, P_SCREEN, P_RESOLVE_ROWS, P_RESOLVE_COLS, P_EXACT_SCAN, P_MATCH_FINALIZE, P_GATHER, P_SOLVE, P_SCORE, P_VFINALIZE, P_NSLOTS };
}  // namespace g2v
