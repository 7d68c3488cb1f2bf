// Host orchestration of the batched TwoWayMatcher (C ABI: gtsfm_match_batch / gtsfm_match_fetch / gtsfm_match_pair).
// Replaces the one-Dask-task-per-pair loop of
// gtsfm/frontend/correspondence_generator/det_desc_correspondence_generator.py:71-81 with chunks of pairs that run as
// a handful of kernels on one stream; nothing returns to the host between the kernels of a chunk.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "ctx.h"

namespace g2v {
cudaError_t launch_match_screen(const CUtensorMap& tm_desc, const CUtensorMap& tm_augq, const CUtensorMap& tm_augt,
                                const PairDesc* pairs, int n_pairs, int n_items, float* rp_v1, float* rp_v2, uint8_t* rp_g,
                                float* colraw, int kpad, int n_sm, cudaStream_t stream);
cudaError_t launch_match_screen_f(const CUtensorMap& tm_desc, const CUtensorMap& tm_augq, const CUtensorMap& tm_augt,
                                  const PairDesc* pairs, int n_pairs, int n_items, float* rp_v1, float* rp_v2, uint8_t* rp_g,
                                  float* colraw, int kpad, int n_sm, cudaStream_t stream);
cudaError_t launch_resolve_f(const StoreView& st, const MatchWork& w, long long total_rows, long long total_cols, int use_ratio, double ratio,
                             cudaStream_t s);
cudaError_t launch_hamming(const StoreView& st, const MatchWork& w, long long total_rows, long long total_cols, int words, cudaStream_t s);
cudaError_t launch_resolve_int(const StoreView& st, const MatchWork& w, int which, long long total, int max_class_rows,
                               cudaStream_t s);
cudaError_t launch_exact_scan(const StoreView& st, const MatchWork& w, int mode, int n_jobs, cudaStream_t s);
cudaError_t launch_exact_tile(const StoreView& st, const MatchWork& w, int n_jobs, long long total_cols, cudaStream_t s);
bool exact_tile_supported(const StoreView& st);
cudaError_t launch_match_finalize(const MatchWork& w, int use_ratio, double ratio, unsigned long long* keys_scratch,
                                  uint2* match_slot, int* match_count, long long* offsets, long long* running, uint2* compact_out,
                                  cudaStream_t s);
int store_ensure_maps(gtsfm_ctx* c);

// Flagged rows / columns per chunk that the exact re-scan can take (merge cells of 16 B each).  A chunk is cut short
// before its rows + columns exceed this, so the flag list can never overflow.
constexpr long long kFlagScratchCap = 1ll << 23;   // also caps a chunk at ~512 C2 pairs: larger chunks measured 1-3 % slower

struct Chunk {
    int p0 = 0, p1 = 0;
    long long rp = 0, cr = 0, rows = 0, cols = 0;
    int items = 0, jobs = 0, max_class_rows = 0, max_hb = 0;
    int fitems = 0;          // work items of the fp16 screening kernel (path-2 pairs)
    long long rtab = 0, ctab = 0;   // offsets of the chunk's block -> pair tables
    int hpairs = 0;                 // binary (HAMMING) pairs
};
}  // namespace g2v

using namespace g2v;

extern "C" {

int gtsfm_match_batch(gtsfm_ctx* c, const int32_t* pairs, int n_pairs, double ratio, int32_t* counts, int64_t* total)
{
    if (!c || n_pairs < 0 || (n_pairs > 0 && !pairs)) return fail(c, GTSFM_E_ARG, "bad pair list");
    cudaSetDevice(c->device);
    const int use_ratio = std::isnan(ratio) ? 0 : 1;
    const int n_img_all = static_cast<int>(c->images.size());
    int max_image = -1;
    for (int p = 0; p < n_pairs; ++p) {
        const int a = pairs[2 * p], b = pairs[2 * p + 1];
        if (a < 0 || b < 0 || a >= n_img_all || b >= n_img_all) return fail(c, GTSFM_E_ARG, "pair " + std::to_string(p) + " names an image that was not uploaded");
        max_image = std::max(max_image, std::max(a, b));
    }
    // uploads run on the copy stream: wait (stream and host) for the last image this batch names, not for later ones
    int r = store_sync_flags(c, max_image);
    if (r) return r;
    if ((r = c->wait_images(max_image, false))) return r;
    c->batch_n = 0;
    c->batch_total = 0;
    c->verify_n = 0; c->ba_n = 0;          // the device match rows of the last verify batch are about to be overwritten
    if (total) *total = 0;
    if ((r = c->wait_fetch())) return r;                     // an asynchronous fetch of the previous batch may still read match_out
    if (n_pairs == 0) return GTSFM_OK;
    const int n_img = static_cast<int>(c->images.size());

    // ---- pair descriptors and chunking
    std::vector<PairDesc>& pds = c->batch_pairs;
    pds.assign(n_pairs, PairDesc{});
    c->batch_img.assign(pairs, pairs + 2 * static_cast<size_t>(n_pairs));
    std::vector<Chunk> chunks;
    Chunk cur;
    long long cur_bytes = 0, out_cap = 0;
    const bool screen_ok = c->kpad == 64 || c->kpad == 128;
    // float descriptors: fp16 tensor-core screening, with or without the ratio test (the resolve step certifies or recomputes
    // the second-nearest distance); a non-positive ratio keeps nothing worth screening for and takes the exhaustive route
    const bool screen_f_ok = (!use_ratio || ratio > 0) && (c->kpad == 64 || c->kpad == 128 || c->kpad == 256) && (c->dim % 4) == 0 &&
                             !c->force_exact_float;
    for (int p = 0; p < n_pairs; ++p) {
        const int a = pairs[2 * p], b = pairs[2 * p + 1];
        if (a < 0 || b < 0 || a >= n_img || b >= n_img) return fail(c, GTSFM_E_ARG, "pair " + std::to_string(p) + " names an image that was not uploaded");
        const ImageRec& qa = c->images[a];
        const ImageRec& tb = c->images[b];
        PairDesc pd{};
        pd.img_q = a; pd.img_t = b;
        pd.q_row0 = qa.row0; pd.t_row0 = tb.row0;
        // twoway_matcher.py:71-72,80-81: an empty side (or a side with only NaN rows) yields no matches
        const bool empty = qa.n == 0 || tb.n == 0 || qa.n_valid == 0 || tb.n_valid == 0;
        pd.nq = empty ? 0 : qa.n; pd.nt = empty ? 0 : tb.n;
        if (!empty && use_ratio && (qa.n_valid < 2 || tb.n_valid < 2))
            return fail(c, GTSFM_E_KNN_UNPACK, "pair " + std::to_string(p) + ": ratio test with fewer than 2 valid descriptors on one side "
                                               "(the reference raises ValueError: not enough values to unpack)");
        pd.nq_pad = (pd.nq + kRowPad - 1) / kRowPad * kRowPad;
        pd.nt_pad = (pd.nt + kRowPad - 1) / kRowPad * kRowPad;
        // the bucketed resolve kernels keep one lane class (nq_pad / 128 rows) in shared memory and use 13-bit half-block ids
        const bool fits = pd.nq_pad / kTileQ <= kMaxClassRows && pd.nt_pad / kHalfT < 8192;
        pd.path = (qa.path == 0 && tb.path == 0 && screen_ok && fits) ? 0 : 1;
        if (pd.path == 1 && !empty && qa.path == 1 && tb.path == 1 && qa.f16_ok && tb.f16_ok && screen_f_ok) pd.path = 2;
        if (qa.path == 3 || tb.path == 3) pd.path = 3;          // a store holds either binary or float32 descriptors
        pd.q_maxn2 = qa.max_norm2; pd.t_maxn2 = tb.max_norm2;
        const long long rp = (pd.path == 0 || pd.path == 2) ? static_cast<long long>(pd.nq_pad) * (pd.nt_pad / kHalfT) : 0;
        const int n_tg = (pd.nq_pad / kTileQ + kFlushTiles - 1) / kFlushTiles;      // path 2: column minima per group of 4 query tiles
        const long long cr = pd.path == 0 ? static_cast<long long>(pd.nt_pad) * 4       // folded column minima: v1, v2, class (+ pad)
                                          : pd.path == 2 ? static_cast<long long>(pd.nt_pad) * n_tg * kTileQ : 0;
        const int items = pd.path == 0 && pd.nq > 0 ? pd.nt_pad / kBlockT : 0;
        const int fitems = pd.path == 2 && pd.nq > 0 ? pd.nt_pad / kBlockT : 0;
        const int jobs = pd.path == 1 && pd.nq > 0 ? (pd.nq + 31) / 32 + (pd.nt + 31) / 32 : 0;
        const long long bytes = rp * 9 + cr * 4 + static_cast<long long>(pd.nq_pad) * (12 + 16 + 8 + 8 + 8) + static_cast<long long>(pd.nt_pad) * (12 + 8 + 8);
        if (cur.p1 > cur.p0 && (cur_bytes + bytes > c->workspace_bytes || cur.rows + cur.cols + pd.nq_pad + pd.nt_pad > kFlagScratchCap)) { chunks.push_back(cur); cur = Chunk{}; cur.p0 = cur.p1 = p; cur_bytes = 0; }
        pd.rp_base = cur.rp; pd.cr_base = cur.cr; pd.out_base = cur.rows; pd.cb_base = cur.cols; pd.item0 = cur.items; pd.job0 = cur.jobs; pd.fitem0 = cur.fitems;
        cur.fitems += fitems;
        if (pd.path == 3 && pd.nq > 0) cur.hpairs += 1;       // after the split test: the pair belongs to the chunk it lands in
        cur.rp += rp; cur.cr += cr; cur.rows += pd.nq_pad; cur.cols += pd.nt_pad; cur.items += items; cur.jobs += jobs;
        cur.p1 = p + 1;
        if (items > 0) { cur.max_class_rows = std::max(cur.max_class_rows, pd.nq_pad / kTileQ); cur.max_hb = std::max(cur.max_hb, pd.nt_pad / kHalfT); }
        cur_bytes += bytes;
        out_cap += pd.nq < pd.nt ? pd.nq : pd.nt;
        pds[p] = pd;
    }
    chunks.push_back(cur);

    // ---- buffers
    long long m_rp = 0, m_cr = 0, m_rows = 0, m_cols = 0, m_rcnt = 0, m_ccnt = 0;
    for (const Chunk& k : chunks) {
        m_rcnt = std::max(m_rcnt, static_cast<long long>(k.p1 - k.p0) * (8 * k.max_hb + 1));
        m_ccnt = std::max(m_ccnt, static_cast<long long>(k.p1 - k.p0) * (kTileQ + 1));
    }
    for (const Chunk& k : chunks) { m_rp = std::max(m_rp, k.rp); m_cr = std::max(m_cr, k.cr); m_rows = std::max(m_rows, k.rows); m_cols = std::max(m_cols, k.cols); }
    cudaStream_t s = c->stream;
    G2V_CUDA(c, c->d_pairs.reserve(sizeof(PairDesc) * n_pairs, false, s));
    G2V_CUDA(c, c->rp_v1.reserve(4 * (m_rp + 1), false, s));
    G2V_CUDA(c, c->rp_v2.reserve(4 * (m_rp + 1), false, s));
    G2V_CUDA(c, c->rp_g.reserve(m_rp + 1, false, s));
    G2V_CUDA(c, c->colraw.reserve(4 * (m_cr + 1), false, s));
    G2V_CUDA(c, c->row_best.reserve(4 * (m_rows + 1), false, s));
    G2V_CUDA(c, c->row_d1.reserve(4 * (m_rows + 1), false, s));
    G2V_CUDA(c, c->row_d2.reserve(4 * (m_rows + 1), false, s));
    G2V_CUDA(c, c->col_best.reserve(4 * (m_cols + 1), false, s));
    G2V_CUDA(c, c->col_d1.reserve(4 * (m_cols + 1), false, s));
    G2V_CUDA(c, c->col_d2.reserve(4 * (m_cols + 1), false, s));
    G2V_CUDA(c, c->row_key.reserve(4 * (m_rows + 16), false, s));
    G2V_CUDA(c, c->col_key.reserve(4 * (m_cols + 16), false, s));
    G2V_CUDA(c, c->row_list.reserve(4 * (m_rows + 16), false, s));
    G2V_CUDA(c, c->col_list.reserve(4 * (m_cols + 16), false, s));
    G2V_CUDA(c, c->row_cnt.reserve(4 * (m_rcnt + 16), false, s));
    G2V_CUDA(c, c->col_cnt.reserve(4 * (m_ccnt + 16), false, s));
    G2V_CUDA(c, c->flag_list.reserve(8 * (m_rows + m_cols + 1), false, s));
    G2V_CUDA(c, c->n_flag.reserve(sizeof(int), false, s));
    bool any_float = false;
    for (const Chunk& k : chunks) any_float |= k.jobs > 0;
    if (any_float) G2V_CUDA(c, c->col_cells.reserve(16 * static_cast<size_t>(m_cols + 1), false, s));
    G2V_CUDA(c, c->flag_scratch.reserve(16 * static_cast<size_t>(kFlagScratchCap), false, s));
    G2V_CUDA(c, c->keys_scratch.reserve(16 * (m_rows + 1), false, s));
    G2V_CUDA(c, c->match_slot.reserve(8 * (m_rows + 1), false, s));
    G2V_CUDA(c, c->match_count.reserve(4 * static_cast<size_t>(n_pairs), false, s));
    G2V_CUDA(c, c->match_offsets.reserve(8 * (static_cast<size_t>(n_pairs) + 1), false, s));
    G2V_CUDA(c, c->match_out.reserve(8 * static_cast<size_t>(out_cap + 1), false, s));
    G2V_CUDA(c, c->d_total.reserve(8, false, s));
    G2V_CUDA(c, c->h_pairs.reserve(sizeof(PairDesc) * n_pairs));
    std::memcpy(c->h_pairs.p, pds.data(), sizeof(PairDesc) * n_pairs);
    G2V_CUDA(c, cudaMemcpyAsync(c->d_pairs.p, c->h_pairs.p, sizeof(PairDesc) * n_pairs, cudaMemcpyHostToDevice, s));
    G2V_CUDA(c, cudaMemsetAsync(c->d_total.p, 0, 8, s));
    if ((r = store_ensure_maps(c))) return r;
    // block -> pair tables of the float screening path (one entry per kRowPad rows / columns, chunk-relative pair index)
    bool any_screen_f = false;
    for (const Chunk& k : chunks) any_screen_f |= k.fitems > 0 || k.hpairs > 0;
    if (any_screen_f) {
        std::vector<int> tab;
        for (Chunk& k : chunks) {
            k.rtab = static_cast<long long>(tab.size());
            for (int p = k.p0; p < k.p1; ++p) tab.insert(tab.end(), pds[p].nq_pad / kRowPad, p - k.p0);
            k.ctab = static_cast<long long>(tab.size());
            for (int p = k.p0; p < k.p1; ++p) tab.insert(tab.end(), pds[p].nt_pad / kRowPad, p - k.p0);
        }
        G2V_CUDA(c, c->blk_tab.reserve(4 * (tab.size() + 1), false, s));
        G2V_CUDA(c, c->h_blk_tab.reserve(4 * (tab.size() + 1)));
        std::memcpy(c->h_blk_tab.p, tab.data(), 4 * tab.size());
        G2V_CUDA(c, cudaMemcpyAsync(c->blk_tab.p, c->h_blk_tab.p, 4 * tab.size(), cudaMemcpyHostToDevice, s));
    }

    const StoreView st = c->view();
    for (const Chunk& k : chunks) {
        MatchWork w{};
        w.pairs = c->d_pairs.as<PairDesc>() + k.p0;
        w.n_pairs = k.p1 - k.p0;
        w.rp_v1 = c->rp_v1.as<float>(); w.rp_v2 = c->rp_v2.as<float>(); w.rp_g = c->rp_g.as<uint8_t>(); w.colraw = c->colraw.as<float>();
        w.row_best = c->row_best.as<int>(); w.row_d1 = c->row_d1.as<float>(); w.row_d2 = c->row_d2.as<float>();
        w.col_best = c->col_best.as<int>(); w.col_d1 = c->col_d1.as<float>(); w.col_d2 = c->col_d2.as<float>();
        w.row_key = c->row_key.as<unsigned int>(); w.col_key = c->col_key.as<unsigned int>();
        w.row_cnt = c->row_cnt.as<int>(); w.col_cnt = c->col_cnt.as<int>(); w.row_list = c->row_list.as<int>(); w.col_list = c->col_list.as<int>();
        w.max_hb = k.max_hb;
        w.row_blk_pair = c->blk_tab.as<int>() + k.rtab; w.col_blk_pair = c->blk_tab.as<int>() + k.ctab;
        w.flag_list = c->flag_list.as<unsigned long long>(); w.n_flag = c->n_flag.as<int>();
        w.flag_scratch = c->flag_scratch.as<unsigned long long>();
        w.col_cells = c->col_cells.as<unsigned long long>();
        w.flag_cap = static_cast<int>(std::min<long long>(k.rows + k.cols, kFlagScratchCap));
        G2V_CUDA(c, cudaMemsetAsync(c->n_flag.p, 0, sizeof(int), s));
        G2V_CUDA(c, cudaMemsetAsync(c->flag_scratch.p, 0xff, 16 * static_cast<size_t>(w.flag_cap), s));
        if (k.items > 0) {
            c->prof_begin(P_SCREEN);
            G2V_CUDA(c, launch_match_screen(c->tm_desc, c->tm_augq, c->tm_augt, w.pairs, w.n_pairs, k.items, c->rp_v1.as<float>(),
                                            c->rp_v2.as<float>(), c->rp_g.as<uint8_t>(), c->colraw.as<float>(), c->kpad, c->n_sm, s));
            c->prof_end();
            c->prof_begin(P_RESOLVE_ROWS);
            G2V_CUDA(c, launch_resolve_int(st, w, 0, k.rows, k.max_class_rows, s));
            c->prof_end();
            c->prof_begin(P_RESOLVE_COLS);
            G2V_CUDA(c, launch_resolve_int(st, w, 1, k.cols, k.max_class_rows, s));
            c->prof_end();
            c->launches += 9;
        }
        if (k.fitems > 0) {
            c->prof_begin(P_SCREEN);
            G2V_CUDA(c, launch_match_screen_f(c->tm_desc_h, c->tm_augq_h, c->tm_augt_h, w.pairs, w.n_pairs, k.fitems, c->rp_v1.as<float>(),
                                              c->rp_v2.as<float>(), c->rp_g.as<uint8_t>(), c->colraw.as<float>(), c->kpad, c->n_sm, s));
            c->prof_end();
            c->prof_begin(P_RESOLVE_ROWS);
            G2V_CUDA(c, launch_resolve_f(st, w, k.rows, k.cols, use_ratio, ratio, s));
            c->prof_end();
            c->launches += 3;
        }
        if (k.items > 0 || k.fitems > 0) {
            // rows / columns the resolve kernels could not decide: exhaustive fp32 scan in cv2's summation order
            c->prof_begin(P_EXACT_SCAN);
            G2V_CUDA(c, launch_exact_scan(st, w, 1, w.flag_cap, s));
            c->prof_end();
            c->launches += 1;
        }
        if (k.hpairs > 0) {
            c->prof_begin(P_EXACT_SCAN);
            G2V_CUDA(c, launch_hamming(st, w, k.rows, k.cols, c->dim <= 32 ? 8 : 16, s));
            c->prof_end();
            c->launches += 2;
        }
        if (k.jobs > 0) {
            c->prof_begin(P_EXACT_SCAN);
            if (exact_tile_supported(st)) { G2V_CUDA(c, launch_exact_tile(st, w, k.jobs, k.cols, s)); c->launches += 1; }
            else G2V_CUDA(c, launch_exact_scan(st, w, 0, k.jobs, s));
            c->prof_end();
            c->launches += 1;
        }
        c->prof_begin(P_MATCH_FINALIZE);
        G2V_CUDA(c, launch_match_finalize(w, use_ratio, ratio, c->keys_scratch.as<unsigned long long>(), c->match_slot.as<uint2>(),
                                          c->match_count.as<int>() + k.p0, c->match_offsets.as<long long>() + k.p0,
                                          c->d_total.as<long long>(), c->match_out.as<uint2>(), s));
        c->prof_end();
        c->launches += 3;
    }
    G2V_CUDA(c, cudaEventRecord(c->match_done, s));
    // ---- counts back to the host (the only synchronisation of the batch)
    G2V_CUDA(c, c->h_small.reserve(4 * static_cast<size_t>(n_pairs) + 16));
    G2V_CUDA(c, cudaMemcpyAsync(c->h_small.p, c->match_count.p, 4 * static_cast<size_t>(n_pairs), cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaStreamSynchronize(s));
    long long sum = 0;
    const int* hc = c->h_small.as<int>();
    c->batch_counts.assign(hc, hc + n_pairs);
    for (int p = 0; p < n_pairs; ++p) { if (counts) counts[p] = hc[p]; sum += hc[p]; }
    c->batch_n = n_pairs;
    c->batch_total = sum;
    if (total) *total = sum;
    return GTSFM_OK;
}

int gtsfm_match_fetch(gtsfm_ctx* c, uint32_t* out)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    if (c->batch_total == 0) return GTSFM_OK;
    if (!out) return fail(c, GTSFM_E_ARG, "out is NULL");
    G2V_CUDA(c, cudaMemcpyAsync(out, c->match_out.p, 8 * static_cast<size_t>(c->batch_total), cudaMemcpyDeviceToHost, c->stream));
    G2V_CUDA(c, cudaStreamSynchronize(c->stream));
    return GTSFM_OK;
}

int gtsfm_match_fetch_async(gtsfm_ctx* c, uint32_t* out)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    if (c->batch_total == 0) return GTSFM_OK;
    if (!out) return fail(c, GTSFM_E_ARG, "out is NULL");
    G2V_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->match_done, 0));
    G2V_CUDA(c, cudaMemcpyAsync(out, c->match_out.p, 8 * static_cast<size_t>(c->batch_total), cudaMemcpyDeviceToHost, c->copy_stream));
    if (!c->fetch_ev) G2V_CUDA(c, cudaEventCreateWithFlags(&c->fetch_ev, cudaEventDisableTiming));
    G2V_CUDA(c, cudaEventRecord(c->fetch_ev, c->copy_stream));
    c->fetch_pending = true;
    return GTSFM_OK;
}

int gtsfm_debug_fetch(gtsfm_ctx* c, int what, void* dst, int64_t bytes)
{
    if (!c || !dst || bytes < 0) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    const DevBuf* b[] = {&c->rp_v1, &c->rp_v2, &c->rp_g, &c->colraw, &c->row_best, &c->row_d1, &c->row_d2,
                         &c->col_best, &c->col_d1, &c->col_d2, &c->n_flag, &c->row_cnt, &c->col_cnt, &c->row_list, &c->col_list};
    if (what < 0 || what > 14) return fail(c, GTSFM_E_ARG, "unknown buffer");
    if (static_cast<size_t>(bytes) > b[what]->cap) return fail(c, GTSFM_E_ARG, "buffer smaller than requested");
    G2V_CUDA(c, cudaMemcpyAsync(dst, b[what]->p, bytes, cudaMemcpyDeviceToHost, c->stream));
    G2V_CUDA(c, cudaStreamSynchronize(c->stream));
    return GTSFM_OK;
}

int gtsfm_match_pair_u8(gtsfm_ctx* c, const uint8_t* desc1, int n1, const uint8_t* desc2, int n2, int nbytes, double ratio, uint32_t* out,
                        int32_t* count)
{
    if (!c || !count) return fail(c, GTSFM_E_ARG, "bad arguments");
    *count = 0;
    int r;
    if ((r = gtsfm_store_clear(c))) return r;
    if (n1 <= 0 || n2 <= 0 || nbytes <= 0) return GTSFM_OK;                  // twoway_matcher.py:71-72
    if ((r = gtsfm_image_upload_u8(c, 0, desc1, n1, nbytes, nullptr, nullptr))) return r;
    if ((r = gtsfm_image_upload_u8(c, 1, desc2, n2, nbytes, nullptr, nullptr))) return r;
    const int32_t pr[2] = {0, 1};
    int64_t total = 0;
    if ((r = gtsfm_match_batch(c, pr, 1, ratio, count, &total))) return r;
    return gtsfm_match_fetch(c, out);
}

int gtsfm_match_pair(gtsfm_ctx* c, const float* desc1, int n1, const float* desc2, int n2, int dim, double ratio, uint32_t* out,
                     int32_t* count)
{
    if (!c || !count) return fail(c, GTSFM_E_ARG, "bad arguments");
    *count = 0;
    int r;
    if ((r = gtsfm_store_clear(c))) return r;
    if (n1 <= 0 || n2 <= 0 || dim <= 0) return GTSFM_OK;                     // twoway_matcher.py:71-72
    if ((r = gtsfm_image_upload(c, 0, desc1, n1, dim, nullptr, nullptr))) return r;
    if ((r = gtsfm_image_upload(c, 1, desc2, n2, dim, nullptr, nullptr))) return r;
    const int32_t pr[2] = {0, 1};
    int64_t total = 0;
    if ((r = gtsfm_match_batch(c, pr, 1, ratio, count, &total))) return r;
    return gtsfm_match_fetch(c, out);
}

}  // extern "C"
