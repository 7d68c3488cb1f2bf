// Context, descriptor store and image upload (C ABI: gtsfm_ctx_*, gtsfm_store_clear, gtsfm_image_upload).
#include "ctx.h"

#include <algorithm>
#include <cstring>

namespace g2v {

struct PrepareArgs {
    const float* src; __nv_bfloat16* bf; uint8_t* u8; __nv_bfloat16* augq; __nv_bfloat16* augt; int* norm_i; uint8_t* valid;
    int* img_flags; __half* h16; __half* augq_h; __half* augt_h; float* norm_f; int n, n_pad, dim, kpad;
};
cudaError_t launch_prepare(const PrepareArgs& a, cudaStream_t s);

static std::string g_create_error;

cudaError_t DevBuf::reserve(size_t bytes, bool keep, cudaStream_t s)
{
    if (bytes <= cap) return cudaSuccess;
    size_t ncap = cap ? cap : 256;
    while (ncap < bytes) ncap = ncap + ncap / 2 + 4096;
    void* np = nullptr;
    cudaError_t e = cudaMalloc(&np, ncap);
    if (e != cudaSuccess) return e;
    if (p) {
        if (keep) { e = cudaMemcpyAsync(np, p, cap, cudaMemcpyDeviceToDevice, s); if (e != cudaSuccess) return e; }
        e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) return e;
        cudaFree(p);
    }
    p = np;
    cap = ncap;
    return cudaSuccess;
}
void DevBuf::release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
cudaError_t PinnedBuf::reserve(size_t bytes)
{
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    size_t ncap = bytes + bytes / 2 + 4096;
    cudaError_t e = cudaMallocHost(&p, ncap);
    if (e == cudaSuccess) cap = ncap;
    return e;
}
void PinnedBuf::release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }

int fail(gtsfm_ctx* c, int code, const std::string& msg)
{
    if (c) c->err = msg; else g_create_error = msg;
    return code;
}
int cuda_fail(gtsfm_ctx* c, cudaError_t e, const char* what)
{
    return fail(c, GTSFM_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

// ---- TMA tensor maps (driver entry point fetched through the runtime: no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
static int make_map(gtsfm_ctx* c, CUtensorMap* m, void* base, int inner, long long rows, int box_inner, int box_rows, CUtensorMapSwizzle sw,
                    CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16)
{
    EncodeTiledFn fn = encode_tiled();
    if (!fn) return fail(c, GTSFM_E_CUDA, "cuTensorMapEncodeTiled entry point not found");
    cuuint64_t gdim[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(rows)};
    cuuint64_t gstr[1] = {static_cast<cuuint64_t>(inner) * 2};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, dt, 2, base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(c, GTSFM_E_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string(static_cast<int>(r)));
    return GTSFM_OK;
}

int store_reserve_rows(gtsfm_ctx* c, long long rows)
{
    if (rows <= c->rows_cap) return GTSFM_OK;
    long long ncap = c->rows_cap ? c->rows_cap : 4096;
    while (ncap < rows) ncap *= 2;
    cudaStream_t s = c->stream;
    // a growing buffer is copied and freed on `s`: no upload copy (copy_stream) may still be writing into the old one
    G2V_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    G2V_CUDA(c, c->desc_bf16.reserve(static_cast<size_t>(ncap) * c->kpad * 2, true, s));
    G2V_CUDA(c, c->desc_u8.reserve(static_cast<size_t>(ncap) * c->kpad, true, s));
    G2V_CUDA(c, c->desc_f32.reserve(static_cast<size_t>(ncap) * c->dim * 4, true, s));
    G2V_CUDA(c, c->augq.reserve(static_cast<size_t>(ncap) * kAugK * 2, true, s));
    G2V_CUDA(c, c->augt.reserve(static_cast<size_t>(ncap) * kAugK * 2, true, s));
    G2V_CUDA(c, c->norm_i.reserve(static_cast<size_t>(ncap) * 4, true, s));
    G2V_CUDA(c, c->valid.reserve(static_cast<size_t>(ncap), true, s));
    G2V_CUDA(c, c->kps.reserve(static_cast<size_t>(ncap) * 16, true, s));
    G2V_CUDA(c, c->desc_h16.reserve(static_cast<size_t>(ncap) * c->kpad * 2, true, s));
    G2V_CUDA(c, c->augq_h.reserve(static_cast<size_t>(ncap) * kAugK * 2, true, s));
    G2V_CUDA(c, c->augt_h.reserve(static_cast<size_t>(ncap) * kAugK * 2, true, s));
    G2V_CUDA(c, c->norm_f.reserve(static_cast<size_t>(ncap) * 4, true, s));
    c->rows_cap = ncap;
    c->maps_valid = false;
    return GTSFM_OK;
}

static int ensure_maps(gtsfm_ctx* c)
{
    if (c->maps_valid) return GTSFM_OK;
    int r;
    if ((r = make_map(c, &c->tm_desc, c->desc_bf16.p, c->kpad, c->rows_cap, 64, kTileQ, CU_TENSOR_MAP_SWIZZLE_128B))) return r;
    if ((r = make_map(c, &c->tm_augq, c->augq.p, kAugK, c->rows_cap, kAugK, kTileQ, CU_TENSOR_MAP_SWIZZLE_32B))) return r;
    if ((r = make_map(c, &c->tm_augt, c->augt.p, kAugK, c->rows_cap, kAugK, kTileQ, CU_TENSOR_MAP_SWIZZLE_32B))) return r;
    const CUtensorMapDataType f16 = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    if ((r = make_map(c, &c->tm_desc_h, c->desc_h16.p, c->kpad, c->rows_cap, 64, kTileQ, CU_TENSOR_MAP_SWIZZLE_128B, f16))) return r;
    if ((r = make_map(c, &c->tm_augq_h, c->augq_h.p, kAugK, c->rows_cap, kAugK, kTileQ, CU_TENSOR_MAP_SWIZZLE_32B, f16))) return r;
    if ((r = make_map(c, &c->tm_augt_h, c->augt_h.p, kAugK, c->rows_cap, kAugK, kTileQ, CU_TENSOR_MAP_SWIZZLE_32B, f16))) return r;
    c->maps_valid = true;
    return GTSFM_OK;
}
int store_ensure_maps(gtsfm_ctx* c) { return ensure_maps(c); }

// Read back the per-image flags written by the prepare kernel (integer-valued? number of valid rows) of images
// 0..max_image (-1: all): waits for the uploads of those images only.
int store_sync_flags(gtsfm_ctx* c, int max_image)
{
    const int n_all = static_cast<int>(c->images.size());
    const int n = (max_image < 0 || max_image >= n_all) ? n_all : max_image + 1;
    if (c->flags_synced >= n) return GTSFM_OK;
    int r = c->wait_images(n - 1, false);
    if (r) return r;
    const int first = c->flags_synced;
    std::vector<int> h(static_cast<size_t>(kImgFlagInts) * (n - first));
    G2V_CUDA(c, cudaMemcpyAsync(h.data(), c->img_flags.as<int>() + static_cast<size_t>(kImgFlagInts) * first, h.size() * sizeof(int),
                                cudaMemcpyDeviceToHost, c->stream));
    G2V_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int i = first; i < n; ++i) {
        const int* f = h.data() + static_cast<size_t>(kImgFlagInts) * (i - first);
        c->images[i].path = (f[0] & 4) ? 3 : (f[0] & 1) ? 1 : 0;
        c->images[i].f16_ok = (f[0] & 2) ? 0 : 1;
        c->images[i].n_valid = f[1];
        std::memcpy(&c->images[i].max_norm2, &f[2], 4);
    }
    c->flags_synced = n;
    return GTSFM_OK;
}

}  // namespace g2v

using namespace g2v;

int gtsfm_ctx::wait_images(int max_image, bool host)
{
    if (max_image < 0 || img_ev.empty()) return GTSFM_OK;
    if (max_image >= static_cast<int>(images.size())) max_image = static_cast<int>(images.size()) - 1;
    cudaEvent_t ev = img_ev[max_image];
    G2V_CUDA(this, cudaStreamWaitEvent(stream, ev, 0));
    if (host) G2V_CUDA(this, cudaEventSynchronize(ev));
    return GTSFM_OK;
}

int gtsfm_ctx::wait_fetch()
{
    if (!fetch_pending) return GTSFM_OK;
    G2V_CUDA(this, cudaEventSynchronize(fetch_ev));
    fetch_pending = false;
    return GTSFM_OK;
}

cudaEvent_t gtsfm_ctx::prof_event()
{
    if (prof_used == prof_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); prof_pool.push_back(e); }
    return prof_pool[prof_used++];
}
void gtsfm_ctx::prof_begin(int slot)
{
    if (!prof_on) return;
    ProfSpan sp{slot, prof_event(), prof_event()};
    cudaEventRecord(sp.a, stream);
    prof_spans.push_back(sp);
}
void gtsfm_ctx::prof_end()
{
    if (!prof_on || prof_spans.empty()) return;
    cudaEventRecord(prof_spans.back().b, stream);
}
int gtsfm_ctx::prof_collect()
{
    if (prof_spans.empty()) return 0;
    cudaStreamSynchronize(stream);
    for (const ProfSpan& sp : prof_spans) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) { prof_ms[sp.slot] += ms; prof_calls[sp.slot] += 1; }
    }
    prof_spans.clear();
    prof_used = 0;
    return 0;
}

g2v::StoreView gtsfm_ctx::view() const
{
    StoreView v;
    v.desc_bf16 = desc_bf16.as<__nv_bfloat16>();
    v.desc_u8 = desc_u8.as<uint8_t>();
    v.desc_f32 = desc_f32.as<float>();
    v.norm_i = norm_i.as<int>();
    v.norm_f = norm_f.as<float>();
    v.valid = valid.as<uint8_t>();
    v.kps = kps.as<double2>();
    v.dim = dim;
    v.kpad = kpad;
    return v;
}

// The flag array grows rarely; growth copies it on `stream`, so nothing may be writing it on the copy stream meanwhile.
static int reserve_img_flags(gtsfm_ctx* c)
{
    const size_t need = (c->images.size() + 1) * kImgFlagInts * sizeof(int);
    if (need <= c->img_flags.cap) return GTSFM_OK;
    G2V_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    G2V_CUDA(c, c->img_flags.reserve(std::max<size_t>(need, 4 * c->img_flags.cap), true, c->stream));
    return GTSFM_OK;
}

// img_ev[image] is recorded behind everything the upload of the newest image enqueued on `s`.
static int record_image_event(gtsfm_ctx* c, cudaStream_t s)
{
    const size_t i = c->images.size() - 1;
    while (c->img_ev.size() <= i) {
        cudaEvent_t e = nullptr;
        G2V_CUDA(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c->img_ev.push_back(e);
    }
    G2V_CUDA(c, cudaEventRecord(c->img_ev[i], s));
    return GTSFM_OK;
}

extern "C" {

int gtsfm_ctx_create(int device, gtsfm_ctx** out)
{
    if (!out) return fail(nullptr, GTSFM_E_ARG, "out is NULL");
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) return fail(nullptr, GTSFM_E_NO_DEVICE, "no CUDA device (the B200 path has no CPU fallback)");
    if (device < 0 || device >= n) return fail(nullptr, GTSFM_E_ARG, "device index out of range");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail(nullptr, GTSFM_E_CUDA, "cudaGetDeviceProperties failed");
    if (prop.major != 10) return fail(nullptr, GTSFM_E_NO_DEVICE, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) + ", the kernels are built for sm_100a only");
    if (cudaSetDevice(device) != cudaSuccess) return fail(nullptr, GTSFM_E_CUDA, "cudaSetDevice failed");
    gtsfm_ctx* c = new gtsfm_ctx();
    c->device = device;
    c->n_sm = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return fail(nullptr, GTSFM_E_CUDA, "cudaStreamCreate failed"); }
    if (cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->match_done, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->upload_ev, cudaEventDisableTiming) != cudaSuccess) {
        delete c;
        return fail(nullptr, GTSFM_E_CUDA, "cudaStreamCreate failed");
    }
    *out = c;
    return GTSFM_OK;
}

void gtsfm_ctx_destroy(gtsfm_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    DevBuf* bufs[] = {&c->desc_h16, &c->augq_h, &c->augt_h, &c->norm_f, &c->desc_bf16, &c->desc_u8, &c->desc_f32, &c->augq, &c->augt, &c->norm_i, &c->valid, &c->kps, &c->img_flags, &c->stage,
                      &c->d_pairs, &c->rp_v1, &c->rp_v2, &c->rp_g, &c->colraw, &c->row_best, &c->row_d1, &c->row_d2, &c->col_best, &c->col_d1,
                      &c->col_d2, &c->flag_list, &c->n_flag, &c->keys_scratch, &c->match_slot, &c->match_count, &c->match_offsets,
                      &c->match_out, &c->d_total, &c->row_key, &c->col_key, &c->row_cnt, &c->col_cnt, &c->row_list, &c->col_list,
                      &c->flag_scratch, &c->col_cells};
    for (DevBuf* b : bufs) b->release();
    for (DevBuf& b : c->v_bufs) b.release();
    c->h_pairs.release();
    c->h_blk_tab.release();
    c->blk_tab.release();
    c->h_small.release();
    c->r_desc.release(); c->r_sim.release(); c->r_tiles.release(); c->r_out.release();
    for (cudaEvent_t e : c->prof_pool) cudaEventDestroy(e);
    for (cudaEvent_t e : c->user_ev) if (e) cudaEventDestroy(e);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->match_done) cudaEventDestroy(c->match_done);
    if (c->upload_ev) cudaEventDestroy(c->upload_ev);
    if (c->xs_ev) cudaEventDestroy(c->xs_ev);
    if (c->fetch_ev) cudaEventDestroy(c->fetch_ev);
    for (cudaEvent_t e : c->img_ev) cudaEventDestroy(e);
    cudaStreamDestroy(c->stream);
    delete c;
}

const char* gtsfm_last_error(const gtsfm_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int gtsfm_ctx_set_workspace(gtsfm_ctx* c, int64_t bytes)
{
    if (!c || bytes < (64ll << 20)) return fail(c, GTSFM_E_ARG, "workspace must be at least 64 MiB");
    c->workspace_bytes = bytes;
    return GTSFM_OK;
}

int64_t gtsfm_kernel_launches(const gtsfm_ctx* c) { return c ? c->launches : 0; }

int gtsfm_ctx_set_float_screening(gtsfm_ctx* c, int on)
{
    if (!c) return GTSFM_E_ARG;
    c->force_exact_float = on == 0;
    return GTSFM_OK;
}

int gtsfm_profile_enable(gtsfm_ctx* c, int on)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    c->prof_collect();
    c->prof_on = on != 0;
    for (int k = 0; k < 16; ++k) { c->prof_ms[k] = 0; c->prof_calls[k] = 0; }
    return GTSFM_OK;
}

int gtsfm_profile_get(gtsfm_ctx* c, double* ms, int64_t* calls, int n_slots)
{
    if (!c || n_slots < 0 || n_slots > 16) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    c->prof_collect();
    for (int k = 0; k < n_slots; ++k) { if (ms) ms[k] = c->prof_ms[k]; if (calls) calls[k] = c->prof_calls[k]; }
    return GTSFM_OK;
}

int gtsfm_event_record(gtsfm_ctx* c, int which)
{
    if (!c || which < 0 || which >= 8) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    if (!c->user_ev[which]) G2V_CUDA(c, cudaEventCreate(&c->user_ev[which]));
    G2V_CUDA(c, cudaEventRecord(c->user_ev[which], c->stream));
    return GTSFM_OK;
}

int gtsfm_event_elapsed_ms(gtsfm_ctx* c, int a, int b, double* ms)
{
    if (!c || !ms || a < 0 || b < 0 || a >= 8 || b >= 8 || !c->user_ev[a] || !c->user_ev[b]) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    G2V_CUDA(c, cudaEventSynchronize(c->user_ev[b]));
    float f = 0;
    G2V_CUDA(c, cudaEventElapsedTime(&f, c->user_ev[a], c->user_ev[b]));
    *ms = f;
    return GTSFM_OK;
}

int gtsfm_store_clear(gtsfm_ctx* c)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    c->images.clear();
    c->rows_used = 0;
    c->dim = 0;
    c->kpad = 0;
    c->binary = -1;
    c->flags_dirty = false;
    c->flags_synced = 0;
    c->maps_valid = false;
    c->batch_n = 0;
    c->batch_total = 0;
    // the next uploads overwrite rows that kernels already enqueued on `stream` may still read
    G2V_CUDA(c, cudaEventRecord(c->upload_ev, c->stream));
    G2V_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->upload_ev, 0));
    return GTSFM_OK;
}

int gtsfm_image_upload(gtsfm_ctx* c, int image_id, const float* desc, int n, int dim, const double* kps, const double* cal)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    if (image_id != static_cast<int>(c->images.size())) return fail(c, GTSFM_E_ARG, "image ids must be dense and uploaded in order");
    if (n < 0 || dim < 0 || (n > 0 && dim > 0 && !desc)) return fail(c, GTSFM_E_ARG, "bad descriptor array");
    if (dim > 4096) return fail(c, GTSFM_E_UNSUPPORTED, "descriptor dimension above 4096");
    if (dim > 0 && n > 0) {
        if (c->binary == 1) return fail(c, GTSFM_E_ARG, "this store holds binary (HAMMING) descriptors");
        c->binary = 0;
    }
    if (c->dim == 0 && dim > 0) {
        if (c->rows_used != 0) return fail(c, GTSFM_E_ARG, "descriptor dimension changed inside one store");
        c->dim = dim;
        c->kpad = (dim + 63) / 64 * 64;
        c->rows_cap = 0;   // row sizes changed: buffers are re-laid out (no image uploaded yet)
    }
    if (dim > 0 && n > 0 && dim != c->dim) return fail(c, GTSFM_E_ARG, "all images of a store must share the descriptor dimension");
    ImageRec rec{};
    rec.row0 = static_cast<int>(c->rows_used);
    rec.n = (dim > 0) ? n : 0;
    rec.n_kps = kps ? n : 0;
    rec.dim = dim;
    rec.path = 1;
    rec.n_valid = 0;
    rec.has_kps = kps != nullptr;
    if (cal) std::memcpy(rec.cal, cal, sizeof(rec.cal)); else { rec.cal[0] = 1; rec.cal[1] = rec.cal[2] = rec.cal[3] = rec.cal[4] = 0; }
    // keypoint-only images (verification of caller-supplied matches) still need store rows for their coordinates
    if (c->dim == 0 && kps && n > 0) { c->dim = 1; c->kpad = 64; c->rows_cap = 0; }
    const int n_rows = ((rec.n > rec.n_kps ? rec.n : rec.n_kps) + kRowPad - 1) / kRowPad * kRowPad;
    rec.n_pad = n_rows;
    int rr = reserve_img_flags(c);
    if (rr) return rr;
    G2V_CUDA(c, cudaMemsetAsync(c->img_flags.as<int>() + kImgFlagInts * c->images.size(), 0, kImgFlagInts * sizeof(int), c->copy_stream));
    if (n_rows > 0) {
        int r = store_reserve_rows(c, c->rows_used + n_rows);
        if (r) return r;
    }
    if (rec.n > 0) {
        float* dst = c->desc_f32.as<float>() + static_cast<long long>(rec.row0) * c->dim;
        // H2D copy AND the prepare kernel on the copy stream: `stream` is not involved, so the kernels of a batch that does not
        // name this image run while it is still arriving (gtsfm_match_batch / gtsfm_verify_batch wait for the event of the
        // last image they name)
        G2V_CUDA(c, cudaMemcpyAsync(dst, desc, static_cast<size_t>(rec.n) * c->dim * 4, cudaMemcpyHostToDevice, c->copy_stream));
        if (kps)
            G2V_CUDA(c, cudaMemcpyAsync(c->kps.as<double2>() + rec.row0, kps, static_cast<size_t>(n) * 16, cudaMemcpyHostToDevice, c->copy_stream));
        PrepareArgs a;
        a.src = dst;
        a.bf = c->desc_bf16.as<__nv_bfloat16>() + static_cast<long long>(rec.row0) * c->kpad;
        a.u8 = c->desc_u8.as<uint8_t>() + static_cast<long long>(rec.row0) * c->kpad;
        a.augq = c->augq.as<__nv_bfloat16>() + static_cast<long long>(rec.row0) * kAugK;
        a.augt = c->augt.as<__nv_bfloat16>() + static_cast<long long>(rec.row0) * kAugK;
        a.norm_i = c->norm_i.as<int>() + rec.row0;
        a.valid = c->valid.as<uint8_t>() + rec.row0;
        a.img_flags = c->img_flags.as<int>() + kImgFlagInts * c->images.size();
        a.h16 = c->desc_h16.as<__half>() + static_cast<long long>(rec.row0) * c->kpad;
        a.augq_h = c->augq_h.as<__half>() + static_cast<long long>(rec.row0) * kAugK;
        a.augt_h = c->augt_h.as<__half>() + static_cast<long long>(rec.row0) * kAugK;
        a.norm_f = c->norm_f.as<float>() + rec.row0;
        a.n = rec.n; a.n_pad = rec.n_pad; a.dim = c->dim; a.kpad = c->kpad;
        G2V_CUDA(c, launch_prepare(a, c->copy_stream));
        ++c->launches;
        c->flags_dirty = true;
    }
    if (kps && n > 0 && rec.n == 0)         // keypoint-only image (the other case copied them above)
        G2V_CUDA(c, cudaMemcpyAsync(c->kps.as<double2>() + rec.row0, kps, static_cast<size_t>(n) * 16, cudaMemcpyHostToDevice, c->copy_stream));
    c->images.push_back(rec);
    c->rows_used += n_rows;
    return record_image_event(c, c->copy_stream);
}

int gtsfm_images_upload(gtsfm_ctx* c, int first_image_id, int n_images, const float* const* desc, const int32_t* n, int dim,
                        const double* const* kps, const double* cal)
{
    if (!c || n_images < 0 || (n_images > 0 && (!desc || !n))) return fail(c, GTSFM_E_ARG, "bad arguments");
    cudaSetDevice(c->device);
    // one growth of the store for the whole list instead of a doubling (copy + synchronise) every few images
    if (dim > 0 && (c->dim == 0 || c->dim == dim) && c->binary != 1 && dim <= 4096) {
        long long rows = c->rows_used;
        for (int i = 0; i < n_images; ++i) rows += (static_cast<long long>(n[i] > 0 ? n[i] : 0) + kRowPad - 1) / kRowPad * kRowPad;
        if (c->dim == 0) { if (c->rows_used == 0) { c->dim = dim; c->kpad = (dim + 63) / 64 * 64; c->rows_cap = 0; } }
        if (c->dim == dim) { int r = store_reserve_rows(c, rows); if (r) return r; }
    }
    for (int i = 0; i < n_images; ++i) {
        int r = gtsfm_image_upload(c, first_image_id + i, desc[i], n[i], dim, kps ? kps[i] : nullptr, cal ? cal + 5 * i : nullptr);
        if (r) return r;
    }
    return GTSFM_OK;
}

int gtsfm_image_upload_u8(gtsfm_ctx* c, int image_id, const uint8_t* desc, int n, int nbytes, const double* kps, const double* cal)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    if (image_id != static_cast<int>(c->images.size())) return fail(c, GTSFM_E_ARG, "image ids must be dense and uploaded in order");
    if (n < 0 || (n > 0 && (nbytes <= 0 || !desc))) return fail(c, GTSFM_E_ARG, "bad descriptor array");
    if (n == 0) {
        // an image without features (np.array([]) in the reference, twoway_matcher.py:71-72): its pairs yield no matches; it
        // occupies no rows and fixes neither the descriptor size nor the kind of the store
        ImageRec rec{};
        rec.row0 = static_cast<int>(c->rows_used);
        rec.path = 3; rec.dim = nbytes > 0 ? nbytes : 0;
        if (cal) std::memcpy(rec.cal, cal, sizeof(rec.cal)); else { rec.cal[0] = 1; rec.cal[1] = rec.cal[2] = rec.cal[3] = rec.cal[4] = 0; }
        { int rr = reserve_img_flags(c); if (rr) return rr; }
        const int flags0[kImgFlagInts] = {4, 0, 0, 0};
        G2V_CUDA(c, cudaMemcpyAsync(c->img_flags.as<int>() + kImgFlagInts * c->images.size(), flags0, sizeof(flags0), cudaMemcpyHostToDevice, c->stream));
        c->images.push_back(rec);
        return record_image_event(c, c->stream);
    }
    if (nbytes > 64) return fail(c, GTSFM_E_UNSUPPORTED, "binary descriptors above 64 bytes");
    if (c->binary == 0) return fail(c, GTSFM_E_ARG, "this store holds float32 (EUCLIDEAN) descriptors");
    if (c->dim == 0) {
        if (c->rows_used != 0) return fail(c, GTSFM_E_ARG, "descriptor dimension changed inside one store");
        c->dim = nbytes; c->kpad = 64; c->rows_cap = 0;
    }
    if (nbytes != c->dim) return fail(c, GTSFM_E_ARG, "all images of a store must share the descriptor size");
    c->binary = 1;
    ImageRec rec{};
    rec.row0 = static_cast<int>(c->rows_used);
    rec.n = n; rec.n_kps = kps ? n : 0; rec.dim = nbytes; rec.path = 3; rec.n_valid = n; rec.has_kps = kps != nullptr;
    if (cal) std::memcpy(rec.cal, cal, sizeof(rec.cal)); else { rec.cal[0] = 1; rec.cal[1] = rec.cal[2] = rec.cal[3] = rec.cal[4] = 0; }
    rec.n_pad = (n + kRowPad - 1) / kRowPad * kRowPad;
    { int rr = reserve_img_flags(c); if (rr) return rr; }
    const int flags[kImgFlagInts] = {4, n, 0, 0};             // bit 2: binary image; number of valid rows (uint8 has no NaN)
    G2V_CUDA(c, cudaMemcpyAsync(c->img_flags.as<int>() + kImgFlagInts * c->images.size(), flags, sizeof(flags), cudaMemcpyHostToDevice, c->stream));
    if (rec.n_pad > 0) {
        int r = store_reserve_rows(c, c->rows_used + rec.n_pad);
        if (r) return r;
        uint8_t* dst = c->desc_u8.as<uint8_t>() + static_cast<long long>(rec.row0) * c->kpad;
        G2V_CUDA(c, cudaMemsetAsync(dst, 0, static_cast<size_t>(rec.n_pad) * c->kpad, c->stream));
        G2V_CUDA(c, cudaMemcpy2DAsync(dst, c->kpad, desc, nbytes, nbytes, n, cudaMemcpyHostToDevice, c->stream));
        G2V_CUDA(c, cudaMemsetAsync(c->valid.as<uint8_t>() + rec.row0, 0, rec.n_pad, c->stream));
        G2V_CUDA(c, cudaMemsetAsync(c->valid.as<uint8_t>() + rec.row0, 1, n, c->stream));
    }
    if (kps && n > 0)
        G2V_CUDA(c, cudaMemcpyAsync(c->kps.as<double2>() + rec.row0, kps, static_cast<size_t>(n) * 16, cudaMemcpyHostToDevice, c->stream));
    c->images.push_back(rec);
    c->rows_used += rec.n_pad;
    return record_image_event(c, c->stream);
}

int gtsfm_host_alloc(void** p, int64_t bytes)
{
    if (!p || bytes < 0) return GTSFM_E_ARG;
    *p = nullptr;
    if (bytes == 0) return GTSFM_OK;
    return cudaHostAlloc(p, static_cast<size_t>(bytes), cudaHostAllocPortable) == cudaSuccess ? GTSFM_OK : GTSFM_E_CUDA;
}

void gtsfm_host_free(void* p)
{
    if (p) cudaFreeHost(p);
}

int gtsfm_sync(gtsfm_ctx* c)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    G2V_CUDA(c, cudaStreamSynchronize(c->stream));
    G2V_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    return GTSFM_OK;
}

}  // extern "C"
