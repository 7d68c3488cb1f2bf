// Device-resident result records of a verify batch, for the multi-GPU result exchange: "per-pair fixed header
// {n_matches, n_inliers, status, R[9], t[3], ratio} plus variable-length index payloads packed with a prefix-sum offset table".
// What is packed is what unzip_two_view_results hands to the back end (gtsfm/runner/gtsfm_runner_base.py:415-440): i2Ri1, i2Ui1,
// v_corr_idxs (= the match rows with inlier mask 1, input order kept, opencv_verifier_base.py:98-100) and the inlier ratio.
//
//   P1 pack_header_kernel   one thread per pair: 128-byte header; number of payload rows = n_inliers for verified pairs, else 0
//   P2 (scan)               exclusive prefix sum of the payload rows over the pairs of the batch (one CTA, chunked) -> offsets
//   P3 pack_payload_kernel  one CTA per pair: stable compaction of the pair's inlier rows to payload[offset[p] ...]
// The block a rank contributes to the NCCL all-gather is [headers | payload], written straight into the caller's device
// buffer (gtsfm_result_pack_copy), so nothing bounces through the host.
#include "ctx.h"
#include "verify_types.h"

namespace g2v {

struct PackedHeader {            // 128 bytes; mirrored by gtsfm_b200/two_view_batch.py HEADER_DTYPE
    long long pair_id;           // caller's id of the pair (position in the global pair list)
    int n_matches, n_inliers, status, pad;
    double R[9], t[3];
    double inlier_ratio;         // n_inliers / n_matches (opencv_verifier_base.py:101), 0 for failed pairs
};
static_assert(sizeof(PackedHeader) == 128, "header layout");

__global__ void pack_header_kernel(const VPair* __restrict__ pairs, const double* __restrict__ R, const double* __restrict__ t,
                                   const int* __restrict__ ninl, const int* __restrict__ status, const long long* __restrict__ pair_ids,
                                   int n_pairs, PackedHeader* __restrict__ hdr, long long* __restrict__ rows)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    PackedHeader h;
    h.pair_id = pair_ids ? pair_ids[p] : p;
    h.n_matches = pairs[p].M;
    h.status = status[p];
    const bool ok = h.status == 0;
    h.n_inliers = ok ? ninl[p] : 0;
    h.pad = 0;
#pragma unroll
    for (int i = 0; i < 9; ++i) h.R[i] = ok ? R[9ll * p + i] : 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) h.t[i] = ok ? t[3ll * p + i] : 0.0;
    h.inlier_ratio = (ok && h.n_matches > 0) ? static_cast<double>(h.n_inliers) / h.n_matches : 0.0;
    hdr[p] = h;
    rows[p] = h.n_inliers;
}

// exclusive prefix sum over n values by ONE CTA of 1024 threads (n is a pair count: <= a few 100k); off[n] = total
__global__ void __launch_bounds__(1024) scan_rows_kernel(const long long* __restrict__ rows, int n, long long* __restrict__ off)
{
    __shared__ long long s_warp[32];
    __shared__ long long s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const long long v = i < n ? rows[i] : 0;
        long long x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const long long y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        if (warp == 0) {
            long long w = s_warp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const long long y = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += y; }
            s_warp[lane] = w;
        }
        __syncthreads();
        const long long before = s_carry + (warp ? s_warp[warp - 1] : 0) + x - v;
        if (i < n) off[i] = before;
        __syncthreads();
        if (tid == 1023) s_carry = before + v;
        __syncthreads();
    }
    if (tid == 0) off[n] = s_carry;
}

// one CTA per pair: rows with mask 1, in order, to payload[off[p] + k]
__global__ void __launch_bounds__(256) pack_payload_kernel(const VPair* __restrict__ pairs, const uint2* __restrict__ match,
                                                           const unsigned char* __restrict__ mask, const PackedHeader* __restrict__ hdr,
                                                           const long long* __restrict__ off, int n_pairs, uint2* __restrict__ payload)
{
    __shared__ int s_warp[8];
    __shared__ int s_base;
    const int p = blockIdx.x;
    if (p >= n_pairs || hdr[p].n_inliers == 0) return;
    const VPair vp = pairs[p];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint2* dst = payload + off[p];
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int i0 = 0; i0 < vp.M; i0 += 256) {
        const int i = i0 + tid;
        const bool in = i < vp.M && mask[vp.m0 + i] == 1;
        const unsigned bal = __ballot_sync(0xffffffffu, in);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int before = s_base;
        for (int w = 0; w < warp; ++w) before += s_warp[w];
        if (in) dst[before + __popc(bal & ((1u << lane) - 1u))] = match[vp.m0 + i];
        __syncthreads();
        if (tid == 0) { int tot = 0; for (int w = 0; w < 8; ++w) tot += s_warp[w]; s_base += tot; }
        __syncthreads();
    }
}

}  // namespace g2v

using namespace g2v;

extern "C" {

int gtsfm_result_pack(gtsfm_ctx* c, int n_pairs, const int64_t* pair_ids, int64_t* n_payload_rows)
{
    if (!c || n_pairs < 0 || !n_payload_rows) return fail(c, GTSFM_E_ARG, "bad arguments");
    cudaSetDevice(c->device);
    *n_payload_rows = 0;
    c->pack_n = 0;
    if (n_pairs == 0) return GTSFM_OK;
    if (c->verify_n != n_pairs) return fail(c, GTSFM_E_ARG, "result_pack needs the pairs of the last verify_batch");
    cudaStream_t s = c->stream;
    DevBuf* vb = c->v_bufs;
    G2V_CUDA(c, vb[24].reserve(sizeof(PackedHeader) * static_cast<size_t>(n_pairs), false, s));
    G2V_CUDA(c, vb[25].reserve(8 * (2 * static_cast<size_t>(n_pairs) + 2), false, s));           // rows (P), offsets (P + 1)
    long long* rows = vb[25].as<long long>();
    long long* off = rows + n_pairs;
    const long long* d_ids = nullptr;
    if (pair_ids) {
        G2V_CUDA(c, vb[26].reserve(8 * static_cast<size_t>(n_pairs), false, s));
        G2V_CUDA(c, cudaMemcpyAsync(vb[26].p, pair_ids, 8 * static_cast<size_t>(n_pairs), cudaMemcpyHostToDevice, s));
        d_ids = vb[26].as<long long>();
    }
    pack_header_kernel<<<(n_pairs + 127) / 128, 128, 0, s>>>(vb[0].as<VPair>(), vb[7].as<double>(), vb[8].as<double>(), vb[10].as<int>(),
                                                             vb[11].as<int>(), d_ids, n_pairs, vb[24].as<PackedHeader>(), rows);
    G2V_CUDA(c, cudaGetLastError());
    scan_rows_kernel<<<1, 1024, 0, s>>>(rows, n_pairs, off);
    G2V_CUDA(c, cudaGetLastError());
    c->launches += 2;
    long long total = 0;
    G2V_CUDA(c, cudaMemcpyAsync(&total, off + n_pairs, 8, cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaStreamSynchronize(s));
    c->pack_n = n_pairs;
    c->pack_rows = total;
    *n_payload_rows = total;
    return GTSFM_OK;
}

int gtsfm_result_pack_copy(gtsfm_ctx* c, void* d_dst, int64_t dst_bytes, int64_t payload_offset_bytes)
{
    if (!c || !d_dst) return fail(c, GTSFM_E_ARG, "bad arguments");
    cudaSetDevice(c->device);
    const int n_pairs = c->pack_n;
    if (n_pairs <= 0 || c->verify_n != n_pairs) return fail(c, GTSFM_E_ARG, "result_pack_copy needs a preceding gtsfm_result_pack");
    const int64_t hdr_bytes = static_cast<int64_t>(sizeof(PackedHeader)) * n_pairs;
    if (payload_offset_bytes < hdr_bytes || (payload_offset_bytes & 7) || payload_offset_bytes + 8 * c->pack_rows > dst_bytes)
        return fail(c, GTSFM_E_ARG, "destination too small: " + std::to_string(hdr_bytes) + " header bytes and " + std::to_string(8 * c->pack_rows) +
                                        " payload bytes at offset " + std::to_string(payload_offset_bytes) + " do not fit " + std::to_string(dst_bytes));
    cudaStream_t s = c->stream;
    DevBuf* vb = c->v_bufs;
    G2V_CUDA(c, cudaMemcpyAsync(d_dst, vb[24].p, static_cast<size_t>(hdr_bytes), cudaMemcpyDeviceToDevice, s));
    const long long* off = vb[25].as<long long>() + n_pairs;
    pack_payload_kernel<<<n_pairs, 256, 0, s>>>(vb[0].as<VPair>(), c->match_out.as<uint2>(), vb[12].as<unsigned char>(), vb[24].as<PackedHeader>(), off,
                                                n_pairs, reinterpret_cast<uint2*>(static_cast<char*>(d_dst) + payload_offset_bytes));
    G2V_CUDA(c, cudaGetLastError());
    c->launches += 1;
    return GTSFM_OK;
}

int gtsfm_stream_wait(gtsfm_ctx* c, void* cuda_stream, int direction)
{
    if (!c) return GTSFM_E_ARG;
    cudaSetDevice(c->device);
    cudaStream_t ext = static_cast<cudaStream_t>(cuda_stream);
    if (!c->xs_ev) G2V_CUDA(c, cudaEventCreateWithFlags(&c->xs_ev, cudaEventDisableTiming));
    if (direction == 0) {            // the caller's stream waits for everything the context has enqueued so far
        G2V_CUDA(c, cudaEventRecord(c->xs_ev, c->stream));
        G2V_CUDA(c, cudaStreamWaitEvent(ext, c->xs_ev, 0));
    } else {                         // the context's stream waits for the caller's stream
        G2V_CUDA(c, cudaEventRecord(c->xs_ev, ext));
        G2V_CUDA(c, cudaStreamWaitEvent(c->stream, c->xs_ev, 0));
    }
    return GTSFM_OK;
}

}  // extern "C"
