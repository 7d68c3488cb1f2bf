// Retriever step in front of the matcher: global-descriptor similarity matrix and per-image top-K
// pair selection of gtsfm/retriever/netvlad_retriever.py:77-149 (blocked `einsum("id,jd->ij")`) and :196-228
// (`pairs_from_score_matrix`: mask, min_score, torch.topk per row).
//
//   R1 similarity_kernel   sim = G G^T in fp32 on the SIMT pipe (N <= 10,000 images, D = 4,096 for NetVLAD: 0.03-0.8 TFLOP,
//                          milliseconds; tf32 / bf16 tensor-core products would change the scores the top-K ranks by).
//                          64 x 64 tile per CTA, K chunks of 16 through shared memory, 4 x 4 accumulators per thread,
//                          accumulated in ascending k in two levels (spans of 128).  Only the tiles the reference computes
//                          are produced: entry (i, j) exists iff block(j) >= block(i) for the caller's blocksize, everything
//                          else stays 0 (netvlad_retriever.py:100-108,150-153).
//   R2 topk_rows_kernel    one warp per image i: candidates are j > i (upper triangle without the diagonal,
//                          netvlad_retriever.py:166-168) with !(score < min_score); num_select rounds of a warp arg-max in the
//                          order (score descending, j ascending), each round restricted to what follows the previous pick.
//                          Rows are re-read from L2 per round (N * 4 B per row), nothing is sorted or stored.
#include <cfloat>
#include <climits>

#include "ctx.h"

namespace g2v {

constexpr int kSimTile = 64;
constexpr int kSimK = 16;
constexpr int kSimSpan = 128;       // k's per first-level partial sum (a multiple of kSimK, power-of-two ratio)

__global__ void __launch_bounds__(256) similarity_kernel(const float* __restrict__ g, int n, int dim, int blocksize,
                                                         const int2* __restrict__ tiles, float* __restrict__ sim)
{
    __shared__ float s_a[kSimK][kSimTile + 4];
    __shared__ float s_b[kSimK][kSimTile + 4];
    const int2 tile = tiles[blockIdx.x];
    const int i0 = tile.x * kSimTile, j0 = tile.y * kSimTile;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    // two-level accumulation: `acc` sums one span of kSimSpan k's, `tot` sums the spans, so the rounding error grows with
    // kSimSpan + dim / kSimSpan instead of dim (a 4,096-term fp32 chain would be ~1e-5 off on a unit-norm diagonal)
    float acc[4][4], tot[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) { acc[a][b] = 0.f; tot[a][b] = 0.f; }
    // loader: thread -> (row r = tid / 4, 4 consecutive k)
    const int lr = threadIdx.x >> 2, lk = (threadIdx.x & 3) * 4;
    for (int k0 = 0; k0 < dim; k0 += kSimK) {
        float va[4] = {0.f, 0.f, 0.f, 0.f}, vb[4] = {0.f, 0.f, 0.f, 0.f};
        const int ia = i0 + lr, jb = j0 + lr;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int k = k0 + lk + u;
            if (k < dim) {
                if (ia < n) va[u] = g[static_cast<long long>(ia) * dim + k];
                if (jb < n) vb[u] = g[static_cast<long long>(jb) * dim + k];
            }
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < 4; ++u) { s_a[lk + u][lr] = va[u]; s_b[lk + u][lr] = vb[u]; }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kSimK; ++k) {
            float a[4], b[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { a[u] = s_a[k][ty * 4 + u]; b[u] = s_b[k][tx * 4 + u]; }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) acc[u][v] = fmaf(a[u], b[v], acc[u][v]);
        }
        if (((k0 / kSimK) & (kSimSpan / kSimK - 1)) == kSimSpan / kSimK - 1) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) { tot[u][v] += acc[u][v]; acc[u][v] = 0.f; }
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[u][v] += tot[u][v];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int i = i0 + ty * 4 + u;
        if (i >= n) continue;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const int j = j0 + tx * 4 + v;
            if (j < n && j / blocksize >= i / blocksize) sim[static_cast<long long>(i) * n + j] = acc[u][v];
        }
    }
}

__global__ void __launch_bounds__(256) topk_rows_kernel(const float* __restrict__ sim, int n, int k_sel, float min_score, int use_min,
                                                        int* __restrict__ out_idx, float* __restrict__ out_score)
{
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    const float* r = sim + static_cast<long long>(row) * n;
    float pv = FLT_MAX;                  // previous pick: everything not after it in (score desc, j asc) is excluded
    int pj = -1;
    for (int sel = 0; sel < k_sel; ++sel) {
        float bv = -FLT_MAX;
        int bj = INT_MAX;
        if (pj != INT_MAX) {
            for (int j = row + 1 + lane; j < n; j += 32) {
                const float v = r[j];
                if (use_min && v < min_score) continue;          // invalid |= scores < min_score (netvlad_retriever.py:218-220)
                if (!(v > -FLT_MAX && v <= FLT_MAX)) continue;     // topk.values.isfinite() (netvlad_retriever.py:223): no NaN / inf
                const bool after = v < pv || (v == pv && j > pj);
                if (!after) continue;
                if (v > bv || (v == bv && j < bj)) { bv = v; bj = j; }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, off);
                const int oj = __shfl_xor_sync(0xffffffffu, bj, off);
                if (ov > bv || (ov == bv && oj < bj)) { bv = ov; bj = oj; }
            }
        }
        if (lane == 0) {
            out_idx[static_cast<long long>(row) * k_sel + sel] = bj == INT_MAX ? -1 : bj;
            out_score[static_cast<long long>(row) * k_sel + sel] = bj == INT_MAX ? -INFINITY : bv;
        }
        pv = bv;
        pj = bj;                         // INT_MAX: the row is exhausted, the remaining slots are filled with -1
    }
}

}  // namespace g2v

using namespace g2v;

extern "C" {

int gtsfm_retriever_similarity(gtsfm_ctx* c, const float* global_desc, int n_images, int dim, int blocksize, float* sim_out)
{
    if (!c || !global_desc || n_images <= 0 || dim <= 0 || blocksize <= 0) return fail(c, GTSFM_E_ARG, "bad arguments");
    if (n_images > 10000) return fail(c, GTSFM_E_ARG, "Cannot construct similarity matrix of this size.");   // MAX_NUM_IMAGES
    cudaSetDevice(c->device);
    cudaStream_t s = c->stream;
    const size_t nn = static_cast<size_t>(n_images) * n_images;
    G2V_CUDA(c, c->r_desc.reserve(sizeof(float) * static_cast<size_t>(n_images) * dim, false, s));
    G2V_CUDA(c, c->r_sim.reserve(sizeof(float) * nn, false, s));
    const int nt = (n_images + kSimTile - 1) / kSimTile;
    std::vector<int2> tiles;
    for (int ti = 0; ti < nt; ++ti)
        for (int tj = 0; tj < nt; ++tj) {
            // the tile holds an entry the reference computes iff its last column's block >= its first row's block
            const int j_last = (tj + 1) * kSimTile - 1 < n_images ? (tj + 1) * kSimTile - 1 : n_images - 1;
            if (j_last / blocksize >= (ti * kSimTile) / blocksize) tiles.push_back(make_int2(ti, tj));
        }
    G2V_CUDA(c, c->r_tiles.reserve(sizeof(int2) * tiles.size(), false, s));
    G2V_CUDA(c, cudaMemcpyAsync(c->r_desc.p, global_desc, sizeof(float) * static_cast<size_t>(n_images) * dim, cudaMemcpyHostToDevice, s));
    G2V_CUDA(c, cudaMemcpyAsync(c->r_tiles.p, tiles.data(), sizeof(int2) * tiles.size(), cudaMemcpyHostToDevice, s));
    G2V_CUDA(c, cudaMemsetAsync(c->r_sim.p, 0, sizeof(float) * nn, s));
    similarity_kernel<<<static_cast<unsigned>(tiles.size()), 256, 0, s>>>(c->r_desc.as<float>(), n_images, dim, blocksize,
                                                                          c->r_tiles.as<int2>(), c->r_sim.as<float>());
    G2V_CUDA(c, cudaGetLastError());
    c->launches += 1;
    c->r_n = n_images;
    if (sim_out) G2V_CUDA(c, cudaMemcpyAsync(sim_out, c->r_sim.p, sizeof(float) * nn, cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaStreamSynchronize(s));        // `tiles` and the caller's buffers are released on return
    return GTSFM_OK;
}

int gtsfm_retriever_topk(gtsfm_ctx* c, const float* sim_in, int n_images, int num_select, double min_score, int32_t* out_idx,
                         float* out_score)
{
    if (!c || n_images <= 0 || num_select <= 0 || !out_idx || !out_score) return fail(c, GTSFM_E_ARG, "bad arguments");
    cudaSetDevice(c->device);
    cudaStream_t s = c->stream;
    const size_t nn = static_cast<size_t>(n_images) * n_images;
    if (sim_in) {
        G2V_CUDA(c, c->r_sim.reserve(sizeof(float) * nn, false, s));
        G2V_CUDA(c, cudaMemcpyAsync(c->r_sim.p, sim_in, sizeof(float) * nn, cudaMemcpyHostToDevice, s));
        c->r_n = n_images;
    } else if (c->r_n != n_images) {
        return fail(c, GTSFM_E_ARG, "retriever_topk without a matrix needs the similarity of the last retriever_similarity call");
    }
    const int k = num_select < n_images ? num_select : n_images;       // netvlad_retriever.py:213-214
    G2V_CUDA(c, c->r_out.reserve(8 * static_cast<size_t>(n_images) * k, false, s));
    int* d_idx = c->r_out.as<int>();
    float* d_score = reinterpret_cast<float*>(d_idx + static_cast<size_t>(n_images) * k);
    const int use_min = min_score == min_score ? 1 : 0;                // NaN = no minimum score
    topk_rows_kernel<<<(n_images + 7) / 8, 256, 0, s>>>(c->r_sim.as<float>(), n_images, k, static_cast<float>(min_score), use_min, d_idx, d_score);
    G2V_CUDA(c, cudaGetLastError());
    c->launches += 1;
    G2V_CUDA(c, cudaMemcpyAsync(out_idx, d_idx, 4 * static_cast<size_t>(n_images) * k, cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaMemcpyAsync(out_score, d_score, 4 * static_cast<size_t>(n_images) * k, cudaMemcpyDeviceToHost, s));
    G2V_CUDA(c, cudaStreamSynchronize(s));
    return GTSFM_OK;
}

}  // extern "C"
