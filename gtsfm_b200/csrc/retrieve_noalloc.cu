// gtsfm_retriever_pairs_noalloc: the retriever step (similarity matrix + per-image top-K, see retrieve.cu) for callers that
// must not allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart
// in the reference.
//
// The regular entry points keep the (n_images, n_images) similarity matrix in a context buffer that grows with cudaMalloc.
// Here that scratch is a file-scope __device__ array of a fixed, modest size (kNoAllocMaxImages^2 floats = 4 MiB, held by
// every GPU that loads the library until the process exits); inputs and outputs are the caller's DEVICE buffers and every
// operation is enqueued on the caller's stream.  The entry point calls no cudaMalloc / cudaFree / allocating library routine
// and never synchronises.  All calls share the one array (per device): it keeps the previous call's contents, so the part a
// call uses is cleared first, and two calls that overlap on the GPU (different streams without an ordering) race.
#include <cfloat>
#include <climits>
#include <mutex>

#include "ctx.h"

namespace g2v {

constexpr int kNoAllocMaxImages = GTSFM_RETRIEVER_NOALLOC_MAX_IMAGES;   // 1024 (include/gtsfm_b200.h)
constexpr int kNaTile = 64;
constexpr int kNaK = 16;
constexpr int kNaSpan = 128;

// the scratch: similarity matrix of the current call, row stride = that call's n_images
static __device__ __align__(256) float g_noalloc_sim[static_cast<size_t>(kNoAllocMaxImages) * kNoAllocMaxImages];

// Same arithmetic as similarity_kernel (retrieve.cu): 64 x 64 tile per CTA, ascending k, two-level accumulation in spans of
// 128, so both entry points give bit-identical scores.  The tile list of the regular path (a host vector copied to a growing
// device buffer) is replaced by a 2-D grid whose CTAs leave when their tile holds no entry the reference computes.
__global__ void __launch_bounds__(256) similarity_noalloc_kernel(const float* __restrict__ g, int n, int dim, int blocksize,
                                                                 float* __restrict__ sim)
{
    __shared__ float s_a[kNaK][kNaTile + 4];
    __shared__ float s_b[kNaK][kNaTile + 4];
    const int i0 = blockIdx.y * kNaTile, j0 = blockIdx.x * kNaTile;
    const int j_last = j0 + kNaTile - 1 < n ? j0 + kNaTile - 1 : n - 1;
    if (j_last / blocksize < i0 / blocksize) return;             // netvlad_retriever.py:100-108: blocks below the diagonal stay 0
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4], tot[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) { acc[a][b] = 0.f; tot[a][b] = 0.f; }
    const int lr = threadIdx.x >> 2, lk = (threadIdx.x & 3) * 4;
    for (int k0 = 0; k0 < dim; k0 += kNaK) {
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
        for (int k = 0; k < kNaK; ++k) {
            float a[4], b[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { a[u] = s_a[k][ty * 4 + u]; b[u] = s_b[k][tx * 4 + u]; }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) acc[u][v] = fmaf(a[u], b[v], acc[u][v]);
        }
        if (((k0 / kNaK) & (kNaSpan / kNaK - 1)) == kNaSpan / kNaK - 1) {
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) { tot[u][v] += acc[u][v]; acc[u][v] = 0.f; }
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int i = i0 + ty * 4 + u;
        if (i >= n) continue;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const int j = j0 + tx * 4 + v;
            if (j < n && j / blocksize >= i / blocksize) sim[static_cast<long long>(i) * n + j] = (acc[u][v] + tot[u][v]);
        }
    }
}

// One warp per image: num_select rounds of a warp arg-max in the order (score descending, j ascending) over j > i
// (netvlad_retriever.py:166-168,213-228), exactly the selection rule of topk_rows_kernel (retrieve.cu).
__global__ void __launch_bounds__(256) topk_noalloc_kernel(const float* __restrict__ sim, int n, int k_sel, float min_score, int use_min,
                                                           int* __restrict__ out_idx, float* __restrict__ out_score)
{
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    const float* r = sim + static_cast<long long>(row) * n;
    float pv = FLT_MAX;
    int pj = -1;
    for (int sel = 0; sel < k_sel; ++sel) {
        float bv = -FLT_MAX;
        int bj = INT_MAX;
        if (pj != INT_MAX) {
            for (int j = row + 1 + lane; j < n; j += 32) {
                const float v = r[j];
                if (use_min && v < min_score) continue;
                if (!(v > -FLT_MAX && v <= FLT_MAX)) continue;
                if (!(v < pv || (v == pv && j > pj))) continue;
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
        pj = bj;
    }
}

// Device address of the scratch array, looked up once per device with that device current.  In host code the array's name
// is only a stand-in (a host-side shadow symbol), never a device address: it must go through cudaGetSymbolAddress.
static std::mutex g_noalloc_mutex;
static float* g_noalloc_ptr[64] = {nullptr};

static int noalloc_scratch(float** out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(nullptr, GTSFM_E_ARG, "gtsfm_retriever_pairs_noalloc: device index above 63");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_sim);
        if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaGetSymbolAddress(g_noalloc_sim)");
        if (!p) return fail(nullptr, GTSFM_E_CUDA, "cudaGetSymbolAddress(g_noalloc_sim) returned a null address");
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return GTSFM_OK;
}

}  // namespace g2v

using namespace g2v;

extern "C" {

int gtsfm_retriever_pairs_noalloc(const float* d_global_desc, int n_images, int dim, int blocksize, int num_select, double min_score,
                                  int32_t* d_out_idx, float* d_out_score, void* cuda_stream)
{
    if (!d_global_desc || !d_out_idx || !d_out_score || n_images <= 0 || dim <= 0 || blocksize <= 0 || num_select <= 0)
        return fail(nullptr, GTSFM_E_ARG, "gtsfm_retriever_pairs_noalloc: bad arguments");
    if (n_images > kNoAllocMaxImages)
        return fail(nullptr, GTSFM_E_ARG, "gtsfm_retriever_pairs_noalloc: " + std::to_string(n_images) + " images exceed the fixed scratch of " +
                                              std::to_string(kNoAllocMaxImages) + " images (use gtsfm_retriever_similarity / gtsfm_retriever_topk)");
    if (num_select > n_images)
        return fail(nullptr, GTSFM_E_ARG, "gtsfm_retriever_pairs_noalloc: num_select exceeds n_images (the outputs are (n_images, num_select))");
    float* sim = nullptr;
    int r = noalloc_scratch(&sim);
    if (r) return r;
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    const size_t nn = static_cast<size_t>(n_images) * n_images;
    // the array still holds the previous call's matrix (possibly of another size): entries below the block diagonal are
    // never written by the kernel and must read 0
    cudaError_t e = cudaMemsetAsync(sim, 0, sizeof(float) * nn, s);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "cudaMemsetAsync(scratch)");
    const unsigned nt = static_cast<unsigned>((n_images + kNaTile - 1) / kNaTile);
    similarity_noalloc_kernel<<<dim3(nt, nt), 256, 0, s>>>(d_global_desc, n_images, dim, blocksize, sim);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "similarity_noalloc_kernel");
    const int use_min = min_score == min_score ? 1 : 0;             // NaN = no minimum score
    topk_noalloc_kernel<<<(n_images + 7) / 8, 256, 0, s>>>(sim, n_images, num_select, static_cast<float>(min_score), use_min, d_out_idx,
                                                           d_out_score);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "topk_noalloc_kernel");
    return GTSFM_OK;
}

}  // extern "C"
