// K2b -- ratio test, mutual check, ordering and compaction of the matches of every pair.
//
// Restates gtsfm/frontend/matcher/twoway_matcher.py:
//   :137      keep a one-way match iff  m1.distance <= ratio * m2.distance  (Python doubles on float32 distances)
//   :117-120  keep (i1, i2) from the 1->2 matches iff the 2->1 match of i2 is i1; cast to uint32
//   :141-142  1->2 matches are ordered by ascending distance with a STABLE sort over the query index
// One CTA per pair: ordered compaction of the surviving rows, bitonic sort of (float32 distance bits << 32 | i1).
#include <cfloat>

#include "store.h"

namespace g2v {

constexpr int kSortSmem = 8192;   // keys sorted in shared memory; larger pairs sort in their global slot

__device__ __forceinline__ bool ratio_ok(float d1, float d2, int use_ratio, double ratio)
{
    if (!use_ratio) return true;
    return static_cast<double>(d1) <= ratio * static_cast<double>(d2);
}

template <typename KeyPtr>
__device__ __forceinline__ void bitonic_sort(KeyPtr keys, int n_pow2)
{
    for (int k = 2; k <= n_pow2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            // one compare-exchange per thread and iteration: pair q couples t (bit j clear) with t | j
            for (int q = threadIdx.x; q < (n_pow2 >> 1); q += blockDim.x) {
                const int t = ((q & ~(j - 1)) << 1) | (q & (j - 1)), x = t | j;
                const unsigned long long a = keys[t], b = keys[x];
                const bool up = (t & k) == 0;
                if ((a > b) == up) { keys[t] = b; keys[x] = a; }
            }
            __syncthreads();
        }
    }
}

// keys_scratch: [sum nq_pad * 2] uint64 (a power of two >= the match count always fits in 2 * nq_pad)
__global__ void __launch_bounds__(512) select_sort_kernel(MatchWork w, int use_ratio, double ratio,
                                                          unsigned long long* keys_scratch, uint2* match_slot, int* match_count)
{
    extern __shared__ unsigned long long skeys[];   // [kSortSmem]
    __shared__ int warp_cnt[16];
    __shared__ int s_total;
    const int p = blockIdx.x;
    const PairDesc pd = w.pairs[p];
    unsigned long long* gkeys = keys_scratch + 2 * pd.out_base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    if (threadIdx.x == 0) s_total = 0;
    __syncthreads();

    // ordered compaction: rows are visited in ascending order so equal-distance rows stay in query-index order
    for (int base = 0; base < pd.nq; base += blockDim.x) {
        const int i = base + threadIdx.x;
        bool keep = false;
        unsigned long long key = 0;
        if (i < pd.nq) {
            const int j = w.row_best[pd.out_base + i];
            if (j >= 0) {
                const float d1 = w.row_d1[pd.out_base + i];
                if (ratio_ok(d1, w.row_d2[pd.out_base + i], use_ratio, ratio) && w.col_best[pd.cb_base + j] == i &&
                    ratio_ok(w.col_d1[pd.cb_base + j], w.col_d2[pd.cb_base + j], use_ratio, ratio)) {
                    keep = true;
                    key = (static_cast<unsigned long long>(__float_as_uint(d1)) << 32) | static_cast<unsigned int>(i);
                }
            }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_cnt[warp] = __popc(bal);
        __syncthreads();
        int off = s_total;
        for (int k = 0; k < warp; ++k) off += warp_cnt[k];
        if (keep) gkeys[off + __popc(bal & ((1u << lane) - 1))] = key;
        __syncthreads();
        if (threadIdx.x == 0) { int s = 0; for (int k = 0; k < nwarp; ++k) s += warp_cnt[k]; s_total += s; }
        __syncthreads();
    }
    const int n = s_total;
    if (threadIdx.x == 0) match_count[p] = n;
    if (n == 0) return;
    int n2 = 1;
    while (n2 < n) n2 <<= 1;
    if (n2 <= kSortSmem) {
        for (int t = threadIdx.x; t < n2; t += blockDim.x) skeys[t] = t < n ? gkeys[t] : ~0ull;
        __syncthreads();
        bitonic_sort(skeys, n2);
        for (int t = threadIdx.x; t < n; t += blockDim.x) {
            const unsigned i = static_cast<unsigned>(skeys[t] & 0xffffffffu);
            match_slot[pd.out_base + t] = make_uint2(i, static_cast<unsigned>(w.row_best[pd.out_base + i]));
        }
    } else {
        for (int t = n + threadIdx.x; t < n2; t += blockDim.x) gkeys[t] = ~0ull;
        __syncthreads();
        bitonic_sort(gkeys, n2);
        for (int t = threadIdx.x; t < n; t += blockDim.x) {
            const unsigned i = static_cast<unsigned>(gkeys[t] & 0xffffffffu);
            match_slot[pd.out_base + t] = make_uint2(i, static_cast<unsigned>(w.row_best[pd.out_base + i]));
        }
    }
}

// exclusive prefix sum of one chunk's per-pair counts on top of the running total of the batch (single CTA)
__global__ void __launch_bounds__(1024) scan_counts_kernel(const int* counts, int n, long long* offsets, long long* running)
{
    __shared__ long long s_part[1024];
    const long long base = *running;
    const int per = (n + blockDim.x - 1) / blockDim.x;
    const int lo = threadIdx.x * per, hi = min(n, lo + per);
    long long s = 0;
    for (int k = lo; k < hi; ++k) s += counts[k];
    s_part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long run = base;
        for (int k = 0; k < static_cast<int>(blockDim.x); ++k) { const long long v = s_part[k]; s_part[k] = run; run += v; }
        offsets[n] = run;
        *running = run;
    }
    __syncthreads();
    long long run = s_part[threadIdx.x];
    for (int k = lo; k < hi; ++k) { offsets[k] = run; run += counts[k]; }
}

// gather every pair's slot into one compact (total, 2) uint32 array
__global__ void compact_matches_kernel(const PairDesc* pairs, const uint2* match_slot, const int* counts, const long long* offsets,
                                       uint2* out)
{
    const int p = blockIdx.x;
    const long long src = pairs[p].out_base, dst = offsets[p];
    for (int t = threadIdx.x; t < counts[p]; t += blockDim.x) out[dst + t] = match_slot[src + t];
}

cudaError_t launch_match_finalize(const MatchWork& w, int use_ratio, double ratio, unsigned long long* keys_scratch,
                                  uint2* match_slot, int* match_count, long long* offsets, long long* running, uint2* compact_out, cudaStream_t s)
{
    if (w.n_pairs <= 0) return cudaSuccess;
    static g2v::DeviceOnce once;
    int& configured = once.cur();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(select_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSortSmem * 8);
        if (e != cudaSuccess) return e;
        configured = 1;
    }
    select_sort_kernel<<<w.n_pairs, 512, kSortSmem * 8, s>>>(w, use_ratio, ratio, keys_scratch, match_slot, match_count);
    scan_counts_kernel<<<1, 1024, 0, s>>>(match_count, w.n_pairs, offsets, running);
    if (compact_out) compact_matches_kernel<<<w.n_pairs, 256, 0, s>>>(w.pairs, match_slot, match_count, offsets, compact_out);
    return cudaGetLastError();
}

}  // namespace g2v
