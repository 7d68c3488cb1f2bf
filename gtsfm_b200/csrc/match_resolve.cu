// K2 -- exact nearest / second-nearest neighbours per query row and per train column, from the screening output.
//
// Restates, on the GPU, what cv::BFMatcher::knnMatch(k=2) returns for each direction (reference call sites
// gtsfm/frontend/matcher/twoway_matcher.py:136,139): the best train index with the LOWEST index among equal float32
// distances, the best distance sqrtf(d^2) and the second-best distance.
//
// Integer path: the screening kernel gives, per query row, the best and second-best GROUP minimum (groups of 16 train
// columns) and, per train column, the minimum of each of the 128 lane classes.  Only the winning group / class is
// recomputed here, exactly, with integer dot products (dp4a) on the uint8 copy of the descriptors.  Whenever the
// runner-up group (class) could tie with the winner in float32-distance space the row (column) is flagged and
// re-scanned exhaustively by exact_scan_kernel (match_exact.cu), so the result never depends on the screening order.
//
// Four kernels per direction (a counting sort of the rows / columns by the descriptor block they must be compared with):
//   merge   : reduces the screening partials of a row (column) to (k1, k2, bucket) and takes a slot in the bucket
//             (one atomicAdd on the bucket counter); coalesced, HBM streaming.
//   prefix  : bucket counters -> exclusive offsets, one CTA per pair.
//   scatter : writes every row (column) into its pair's member list at offset[bucket] + slot.
//   bucket  : one CTA per bucket -- (pair, 128-column half block) for rows, (pair, lane class) for columns.  The CTA
//             stages the <= 128 (class: nq/128) descriptor rows every member of the bucket is compared against in
//             shared memory ONCE and recomputes the candidates of its members from there.  Per pair this reads each
//             descriptor row about once (2 MB) instead of gathering the same rows once per member through L2 (80 MB
//             for 8192 x 8192).
#include <cfloat>

#include "store.h"

namespace g2v {

__device__ __forceinline__ int find_by_out_base(const PairDesc* pairs, int n, long long pos)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].out_base <= pos) lo = mid; else hi = mid - 1;
    }
    return lo;
}
__device__ __forceinline__ int find_by_cb_base(const PairDesc* pairs, int n, long long pos)
{
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].cb_base <= pos) lo = mid; else hi = mid - 1;
    }
    return lo;
}
__device__ __forceinline__ void flag(const MatchWork& w, int pair, int dir, int idx)
{
    const int k = atomicAdd(w.n_flag, 1);
    if (k < w.flag_cap)
        w.flag_list[k] = (static_cast<unsigned long long>(pair) << 33) | (static_cast<unsigned long long>(dir) << 32) |
                         static_cast<unsigned int>(idx);
}

// float32 distance exactly as cv2 produces it for integer-valued descriptors: the squared distance is an exact
// integer below 2^24 in any summation order, then sqrtf (correctly rounded).
__device__ __forceinline__ float dist_from_d2(int d2) { return sqrtf(static_cast<float>(d2)); }
__device__ __forceinline__ float dist_from_half(float h) { return h >= kHuge * 0.25f ? FLT_MAX : sqrtf(h + h); }

// ---------------------------------------------------------------------------------------------------------------
// merge, rows: one thread per query row.  row_d1 <- k1, row_d2 <- k2 (halved squared distances), row_key <- winning
// half block << 3 | winning group | slot in the bucket << 16.
__global__ void __launch_bounds__(128) merge_rows_kernel(StoreView st, MatchWork w, long long total_rows)
{
    const long long pos = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= total_rows) return;
    const int p = find_by_out_base(w.pairs, w.n_pairs, static_cast<long long>(blockIdx.x) * blockDim.x);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0) return;
    const int i = static_cast<int>(pos - pd.out_base);
    if (i >= pd.nq || !st.valid[pd.q_row0 + i]) {
        w.row_key[pos] = 0xffffffffu;
        if (i < pd.nq) { w.row_best[pos] = -1; w.row_d1[pos] = FLT_MAX; w.row_d2[pos] = FLT_MAX; }
        return;
    }
    const int n_hb = pd.nt_pad / kHalfT;
    float k1 = kHuge * 8.0f, k2 = kHuge * 8.0f;
    int hb1 = 0;
    const long long e0 = pd.rp_base + i;
    // the loads of a batch are independent of the running minimum, so they are all in flight together; the group id of
    // the winner is fetched once at the end (a load inside the compare would serialise on every new minimum)
#pragma unroll 8
    for (int hb = 0; hb < n_hb; ++hb) {
        const long long e = e0 + static_cast<long long>(hb) * pd.nq_pad;
        const float a = __ldg(w.rp_v1 + e), b = __ldg(w.rp_v2 + e);
        const bool lt = a < k1;                                   // strict: the lowest half block wins ties
        k2 = lt ? fminf(k1, b) : fminf(k2, a);
        hb1 = lt ? hb : hb1;
        k1 = fminf(k1, a);
    }
    const int g1 = w.rp_g[e0 + static_cast<long long>(hb1) * pd.nq_pad];
    w.row_d1[pos] = k1;
    w.row_d2[pos] = k2;
    const unsigned slot = static_cast<unsigned>(atomicAdd(w.row_cnt + static_cast<long long>(p) * (8 * w.max_hb + 1) + ((hb1 << 3) | g1), 1));
    w.row_key[pos] = static_cast<unsigned>((hb1 << 3) | g1) | (slot << 16);
}

// merge, columns: the screening kernel already folded the 128 lane classes of every column (match_screen.cu), so this is a
// 12-byte read per column: col_d1 <- k1, col_d2 <- k2, col_key <- winning lane class | slot << 8.  Every CTA lies inside one
// pair because column ranges start on multiples of 256, so the pair lookup is done once per CTA.
constexpr int kColsPerCta = 256;
__global__ void __launch_bounds__(256) merge_cols_kernel(StoreView st, MatchWork w, long long total_cols)
{
    const long long pos0 = static_cast<long long>(blockIdx.x) * kColsPerCta;
    const long long pos = pos0 + threadIdx.x;
    if (pos >= total_cols) return;
    const int p = find_by_cb_base(w.pairs, w.n_pairs, pos0);
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0) return;
    const int j = static_cast<int>(pos - pd.cb_base);
    if (j >= pd.nt_pad) return;
    const bool live = j < pd.nt && st.valid[pd.t_row0 + j];
    if (live) {
        const float* cf = w.colraw + pd.cr_base + j;
        const int c1 = __float_as_int(__ldg(cf + 2 * static_cast<long long>(pd.nt_pad)));
        w.col_d1[pos] = __ldg(cf);
        w.col_d2[pos] = __ldg(cf + pd.nt_pad);
        const unsigned slot = static_cast<unsigned>(atomicAdd(w.col_cnt + static_cast<long long>(p) * (kTileQ + 1) + c1, 1));
        w.col_key[pos] = static_cast<unsigned>(c1) | (slot << 8);
    } else {
        w.col_key[pos] = 0xffffffffu;
        if (j < pd.nt) { w.col_best[pos] = -1; w.col_d1[pos] = FLT_MAX; w.col_d2[pos] = FLT_MAX; }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// prefix: cnt[p][0 .. stride - 2] (bucket sizes) -> exclusive offsets in place, cnt[p][stride - 1] <- the total.
__global__ void __launch_bounds__(128) prefix_counts_kernel(int* cnt, int stride)
{
    __shared__ int s_w[4];
    __shared__ int s_carry;
    int* c = cnt + static_cast<long long>(blockIdx.x) * stride;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < stride; base += 128) {
        const int k = base + tid;
        const int v = (k < stride - 1) ? c[k] : 0;
        int inc = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int o = __shfl_up_sync(0xffffffffu, inc, off); if (lane >= off) inc += o; }
        if (lane == 31) s_w[warp] = inc;
        __syncthreads();
        int before = s_carry;
        for (int w2 = 0; w2 < warp; ++w2) before += s_w[w2];
        if (k < stride) c[k] = before + inc - v;
        __syncthreads();
        if (tid == 127) s_carry = before + inc;
        __syncthreads();
    }
}

// scatter: member lists.  Rows: list[out_base + offset[half block << 3 | group] + slot] = row;
// columns: list[cb_base + offset[class] + slot] = column.
__global__ void __launch_bounds__(256) scatter_rows_kernel(MatchWork w, long long total_rows)
{
    const long long pos = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= total_rows) return;
    const int p = find_by_out_base(w.pairs, w.n_pairs, static_cast<long long>(blockIdx.x) * blockDim.x);
    if (w.pairs[p].path != 0) return;
    const unsigned key = w.row_key[pos];
    if (key == 0xffffffffu) return;
    const long long ob = w.pairs[p].out_base;
    const int off = w.row_cnt[static_cast<long long>(p) * (8 * w.max_hb + 1) + static_cast<int>(key & 0xffffu)];
    w.row_list[ob + off + (key >> 16)] = static_cast<int>(pos - ob);
}
__global__ void __launch_bounds__(256) scatter_cols_kernel(MatchWork w, long long total_cols)
{
    const long long pos = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= total_cols) return;
    const int p = find_by_cb_base(w.pairs, w.n_pairs, static_cast<long long>(blockIdx.x) * blockDim.x);
    if (w.pairs[p].path != 0) return;
    const unsigned key = w.col_key[pos];
    if (key == 0xffffffffu) return;
    const long long cb = w.pairs[p].cb_base;
    const int off = w.col_cnt[static_cast<long long>(p) * (kTileQ + 1) + static_cast<int>(key & 0xffu)];
    w.col_list[cb + off + (key >> 8)] = static_cast<int>(pos - cb);
}

// ---------------------------------------------------------------------------------------------------------------
// Exact integer dot products of the bucket kernels: warp-level IMMA, D(16 x 8, s32) += A(16 x 32, u8) * B(32 x 8, u8).
// A rows = members of the bucket (gathered from the store straight into the fragment registers), B columns = the staged
// descriptor rows they are compared with, so a staged row read from shared memory serves 16 members at once (a thread
// per member would read it once per member and leave the kernel bound by shared-memory bandwidth).
// Fragment layout (g = lane >> 2, q = lane & 3): a0 / a2 = row g, a1 / a3 = row g + 8, b0 / b1 = column g; the k index
// of a register only has to agree between A and B, so lane q simply takes bytes [64 s + 16 q, +16) of its rows: one
// 16-byte load feeds two k-steps.  c0, c1 = (row g, columns 2q, 2q + 1), c2, c3 = (row g + 8, same columns).
__device__ __forceinline__ void imma_u8(int (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// staged row stride in 16-byte units, = 4 (mod 8): the eight lanes of a quarter warp then read eight different 16-byte
// bank groups when they fetch their B fragments
template <int N16> struct Staged { static constexpr int kStride = (N16 % 8 == 4) ? N16 : N16 + 4; };

// Running (best, second best, index of the best) over integer squared distances.  sqrtf is strictly increasing on the
// integers below 2^22 (neighbouring square roots are more than one float32 ulp apart), so the order of the float32
// distances cv2 compares is the order of the integers; candidates are visited by ascending index and '<' is strict, so
// the lowest index wins ties.  A best distance of 2^22 or more makes the caller flag the entry for the exhaustive scan.
constexpr int kSqrtMonotoneBelow = 1 << 22;
struct Top2 {
    int b1, b2, i1;
    __device__ __forceinline__ void push(int d2, int idx)
    {
        const bool lt = d2 < b1;
        b2 = lt ? b1 : min(b2, d2);
        i1 = lt ? idx : i1;
        b1 = min(b1, d2);
    }
    // merge with the partner lane `off` lanes away (all lanes of the warp call this)
    __device__ __forceinline__ void merge_xor(int off)
    {
        const int o1 = __shfl_xor_sync(0xffffffffu, b1, off), o2 = __shfl_xor_sync(0xffffffffu, b2, off);
        const int oi = __shfl_xor_sync(0xffffffffu, i1, off);
        const bool take = (o1 < b1) || (o1 == b1 && oi < i1);
        const int lose = take ? b1 : o1;
        b2 = min(min(b2, o2), lose);
        if (take) { b1 = o1; i1 = oi; }
    }
};
__device__ __forceinline__ float dist_or_max(int d2) { return d2 == INT_MAX ? FLT_MAX : dist_from_d2(d2); }

// bucket, rows: one CTA per (pair, half block of 128 train columns).  The members of the half block are sorted by group;
// a warp takes 16 members of one group at a time and multiplies them with the group's 16 staged train rows.
constexpr int kBucketRowThreads = 512;   // 8 warps: a half block has ~12 tiles of 16 members, i.e. two rounds of a gather-latency-bound chain
template <int N16>
__global__ void __launch_bounds__(kBucketRowThreads, 2) bucket_rows_kernel(StoreView st, MatchWork w)
{
    constexpr int kStride = Staged<N16>::kStride;
    __shared__ uint4 s_t[kHalfT * kStride];
    __shared__ int s_norm[kHalfT];
    __shared__ uint8_t s_valid[kHalfT];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, q = lane & 3;
    const int p = blockIdx.x / w.max_hb, hb = blockIdx.x - p * w.max_hb;
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0 || pd.nq == 0 || hb * kHalfT >= pd.nt) return;   // padding columns are never a winner
    const int* off = w.row_cnt + static_cast<long long>(p) * (8 * w.max_hb + 1) + 8 * hb;
    int o[kHalfT / kGroup + 1];
#pragma unroll
    for (int g = 0; g <= kHalfT / kGroup; ++g) o[g] = off[g];
    if (o[kHalfT / kGroup] == o[0]) return;
    int tiles = 0;
#pragma unroll
    for (int g = 0; g < kHalfT / kGroup; ++g) tiles += (o[g + 1] - o[g] + 15) >> 4;

    // Registers of one tile of 16 members: everything that comes from global memory.  The first tile of every warp is
    // fetched BEFORE the train rows are staged, so its gather latency hides behind the staging loads and the barrier
    // (both were ~28 % of the stall samples each).
    int g = 0, n_g = 0, i_lo = 0, i_hi = 0, n_lo = 0, n_hi = 0;
    bool act_lo = false, act_hi = false;
    long long pos = 0;
    float k1 = 0.f, k2 = 0.f;
    uint4 L[N16 / 4], H[N16 / 4];
    auto fetch = [&](int tile) {
        int local = tile, first = o[0];
        g = 0; n_g = o[1] - o[0];
#pragma unroll
        for (int gg = 0; gg < kHalfT / kGroup; ++gg) {
            const int tg = (o[gg + 1] - o[gg] + 15) >> 4;
            const bool here = local >= 0 && local < tg;
            if (here) { g = gg; first = o[gg] + 16 * local; n_g = o[gg + 1] - first; }
            local -= here ? (1 << 30) : tg;
        }
        const int* list = w.row_list + pd.out_base + first;
        act_lo = gid < n_g; act_hi = gid + 8 < n_g;
        i_lo = act_lo ? list[gid] : list[0]; i_hi = act_hi ? list[gid + 8] : i_lo;
        const uint4* r_lo = reinterpret_cast<const uint4*>(st.desc_u8) + static_cast<long long>(pd.q_row0 + i_lo) * N16 + q;
        const uint4* r_hi = reinterpret_cast<const uint4*>(st.desc_u8) + static_cast<long long>(pd.q_row0 + i_hi) * N16 + q;
#pragma unroll
        for (int k = 0; k < N16 / 4; ++k) { L[k] = __ldg(r_lo + 4 * k); H[k] = __ldg(r_hi + 4 * k); }
        n_lo = st.norm_i[pd.q_row0 + i_lo]; n_hi = st.norm_i[pd.q_row0 + i_hi];
        pos = pd.out_base + ((q & 1) ? i_hi : i_lo);                       // lane q = 0 finishes row g, lane q = 1 row g + 8
        k1 = w.row_d1[pos]; k2 = w.row_d2[pos];                            // screened values
    };
    if (warp < tiles) fetch(warp);

    const long long trow0 = static_cast<long long>(pd.t_row0) + hb * kHalfT;
    const uint4* tsrc = reinterpret_cast<const uint4*>(st.desc_u8) + trow0 * N16;
#pragma unroll
    for (int it = 0; it < kHalfT * N16 / kBucketRowThreads; ++it) {         // all loads of the block in flight at once
        const int idx = tid + it * kBucketRowThreads;
        const int r = idx / N16, k = idx % N16;
        s_t[r * kStride + k] = __ldg(tsrc + idx);
    }
    if (tid < kHalfT) {
        s_norm[tid] = st.norm_i[trow0 + tid];
        s_valid[tid] = (hb * kHalfT + tid < pd.nt) ? st.valid[trow0 + tid] : 0;
    }
    __syncthreads();

    for (int tile = warp; tile < tiles; tile += kBucketRowThreads / 32) {
        if (tile != warp) fetch(tile);
        Top2 tl{INT_MAX, INT_MAX, INT_MAX}, th{INT_MAX, INT_MAX, INT_MAX};
#pragma unroll
        for (int nt = 0; nt < kGroup / 8; ++nt) {
            int acc[4] = {0, 0, 0, 0};
            const uint4* brow = s_t + (g * kGroup + nt * 8 + gid) * kStride + q;
#pragma unroll
            for (int k = 0; k < N16 / 4; ++k) {
                const uint4 B = brow[4 * k];
                imma_u8(acc, L[k].x, H[k].x, L[k].y, H[k].y, B.x, B.y);
                imma_u8(acc, L[k].z, H[k].z, L[k].w, H[k].w, B.z, B.w);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int r = g * kGroup + nt * 8 + 2 * q + u;
                if (s_valid[r]) {
                    const int j = hb * kHalfT + r, nn = s_norm[r];
                    tl.push(n_lo + nn - 2 * acc[u], j);
                    th.push(n_hi + nn - 2 * acc[2 + u], j);
                }
            }
        }
        tl.merge_xor(1); tl.merge_xor(2);
        th.merge_xor(1); th.merge_xor(2);
        if ((q == 0 && act_lo) || (q == 1 && act_hi)) {
            const Top2 t = (q == 0) ? tl : th;
            const int i = (q == 0) ? i_lo : i_hi;
            const float f1 = dist_or_max(t.b1), f2 = dist_or_max(t.b2), f2o = dist_from_half(k2);
            // Ambiguous when another group may hold an equal (float32) distance at a lower index, or when the
            // recomputed minimum does not reproduce the screened value (never expected; keeps the result
            // independent of screening).
            const bool ambiguous = (t.i1 == INT_MAX) || (f2o <= f1) || (static_cast<float>(t.b1) != k1 + k1) || (t.b1 >= kSqrtMonotoneBelow);
            if (ambiguous) flag(w, p, 0, i);
            w.row_best[pos] = (t.i1 == INT_MAX) ? -1 : t.i1;
            w.row_d1[pos] = f1;
            w.row_d2[pos] = fminf(f2, f2o);
        }
    }
}

// bucket, columns: one CTA per (pair, lane class c).  The class rows c, c + 128, ... of the query image are staged in
// shared memory (padded to a multiple of eight rows); a warp takes 16 member columns at a time.
template <int N16>
__global__ void __launch_bounds__(128) bucket_cols_kernel(StoreView st, MatchWork w)
{
    constexpr int kStride = Staged<N16>::kStride;
    extern __shared__ uint4 s_dyn[];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, q = lane & 3;
    const int p = blockIdx.x >> 7, c = blockIdx.x & 127;
    const PairDesc pd = w.pairs[p];
    if (pd.path != 0 || pd.nq == 0 || pd.nt == 0 || c >= pd.nq) return;
    const int* off = w.col_cnt + static_cast<long long>(p) * (kTileQ + 1) + c;
    const int start = off[0], n = off[1] - start;
    if (n <= 0) return;
    const int n_cls = pd.nq_pad / kTileQ, n_cls8 = (n_cls + 7) & ~7;
    uint4* s_q = s_dyn;
    int* s_norm = reinterpret_cast<int*>(s_dyn + n_cls8 * kStride);
    uint8_t* s_valid = reinterpret_cast<uint8_t*>(s_norm + n_cls8);
    // registers of one tile of 16 member columns: everything that comes from global memory.  A warp's first tile is fetched
    // BEFORE the class rows are staged, so the two dependent gathers (member list -> descriptor rows) hide behind the staging
    // loads and the barrier, as in bucket_rows
    const int* list = w.col_list + pd.cb_base + start;
    bool act_lo = false, act_hi = false;
    int j_lo = 0, j_hi = 0, n_lo = 0, n_hi = 0;
    long long pos = 0;
    float k1 = 0.f, k2 = 0.f;
    uint4 L[N16 / 4], H[N16 / 4];
    auto fetch = [&](int first) {
        act_lo = first + gid < n; act_hi = first + gid + 8 < n;
        j_lo = act_lo ? list[first + gid] : list[first]; j_hi = act_hi ? list[first + gid + 8] : j_lo;
        const uint4* r_lo = reinterpret_cast<const uint4*>(st.desc_u8) + static_cast<long long>(pd.t_row0 + j_lo) * N16 + q;
        const uint4* r_hi = reinterpret_cast<const uint4*>(st.desc_u8) + static_cast<long long>(pd.t_row0 + j_hi) * N16 + q;
#pragma unroll
        for (int k = 0; k < N16 / 4; ++k) { L[k] = __ldg(r_lo + 4 * k); H[k] = __ldg(r_hi + 4 * k); }
        n_lo = st.norm_i[pd.t_row0 + j_lo]; n_hi = st.norm_i[pd.t_row0 + j_hi];
        pos = pd.cb_base + ((q & 1) ? j_hi : j_lo);
        k1 = w.col_d1[pos]; k2 = w.col_d2[pos];
    };
    if (16 * warp < n) fetch(16 * warp);
    const uint4* qsrc = reinterpret_cast<const uint4*>(st.desc_u8) + (static_cast<long long>(pd.q_row0) + c) * N16;
    for (int idx = tid; idx < n_cls8 * N16; idx += 128) {
        const int t = idx / N16, k = idx % N16;
        s_q[t * kStride + k] = (t < n_cls) ? __ldg(qsrc + static_cast<long long>(t) * kTileQ * N16 + k) : make_uint4(0u, 0u, 0u, 0u);
    }
    for (int t = tid; t < n_cls8; t += 128) {
        const bool real = t < n_cls && c + t * kTileQ < pd.nq;
        const long long row = static_cast<long long>(pd.q_row0) + c + t * kTileQ;
        s_norm[t] = real ? st.norm_i[row] : 0;
        s_valid[t] = real ? st.valid[row] : 0;
    }
    __syncthreads();

    for (int first = 16 * warp; first < n; first += 64) {
        if (first != 16 * warp) fetch(first);
        Top2 tl{INT_MAX, INT_MAX, INT_MAX}, th{INT_MAX, INT_MAX, INT_MAX};
#pragma unroll 2
        for (int nt = 0; nt < n_cls8 / 8; ++nt) {
            int acc[4] = {0, 0, 0, 0};
            const uint4* brow = s_q + (nt * 8 + gid) * kStride + q;
#pragma unroll
            for (int k = 0; k < N16 / 4; ++k) {
                const uint4 B = brow[4 * k];
                imma_u8(acc, L[k].x, H[k].x, L[k].y, H[k].y, B.x, B.y);
                imma_u8(acc, L[k].z, H[k].z, L[k].w, H[k].w, B.z, B.w);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int t = nt * 8 + 2 * q + u;
                if (s_valid[t]) {
                    const int i = c + t * kTileQ, nn = s_norm[t];
                    tl.push(n_lo + nn - 2 * acc[u], i);
                    th.push(n_hi + nn - 2 * acc[2 + u], i);
                }
            }
        }
        tl.merge_xor(1); tl.merge_xor(2);
        th.merge_xor(1); th.merge_xor(2);
        if ((q == 0 && act_lo) || (q == 1 && act_hi)) {
            const Top2 t2 = (q == 0) ? tl : th;
            const int j = (q == 0) ? j_lo : j_hi;
            const float f1 = dist_or_max(t2.b1), f2 = dist_or_max(t2.b2), f2o = dist_from_half(k2);
            const bool ambiguous = (t2.i1 == INT_MAX) || (f2o <= f1) || (static_cast<float>(t2.b1) != k1 + k1) || (t2.b1 >= kSqrtMonotoneBelow);
            if (ambiguous) flag(w, p, 1, j);
            w.col_best[pos] = (t2.i1 == INT_MAX) ? -1 : t2.i1;
            w.col_d1[pos] = f1;
            w.col_d2[pos] = fminf(f2, f2o);
        }
    }
}

template <int N16> static cudaError_t launch_bucket_cols(const StoreView& st, const MatchWork& w, int max_class_rows, cudaStream_t s)
{
    static g2v::DeviceOnce once;
    int& configured = once.cur();
    const int smem = ((max_class_rows + 7) & ~7) * (Staged<N16>::kStride * 16 + 5) + 16;
    if (smem > configured) {
        const int want = kMaxClassRows * (Staged<N16>::kStride * 16 + 5) + 16;
        cudaError_t e = cudaFuncSetAttribute(bucket_cols_kernel<N16>, cudaFuncAttributeMaxDynamicSharedMemorySize, want);
        if (e != cudaSuccess) return e;
        configured = want;
    }
    bucket_cols_kernel<N16><<<static_cast<unsigned>(w.n_pairs) * kTileQ, 128, smem, s>>>(st, w);
    return cudaGetLastError();
}

// which: 0 = query rows, 1 = train columns.
cudaError_t launch_resolve_int(const StoreView& st, const MatchWork& w, int which, long long total, int max_class_rows, cudaStream_t s)
{
    if (total <= 0 || w.n_pairs <= 0 || w.max_hb <= 0) return cudaSuccess;
    if (st.kpad != 64 && st.kpad != 128) return cudaErrorInvalidValue;
    cudaError_t e;
    if (which == 0) {
        const int stride = 8 * w.max_hb + 1;
        if ((e = cudaMemsetAsync(w.row_cnt, 0, sizeof(int) * static_cast<size_t>(w.n_pairs) * stride, s)) != cudaSuccess) return e;
        merge_rows_kernel<<<static_cast<unsigned>((total + 127) / 128), 128, 0, s>>>(st, w, total);
        prefix_counts_kernel<<<w.n_pairs, 128, 0, s>>>(w.row_cnt, stride);
        scatter_rows_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s>>>(w, total);
        const unsigned grid = static_cast<unsigned>(w.n_pairs) * w.max_hb;
        if (st.kpad == 64) bucket_rows_kernel<4><<<grid, kBucketRowThreads, 0, s>>>(st, w);
        else bucket_rows_kernel<8><<<grid, kBucketRowThreads, 0, s>>>(st, w);
    } else {
        const int stride = kTileQ + 1;
        if ((e = cudaMemsetAsync(w.col_cnt, 0, sizeof(int) * static_cast<size_t>(w.n_pairs) * stride, s)) != cudaSuccess) return e;
        merge_cols_kernel<<<static_cast<unsigned>((total + kColsPerCta - 1) / kColsPerCta), 256, 0, s>>>(st, w, total);
        prefix_counts_kernel<<<w.n_pairs, 128, 0, s>>>(w.col_cnt, stride);
        scatter_cols_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s>>>(w, total);
        e = st.kpad == 64 ? launch_bucket_cols<4>(st, w, max_class_rows, s) : launch_bucket_cols<8>(st, w, max_class_rows, s);
        if (e != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

}  // namespace g2v
