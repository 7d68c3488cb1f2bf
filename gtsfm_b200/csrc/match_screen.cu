// K1 -- descriptor-distance screening on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a only.
//
// Replaces the arithmetic of cv::BFMatcher::knnMatch / match as called by the reference at
// gtsfm/frontend/matcher/twoway_matcher.py:136,139 (both directions, one distance matrix).
//
// For a pair (image 1 = query rows q, image 2 = train rows t) the accumulator of the MMA holds
//     acc[q][t] = (|q|^2 + |t|^2) / 2 - q.t  =  d^2(q,t) / 2
// exactly (integer-valued SIFT-like descriptors <= 255 are exact in bf16, products are exact in fp32 and all partial
// sums stay below 2^24).  The norms are folded in by one extra K step of 16 bf16 ("aug" columns) so the epilogue
// never touches per-row/column constants; the main K steps run with the B operand negated.
//
// Work item = (pair, block of 256 train rows).  The train block stays in shared memory, the query image streams
// through a TMA ring in tiles of 128 rows.  Epilogue threads own one TMEM lane (= query row) and 128 train columns:
//   * query direction: minimum of every group of 16 consecutive train columns (FMNMX3 chains), then best / second-best
//     group of the 8 groups -> row partial (v1, v2, g1) per (query row, 128-column half block);
//   * train direction: running minimum per column over all query tiles (1 FMNMX per element), i.e. the minimum of
//     each "lane class" {q : q mod 128 == lane}; at the end of the item the 128 classes of a column are folded on the SM
//     (CREDUX.MIN.F32 + ballot per warp, shared memory across the four warps) into (smallest, second smallest, winning
//     class) and written as colraw[{v1, v2, class}][t].
// The exact best / second-best and their indices are recovered by the resolve kernels, which recompute only the
// winning group / lane class exactly (match_resolve.cu).  Nothing here is approximate for the integer path.
#include "ptx.cuh"
#include "types.h"

namespace g2v {

constexpr int kStages = 4;
constexpr int kThreads = 384;          // warp 0 TMA, warp 1 MMA, warp 2 TMEM alloc, warp 3 idle, warps 4..11 epilogue
constexpr int kEpiWarp0 = 4;
constexpr uint32_t kTmemCols = 512;    // two 128x256 fp32 accumulators

template <int KCH> struct ScreenSmem {
    static constexpr int kChunkQ = kTileQ * 128;            // bytes of one 64-wide K chunk of a query tile
    static constexpr int kChunkT = kBlockT * 128;
    static constexpr int kAugQ = kTileQ * kAugK * 2;
    static constexpr int kAugT = kBlockT * kAugK * 2;
    static constexpr int kTBytes = KCH * kChunkT + kAugT;
    static constexpr int kQBytes = KCH * kChunkQ + kAugQ;
    static constexpr int kOffT = 0;
    static constexpr int kOffTAug = KCH * kChunkT;
    static constexpr int kOffQ = kTBytes;
    static constexpr int kOffBar = kTBytes + kStages * kQBytes;
    static constexpr int kOffFold = kOffBar + 256;          // column fold: [2 halves][4 warps][128 columns] (v1, v2) + class bytes
    static constexpr int kFoldBytes = 2 * 4 * kHalfT * (8 + 1);
    static constexpr int kTotal = kOffFold + kFoldBytes + 1024;   // + alignment slack
};

struct ScreenArgs {
    const PairDesc* pairs;
    int n_pairs;
    int n_items;
    float* rp_v1;
    float* rp_v2;
    uint8_t* rp_g;
    float* colraw;
};

__device__ __forceinline__ int find_pair(const PairDesc* pairs, int n_pairs, int item)
{
    int lo = 0, hi = n_pairs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].item0 <= item) lo = mid; else hi = mid - 1;
    }
    return lo;
}

template <int KCH>
__global__ void __launch_bounds__(kThreads, 1)
match_screen_kernel(const __grid_constant__ CUtensorMap tm_desc, const __grid_constant__ CUtensorMap tm_augq,
                    const __grid_constant__ CUtensorMap tm_augt, const ScreenArgs args)
{
    using L = ScreenSmem<KCH>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kOffBar);
    uint64_t* q_full = bars;                 // [kStages]
    uint64_t* q_empty = bars + kStages;      // [kStages]
    uint64_t* acc_full = bars + 2 * kStages; // [2]
    uint64_t* acc_empty = acc_full + 2;      // [2]
    uint64_t* t_full = acc_empty + 2;
    uint64_t* t_empty = t_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm_desc);
        ptx::prefetch_tmap(&tm_augq);
        ptx::prefetch_tmap(&tm_augt);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) { ptx::mbar_init(&q_full[s], 1); ptx::mbar_init(&q_empty[s], 1); }
        for (int a = 0; a < 2; ++a) { ptx::mbar_init(&acc_full[a], 1); ptx::mbar_init(&acc_empty[a], 8); }
        ptx::mbar_init(t_full, 1);
        ptx::mbar_init(t_empty, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 2) {
        ptx::tmem_alloc(tmem_slot, kTmemCols);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < kEpiWarp0) {
    ptx::setmaxnreg_dec<40>();
    if (warp == 0) {
        // ===================================================== TMA producer (one elected lane)
        if (ptx::elect_one()) {
            uint32_t qcnt = 0, it = 0;
            for (int item = blockIdx.x; item < args.n_items; item += gridDim.x, ++it) {
                const int p = find_pair(args.pairs, args.n_pairs, item);
                const PairDesc pd = args.pairs[p];
                const int trow = pd.t_row0 + (item - pd.item0) * kBlockT;
                ptx::mbar_wait(t_empty, (it & 1) ^ 1);
                ptx::mbar_arrive_expect_tx(t_full, L::kTBytes);
#pragma unroll
                for (int kc = 0; kc < KCH; ++kc) {
                    ptx::tma_load_2d(smem + L::kOffT + kc * L::kChunkT, &tm_desc, kc * 64, trow, t_full);
                    ptx::tma_load_2d(smem + L::kOffT + kc * L::kChunkT + L::kChunkQ, &tm_desc, kc * 64, trow + 128, t_full);
                }
                ptx::tma_load_2d(smem + L::kOffTAug, &tm_augt, 0, trow, t_full);
                ptx::tma_load_2d(smem + L::kOffTAug + L::kAugQ, &tm_augt, 0, trow + 128, t_full);
                const int ntiles = pd.nq_pad / kTileQ;
                for (int t = 0; t < ntiles; ++t, ++qcnt) {
                    const uint32_t s = qcnt % kStages, ph = (qcnt / kStages) & 1;
                    ptx::mbar_wait(&q_empty[s], ph ^ 1);
                    uint8_t* qs = smem + L::kOffQ + s * L::kQBytes;
                    ptx::mbar_arrive_expect_tx(&q_full[s], L::kQBytes);
                    const int qrow = pd.q_row0 + t * kTileQ;
#pragma unroll
                    for (int kc = 0; kc < KCH; ++kc) ptx::tma_load_2d(qs + kc * L::kChunkQ, &tm_desc, kc * 64, qrow, &q_full[s]);
                    ptx::tma_load_2d(qs + KCH * L::kChunkQ, &tm_augq, 0, qrow, &q_full[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one elected lane)
        if (ptx::elect_one()) {
            constexpr uint32_t idesc_main = ptx::idesc_bf16(kTileQ, kBlockT, true);    // D -= Q . T^T
            constexpr uint32_t idesc_aug = ptx::idesc_bf16(kTileQ, kBlockT, false);    // D += (|q|^2 + |t|^2) / 2
            const uint32_t sbase = ptx::smem_u32(smem);
            uint32_t qcnt = 0, acnt = 0, it = 0;
            for (int item = blockIdx.x; item < args.n_items; item += gridDim.x, ++it) {
                const int p = find_pair(args.pairs, args.n_pairs, item);
                const int ntiles = args.pairs[p].nq_pad / kTileQ;
                ptx::mbar_wait(t_full, it & 1);
                for (int t = 0; t < ntiles; ++t, ++qcnt, ++acnt) {
                    const uint32_t s = qcnt % kStages, ph = (qcnt / kStages) & 1;
                    const uint32_t a = acnt & 1, aph = (acnt >> 1) & 1;
                    ptx::mbar_wait(&q_full[s], ph);
                    ptx::mbar_wait(&acc_empty[a], aph ^ 1);
                    ptx::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + a * kBlockT;
                    const uint32_t qaddr = sbase + L::kOffQ + s * L::kQBytes;
#pragma unroll
                    for (int kc = 0; kc < KCH; ++kc) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t da = ptx::smem_desc(qaddr + kc * L::kChunkQ + ks * 32, 1024, ptx::kLayoutSw128);
                            const uint64_t db = ptx::smem_desc(sbase + L::kOffT + kc * L::kChunkT + ks * 32, 1024, ptx::kLayoutSw128);
                            ptx::umma_bf16(d_tmem, da, db, idesc_main, (kc | ks) != 0);
                        }
                    }
                    {
                        const uint64_t da = ptx::smem_desc(qaddr + KCH * L::kChunkQ, 256, ptx::kLayoutSw32);
                        const uint64_t db = ptx::smem_desc(sbase + L::kOffTAug, 256, ptx::kLayoutSw32);
                        ptx::umma_bf16(d_tmem, da, db, idesc_aug, 1);
                    }
                    ptx::umma_commit(&q_empty[s]);
                    ptx::umma_commit(&acc_full[a]);
                }
                ptx::umma_commit(t_empty);
            }
        }
    }
    } else {
        ptx::setmaxnreg_inc<232>();
        // ===================================================== epilogue: 8 warps, lane = query row, 128 train columns each
        const int ew = warp - kEpiWarp0;
        const int half = ew >> 2;                       // which 128-column half of the 256-column accumulator
        const int lane_base = (warp & 3) * 32;          // TMEM lanes this warp may access
        const int m_local = lane_base + lane;
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(lane_base) << 16) + half * kHalfT;
        uint32_t acnt = 0;
        float cmin[kHalfT];
        for (int item = blockIdx.x; item < args.n_items; item += gridDim.x) {
            const int p = find_pair(args.pairs, args.n_pairs, item);
            const int blk = item - args.pairs[p].item0;
            const int nq_pad = args.pairs[p].nq_pad;
            const int ntiles = nq_pad / kTileQ;
            long long e = args.pairs[p].rp_base + static_cast<long long>(blk * 2 + half) * nq_pad + m_local;
#pragma unroll
            for (int i = 0; i < kHalfT; ++i) cmin[i] = kHuge * 4.0f;
            for (int t = 0; t < ntiles; ++t, ++acnt) {
                const uint32_t a = acnt & 1, aph = (acnt >> 1) & 1;
                ptx::mbar_wait(&acc_full[a], aph);
                ptx::tc_fence_after();
                float v1 = kHuge * 4.0f, v2 = kHuge * 4.0f;
                int g1 = 0;
#pragma unroll
                for (int c = 0; c < kHalfT / kGroup; ++c) {
                    float v[kGroup];
                    ptx::tmem_ld_32x32b_x16(t_lane + a * kBlockT + c * kGroup, v);
#pragma unroll
                    for (int i = 0; i < kGroup; ++i) cmin[c * kGroup + i] = fminf(cmin[c * kGroup + i], v[i]);
                    float m = fminf(fminf(v[0], v[1]), v[2]);
#pragma unroll
                    for (int i = 3; i < 15; i += 2) m = fminf(fminf(m, v[i]), v[i + 1]);
                    m = fminf(m, v[15]);
                    const bool lt = m < v1;                      // strict: the lowest group id wins ties
                    v2 = fminf(v2, fmaxf(v1, m));
                    g1 = lt ? c : g1;
                    v1 = fminf(v1, m);
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&acc_empty[a]);
                args.rp_v1[e] = v1;
                args.rp_v2[e] = v2;
                args.rp_g[e] = static_cast<uint8_t>(g1);
                e += kTileQ;
            }
            // ---- fold the column minima over the 128 lane classes before they leave the SM: per column the smallest value,
            // the second smallest and the lowest class that holds the smallest (ties as merge_cols resolved them: ascending
            // class order, strict <).  Inside a warp two CREDUX.MIN.F32 and a ballot per column; the four warps of a half meet
            // in shared memory and thread m_local finishes column m_local.  9 bytes per column instead of 512.
            uint2* s_fold = reinterpret_cast<uint2*>(smem + L::kOffFold) + (half * 4) * kHalfT;
            uint8_t* s_fcls = smem + L::kOffFold + 2 * 4 * kHalfT * 8 + (half * 4) * kHalfT;
            const int wq = warp & 3;
#pragma unroll
            for (int i = 0; i < kHalfT; ++i) {
                const float v = cmin[i];
                const float mn = ptx::redux_min_f32(v);
                const int l1 = __ffs(__ballot_sync(0xffffffffu, v == mn)) - 1;
                const float mn2 = ptx::redux_min_f32(lane == l1 ? __int_as_float(0x7f800000) : v);
                if (lane == (i & 31)) {
                    s_fold[wq * kHalfT + i] = make_uint2(__float_as_uint(mn), __float_as_uint(mn2));
                    s_fcls[wq * kHalfT + i] = static_cast<uint8_t>(lane_base + l1);
                }
            }
            ptx::named_bar_sync(1 + half, kHalfT);
            {
                float k1 = kHuge * 8.0f, k2 = kHuge * 8.0f;
                int c1 = 0;
#pragma unroll
                for (int w4 = 0; w4 < 4; ++w4) {
                    const uint2 e2 = s_fold[w4 * kHalfT + m_local];
                    const float a = __uint_as_float(e2.x), b2 = __uint_as_float(e2.y);
                    const bool lt = a < k1;                      // strict: the lowest class wins ties
                    k2 = lt ? fminf(k1, b2) : fminf(k2, a);
                    c1 = lt ? s_fcls[w4 * kHalfT + m_local] : c1;
                    k1 = fminf(k1, a);
                }
                const int nt_pad = args.pairs[p].nt_pad;
                float* cf = args.colraw + args.pairs[p].cr_base + blk * kBlockT + half * kHalfT + m_local;
                cf[0] = k1;
                cf[nt_pad] = k2;
                cf[2 * static_cast<long long>(nt_pad)] = __int_as_float(c1);
            }
            ptx::named_bar_sync(1 + half, kHalfT);               // the fold buffer is rewritten at the end of the next item
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

template <int KCH>
static cudaError_t launch_screen_t(const CUtensorMap& tm_desc, const CUtensorMap& tm_augq, const CUtensorMap& tm_augt,
                                   const ScreenArgs& args, int n_sm, cudaStream_t stream)
{
    using L = ScreenSmem<KCH>;
    static g2v::DeviceOnce once;
    int& configured = once.cur();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(match_screen_kernel<KCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
        if (e != cudaSuccess) return e;
        configured = 1;
    }
    const int grid = args.n_items < n_sm ? args.n_items : n_sm;
    if (grid <= 0) return cudaSuccess;
    match_screen_kernel<KCH><<<grid, kThreads, L::kTotal, stream>>>(tm_desc, tm_augq, tm_augt, args);
    return cudaGetLastError();
}

// kpad = descriptor dimension padded to a multiple of 64 (64 or 128 supported by the stationary-block kernel).
cudaError_t launch_match_screen(const CUtensorMap& tm_desc, const CUtensorMap& tm_augq, const CUtensorMap& tm_augt,
                                const PairDesc* pairs, int n_pairs, int n_items, float* rp_v1, float* rp_v2, uint8_t* rp_g,
                                float* colraw, int kpad, int n_sm, cudaStream_t stream)
{
    ScreenArgs a{pairs, n_pairs, n_items, rp_v1, rp_v2, rp_g, colraw};
    if (kpad == 64) return launch_screen_t<1>(tm_desc, tm_augq, tm_augt, a, n_sm, stream);
    if (kpad == 128) return launch_screen_t<2>(tm_desc, tm_augq, tm_augt, a, n_sm, stream);
    return cudaErrorInvalidValue;
}

}  // namespace g2v
