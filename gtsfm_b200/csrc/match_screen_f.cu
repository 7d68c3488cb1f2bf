// K1f -- descriptor-distance screening of FLOAT descriptors on the 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
// Same work decomposition and fused epilogue as match_screen.cu (work item = pair x 256 stationary train rows, query tiles
// of 128 rows streamed by TMA, row partials per 128-column half block and column minima per lane class), but
//   * the operands are the fp16 copies of the descriptors (11-bit significands: |q.t - q^.t^| <= 2^-10 |q||t|), so the
//     accumulator holds d^2 / 2 only up to the error bound the resolve kernels apply (match_resolve_f.cu);
//   * descriptors of up to 256 dimensions: the train block takes KCH x 32 KB of shared memory, so the query tiles stream
//     through a ring of 64-column K CHUNKS (16 KB each, KCH + 1 slots) instead of whole tiles.
// cv2's arithmetic (twoway_matcher.py:136,139 -> cv::batchDistance) is reproduced by the exact recompute of the winning
// group / lane class and, for ambiguous rows, by the exhaustive fp32 scan.
#include "ptx.cuh"
#include "types.h"

namespace g2v {

constexpr int kThreadsF = 384;          // warp 0 TMA, warp 1 MMA, warp 2 TMEM alloc, warp 3 idle, warps 4..11 epilogue
constexpr int kEpiWarp0F = 4;
constexpr uint32_t kTmemColsF = 512;    // two 128x256 fp32 accumulators
// kGroupF (train columns per group of the row partials) and kFlushTiles (query tiles per column-minimum flush): types.h

template <int KCH> struct ScreenSmemF {
    static constexpr int kSlots = KCH + 1;                  // ring of 64-wide K chunks of query tiles (see the aug hazard note)
    static constexpr int kChunkQ = kTileQ * 128;            // bytes of one 64-wide K chunk of a query tile
    static constexpr int kChunkT = kBlockT * 128;
    static constexpr int kAugQ = kTileQ * kAugK * 2;
    static constexpr int kAugT = kBlockT * kAugK * 2;
    static constexpr int kTBytes = KCH * kChunkT + kAugT;
    static constexpr int kOffT = 0;
    static constexpr int kOffTAug = KCH * kChunkT;
    static constexpr int kOffQ = kTBytes;
    static constexpr int kOffQAug = kOffQ + kSlots * kChunkQ;   // two aug buffers (tile parity)
    static constexpr int kOffBar = kOffQAug + 2 * kAugQ;
    static constexpr int kTotal = kOffBar + 256 + 1024;     // barriers + alignment slack
};

struct ScreenArgsF {
    const PairDesc* pairs;
    int n_pairs;
    int n_items;
    float* rp_v1;
    float* rp_v2;
    uint8_t* rp_g;
    float* colraw;
};

__device__ __forceinline__ int find_pair_f(const PairDesc* pairs, int n_pairs, int item)
{
    int lo = 0, hi = n_pairs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].fitem0 <= item) lo = mid; else hi = mid - 1;
    }
    return lo;
}

template <int KCH>
__global__ void __launch_bounds__(kThreadsF, 1)
match_screen_f_kernel(const __grid_constant__ CUtensorMap tm_desc, const __grid_constant__ CUtensorMap tm_augq,
                    const __grid_constant__ CUtensorMap tm_augt, const ScreenArgsF args)
{
    using L = ScreenSmemF<KCH>;
    constexpr int NS = L::kSlots;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kOffBar);
    uint64_t* q_full = bars;                 // [NS]
    uint64_t* q_empty = bars + NS;           // [NS]
    uint64_t* acc_full = bars + 2 * NS;      // [2]
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
        for (int s = 0; s < NS; ++s) { ptx::mbar_init(&q_full[s], 1); ptx::mbar_init(&q_empty[s], 1); }
        for (int a = 0; a < 2; ++a) { ptx::mbar_init(&acc_full[a], 1); ptx::mbar_init(&acc_empty[a], 8); }
        ptx::mbar_init(t_full, 1);
        ptx::mbar_init(t_empty, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 2) {
        ptx::tmem_alloc(tmem_slot, kTmemColsF);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < kEpiWarp0F) {
    ptx::setmaxnreg_dec<40>();
    if (warp == 0) {
        // ===================================================== TMA producer (one elected lane)
        if (ptx::elect_one()) {
            uint32_t qcnt = 0, tcnt = 0, it = 0;
            for (int item = blockIdx.x; item < args.n_items; item += gridDim.x, ++it) {
                const int p = find_pair_f(args.pairs, args.n_pairs, item);
                const PairDesc pd = args.pairs[p];
                const int trow = pd.t_row0 + (item - pd.fitem0) * kBlockT;
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
                for (int t = 0; t < ntiles; ++t, ++tcnt) {
                    const int qrow = pd.q_row0 + t * kTileQ;
#pragma unroll
                    for (int kc = 0; kc < KCH; ++kc, ++qcnt) {
                        const uint32_t s = qcnt % NS, ph = (qcnt / NS) & 1;
                        ptx::mbar_wait(&q_empty[s], ph ^ 1);
                        // the aug rows of the tile ride on its first chunk.  Hazard: aug buffer (tcnt & 1) was last read by
                        // the aug MMA of tile tcnt - 2; the slot waited for here was released by chunk qcnt - NS >= the last
                        // chunk of tile tcnt - 2 (NS <= KCH + 1), whose commit follows that aug MMA.
                        ptx::mbar_arrive_expect_tx(&q_full[s], L::kChunkQ + (kc == 0 ? L::kAugQ : 0));
                        ptx::tma_load_2d(smem + L::kOffQ + s * L::kChunkQ, &tm_desc, kc * 64, qrow, &q_full[s]);
                        if (kc == 0) ptx::tma_load_2d(smem + L::kOffQAug + (tcnt & 1) * L::kAugQ, &tm_augq, 0, qrow, &q_full[s]);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one elected lane)
        if (ptx::elect_one()) {
            constexpr uint32_t idesc_main = ptx::idesc_f16(kTileQ, kBlockT, true);    // D -= Q . T^T
            constexpr uint32_t idesc_aug = ptx::idesc_f16(kTileQ, kBlockT, false);    // D += (|q|^2 + |t|^2) / 2
            const uint32_t sbase = ptx::smem_u32(smem);
            uint32_t qcnt = 0, acnt = 0, it = 0;
            for (int item = blockIdx.x; item < args.n_items; item += gridDim.x, ++it) {
                const int p = find_pair_f(args.pairs, args.n_pairs, item);
                const int ntiles = args.pairs[p].nq_pad / kTileQ;
                ptx::mbar_wait(t_full, it & 1);
                for (int t = 0; t < ntiles; ++t, ++acnt) {
                    const uint32_t a = acnt & 1, aph = (acnt >> 1) & 1;
                    ptx::mbar_wait(&acc_empty[a], aph ^ 1);
                    const uint32_t d_tmem = tmem_base + a * kBlockT;
#pragma unroll
                    for (int kc = 0; kc < KCH; ++kc, ++qcnt) {
                        const uint32_t s = qcnt % NS, ph = (qcnt / NS) & 1;
                        ptx::mbar_wait(&q_full[s], ph);
                        ptx::tc_fence_after();
                        const uint32_t qaddr = sbase + L::kOffQ + s * L::kChunkQ;
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t da = ptx::smem_desc(qaddr + ks * 32, 1024, ptx::kLayoutSw128);
                            const uint64_t db = ptx::smem_desc(sbase + L::kOffT + kc * L::kChunkT + ks * 32, 1024, ptx::kLayoutSw128);
                            ptx::umma_bf16(d_tmem, da, db, idesc_main, (kc | ks) != 0);
                        }
                        if (kc == KCH - 1) {
                            const uint64_t da = ptx::smem_desc(sbase + L::kOffQAug + (acnt & 1) * L::kAugQ, 256, ptx::kLayoutSw32);
                            const uint64_t db = ptx::smem_desc(sbase + L::kOffTAug, 256, ptx::kLayoutSw32);
                            ptx::umma_bf16(d_tmem, da, db, idesc_aug, 1);
                        }
                        ptx::umma_commit(&q_empty[s]);
                    }
                    ptx::umma_commit(&acc_full[a]);
                }
                ptx::umma_commit(t_empty);
            }
        }
    }
    } else {
        ptx::setmaxnreg_inc<232>();
        // ===================================================== epilogue: 8 warps, lane = query row, 128 train columns each
        const int ew = warp - kEpiWarp0F;
        const int half = ew >> 2;                       // which 128-column half of the 256-column accumulator
        const int lane_base = (warp & 3) * 32;          // TMEM lanes this warp may access
        const int m_local = lane_base + lane;
        const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(lane_base) << 16) + half * kHalfT;
        uint32_t acnt = 0;
        float cmin[kHalfT];
        for (int item = blockIdx.x; item < args.n_items; item += gridDim.x) {
            const int p = find_pair_f(args.pairs, args.n_pairs, item);
            const int blk = item - args.pairs[p].fitem0;
            const int nq_pad = args.pairs[p].nq_pad;
            const int ntiles = nq_pad / kTileQ;
            long long e = args.pairs[p].rp_base + static_cast<long long>(blk * 2 + half) * nq_pad + m_local;
#pragma unroll
            for (int i = 0; i < kHalfT; ++i) cmin[i] = kHuge * 4.0f;
            const int n_tg = (ntiles + kFlushTiles - 1) / kFlushTiles;
            float* cr = args.colraw + args.pairs[p].cr_base +
                        (static_cast<long long>(blk) * kBlockT + half * kHalfT) * (static_cast<long long>(n_tg) * kTileQ) + m_local;
            for (int t = 0; t < ntiles; ++t, ++acnt) {
                const uint32_t a = acnt & 1, aph = (acnt >> 1) & 1;
                ptx::mbar_wait(&acc_full[a], aph);
                ptx::tc_fence_after();
                float v1 = kHuge * 4.0f, v2 = kHuge * 4.0f;
                int g1 = 0;
#pragma unroll
                for (int c = 0; c < kHalfT / 16; ++c) {
                    float v[16];
                    ptx::tmem_ld_32x32b_x16(t_lane + a * kBlockT + c * 16, v);
#pragma unroll
                    for (int i = 0; i < 16; ++i) cmin[c * 16 + i] = fminf(cmin[c * 16 + i], v[i]);
#pragma unroll
                    for (int sg = 0; sg < 16 / kGroupF; ++sg) {
                        const float m = fminf(fminf(fminf(v[4 * sg], v[4 * sg + 1]), v[4 * sg + 2]), v[4 * sg + 3]);
                        const bool lt = m < v1;                      // strict: the lowest group id wins ties
                        v2 = fminf(v2, fmaxf(v1, m));
                        g1 = lt ? c * (16 / kGroupF) + sg : g1;
                        v1 = fminf(v1, m);
                    }
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&acc_empty[a]);
                args.rp_v1[e] = v1;
                args.rp_v2[e] = v2;
                args.rp_g[e] = static_cast<uint8_t>(g1);
                e += kTileQ;
                // column minima leave the registers every kFlushTiles query tiles: a lane class then holds at most
                // kFlushTiles rows {(kFlushTiles tg + j) 128 + lane}, which is what the resolve kernel recomputes exactly
                if ((t % kFlushTiles) == kFlushTiles - 1 || t == ntiles - 1) {
                    float* dst = cr + static_cast<long long>(t / kFlushTiles) * kTileQ;
#pragma unroll
                    for (int i = 0; i < kHalfT; ++i) {
                        dst[static_cast<long long>(i) * n_tg * kTileQ] = cmin[i];
                        cmin[i] = kHuge * 4.0f;
                    }
                }
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) ptx::tmem_dealloc(tmem_base, kTmemColsF);
}

template <int KCH>
static cudaError_t launch_screen_f_t(const CUtensorMap& tm_desc, const CUtensorMap& tm_augq, const CUtensorMap& tm_augt,
                                     const ScreenArgsF& args, int n_sm, cudaStream_t stream)
{
    using L = ScreenSmemF<KCH>;
    static g2v::DeviceOnce once;
    int& configured = once.cur();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(match_screen_f_kernel<KCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal);
        if (e != cudaSuccess) return e;
        configured = 1;
    }
    const int grid = args.n_items < n_sm ? args.n_items : n_sm;
    if (grid <= 0) return cudaSuccess;
    match_screen_f_kernel<KCH><<<grid, kThreadsF, L::kTotal, stream>>>(tm_desc, tm_augq, tm_augt, args);
    return cudaGetLastError();
}

// kpad = descriptor dimension padded to a multiple of 64: 64, 128 or 256.
cudaError_t launch_match_screen_f(const CUtensorMap& tm_desc, const CUtensorMap& tm_augq, const CUtensorMap& tm_augt,
                                  const PairDesc* pairs, int n_pairs, int n_items, float* rp_v1, float* rp_v2, uint8_t* rp_g,
                                  float* colraw, int kpad, int n_sm, cudaStream_t stream)
{
    ScreenArgsF a{pairs, n_pairs, n_items, rp_v1, rp_v2, rp_g, colraw};
    if (kpad == 64) return launch_screen_f_t<1>(tm_desc, tm_augq, tm_augt, a, n_sm, stream);
    if (kpad == 128) return launch_screen_f_t<2>(tm_desc, tm_augq, tm_augt, a, n_sm, stream);
    if (kpad == 256) return launch_screen_f_t<4>(tm_desc, tm_augq, tm_augt, a, n_sm, stream);
    return cudaErrorInvalidValue;
}

}  // namespace g2v
