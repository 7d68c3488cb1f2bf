// K3-K7 -- batched essential-matrix RANSAC of the reference's Ransac verifier
// (gtsfm/frontend/verifier/ransac.py:52-82, opencv_verifier_base.py:45-109, utils/verification.py:52-94).
//
//   K3 gather_normalize   match rows -> normalised correspondences (fp64 master + fp32 copy)   features.py:40-50
//   K4 solve              one 5-point hypothesis per thread, fp64                              (cv2 USAC's minimal solver)
//   K5 score              MSAC cost sum(min(sampson^2, thr^2)) of every candidate over all correspondences, fp32,
//                         correspondences staged in shared memory, one candidate per warp, warp-shuffle reduction
//   K6 finalize           best candidates re-solved and re-scored in fp64, Levenberg-Marquardt local optimisation,
//                         final mask sampson^2 < thr^2 (fp64)                                  verification.py:170-214
//   K7 (inside finalize)  pose recovery: SVD of E, four (R,t) candidates, DLT cheirality vote   verification.py:81
#include <cfloat>

#include "store.h"
#include "verify_block.cuh"
#include "verify_fivept.cuh"
#include "verify_types.h"

namespace g2v {

// ---------------------------------------------------------------------------------------------------------------- K3
__global__ void __launch_bounds__(256) gather_normalize_kernel(StoreView st, VerifyBufs b, int n_pairs)
{
    const int p = blockIdx.y;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < vp.M; i += gridDim.x * blockDim.x) {
        const uint2 m = static_cast<const uint2*>(b.match)[vp.m0 + i];
        const double2 k1 = st.kps[vp.row1 + m.x], k2 = st.kps[vp.row2 + m.y];
        double x1, y1, x2, y2;
        vm::calibrate(k1.x, k1.y, vp.cal1, &x1, &y1);
        vm::calibrate(k2.x, k2.y, vp.cal2, &x2, &y2);
        static_cast<double4*>(b.corr_d)[vp.m0 + i] = make_double4(x1, y1, x2, y2);
        static_cast<float4*>(b.corr_f)[vp.m0 + i] = make_float4(static_cast<float>(x1), static_cast<float>(y1), static_cast<float>(x2), static_cast<float>(y2));
    }
}

__device__ __forceinline__ void hypothesis_samples(const VerifyBufs& b, const VerifyDims& d, int p, int h, const VPair& vp, int* s)
{
    if (b.samples) {
        const int* src = b.samples + (static_cast<long long>(p) * d.H + h) * 5;
        for (int k = 0; k < 5; ++k) s[k] = src[k];
    } else {
        sample5(d.seed, vp.key, static_cast<uint32_t>(h), static_cast<uint32_t>(vp.M), s);
    }
}

// phase A of hypothesis h of pair p: returns the number of real roots and leaves the hand-off block in the scratch
template <class W>
__device__ int isolate_hypothesis(const VerifyBufs& b, const VerifyDims& d, int p, int h, const W& wk, unsigned* mid)
{
    const VPair& vp = b.pairs[p];
    int s[5];
    hypothesis_samples(b, d, p, h, vp, s);
    const double4* corr = static_cast<const double4*>(b.corr_d) + vp.m0;
    int nr;
    {
        double x1[5], y1[5], x2[5], y2[5];
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            s[k] = min(max(s[k], 0), vp.M - 1);
            const double4 c = corr[s[k]];
            x1[k] = c.x; y1[k] = c.y; x2[k] = c.z; y2[k] = c.w;
        }
        nr = vm::fivept_isolate(x1, y1, x2, y2, wk, mid);
    }
    if (nr > 0) {
        // the sample correspondences of the hand-off block are re-read (cache hits) rather than kept alive in
        // registers through the whole solve
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const double4 c = corr[s[k]];
            wk(vm::kOffS + k) = c.x; wk(vm::kOffS + 5 + k) = c.y; wk(vm::kOffS + 10 + k) = c.z; wk(vm::kOffS + 15 + k) = c.w;
        }
    }
    return nr;
}

// ---------------------------------------------------------------------------------------------------------------- K4
// Persistent CTAs over blocks of kSolveThreads hypotheses.  Phase A runs one hypothesis per thread and leaves its state in
// shared memory ([field][thread], conflict-free); the real roots of the block are then compacted so that phase B
// (root refinement, back-substitution, polish) runs one ROOT per thread with full warps, whatever the number of
// real roots each hypothesis has.  Slot r of a hypothesis holds the essential matrix of its r-th root (ascending z),
// NaN when the root gave no finite matrix.
constexpr int kSolveThreads = 64;
constexpr int kSolveSmem = vm::kWorkDoubles * kSolveThreads * 8;
static_assert(vm::kOffXyz >= vm::kOffP + 11 && vm::kOffXyz + 3 * vm::kMaxCand <= vm::kWorkDoubles,
              "the parked (x, y, z) of every root must fit the scratch behind the hand-off polynomial");

__global__ void __launch_bounds__(kSolveThreads, 2) solve_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    extern __shared__ double s_h[];                       // [kWorkDoubles][kSolveThreads]: scratch, then hand-off state
    __shared__ unsigned s_mid[kSolveThreads];
    __shared__ unsigned s_valid[kSolveThreads];           // bit r: root r of the thread's hypothesis gave a valid model
    __shared__ unsigned short s_map2[kSolveThreads * vm::kMaxCand];
    __shared__ int s_n2;
    __shared__ int s_off[kSolveThreads + 1];
    __shared__ unsigned short s_map[kSolveThreads * vm::kMaxCand];
    __shared__ int s_wsum[kSolveThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long total = static_cast<long long>(n_pairs) * d.H;
    const long long n_blocks = (total + kSolveThreads - 1) / kSolveThreads;
    for (long long blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        const long long g = blk * kSolveThreads + tid;
        int nr = 0;
        s_valid[tid] = 0u;
        if (tid == 0) s_n2 = 0;
        if (g < total) {
            const int p = static_cast<int>(g / d.H), h = static_cast<int>(g - static_cast<long long>(p) * d.H);
            if (b.pairs[p].status == 0) {
                unsigned mid = 0;
                nr = isolate_hypothesis(b, d, p, h, vm::Work<kSolveThreads>{s_h + tid}, &mid);
                s_mid[tid] = mid;
            }
        }
        // exclusive prefix sum of nr over the block
        int inc = nr;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, off); if (lane >= off) inc += v; }
        if (lane == 31) s_wsum[warp] = inc;
        __syncthreads();
        int base = 0;
        for (int w2 = 0; w2 < warp; ++w2) base += s_wsum[w2];
        const int excl = base + inc - nr;
        s_off[tid] = excl;
        if (tid == kSolveThreads - 1) s_off[kSolveThreads] = excl + nr;
        for (int r = 0; r < nr; ++r) s_map[excl + r] = static_cast<unsigned short>((tid << 4) | r);
        __syncthreads();
        const int n_roots = s_off[kSolveThreads];
        if (tid == 0 && b.stats) {
            const long long left = total - blk * kSolveThreads;
            atomicAdd(b.stats + 1, static_cast<unsigned long long>(left < kSolveThreads ? left : kSolveThreads));
        }
        // phase B1: one root per thread -> (x, y, z) and the validity of the unpolished model; survivors are compacted
        // once more (their xyz parked in the unused tail of the scratch) so that the polish runs on full warps
        double* s_xyz = s_h + vm::kOffXyz * kSolveThreads;
        for (int idx = tid; idx < n_roots; idx += kSolveThreads) {
            const int code = s_map[idx], j = code >> 4, r = code & 15;
            double xyz[3];
            if (vm::fivept_root_xyz(vm::HypView<kSolveThreads>{s_h + j}, s_mid[j], r, xyz)) {
                const int pos = atomicAdd(&s_n2, 1);
                s_map2[pos] = static_cast<unsigned short>(code);
                s_xyz[3 * pos] = xyz[0]; s_xyz[3 * pos + 1] = xyz[1]; s_xyz[3 * pos + 2] = xyz[2];
            }
        }
        __syncthreads();
        const int n_ok = s_n2;
        // phase B2: polish + final model of the survivors
        for (int idx = tid; idx < n_ok; idx += kSolveThreads) {
            const int code = s_map2[idx], j = code >> 4, r = code & 15;
            double xyz[3] = {s_xyz[3 * idx], s_xyz[3 * idx + 1], s_xyz[3 * idx + 2]}, E[9];
            if (vm::fivept_root_polish(vm::HypView<kSolveThreads>{s_h + j}, xyz, E)) {
                float* out = b.E32 + ((blk * kSolveThreads + j) * vm::kMaxCand + r) * 9;
#pragma unroll
                for (int e = 0; e < 9; ++e) out[e] = static_cast<float>(E[e]);
                atomicOr(&s_valid[j], 1u << r);
            }
        }
        __syncthreads();
        if (g < total) b.vmask[g] = static_cast<unsigned short>(s_valid[tid]);
        if (b.stats) {
            int nv = __popc(s_valid[tid]);
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) nv += __shfl_xor_sync(0xffffffffu, nv, off);
            if (lane == 0 && nv) atomicAdd(b.stats, static_cast<unsigned long long>(nv));
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------- K5
constexpr int kScoreHyp = 128;     // hypotheses per CTA (the staged correspondences are shared by all their candidates)
constexpr int kScoreTile = 4096;   // correspondences staged per shared-memory tile (64 KB)
constexpr int kScoreMaxCand = kScoreHyp * vm::kMaxCand;
constexpr int kScoreCheck = 256;   // packed pairs (512 correspondences) between two early-exit checks

// packed fp32x2 arithmetic (FFMA2 / FMUL2 / FADD2 on sm_100a): two correspondences per lane and instruction
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk2(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float rcp_fast(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// Shared-memory tile: correspondences 2q and 2q + 1 are stored as A[q] = (x1a, x1b, y1a, y1b), B[q] = (x2a, x2b, y2a, y2b)
// so that one LDS.128 yields two packed operands.
__global__ void __launch_bounds__(256) score_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    extern __shared__ float4 s_corr[];                 // [kScoreTile / 2] A, then [kScoreTile / 2] B
    __shared__ float s_cost[kScoreMaxCand];
    __shared__ unsigned short s_cand[kScoreMaxCand];   // (local hypothesis << 4) | candidate
    __shared__ unsigned char s_dead[kScoreMaxCand];    // candidate left the race (partial cost above the pair's best)
    __shared__ unsigned long long s_exec;
    __shared__ int s_n, s_next;
    __shared__ int s_wcnt[kScoreHyp / 32];
    const int p = blockIdx.y;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    if (vp.status != 0) return;
    const int h0 = blockIdx.x * kScoreHyp;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // candidate list of the CTA's hypotheses, in ascending (hypothesis, root) order: one hypothesis per thread
    unsigned vm_bits = 0;
    int incl = 0;
    if (threadIdx.x < kScoreHyp) {
        vm_bits = (h0 + threadIdx.x < d.H) ? b.vmask[static_cast<long long>(p) * d.H + h0 + threadIdx.x] : 0u;
        incl = __popc(vm_bits);
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += v; }
        if (lane == 31) s_wcnt[warp] = incl;
    }
    for (int k = threadIdx.x; k < kScoreMaxCand; k += blockDim.x) { s_cost[k] = 0.f; s_dead[k] = 0; }
    if (threadIdx.x == 0) { s_exec = 0ull; s_next = 0; }
    __syncthreads();
    if (threadIdx.x < kScoreHyp) {
        int pos = incl - __popc(vm_bits);
        for (int w2 = 0; w2 < warp; ++w2) pos += s_wcnt[w2];
        unsigned m = vm_bits;
        while (m) { const int c = __ffs(m) - 1; m &= m - 1; s_cand[pos++] = static_cast<unsigned short>((threadIdx.x << 4) | c); }
        if (threadIdx.x == kScoreHyp - 1) s_n = pos;
    }
    __syncthreads();
    const int n = s_n;
    if (n == 0) return;
    if (threadIdx.x == 0 && b.stats) atomicAdd(b.stats + 2, static_cast<unsigned long long>(n) * static_cast<unsigned long long>(vp.M));
    const float thr2 = static_cast<float>(vp.thr2);
    const float q_inv = thr2 * static_cast<float>(vp.M) * (1.0f / 1073741824.0f), q_scale = 1.0f / q_inv;
    const float* Ebase = b.E32 + (static_cast<long long>(p) * d.H + h0) * (9 * vm::kMaxCand);
    const volatile float* best = reinterpret_cast<const volatile float*>(b.best_cost + p);
    float4* sA = s_corr;
    float4* sB = s_corr + kScoreTile / 2;
    for (int t0 = 0; t0 < vp.M; t0 += kScoreTile) {
        const int tn = min(kScoreTile, vp.M - t0);
        const int nq = tn >> 1;                        // complete pairs of correspondences
        const bool last_tile = t0 + kScoreTile >= vp.M;
        __syncthreads();
        for (int q = threadIdx.x; q < (tn + 1) >> 1; q += blockDim.x) {
            const float4 ca = static_cast<const float4*>(b.corr_f)[vp.m0 + t0 + 2 * q];
            const float4 cb = (2 * q + 1 < tn) ? static_cast<const float4*>(b.corr_f)[vp.m0 + t0 + 2 * q + 1] : ca;
            sA[q] = make_float4(ca.x, cb.x, ca.y, cb.y);
            sB[q] = make_float4(ca.z, cb.z, ca.w, cb.w);
        }
        if (threadIdx.x == 0) s_next = 0;
        __syncthreads();
        // One candidate's state while a warp scores it over the staged tile
        struct Cand { f32x2 p[9]; float e[9]; float before, acc0, acc1; int ci, done; bool dead; };
        auto load_cand = [&](Cand& c, int ci) {
            const int code = s_cand[ci];
            const float* E = Ebase + ((code >> 4) * vm::kMaxCand + (code & 15)) * 9;
#pragma unroll
            for (int k = 0; k < 9; ++k) { c.e[k] = __ldg(E + k); c.p[k] = pk2(c.e[k], c.e[k]); }
            c.before = s_cost[ci];                     // cost over the previous tiles
            c.acc0 = c.acc1 = 0.f; c.ci = ci; c.done = 0; c.dead = false;
        };
        // squared Sampson distances of the packed pair (A, B) under candidate c, clamped and added to its two accumulators
        auto eval = [&](Cand& c, const ulonglong2& A, const ulonglong2& B) {
            const f32x2 l0 = fma2(c.p[0], A.x, fma2(c.p[1], A.y, c.p[2])), l1 = fma2(c.p[3], A.x, fma2(c.p[4], A.y, c.p[5])),
                        l2 = fma2(c.p[6], A.x, fma2(c.p[7], A.y, c.p[8]));
            const f32x2 m0 = fma2(c.p[0], B.x, fma2(c.p[3], B.y, c.p[6])), m1 = fma2(c.p[1], B.x, fma2(c.p[4], B.y, c.p[7]));
            const f32x2 num = fma2(B.x, l0, fma2(B.y, l1, l2));
            const f32x2 den = fma2(l0, l0, fma2(l1, l1, fma2(m0, m0, mul2(m1, m1))));
            float d0, d1;
            upk2(den, d0, d1);
            const f32x2 sq = mul2(mul2(num, num), pk2(rcp_fast(d0), rcp_fast(d1)));
            float s0, s1;
            upk2(sq, s0, s1);
            c.acc0 += fminf(s0, thr2);                 // NaN -> thr2: counts as an outlier
            c.acc1 += fminf(s1, thr2);
        };
        // Early exit (exact): the MSAC cost only grows, so a candidate whose partial cost already exceeds the lowest COMPLETE
        // cost of the pair (by more than the fp32 accumulation noise) cannot be the minimum.  The warp total of the partial
        // costs is taken in fixed point (one REDUX instead of a five-step shuffle tree): every lane's share is below thr2 * M,
        // so 2^30 / (thr2 * M) per unit cannot overflow; the rounding (< 2^-25 of the full scale) is far inside the 1.001 margin.
        auto check = [&](Cand& c, int c1, float bestv) {
            c.done = c1;
            const int part_q = __float2int_rn((c.acc0 + c.acc1) * q_scale);
            const float part = static_cast<float>(__reduce_add_sync(0xffffffffu, part_q)) * q_inv;
            if (c.before + part > bestv * 1.001f) c.dead = true;
        };
        auto run_single = [&](Cand& c, int from) {
            for (int c0 = from; c0 < nq && !c.dead; c0 += kScoreCheck) {
                const int c1 = min(nq, c0 + kScoreCheck);
                const float bestv = *best;             // issued here, consumed after the segment: latency hidden
#pragma unroll 2
                for (int k = c0 + lane; k < c1; k += 32) {
                    const ulonglong2 A = *reinterpret_cast<const ulonglong2*>(sA + k);   // x1 pair, y1 pair
                    const ulonglong2 B = *reinterpret_cast<const ulonglong2*>(sB + k);   // x2 pair, y2 pair
                    eval(c, A, B);
                }
                check(c, c1, bestv);
            }
        };
        auto finish = [&](Cand& c) {
            float acc = c.acc0 + c.acc1;
            if (!c.dead && (tn & 1) && lane == 0) {    // odd tail: first half of the last (duplicated) pair
                const float4 a4 = sA[nq], b4 = sB[nq];
                const float x1 = a4.x, y1 = a4.z, x2 = b4.x, y2 = b4.z;
                const float* e = c.e;
                const float l0 = fmaf(e[0], x1, fmaf(e[1], y1, e[2])), l1 = fmaf(e[3], x1, fmaf(e[4], y1, e[5])), l2 = fmaf(e[6], x1, fmaf(e[7], y1, e[8]));
                const float m0 = fmaf(e[0], x2, fmaf(e[3], y2, e[6])), m1 = fmaf(e[1], x2, fmaf(e[4], y2, e[7]));
                const float num = fmaf(x2, l0, fmaf(y2, l1, l2));
                const float den = fmaf(l0, l0, fmaf(l1, l1, fmaf(m0, m0, m1 * m1)));
                acc += fminf(num * num * rcp_fast(den), thr2);
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
            if (lane == 0) {
                const float total = c.before + acc;
                s_cost[c.ci] = total;
                if (c.dead) s_dead[c.ci] = 1;
                else if (last_tile) atomicMin(b.best_cost + p, __float_as_uint(total));
                atomicAdd(&s_exec, static_cast<unsigned long long>(2 * c.done + (c.dead ? 0 : (tn & 1))));
            }
        };
        for (;;) {
            // candidates are handed out dynamically, two at a time: early exits make their cost very uneven.  A warp scores
            // its two candidates in ONE pass over the tile -- every LDS.128 feeds both (the shared-memory pipe was the limiter
            // with one candidate per pass) -- and each keeps its own accumulators, checks and exit, so a candidate's cost does
            // not depend on its partner.
            int ci = 0;
            if (lane == 0) ci = atomicAdd(&s_next, 2);
            ci = __shfl_sync(0xffffffffu, ci, 0);
            if (ci >= n) break;
            const bool liveA = !s_dead[ci], liveB = ci + 1 < n && !s_dead[ci + 1];
            if (!liveA && !liveB) continue;
            if (liveA != liveB) {
                Cand c;
                load_cand(c, liveA ? ci : ci + 1);
                run_single(c, 0);
                finish(c);
                continue;
            }
            Cand ca, cb;
            load_cand(ca, ci);
            load_cand(cb, ci + 1);
            for (int c0 = 0; c0 < nq; c0 += kScoreCheck) {
                const int c1 = min(nq, c0 + kScoreCheck);
                const float bestv = *best;
#pragma unroll 2
                for (int k = c0 + lane; k < c1; k += 32) {
                    const ulonglong2 A = *reinterpret_cast<const ulonglong2*>(sA + k);
                    const ulonglong2 B = *reinterpret_cast<const ulonglong2*>(sB + k);
                    eval(ca, A, B);
                    eval(cb, A, B);
                }
                check(ca, c1, bestv);
                check(cb, c1, bestv);
                if (ca.dead || cb.dead) {              // the survivor (if any) goes on alone
                    if (!ca.dead) run_single(ca, c1);
                    if (!cb.dead) run_single(cb, c1);
                    break;
                }
            }
            finish(ca);
            finish(cb);
        }
    }
    __syncthreads();
    for (int ci = threadIdx.x; ci < n; ci += blockDim.x) {
        const int code = s_cand[ci];
        b.cost[(static_cast<long long>(p) * d.H + h0 + (code >> 4)) * vm::kMaxCand + (code & 15)] = s_cost[ci];
    }
    if (threadIdx.x == 0 && b.stats) atomicAdd(b.stats + 3, s_exec);
}

// ---------------------------------------------------------------------------------------------------------------- K6/K7
constexpr int kTopK = 8;            // candidates re-scored in fp64 (fp32 scoring only has to rank the best few)
// kFinThreads, block_sum and recover_pose_block: verify_block.cuh

__device__ inline void skew_times(const double* t, const double* R, double* E)   // E = [t]x R
{
    for (int j = 0; j < 3; ++j) {
        E[j] = -t[2] * R[3 + j] + t[1] * R[6 + j];
        E[3 + j] = t[2] * R[j] - t[0] * R[6 + j];
        E[6 + j] = -t[1] * R[j] + t[0] * R[3 + j];
    }
}
__device__ inline void rodrigues(const double* w, double* R)
{
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
    double a, b;
    if (th < 1e-8) { a = 1.0 - th2 / 6.0; b = 0.5 - th2 / 24.0; } else { a = sin(th) / th; b = (1.0 - cos(th)) / th2; }
    const double K[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
    double K2[9];
    vm::mat3_mul(K, K, K2);
    for (int i = 0; i < 9; ++i) R[i] = ((i % 4) == 0 ? 1.0 : 0.0) + a * K[i] + b * K2[i];
}
__device__ inline void tangent_basis(const double* t, double* b1, double* b2)
{
    int k = 0;
    if (fabs(t[1]) < fabs(t[k])) k = 1;
    if (fabs(t[2]) < fabs(t[k])) k = 2;
    double e[3] = {0, 0, 0};
    e[k] = 1.0;
    vm::cross3(t, e, b1);
    const double n = vm::norm3(b1);
    for (int i = 0; i < 3; ++i) b1[i] /= n;
    vm::cross3(t, b1, b2);
}
// (JtJ + lambda diag(JtJ)) d = -Jtr by Cholesky; false when not positive definite
__device__ inline bool solve_lm5(const double* JtJ, const double* Jtr, double lambda, double* d)
{
    double L[25];
    for (int i = 0; i < 25; ++i) L[i] = JtJ[i];
    for (int i = 0; i < 5; ++i) L[6 * i] += lambda * JtJ[6 * i] + 1e-300;
    for (int j = 0; j < 5; ++j) {
        double s = L[6 * j];
        for (int k = 0; k < j; ++k) s -= L[5 * j + k] * L[5 * j + k];
        if (!(s > 0)) return false;
        const double ljj = sqrt(s);
        L[6 * j] = ljj;
        for (int i = j + 1; i < 5; ++i) {
            double v = L[5 * i + j];
            for (int k = 0; k < j; ++k) v -= L[5 * i + k] * L[5 * j + k];
            L[5 * i + j] = v / ljj;
        }
    }
    double y[5];
    for (int i = 0; i < 5; ++i) { double v = -Jtr[i]; for (int k = 0; k < i; ++k) v -= L[5 * i + k] * y[k]; y[i] = v / L[6 * i]; }
    for (int i = 4; i >= 0; --i) { double v = y[i]; for (int k = i + 1; k < 5; ++k) v -= L[5 * k + i] * d[k]; d[i] = v / L[6 * i]; }
    return true;
}

// One pass over all correspondences at model (R, t) in shared memory: s_out = {cost, #inliers, Jtr[5], JtJ upper[15]}
// of the signed Sampson residuals of the inliers; parameters: R <- R Exp(w), t <- normalize(t + b1 a + b2 b).
__device__ void lo_pass(const VerifyBufs& b, const VPair& vp, const double* s_R, const double* s_t, double* s_E, double* s_dE,
                        double (*s_red)[24], double* s_out, double4 (*s_q)[64])
{
    if (threadIdx.x == 0) {
        double b1[3], b2[3];
        skew_times(s_t, s_R, s_E);
        tangent_basis(s_t, b1, b2);
        for (int i = 0; i < 3; ++i) {
            s_dE[3 * i] = 0;                    s_dE[3 * i + 1] = s_E[3 * i + 2];       s_dE[3 * i + 2] = -s_E[3 * i + 1];
            s_dE[9 + 3 * i] = -s_E[3 * i + 2];  s_dE[9 + 3 * i + 1] = 0;                s_dE[9 + 3 * i + 2] = s_E[3 * i];
            s_dE[18 + 3 * i] = s_E[3 * i + 1];  s_dE[18 + 3 * i + 1] = -s_E[3 * i];     s_dE[18 + 3 * i + 2] = 0;
        }
        skew_times(b1, s_R, s_dE + 27);
        skew_times(b2, s_R, s_dE + 36);
    }
    __syncthreads();
    double E[9];
    for (int i = 0; i < 9; ++i) E[i] = s_E[i];
    double acc[22];
    for (int i = 0; i < 22; ++i) acc[i] = 0;
    // The Jacobian rows are only needed for the inliers (about half of the correspondences, scattered), and they are
    // ~85 % of the fp64 work of a pass: every warp first classifies 32 correspondences (residual only), queues its inliers
    // in shared memory and runs the Jacobian part whenever 32 of them are waiting, i.e. on full warps.
    auto inlier_terms = [&](const double4& c) {
        const double a1 = c.x, c1 = c.y, a2 = c.z, c2 = c.w;
        const double l0 = E[0] * a1 + E[1] * c1 + E[2], l1 = E[3] * a1 + E[4] * c1 + E[5], l2 = E[6] * a1 + E[7] * c1 + E[8];
        const double m0 = E[0] * a2 + E[3] * c2 + E[6], m1 = E[1] * a2 + E[4] * c2 + E[7];
        const double num = a2 * l0 + c2 * l1 + l2, den = l0 * l0 + l1 * l1 + m0 * m0 + m1 * m1;
        const double isd = 1.0 / sqrt(den), r = num * isd, q = r / den;
        acc[0] += num * num / den;
        acc[1] += 1.0;
        double J[5];
#pragma unroll
        for (int pz = 0; pz < 5; ++pz) {
            const double* D = s_dE + 9 * pz;
            const double dl0 = D[0] * a1 + D[1] * c1 + D[2], dl1 = D[3] * a1 + D[4] * c1 + D[5], dl2 = D[6] * a1 + D[7] * c1 + D[8];
            const double dm0 = D[0] * a2 + D[3] * c2 + D[6], dm1 = D[1] * a2 + D[4] * c2 + D[7];
            const double dnum = a2 * dl0 + c2 * dl1 + dl2;
            J[pz] = dnum * isd - q * (l0 * dl0 + l1 * dl1 + m0 * dm0 + m1 * dm1);
        }
        int k = 7;
#pragma unroll
        for (int pz = 0; pz < 5; ++pz) {
            acc[2 + pz] += J[pz] * r;
#pragma unroll
            for (int u = pz; u < 5; ++u) acc[k++] += J[pz] * J[u];
        }
    };
    const int lane = threadIdx.x & 31;
    double4* q = s_q[threadIdx.x >> 5];
    int qn = 0;                                             // warp-uniform: inliers waiting in the queue (< 32 between rounds)
    // the next row's correspondence is requested before the current one is used: the rows of a pair are re-read from L2 on
    // every pass (211 KB of double4 per C2 pair against four CTAs' worth of L1), and a dependent load per iteration left the
    // warps waiting on it (long scoreboard was the first stall reason)
    const double4* corr = static_cast<const double4*>(b.corr_d) + vp.m0;
    double4 cn = make_double4(0, 0, 0, 0);
    if (static_cast<int>(threadIdx.x) < vp.M) cn = corr[threadIdx.x];
    for (int i0 = threadIdx.x - lane; i0 < vp.M; i0 += kFinThreads) {
        const int i = i0 + lane;
        bool in = false;
        const double4 c = cn;
        if (i + kFinThreads < vp.M) cn = corr[i + kFinThreads];
        if (i < vp.M) {
            in = vm::sampson_sq(E, c.x, c.y, c.z, c.w) < vp.thr2;
            if (!in) acc[0] += vp.thr2;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, in);
        if (in) q[qn + __popc(bal & ((1u << lane) - 1u))] = c;
        qn += __popc(bal);
        __syncwarp();
        if (qn >= 32) {
            qn -= 32;
            const double4 cc = q[qn + lane];
            __syncwarp();
            inlier_terms(cc);
        }
    }
    if (lane < qn) inlier_terms(q[lane]);
    block_sum<22>(acc, s_red, s_out);
}

// K6a: the kTopK best candidates of a pair by fp32 MSAC cost (ties: lowest candidate id = h * 10 + root).
__global__ void __launch_bounds__(256) topk_select_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    __shared__ int s_topi[kTopK];
    __shared__ float s_wv[8];
    __shared__ int s_wi[8];
    const int p = blockIdx.x;
    if (p >= n_pairs) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < kTopK) b.topi[p * kTopK + tid] = INT_MAX;
    if (b.pairs[p].status != 0) return;
    const int n_slots = d.H * vm::kMaxCand;
    const float* cost = b.cost + static_cast<long long>(p) * n_slots;
    const unsigned short* vmask = b.vmask + static_cast<long long>(p) * d.H;
    for (int k = 0; k < kTopK; ++k) {
        float bv = FLT_MAX;
        int bi = INT_MAX;
        for (int h = tid; h < d.H; h += 256) {             // ascending scan: first minimum = lowest id
            unsigned m = vmask[h];
            while (m) {
                const int c = __ffs(m) - 1;
                m &= m - 1;
                const int sidx = h * vm::kMaxCand + c;
                bool used = false;
                for (int j = 0; j < k; ++j) used |= (s_topi[j] == sidx);
                const float v = cost[sidx];
                if (!used && (v < bv)) { bv = v; bi = sidx; }
            }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, off);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (ov < bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_wv[warp] = bv; s_wi[warp] = bi; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < 8; ++w)
                if (s_wv[w] < bv || (s_wv[w] == bv && s_wi[w] < bi)) { bv = s_wv[w]; bi = s_wi[w]; }
            s_topi[k] = bi;
            b.topi[p * kTopK + k] = bi;
        }
        __syncthreads();
        if (s_topi[k] == INT_MAX) break;
    }
}

// K6b: re-solve the selected hypotheses in fp64, one (pair, candidate) per thread with the solver scratch in shared
// memory as in the solve kernel.  The wanted root is the one whose matrix reproduces the fp32 copy.
__global__ void __launch_bounds__(kSolveThreads, 2) topk_solve_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    extern __shared__ double s_h[];
    const long long g = static_cast<long long>(blockIdx.x) * kSolveThreads + threadIdx.x;
    if (g >= static_cast<long long>(n_pairs) * kTopK) return;
    const int p = static_cast<int>(g / kTopK);
    double Eb[9];
    for (int e = 0; e < 9; ++e) Eb[e] = nan("");
    const int sidx = b.topi[g];
    if (sidx != INT_MAX && b.pairs[p].status == 0) {
        const int h = sidx / vm::kMaxCand, c = sidx - h * vm::kMaxCand;
        const float* e32 = b.E32 + (static_cast<long long>(p) * d.H) * (9 * vm::kMaxCand) + static_cast<long long>(sidx) * 9;
        unsigned mid = 0;
        const vm::Work<kSolveThreads> wk{s_h + threadIdx.x};
        const int nr = isolate_hypothesis(b, d, p, h, wk, &mid);
        const vm::HypView<kSolveThreads> hv{s_h + threadIdx.x};
        double bestdot = -1.0, E[9];
        // slot c of E32 is root c of the same deterministic isolation; accept it when it reproduces the fp32 copy,
        // otherwise (a root count that differs in the last bit) take the closest of all roots
        if (c < nr && vm::fivept_root(hv, mid, c, E)) {
            double dot = 0;
            for (int e = 0; e < 9; ++e) dot += E[e] * static_cast<double>(e32[e]);
            if (fabs(dot) > 0.999999) { bestdot = 2.0; for (int e = 0; e < 9; ++e) Eb[e] = E[e]; }
        }
        for (int r = 0; r < nr && bestdot < 2.0; ++r) {
            if (!vm::fivept_root(hv, mid, r, E)) continue;
            double dot = 0;
            for (int e = 0; e < 9; ++e) dot += E[e] * static_cast<double>(e32[e]);
            if (fabs(dot) > bestdot && fabs(dot) < 2.0) { bestdot = fabs(dot); for (int e = 0; e < 9; ++e) Eb[e] = E[e]; }
        }
    }
    for (int e = 0; e < 9; ++e) b.Ek[g * 9 + e] = Eb[e];
}

__global__ void __launch_bounds__(kFinThreads, 4) finalize_kernel(VerifyBufs b, VerifyDims d, int n_pairs)
{
    __shared__ double s_red[kFinThreads / 32][24];
    __shared__ double s_out[24];
    __shared__ double s_Ek[kTopK][9];
    __shared__ double s_R[9], s_t[3], s_E[9], s_dE[45];
    __shared__ double s_pose[21];
    __shared__ double4 s_q[kFinThreads / 32][64];      // lo_pass: per-warp queue of inliers awaiting their Jacobian rows
    __shared__ int s_topi[kTopK];
    __shared__ int s_ctrl;

    const int p = blockIdx.x;
    if (p >= n_pairs) return;
    const VPair& vp = b.pairs[p];
    const int tid = threadIdx.x;
    for (int i = tid; i < vp.M; i += kFinThreads) b.mask[vp.m0 + i] = 0;
    if (tid < 9) { b.R[p * 9 + tid] = 0; b.E[p * 9 + tid] = 0; }
    if (tid < 3) b.t[p * 3 + tid] = 0;
    if (tid == 0) { b.ninl[p] = 0; b.status[p] = vp.status; }
    if (vp.status != 0) return;

    // ---- (a, b) the kTopK best candidates by fp32 cost and their fp64 matrices (topk_select_kernel, topk_solve_kernel)
    if (tid < kTopK) s_topi[tid] = b.topi[p * kTopK + tid];
    if (tid < kTopK * 9) (&s_Ek[0][0])[tid] = b.Ek[static_cast<long long>(p) * kTopK * 9 + tid];
    __syncthreads();
    int n_top = 0;
    while (n_top < kTopK && s_topi[n_top] != INT_MAX) ++n_top;
    if (n_top == 0) { if (tid == 0) b.status[p] = 2; return; }

    // ---- (c) fp64 MSAC cost of the kTopK candidates, then the minimum (ties: lowest id)
    {
        double acc[kTopK];
#pragma unroll
        for (int k = 0; k < kTopK; ++k) acc[k] = 0;
        const double4* corr = static_cast<const double4*>(b.corr_d) + vp.m0;
        double4 cn = make_double4(0, 0, 0, 0);
        if (tid < vp.M) cn = corr[tid];
        for (int i = tid; i < vp.M; i += kFinThreads) {
            const double4 c = cn;                           // software prefetch, as in lo_pass
            if (i + kFinThreads < vp.M) cn = corr[i + kFinThreads];
#pragma unroll
            for (int k = 0; k < kTopK; ++k) {
                if (k < n_top) {
                    const double s = vm::sampson_sq(s_Ek[k], c.x, c.y, c.z, c.w);
                    acc[k] += (s < vp.thr2) ? s : vp.thr2;
                }
            }
        }
        block_sum<kTopK>(acc, s_red, s_out);
    }
    if (tid == 0) {
        int best = 0;
        for (int k = 1; k < n_top; ++k)
            if (s_out[k] < s_out[best] || (s_out[k] == s_out[best] && s_topi[k] < s_topi[best])) best = k;
        for (int e = 0; e < 9; ++e) s_E[e] = s_Ek[best][e];
    }
    __syncthreads();

    // ---- (d) local optimisation: Levenberg-Marquardt on (R, t) over the Sampson residuals of the current inliers
    if (d.lo_iters > 0) {
        double R[9], t[3], JtJ[25], Jtr[5], cst = 0, lambda = 1e-4;
        int ninl = 0, it = 0;
        if (tid == 0) {
            // canonical start (independent of the sign of the minimal model and of the order an SVD returns the twisted
            // pair in): the rotation with the larger trace, t signed so that [t]x R reproduces the model
            double R1[9], R2[9], Et[9], dot = 0;
            vm::decompose_essential(s_E, R1, R2, t);
            const bool first = R1[0] + R1[4] + R1[8] >= R2[0] + R2[4] + R2[8];
            for (int i = 0; i < 9; ++i) R[i] = first ? R1[i] : R2[i];
            skew_times(t, R, Et);
            for (int i = 0; i < 9; ++i) dot += Et[i] * s_E[i];
            if (dot < 0) { t[0] = -t[0]; t[1] = -t[1]; t[2] = -t[2]; }
            for (int i = 0; i < 9; ++i) s_R[i] = R[i];
            for (int i = 0; i < 3; ++i) s_t[i] = t[i];
        }
        __syncthreads();
        lo_pass(b, vp, s_R, s_t, s_E, s_dE, s_red, s_out, s_q);
        auto unpack = [&](double* jtj, double* jtr) {
            for (int i = 0; i < 5; ++i) jtr[i] = s_out[2 + i];
            int k = 7;
            for (int i = 0; i < 5; ++i) for (int u = i; u < 5; ++u) { jtj[5 * i + u] = s_out[k]; jtj[5 * u + i] = s_out[k]; ++k; }
        };
        if (tid == 0) { cst = s_out[0]; ninl = static_cast<int>(s_out[1]); unpack(JtJ, Jtr); }
        for (;;) {
            if (tid == 0) {
                bool done = false;
                for (;;) {
                    if (it >= d.lo_iters || ninl < 5) { done = true; break; }
                    double dl[5];
                    if (!solve_lm5(JtJ, Jtr, lambda, dl)) { lambda *= 10.0; ++it; if (lambda > 1e4) { done = true; break; } continue; }
                    double dR[9], b1[3], b2[3], tn[3];
                    rodrigues(dl, dR);
                    vm::mat3_mul(R, dR, s_R);
                    tangent_basis(t, b1, b2);
                    for (int i = 0; i < 3; ++i) tn[i] = t[i] + b1[i] * dl[3] + b2[i] * dl[4];
                    const double nn = vm::norm3(tn);
                    for (int i = 0; i < 3; ++i) s_t[i] = tn[i] / nn;
                    break;
                }
                s_ctrl = done ? 1 : 0;
            }
            __syncthreads();
            if (s_ctrl) break;
            lo_pass(b, vp, s_R, s_t, s_E, s_dE, s_red, s_out, s_q);
            if (tid == 0) {
                const double cn = s_out[0];
                if (cn < cst) {
                    const double gain = cst - cn;
                    const bool stop = !(gain > 1e-10 * cst);
                    for (int i = 0; i < 9; ++i) R[i] = s_R[i];
                    for (int i = 0; i < 3; ++i) t[i] = s_t[i];
                    unpack(JtJ, Jtr);
                    cst = cn;
                    ninl = static_cast<int>(s_out[1]);
                    lambda *= 0.1;
                    if (lambda < 1e-9) lambda = 1e-9;
                    ++it;
                    if (stop) it = d.lo_iters;
                } else {
                    lambda *= 10.0;
                    ++it;
                    if (lambda > 1e4) it = d.lo_iters;
                }
            }
        }
        if (tid == 0) skew_times(t, R, s_E);
        __syncthreads();
    }

    // ---- (e) final mask: sampson^2 < thr^2 (strict), fp64
    double E[9];
    for (int i = 0; i < 9; ++i) E[i] = s_E[i];
    {
        double cnt[1] = {0};
        for (int i = tid; i < vp.M; i += kFinThreads) {
            const double4 c = static_cast<const double4*>(b.corr_d)[vp.m0 + i];
            const bool in = vm::sampson_sq(E, c.x, c.y, c.z, c.w) < vp.thr2;
            b.mask[vp.m0 + i] = in ? 1 : 0;
            cnt[0] += in ? 1.0 : 0.0;
        }
        block_sum<1>(cnt, s_red, s_out);
    }
    const int n_in = static_cast<int>(s_out[0]);
    __syncthreads();
    if (tid < 9) b.E[p * 9 + tid] = E[tid];
    if (tid == 0) b.ninl[p] = n_in;
    if (n_in == 0) { if (tid == 0) b.status[p] = 2; return; }

    // ---- (f) pose recovery (cv2.recoverPose): 4 candidates, cheirality vote over the inliers
    recover_pose_block(b, vp, p, E, s_pose, s_red, s_out, s_q);
}

// ---------------------------------------------------------------------------------------------------------------- launch
// step: 0 gather/normalise, 1 solve, 2 score, 3 finalize
cudaError_t launch_verify_step(int step, const StoreView& st, const VerifyBufs& b, const VerifyDims& d, int n_pairs, int max_m, cudaStream_t s)
{
    if (n_pairs <= 0) return cudaSuccess;
    static g2v::DeviceOnce once;
    int& configured = once.cur();
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kScoreTile * 16);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSolveSmem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(topk_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSolveSmem);
        if (e != cudaSuccess) return e;
        configured = 1;
    }
    if (step == 0) {
        const int gx = max_m > 0 ? (max_m + 255) / 256 : 1;
        gather_normalize_kernel<<<dim3(gx, n_pairs), 256, 0, s>>>(st, b, n_pairs);
    } else if (step == 1) {
        const long long nb = (static_cast<long long>(n_pairs) * d.H + kSolveThreads - 1) / kSolveThreads;
        int n_sm = 148, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
        const long long grid = nb < 2LL * n_sm ? nb : 2LL * n_sm;
        solve_kernel<<<static_cast<unsigned>(grid), kSolveThreads, kSolveSmem, s>>>(b, d, n_pairs);
    } else if (step == 2) {
        score_kernel<<<dim3((d.H + kScoreHyp - 1) / kScoreHyp, n_pairs), 256, kScoreTile * 16, s>>>(b, d, n_pairs);
    } else {
        topk_select_kernel<<<n_pairs, 256, 0, s>>>(b, d, n_pairs);
        const unsigned nb = static_cast<unsigned>((static_cast<long long>(n_pairs) * kTopK + kSolveThreads - 1) / kSolveThreads);
        topk_solve_kernel<<<nb, kSolveThreads, kSolveSmem, s>>>(b, d, n_pairs);
        finalize_kernel<<<n_pairs, kFinThreads, 0, s>>>(b, d, n_pairs);
    }
    return cudaGetLastError();
}

}  // namespace g2v
