// Block-level pieces shared by the essential-matrix (verify_kernels.cu) and fundamental-matrix (verify_fmat.cu)
// finalize kernels: a deterministic block sum and the cv2.recoverPose restatement (gtsfm/utils/verification.py:81).
#pragma once
#include "verify_math.cuh"
#include "verify_types.h"

namespace g2v {

constexpr int kFinThreads = 128;

// deterministic block sum of NV doubles per thread: warp shuffle tree, then warps added in order by thread 0
template <int NV> __device__ void block_sum(double* v, double (*s_red)[24], double* s_out)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        double x = v[k];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(0xffffffffu, x, off);
        if (lane == 0) s_red[warp][k] = x;
    }
    __syncthreads();
    if (threadIdx.x < NV) {
        double s = 0;
        for (int w = 0; w < kFinThreads / 32; ++w) s += s_red[w][threadIdx.x];
        s_out[threadIdx.x] = s;
    }
    __syncthreads();
}

// Pose recovery of one pair by a CTA of kFinThreads threads (cv2.recoverPose): SVD of E -> {R1, R2} x {+t, -t}, DLT
// triangulation of every inlier under each candidate, cheirality vote (0 < z < 50 in both views), first maximum in the
// order (R1,t), (R2,t), (R1,-t), (R2,-t).  The inliers are the rows whose mask byte is 1; every thread reads the bytes it
// wrote itself (same stride), so no barrier is needed between the mask pass and this one.
// s_pose: 21 doubles of shared memory (R1, R2, t); s_q: a 64-entry queue per warp.
__device__ inline void recover_pose_block(const VerifyBufs& b, const VPair& vp, int p, const double* E, double* s_pose,
                                          double (*s_red)[24], double* s_out, double4 (*s_q)[64])
{
    const int tid = threadIdx.x;
    double* s_R1 = s_pose;
    double* s_R2 = s_pose + 9;
    double* s_tt = s_pose + 18;
    if (tid == 0) vm::decompose_essential(E, s_R1, s_R2, s_tt);
    __syncthreads();
    {
        double good[4] = {0, 0, 0, 0};
        auto vote = [&](const double4& c) {
#pragma unroll 1
            for (int cand = 0; cand < 4; ++cand) {
                const double* Rc = (cand & 1) ? s_R2 : s_R1;
                const double sg = (cand & 2) ? -1.0 : 1.0;
                double P1[12];
                for (int r = 0; r < 3; ++r) { P1[4 * r] = Rc[3 * r]; P1[4 * r + 1] = Rc[3 * r + 1]; P1[4 * r + 2] = Rc[3 * r + 2]; P1[4 * r + 3] = sg * s_tt[r]; }
                good[cand] += vm::cheirality_ok(P1, c.x, c.y, c.z, c.w) ? 1.0 : 0.0;
            }
        };
        // the inliers are scattered: every warp queues the ones among its next 32 rows and triangulates on full warps
        const int lane = tid & 31;
        double4* q = s_q[tid >> 5];
        int qn = 0;
        for (int i0 = tid - lane; i0 < vp.M; i0 += kFinThreads) {
            const int i = i0 + lane;
            const bool in = i < vp.M && b.mask[vp.m0 + i];
            const unsigned bal = __ballot_sync(0xffffffffu, in);
            if (in) q[qn + __popc(bal & ((1u << lane) - 1u))] = static_cast<const double4*>(b.corr_d)[vp.m0 + i];
            qn += __popc(bal);
            __syncwarp();
            if (qn >= 32) {
                qn -= 32;
                const double4 c = q[qn + lane];
                __syncwarp();
                vote(c);
            }
        }
        if (lane < qn) vote(q[lane]);
        block_sum<4>(good, s_red, s_out);
    }
    if (tid == 0) {
        const double g0 = s_out[0], g1 = s_out[1], g2 = s_out[2], g3 = s_out[3];
        int best;
        if (g0 >= g1 && g0 >= g2 && g0 >= g3) best = 0;
        else if (g1 >= g0 && g1 >= g2 && g1 >= g3) best = 1;
        else if (g2 >= g0 && g2 >= g1 && g2 >= g3) best = 2;
        else best = 3;
        const double* Rb = (best & 1) ? s_R2 : s_R1;
        const double sg = (best & 2) ? -1.0 : 1.0;
        for (int i = 0; i < 9; ++i) b.R[p * 9 + i] = Rb[i];
        for (int i = 0; i < 3; ++i) b.t[p * 3 + i] = sg * s_tt[i];
    }
}

}  // namespace g2v
