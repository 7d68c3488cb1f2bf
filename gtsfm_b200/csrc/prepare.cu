// Image upload preprocessing: one pass over the caller's float32 descriptors builds everything the matching kernels
// read -- the bf16 operand of the tensor-core GEMM, the uint8 exact copy, the integer squared norms, the two "aug"
// K-step rows that fold (|q|^2 + |t|^2) / 2 into the accumulator, and the NaN-row mask of
// gtsfm/frontend/matcher/twoway_matcher.py:75-76.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "store.h"

namespace g2v {

struct PrepareArgs {
    const float* src;       // [n][dim] float32 as uploaded
    __nv_bfloat16* bf;      // [n_pad][kpad]
    uint8_t* u8;            // [n_pad][kpad]
    __nv_bfloat16* augq;    // [n_pad][16]
    __nv_bfloat16* augt;    // [n_pad][16]
    int* norm_i;            // [n_pad]
    uint8_t* valid;         // [n_pad]
    int* img_flags;         // [0] |= 1 if some valid row is not integer-valued in [0,255] or its norm is >= 2^23,
                            //     |= 2 if some valid row has |x|^2 > kF16MaxNorm2 (not eligible for the fp16 screening)
                            // [1] += number of valid rows; [2] = max |x|^2 over the valid rows (float bits)
    __half* h16;            // [n_pad][kpad]  fp16 copy: operand of the float-path screening GEMM
    __half* augq_h;         // [n_pad][16]    fp16 "aug" rows of the float path
    __half* augt_h;         // [n_pad][16]
    float* norm_f;          // [n_pad]        |x|^2 in float32
    int n, n_pad, dim, kpad;
};

// one warp per row
__global__ void __launch_bounds__(256) prepare_rows_kernel(PrepareArgs a)
{
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= a.n_pad) return;
    bool has_nan = false, nonint = false;
    float nrm = 0.f;
    if (row < a.n) {
        for (int c = lane; c < a.dim; c += 32) {
            const float v = a.src[static_cast<long long>(row) * a.dim + c];
            has_nan |= isnan(v);
            nonint |= !(v >= 0.f && v <= 255.f && v == rintf(v));
            nrm += v * v;
        }
    }
    has_nan = __any_sync(0xffffffffu, has_nan);
    nonint = __any_sync(0xffffffffu, nonint);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, off);
    const bool ok = row < a.n && !has_nan;
    // integer path needs every squared distance and partial sum below 2^24: require |x|^2 < 2^23 per row
    if (ok && (nonint || nrm >= 8388608.f)) nonint = true;
    for (int c = lane; c < a.kpad; c += 32) {
        const float v = (ok && c < a.dim) ? a.src[static_cast<long long>(row) * a.dim + c] : 0.f;
        a.bf[static_cast<long long>(row) * a.kpad + c] = __float2bfloat16_rn(v);
        a.h16[static_cast<long long>(row) * a.kpad + c] = __float2half_rn(v);
        a.u8[static_cast<long long>(row) * a.kpad + c] = (ok && !nonint) ? static_cast<uint8_t>(v) : 0;
    }
    if (lane < 16) {
        // |x|^2 split into three bf16-exact pieces (8 significant bits each for the integer path; a 3-term bf16
        // expansion otherwise).  Rows that must never be matched get 2^100.
        float p0, p1, p2;
        if (!ok) { p0 = 0.f; p1 = 0.f; p2 = kHuge; }
        else if (!nonint) {
            const int ni = static_cast<int>(nrm);
            p0 = static_cast<float>(ni & 0xff); p1 = static_cast<float>(ni & 0xff00); p2 = static_cast<float>(ni & 0x7fff0000);
        } else {
            p2 = __bfloat162float(__float2bfloat16_rn(nrm));
            p1 = __bfloat162float(__float2bfloat16_rn(nrm - p2));
            p0 = __bfloat162float(__float2bfloat16_rn(nrm - p2 - p1));
        }
        const float pieces[3] = {p0, p1, p2};
        float q = 0.f, t = 0.f;
        if (lane < 3) { q = pieces[lane]; t = 0.5f; }
        else if (lane < 6) { q = 0.5f; t = pieces[lane - 3]; }
        a.augq[static_cast<long long>(row) * kAugK + lane] = __float2bfloat16_rn(q);
        a.augt[static_cast<long long>(row) * kAugK + lane] = __float2bfloat16_rn(t);
        // float path: |x|^2 as a 3-term fp16 expansion (exact to 2^-33 |x|^2 for |x|^2 <= kF16MaxNorm2); rows that must
        // never be matched get kF16Never, above every d^2 / 2 an eligible pair can reach
        float h0, h1, h2;
        if (!ok || nrm > kF16MaxNorm2) { h0 = 0.f; h1 = 0.f; h2 = kF16Never; }
        else {
            h2 = __half2float(__float2half_rn(nrm));
            h1 = __half2float(__float2half_rn(nrm - h2));
            h0 = __half2float(__float2half_rn(nrm - h2 - h1));
        }
        const float hp[3] = {h0, h1, h2};
        float qh = 0.f, th = 0.f;
        if (lane < 3) { qh = hp[lane]; th = 0.5f; }
        else if (lane < 6) { qh = 0.5f; th = hp[lane - 3]; }
        a.augq_h[static_cast<long long>(row) * kAugK + lane] = __float2half_rn(qh);
        a.augt_h[static_cast<long long>(row) * kAugK + lane] = __float2half_rn(th);
    }
    if (lane == 0) {
        a.norm_i[row] = (ok && !nonint) ? static_cast<int>(nrm) : 0;
        a.valid[row] = ok ? 1 : 0;
        a.norm_f[row] = ok ? nrm : 0.f;
        if (ok && !(nrm <= kF16MaxNorm2)) atomicOr(&a.img_flags[0], 2);
        if (ok && nrm <= kF16MaxNorm2) atomicMax(&a.img_flags[2], __float_as_int(nrm));
        if (ok && nonint) atomicOr(&a.img_flags[0], 1);
        if (ok) atomicAdd(&a.img_flags[1], 1);
    }
}

cudaError_t launch_prepare(const PrepareArgs& a, cudaStream_t s)
{
    prepare_rows_kernel<<<(a.n_pad + 7) / 8, 256, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace g2v
