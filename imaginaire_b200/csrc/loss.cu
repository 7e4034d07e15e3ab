// Mean-absolute-difference loss on bf16 feature maps, forward and backward in one pass each.
//
// ref: the perceptual and feature-matching losses of the SPADE trainer compare feature maps with
//      nn.L1Loss: imaginaire/losses/perceptual.py:66-67, :130-131 (five VGG-19 levels per image
//      pair), imaginaire/losses/feature_matching.py:12, :36-38 (every discriminator level).
//
// On the bf16 feature maps of this build the ATen expression
//     l1_loss(a.float(), b.float())
// is two up-casts, sub, abs, mean forward and sgn, mul, div, down-cast backward: ~60 bytes of HBM
// traffic per element in ten launches per level.  Here: forward reads a and b once (4 B per
// element), backward reads them again and writes the bf16 gradient (6 B per element).  The
// arithmetic is the same: |float(a) - float(b)| accumulated in fp32; gradient sign(a - b) * (upstream / n)
// rounded to bf16 once (sign(0) = 0, like torch).
//
// The tensors are treated as flat arrays: the caller guarantees that a, b (and the gradient) are
// dense and share one memory layout.
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

constexpr int kThreadsL = 256;
constexpr int kUnroll = 4;

__device__ __forceinline__ float abs_diff8(const uint4& a, const uint4& b) {
    float s = 0.0f;
    s += fabsf(imx::bf16_lo(a.x) - imx::bf16_lo(b.x)); s += fabsf(imx::bf16_hi(a.x) - imx::bf16_hi(b.x));
    s += fabsf(imx::bf16_lo(a.y) - imx::bf16_lo(b.y)); s += fabsf(imx::bf16_hi(a.y) - imx::bf16_hi(b.y));
    s += fabsf(imx::bf16_lo(a.z) - imx::bf16_lo(b.z)); s += fabsf(imx::bf16_hi(a.z) - imx::bf16_hi(b.z));
    s += fabsf(imx::bf16_lo(a.w) - imx::bf16_lo(b.w)); s += fabsf(imx::bf16_hi(a.w) - imx::bf16_hi(b.w));
    return s;
}

// *sum += sum_i |a[i] - b[i]|.  n8 = number of 8-element vectors; the n - 8 * n8 tail elements are
// added by one thread.  The minimum-CTA launch bound keeps the 2 * kUnroll loads in distinct
// registers, i.e. really in flight (see spade_norm.cu).
__global__ void __launch_bounds__(kThreadsL, 4)
l1_sum_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, long n8, long n, float* __restrict__ sum) {
    __shared__ float warp_sums[kThreadsL / 32];
    const long stride = (long)gridDim.x * kThreadsL;
    long i = (long)blockIdx.x * kThreadsL + threadIdx.x;
    float s = 0.0f;
    for (; i + (kUnroll - 1) * stride < n8; i += kUnroll * stride) {
        uint4 va[kUnroll], vb[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            va[u] = imx::ldg_stream(a + i + u * stride);
            vb[u] = imx::ldg_stream(b + i + u * stride);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) s += abs_diff8(va[u], vb[u]);
    }
    for (; i < n8; i += stride) s += abs_diff8(imx::ldg_stream(a + i), imx::ldg_stream(b + i));
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const uint16_t* ta = reinterpret_cast<const uint16_t*>(a);
        const uint16_t* tb = reinterpret_cast<const uint16_t*>(b);
        for (long k = n8 * 8; k < n; ++k)
            s += fabsf(__uint_as_float((uint32_t)ta[k] << 16) - __uint_as_float((uint32_t)tb[k] << 16));
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < kThreadsL / 32; ++w) t += warp_sums[w];
        atomicAdd(sum, t);
    }
}

__device__ __forceinline__ float sgn_scaled(float d, float s) {
    return d > 0.0f ? s : (d < 0.0f ? -s : 0.0f);
}

__device__ __forceinline__ uint32_t grad2(uint32_t a, uint32_t b, float s) {
    return imx::pack_bf16(sgn_scaled(imx::bf16_lo(a) - imx::bf16_lo(b), s), sgn_scaled(imx::bf16_hi(a) - imx::bf16_hi(b), s));
}

// grad[i] = sign(a[i] - b[i]) * ((*upstream) / denom)   (a division, like autograd's mean backward)
__global__ void __launch_bounds__(kThreadsL, 4)
l1_grad_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ grad, long n8, long n,
               const float* __restrict__ upstream, float denom) {
    const float s = __fdiv_rn(__ldg(upstream), denom);
    const long stride = (long)gridDim.x * kThreadsL;
    long i = (long)blockIdx.x * kThreadsL + threadIdx.x;
    for (; i + (kUnroll - 1) * stride < n8; i += kUnroll * stride) {
        uint4 va[kUnroll], vb[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            va[u] = imx::ldg_stream(a + i + u * stride);
            vb[u] = imx::ldg_stream(b + i + u * stride);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
            imx::stg_stream(grad + i + u * stride, make_uint4(grad2(va[u].x, vb[u].x, s), grad2(va[u].y, vb[u].y, s),
                                                              grad2(va[u].z, vb[u].z, s), grad2(va[u].w, vb[u].w, s)));
    }
    for (; i < n8; i += stride) {
        const uint4 va = imx::ldg_stream(a + i), vb = imx::ldg_stream(b + i);
        imx::stg_stream(grad + i, make_uint4(grad2(va.x, vb.x, s), grad2(va.y, vb.y, s), grad2(va.z, vb.z, s), grad2(va.w, vb.w, s)));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const uint16_t* ta = reinterpret_cast<const uint16_t*>(a);
        const uint16_t* tb = reinterpret_cast<const uint16_t*>(b);
        uint16_t* tg = reinterpret_cast<uint16_t*>(grad);
        for (long k = n8 * 8; k < n; ++k) {
            const float d = __uint_as_float((uint32_t)ta[k] << 16) - __uint_as_float((uint32_t)tb[k] << 16);
            tg[k] = (uint16_t)(imx::pack_bf16(sgn_scaled(d, s), 0.0f) & 0xffffu);
        }
    }
}

// Few, long-lived CTAs: every CTA of the forward ends with one atomic on the same address.
inline int l1_blocks(long n8) {
    long b = (n8 + (long)kThreadsL * kUnroll - 1) / ((long)kThreadsL * kUnroll);
    const long cap = 2L * imx::kNumSMs;
    if (b > cap) b = cap;
    return (int)(b < 1 ? 1 : b);
}

// ------------------------------------------------------------------ per-pixel channel dot product
// out[p] = sum_c a[p][c] * b[p][c] for NHWC bf16 tensors (the projection term of the FPSE
// discriminator: (segmap embedding * feature).sum(dim=1), imaginaire/discriminators/fpse.py:125-127).
// Through ATen this is a materialised product tensor plus a reduction forward and two broadcast
// products backward.  One warp per pixel; fp32 accumulation, one rounding.
__device__ __forceinline__ float dot8(const uint4& a, const uint4& b) {
    float s = imx::bf16_lo(a.x) * imx::bf16_lo(b.x);
    s = fmaf(imx::bf16_hi(a.x), imx::bf16_hi(b.x), s);
    s = fmaf(imx::bf16_lo(a.y), imx::bf16_lo(b.y), s); s = fmaf(imx::bf16_hi(a.y), imx::bf16_hi(b.y), s);
    s = fmaf(imx::bf16_lo(a.z), imx::bf16_lo(b.z), s); s = fmaf(imx::bf16_hi(a.z), imx::bf16_hi(b.z), s);
    s = fmaf(imx::bf16_lo(a.w), imx::bf16_lo(b.w), s); s = fmaf(imx::bf16_hi(a.w), imx::bf16_hi(b.w), s);
    return s;
}

__global__ void __launch_bounds__(kThreadsL)
channel_dot_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint16_t* __restrict__ out, long pixels, int cg) {
    const int lane = threadIdx.x & 31;
    const long warps = (long)gridDim.x * (kThreadsL / 32);
    for (long p = (long)blockIdx.x * (kThreadsL / 32) + (threadIdx.x >> 5); p < pixels; p += warps) {
        float s = 0.0f;
        for (int c8 = lane; c8 < cg; c8 += 32) s += dot8(imx::ldg_stream(a + p * cg + c8), imx::ldg_stream(b + p * cg + c8));
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
        if (lane == 0) out[p] = (uint16_t)(imx::pack_bf16(s, 0.0f) & 0xffffu);
    }
}

__device__ __forceinline__ uint32_t scale2(uint32_t v, float g) {
    return imx::pack_bf16(imx::bf16_lo(v) * g, imx::bf16_hi(v) * g);
}

// da[p][c] = g[p] * b[p][c],  db[p][c] = g[p] * a[p][c]   (either output may be NULL)
__global__ void __launch_bounds__(kThreadsL)
channel_dot_bwd_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const uint16_t* __restrict__ g,
                       uint4* __restrict__ da, uint4* __restrict__ db, long pixels, int cg) {
    const long total = pixels * cg;
    for (long i = (long)blockIdx.x * kThreadsL + threadIdx.x; i < total; i += (long)gridDim.x * kThreadsL) {
        const float gv = __uint_as_float((uint32_t)g[i / cg] << 16);
        if (da) {
            const uint4 v = imx::ldg_stream(b + i);
            imx::stg_stream(da + i, make_uint4(scale2(v.x, gv), scale2(v.y, gv), scale2(v.z, gv), scale2(v.w, gv)));
        }
        if (db) {
            const uint4 v = imx::ldg_stream(a + i);
            imx::stg_stream(db + i, make_uint4(scale2(v.x, gv), scale2(v.y, gv), scale2(v.z, gv), scale2(v.w, gv)));
        }
    }
}

}  // namespace

extern "C" int imx_l1_abs_sum_bf16(const imx_bf16* a, const imx_bf16* b, int64_t n, float* sum, imx_stream_t stream) {
    IMX_REQUIRE(n >= 0, "l1_abs_sum: negative size");
    if (n == 0) return IMX_OK;
    IMX_REQUIRE(a && b && sum, "l1_abs_sum: null pointer");
    IMX_REQUIRE(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0,
                "l1_abs_sum: a and b must be 16-byte aligned");
    const long n8 = n / 8;
    l1_sum_kernel<<<l1_blocks(n8), kThreadsL, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), n8, (long)n, sum);
    return imx::check_launch("l1_abs_sum");
}

extern "C" int imx_l1_abs_grad_bf16(const imx_bf16* a, const imx_bf16* b, imx_bf16* grad, int64_t n,
                                    const float* upstream, float denom, imx_stream_t stream) {
    IMX_REQUIRE(n >= 0, "l1_abs_grad: negative size");
    if (n == 0) return IMX_OK;
    IMX_REQUIRE(a && b && grad && upstream, "l1_abs_grad: null pointer");
    IMX_REQUIRE(denom != 0.0f, "l1_abs_grad: zero denominator");
    IMX_REQUIRE(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(grad)) & 15) == 0,
                "l1_abs_grad: a, b and grad must be 16-byte aligned");
    const long n8 = n / 8;
    long blocks = (n8 + (long)kThreadsL * kUnroll - 1) / ((long)kThreadsL * kUnroll);
    const long cap = 4L * imx::kNumSMs;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    l1_grad_kernel<<<(int)blocks, kThreadsL, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), reinterpret_cast<uint4*>(grad), n8, (long)n,
        upstream, denom);
    return imx::check_launch("l1_abs_grad");
}

extern "C" int imx_channel_dot_nhwc_bf16(const imx_bf16* a, const imx_bf16* b, imx_bf16* out, int64_t pixels, int C,
                                         imx_stream_t stream) {
    IMX_REQUIRE(pixels >= 0 && C >= 0 && C % 8 == 0, "channel_dot: bad size (C %% 8 must be 0)");
    if (pixels == 0) return IMX_OK;
    IMX_REQUIRE(a && b && out, "channel_dot: null pointer");
    long blocks = (pixels + kThreadsL / 32 - 1) / (kThreadsL / 32);
    if (blocks > 16L * imx::kNumSMs) blocks = 16L * imx::kNumSMs;
    channel_dot_kernel<<<(int)blocks, kThreadsL, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), reinterpret_cast<uint16_t*>(out), (long)pixels, C / 8);
    return imx::check_launch("channel_dot");
}

extern "C" int imx_channel_dot_bwd_nhwc_bf16(const imx_bf16* a, const imx_bf16* b, const imx_bf16* grad_out,
                                             imx_bf16* grad_a, imx_bf16* grad_b, int64_t pixels, int C, imx_stream_t stream) {
    IMX_REQUIRE(pixels >= 0 && C >= 0 && C % 8 == 0, "channel_dot_bwd: bad size (C %% 8 must be 0)");
    if (pixels * C == 0 || (!grad_a && !grad_b)) return IMX_OK;
    IMX_REQUIRE(a && b && grad_out, "channel_dot_bwd: null pointer");
    const long total = (long)pixels * (C / 8);
    long blocks = (total + kThreadsL - 1) / kThreadsL;
    if (blocks > 16L * imx::kNumSMs) blocks = 16L * imx::kNumSMs;
    channel_dot_bwd_kernel<<<(int)blocks, kThreadsL, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(a), reinterpret_cast<const uint4*>(b), reinterpret_cast<const uint16_t*>(grad_out),
        reinterpret_cast<uint4*>(grad_a), reinterpret_cast<uint4*>(grad_b), (long)pixels, C / 8);
    return imx::check_launch("channel_dot_bwd");
}

