// Normalisation statistics and the fused normalise * (1+gamma) + beta -> activation
// kernels of the SPADE block.
//
// ref: SpatiallyAdaptiveNorm.forward  imaginaire/layers/activation_norm.py:211-234
//      AdaptiveNorm.forward           imaginaire/layers/activation_norm.py:79-106
//      inner norm factory             imaginaire/layers/activation_norm.py:377-432
//      LeakyReLU(0.2)                 imaginaire/layers/nonlinearity.py:22
//
// All tensors are NHWC bf16; every thread owns one fixed group of 8 channels (one
// 128-bit load) and walks down the pixel rows, so per-channel parameters live in
// registers and every warp-level access is a contiguous 512-byte row segment.
// HBM-bound: algorithmic bytes are 2*NHWC (stats: read x) and
// 2*NHWC*(2 + 2*[gamma/beta materialised]) (modulate: read x, gamma, beta, write y).
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

using imx::bf16_hi;
using imx::bf16_lo;
using imx::pack_bf16;

constexpr int kThreads = 256;

struct RowMap {
    int cg;    // 8-channel groups per pixel (C / 8)
    int tpr;   // threads per pixel row actually used
    int rows;  // pixel rows covered per block iteration
};

inline RowMap make_rowmap(int C) {
    RowMap m;
    m.cg = C / 8;
    m.tpr = m.cg < kThreads ? m.cg : kThreads;
    m.rows = kThreads / m.tpr;
    return m;
}

// Split `np` pixels of one group over blocks.  `per_sm` CTAs per SM in total: the reducing
// kernels end with 2C atomics per CTA onto the same addresses, and atomics on one address
// serialise in L2 (measured ~20 ns each: a 592-CTA grid spends ~12 us in that chain), so they
// want few, long-lived CTAs with many loads in flight per thread; the pure streaming kernels
// take more.  (Adding the CTAs' sums through distributed shared memory in clusters of 8 before
// the atomics was tried and is slower: 37 clusters of 2-CTA/SM kernels do not fit one wave.)
inline int pick_blocks(long np, int rows, int groups, int per_sm = 2) {
    long iters = (np + rows - 1) / rows;
    long want = ((long)per_sm * imx::kNumSMs + groups - 1) / groups;
    long b = iters < want ? iters : want;
    return (int)(b < 1 ? 1 : b);
}

// Statistics kernel: 3 CTAs per SM when the tensor is large (the atomic tail is then small beside
// the streaming time), 2 otherwise.
inline int stats_ctas_per_sm(long elems) { return elems * 2 >= (64L << 20) ? 3 : 2; }

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
    f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
    f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}
__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// ------------------------------------------------------------------ statistics

// grid = (blocks, groups); group g covers pixels [g*HWg, (g+1)*HWg).
// __launch_bounds__(kThreads, 4): with no minimum-CTA hint ptxas minimises registers (44) by
// recycling the destination registers of the eight loads, which leaves ~3 of them in flight; with
// the hint it keeps all eight in distinct registers (64) and the kernel goes from 0.57 to 0.82 of
// the HBM peak.
template <bool kWithSq>
__global__ void __launch_bounds__(kThreads, 4)
stats_kernel(const uint4* __restrict__ x, float* __restrict__ sum, float* __restrict__ sumsq,
             long HWg, int C, int cg, int tpr, int rows, int row_stride_c8) {
    constexpr int NV = kWithSq ? 16 : 8;
    __shared__ float red[kThreads][NV + 1];
    const int tid = threadIdx.x;
    const int ci = tid % tpr, r = tid / tpr;
    const bool active = r < rows;
    const int g = blockIdx.y;
    // block-interleaved pixel rows inside the group (see modulate_kernel)
    const long p0 = (long)blockIdx.x * rows;
    const long p1 = HWg;
    const long pstep = (long)rows * gridDim.x;
    const uint4* base = x + (size_t)g * HWg * row_stride_c8;
    for (int c8b = 0; c8b < cg; c8b += tpr) {   // block-uniform trip count (barriers inside)
        const int c8 = c8b + ci;
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (active && c8 < cg) {
            long p = p0 + r;
            // 8 independent 128-bit loads in flight per thread
            for (; p + 7L * pstep < p1; p += 8L * pstep) {
                uint4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = imx::ldg_stream(base + (size_t)(p + (long)u * pstep) * row_stride_c8 + c8);
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    float f[8];
                    unpack8(v[u], f);
#pragma unroll
                    for (int k = 0; k < 8; ++k) { s[k] += f[k]; if (kWithSq) q[k] = fmaf(f[k], f[k], q[k]); }
                }
            }
            for (; p < p1; p += pstep) {
                float f[8];
                unpack8(imx::ldg_stream(base + (size_t)p * row_stride_c8 + c8), f);
#pragma unroll
                for (int k = 0; k < 8; ++k) { s[k] += f[k]; if (kWithSq) q[k] = fmaf(f[k], f[k], q[k]); }
            }
        }
        // reduce the `rows` partial sums that share this channel group, then one vector atomic per CTA and channel group
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            red[tid][k] = s[k];
            if (kWithSq) red[tid][8 + k] = q[k];
        }
        __syncthreads();
        float acc[NV] = {};
        if (r == 0) {
#pragma unroll
            for (int k = 0; k < NV; ++k) acc[k] = red[tid][k];
            for (int rr = 1; rr < rows; ++rr)
#pragma unroll
                for (int k = 0; k < NV; ++k) acc[k] += red[tid + rr * tpr][k];
        }
        if (r == 0 && c8 < cg) {
            float* dst = sum + (size_t)g * C + c8 * 8;
            imx::red_add_v4(dst, acc[0], acc[1], acc[2], acc[3]);
            imx::red_add_v4(dst + 4, acc[4], acc[5], acc[6], acc[7]);
            if (kWithSq) {
                float* dq = sumsq + (size_t)g * C + c8 * 8;
                imx::red_add_v4(dq, acc[NV - 8], acc[NV - 7], acc[NV - 6], acc[NV - 5]);
                imx::red_add_v4(dq + 4, acc[NV - 4], acc[NV - 3], acc[NV - 2], acc[NV - 1]);
            }
        }
    }
}

// dx = dy * act'(y) for an activation that was fused into a conv epilogue (y = act(pre)), plus
// the per-channel sum of dx (the conv's bias gradient) in the same pass.
__global__ void __launch_bounds__(kThreads, 4)
act_bwd_sum_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y, uint4* __restrict__ dx,
                   float* __restrict__ sum, long NP, int cg, int tpr, int rows, float slope) {
    __shared__ float red[kThreads][9];
    const int tid = threadIdx.x;
    const int ci = tid % tpr, r = tid / tpr;
    const bool active = r < rows;
    const long p0 = (long)blockIdx.x * rows;   // block-interleaved pixel rows (see modulate_kernel)
    const long p1 = NP;
    const long pstep = (long)rows * gridDim.x;
    for (int c8b = 0; c8b < cg; c8b += tpr) {   // block-uniform trip count (barriers inside)
        const int c8 = c8b + ci;
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (active && c8 < cg) {
            auto one = [&](const uint4 gv, const uint4 yv, long p) {
                float g[8], yy[8];
                unpack8(gv, g);
                unpack8(yv, yy);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    g[k] = yy[k] > 0.0f ? g[k] : g[k] * slope;
                }
                const uint4 o = pack8(g);
                imx::stg_stream(dx + (size_t)p * cg + c8, o);
                // the sum is taken over the values as stored, like a separate pass over dx would
                unpack8(o, g);
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] += g[k];
            };
            long p = p0 + r;
            for (; p + 3L * pstep < p1; p += 4L * pstep) {
                uint4 gv[4], yv[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    gv[u] = imx::ldg_stream(dy + (size_t)(p + (long)u * pstep) * cg + c8);
                    yv[u] = imx::ldg_stream(y + (size_t)(p + (long)u * pstep) * cg + c8);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) one(gv[u], yv[u], p + (long)u * pstep);
            }
            for (; p < p1; p += pstep)
                one(imx::ldg_stream(dy + (size_t)p * cg + c8), imx::ldg_stream(y + (size_t)p * cg + c8), p);
        }
        if (sum) {
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 8; ++k) red[tid][k] = s[k];
            __syncthreads();
            float acc[8] = {};
            if (r == 0) {
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = red[tid][k];
                for (int rr = 1; rr < rows; ++rr)
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[k] += red[tid + rr * tpr][k];
            }
            if (r == 0 && c8 < cg) {
                float* dst = sum + c8 * 8;
                imx::red_add_v4(dst, acc[0], acc[1], acc[2], acc[3]);
                imx::red_add_v4(dst + 4, acc[4], acc[5], acc[6], acc[7]);
            }
        }
    }
}

__global__ void finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sumsq,
                                float count, float eps, float* __restrict__ mean,
                                float* __restrict__ rstd, float* running_mean, float* running_var,
                                float momentum, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float m = sum[i] / count;
    const float var = fmaxf(sumsq[i] / count - m * m, 0.0f);
    mean[i] = m;
    rstd[i] = rsqrtf(var + eps);
    if (running_mean) {
        const float unbiased = count > 1.0f ? var * (count / (count - 1.0f)) : var;
        running_mean[i] = (1.0f - momentum) * running_mean[i] + momentum * m;
        running_var[i] = (1.0f - momentum) * running_var[i] + momentum * unbiased;
    }
}

// --------------------------------------------------------------- modulation

// The activation is carried as its negative-side slope (LeakyReLU 0.2, ReLU 0, none 1), so the
// per-element code is select / min / max instructions instead of a branch on the activation type.
inline float act_slope(int act) { return act == 1 ? 0.2f : (act == 2 ? 0.0f : 1.0f); }
__device__ __forceinline__ float act_fwd(float v, float slope) {
    return fmaxf(v, 0.0f) + slope * fminf(v, 0.0f);
}
__device__ __forceinline__ float act_grad(float pre, float slope) {
    return pre > 0.0f ? 1.0f : slope;
}

struct ModArgs {
    const uint4* x;
    const uint4* dy;
    const float* mean;
    const float* rstd;
    const float* w;
    const float* b;
    const void* gamma;
    const void* beta;
    int N, H, W, C;
    int per_instance, gb_mode, up2x;
    float slope;    // activation as its negative-side slope (act_slope)
    float* gbsum;   // backward pass 1, gb_mode 0: optional (2, C) per-channel sums of dgamma / dbeta
};

// Per-thread channel parameters: xn = x * rs + mr (inner norm), xh = xn * w + b (affine).
__device__ __forceinline__ void load_norm_params(const ModArgs& a, int n, int c, float (&rs)[8], float (&mr)[8],
                                                 float (&ww)[8], float (&bb)[8]) {
    float m[8];
    const int so = (a.per_instance ? n * a.C : 0) + c;
    load8f(a.mean + so, m);
    load8f(a.rstd + so, rs);
    if (a.w) load8f(a.w + c, ww);
    if (a.b) load8f(a.b + c, bb);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        mr[k] = -m[k] * rs[k];
        if (!a.w) ww[k] = 1.0f;
        if (!a.b) bb[k] = 0.0f;
    }
}

// grid = (blocks, N).  Each block walks a pixel range of one sample.
// kAffine: the inner norm has a learned weight / bias (false for SPADE's parameter-free norm, which
// then carries 16 fewer live registers per thread).
template <int kMode, bool kAffine>  // 0: forward   1: backward pass 1 (dgamma/dbeta + reductions)   2: backward pass 2 (dx)
__global__ void __launch_bounds__(kThreads, 2)
modulate_kernel(ModArgs a, uint4* __restrict__ out0, void* __restrict__ out1, void* __restrict__ out2,
                float* __restrict__ red, float inv_count, int use_batch_stats,
                int cg, int tpr, int rows) {
    __shared__ float sred[kMode == 1 ? kThreads : 1][17];
    const int tid = threadIdx.x;
    const int ci = tid % tpr, r = tid / tpr;
    const bool active = r < rows;
    const int n = blockIdx.y;
    const long HW = (long)a.H * a.W;
    // Block-interleaved pixel rows (block b takes rows b, b + gridDim.x, ... of `rows` pixels): at any
    // moment the resident CTAs together read one compact window of each tensor, which keeps more
    // DRAM pages open than giving every CTA its own distant contiguous range (1-6 % faster).
    const long p0 = (long)blockIdx.x * rows;
    const long p1 = HW;
    const long pstep = (long)rows * gridDim.x;
    const int Wx = a.up2x ? a.W / 2 : a.W;
    const long HWx = a.up2x ? HW / 4 : HW;
    const uint4* xb = a.x + (size_t)n * HWx * cg;
    const size_t ob = (size_t)n * HW * cg;
    const bool want_gbsum = kMode == 1 && (a.gb_mode == 1 || a.gbsum != nullptr);

    for (int c8b = 0; c8b < cg; c8b += tpr) {   // block-uniform trip count (barriers inside)
        const int c8 = c8b + ci < cg ? c8b + ci : cg - 1;   // surplus threads shadow the last group, write nothing
        const bool own = c8b + ci < cg;
        const int c = c8 * 8;
        float rs[8], mr[8], ww[8], bb[8];
        load_norm_params(a, n, c, rs, mr, ww, bb);
        float gv[8], bv[8];
        if (a.gb_mode == 1) {
            load8f(reinterpret_cast<const float*>(a.gamma) + (size_t)n * a.C + c, gv);
            load8f(reinterpret_cast<const float*>(a.beta) + (size_t)n * a.C + c, bv);
        }
        float m1[8], m2[8];  // pass 2: mean(g), mean(g * xn) in xhat-affine units
        if (kMode == 2) {
            if (use_batch_stats) {
                const int so = (a.per_instance ? n * a.C : 0) + c;
                const int G = a.per_instance ? a.N : 1;
                load8f(red + so, m1);
                load8f(red + (size_t)G * a.C + so, m2);
#pragma unroll
                for (int k = 0; k < 8; ++k) { m1[k] *= inv_count; m2[k] *= inv_count; }
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) { m1[k] = 0.0f; m2[k] = 0.0f; }
            }
        }
        float r1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, r2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        float ag[8] = {0, 0, 0, 0, 0, 0, 0, 0}, ab[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (active && own) {
            // Raw 128-bit loads of one pixel row.  The next row's loads are issued before the
            // current row is processed (register double buffering): twice the bytes in flight
            // per thread, which is what a 2-CTA/SM kernel needs to approach HBM bandwidth.
            struct Raw { uint4 x, g, b, dy; };
            auto load_raw = [&](long p) {
                Raw q;
                long px = p;
                if (a.up2x) {
                    const int yy = (int)(p / a.W), xx = (int)(p - (long)yy * a.W);
                    px = (long)(yy >> 1) * Wx + (xx >> 1);
                }
                q.x = a.up2x ? __ldg(xb + (size_t)px * cg + c8) : imx::ldg_stream(xb + (size_t)px * cg + c8);
                if (a.gb_mode == 0) {
                    q.g = imx::ldg_stream(reinterpret_cast<const uint4*>(a.gamma) + ob + (size_t)p * cg + c8);
                    q.b = imx::ldg_stream(reinterpret_cast<const uint4*>(a.beta) + ob + (size_t)p * cg + c8);
                } else {
                    q.g = make_uint4(0, 0, 0, 0);
                    q.b = make_uint4(0, 0, 0, 0);
                }
                if (kMode != 0) q.dy = imx::ldg_stream(a.dy + ob + (size_t)p * cg + c8);
                else q.dy = make_uint4(0, 0, 0, 0);
                return q;
            };
            Raw cur = load_raw(p0 + r < p1 ? p0 + r : p1 - 1);
            for (long p = p0 + r; p < p1; p += pstep) {
                const long pn = p + pstep;
                const Raw nxt = load_raw(pn < p1 ? pn : p);   // tail: harmless re-read of the current row
                float xv[8], g[8], be[8];
                unpack8(cur.x, xv);
                if (a.gb_mode == 0) {
                    unpack8(cur.g, g);
                    unpack8(cur.b, be);
                } else if (a.gb_mode == 1) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) { g[k] = gv[k]; be[k] = bv[k]; }
                } else {
#pragma unroll
                    for (int k = 0; k < 8; ++k) { g[k] = 0.0f; be[k] = 0.0f; }
                }
                if (kMode == 0) {
                    float y[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float xn = fmaf(xv[k], rs[k], mr[k]);
                        const float xh = kAffine ? fmaf(xn, ww[k], bb[k]) : xn;
                        y[k] = act_fwd(fmaf(xh, 1.0f + g[k], be[k]), a.slope);
                    }
                    imx::stg_stream(out0 + ob + (size_t)p * cg + c8, pack8(y));
                } else {
                    float dyv[8];
                    unpack8(cur.dy, dyv);
                    float o0[8], o1[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float xn = fmaf(xv[k], rs[k], mr[k]);       // normalised x
                        const float xh = kAffine ? fmaf(xn, ww[k], bb[k]) : xn;   // + inner affine
                        const float pre = fmaf(xh, 1.0f + g[k], be[k]);
                        const float dpre = dyv[k] * act_grad(pre, a.slope);
                        const float dxh = dpre * (1.0f + g[k]);           // grad wrt xh
                        if (kMode == 1) {
                            o0[k] = dpre * xh;   // dgamma
                            o1[k] = dpre;        // dbeta
                            r1[k] += dxh;                  // -> dbias of the inner norm
                            r2[k] = fmaf(dxh, xn, r2[k]);  // -> dweight of the inner norm
                            if (want_gbsum) { ag[k] += o0[k]; ab[k] += o1[k]; }
                        } else {
                            // dx = rstd * w * (dxh - mean(dxh) - xn * mean(dxh * xn))
                            o0[k] = (kAffine ? rs[k] * ww[k] : rs[k]) * (dxh - m1[k] - xn * m2[k]);
                        }
                    }
                    if (kMode == 1) {
                        if (a.gb_mode == 0) {
                            imx::stg_stream(reinterpret_cast<uint4*>(out1) + ob + (size_t)p * cg + c8, pack8(o0));
                            imx::stg_stream(reinterpret_cast<uint4*>(out2) + ob + (size_t)p * cg + c8, pack8(o1));
                        }
                    } else {
                        imx::stg_stream(out0 + ob + (size_t)p * cg + c8, pack8(o0));
                    }
                }
                cur = nxt;
            }
        }
        if (kMode == 1) {
            // block-level reduction of r1, r2 (and the dgamma / dbeta sums)
            for (int pass = 0; pass < (want_gbsum ? 2 : 1); ++pass) {
                __syncthreads();
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    sred[tid][k] = pass == 0 ? r1[k] : ag[k];
                    sred[tid][8 + k] = pass == 0 ? r2[k] : ab[k];
                }
                __syncthreads();
                float acc[16] = {};
                if (r == 0) {
#pragma unroll
                    for (int k = 0; k < 16; ++k) acc[k] = sred[tid][k];
                    for (int rr = 1; rr < rows; ++rr)
#pragma unroll
                        for (int k = 0; k < 16; ++k) acc[k] += sred[tid + rr * tpr][k];
                }
                if (r == 0 && own) {
                    if (pass == 0) {
                        const int G = a.per_instance ? a.N : 1;
                        const int so = (a.per_instance ? n * a.C : 0) + c;
                        float* d0 = red + so;
                        float* d1 = red + (size_t)G * a.C + so;
                        imx::red_add_v4(d0, acc[0], acc[1], acc[2], acc[3]);
                        imx::red_add_v4(d0 + 4, acc[4], acc[5], acc[6], acc[7]);
                        imx::red_add_v4(d1, acc[8], acc[9], acc[10], acc[11]);
                        imx::red_add_v4(d1 + 4, acc[12], acc[13], acc[14], acc[15]);
                    } else if (a.gb_mode == 0) {
                        // sums over all pixels of the batch = bias gradients of the gamma / beta convs
                        float* dg = a.gbsum + c;
                        float* db = a.gbsum + a.C + c;
                        imx::red_add_v4(dg, acc[0], acc[1], acc[2], acc[3]);
                        imx::red_add_v4(dg + 4, acc[4], acc[5], acc[6], acc[7]);
                        imx::red_add_v4(db, acc[8], acc[9], acc[10], acc[11]);
                        imx::red_add_v4(db + 4, acc[12], acc[13], acc[14], acc[15]);
                    } else {
                        float* dg = reinterpret_cast<float*>(out1) + (size_t)n * a.C + c;
                        float* db = reinterpret_cast<float*>(out2) + (size_t)n * a.C + c;
#pragma unroll
                        for (int k = 0; k < 8; ++k) { atomicAdd(dg + k, acc[k]); atomicAdd(db + k, acc[8 + k]); }
                    }
                }
            }
        }
    }
}

int check_mod_args(const char* who, const void* x, const float* mean, const float* rstd,
                   const void* gamma, const void* beta, int N, int H, int W, int C,
                   int gb_mode, int act, int up2x) {
    IMX_REQUIRE(x && mean && rstd, "%s: null pointer", who);
    IMX_REQUIRE(N >= 0 && H >= 0 && W >= 0 && C >= 0, "%s: negative size", who);
    IMX_REQUIRE(C % 8 == 0, "%s: C (%d) must be a multiple of 8", who, C);
    IMX_REQUIRE(gb_mode >= 0 && gb_mode <= 2, "%s: bad gb_mode %d", who, gb_mode);
    IMX_REQUIRE(gb_mode == 2 || (gamma && beta), "%s: gamma/beta missing", who);
    IMX_REQUIRE(act >= 0 && act <= 2, "%s: bad act %d", who, act);
    IMX_REQUIRE(!up2x || (H % 2 == 0 && W % 2 == 0), "%s: up2x needs even H, W", who);
    IMX_REQUIRE(N <= 65535, "%s: batch too large", who);
    return IMX_OK;
}

}  // namespace

extern "C" int imx_bn_stats_nhwc_bf16(const imx_bf16* x, float* sum, float* sumsq,
                                      int N, int HW, int C, int per_instance, imx_stream_t stream) {
    IMX_REQUIRE(x && sum && sumsq, "bn_stats: null pointer");
    IMX_REQUIRE(N >= 0 && HW >= 0 && C >= 0, "bn_stats: negative size");
    IMX_REQUIRE(C % 8 == 0, "bn_stats: C (%d) must be a multiple of 8", C);
    if ((long)N * HW * C == 0) return IMX_OK;
    const RowMap m = make_rowmap(C);
    const int groups = per_instance ? N : 1;
    IMX_REQUIRE(groups <= 65535, "bn_stats: batch too large");
    const long HWg = per_instance ? HW : (long)N * HW;
    dim3 grid(pick_blocks(HWg, m.rows, groups, stats_ctas_per_sm((long)N * HW * C)), groups);
    stats_kernel<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(x), sum, sumsq, HWg, C, m.cg, m.tpr, m.rows, m.cg);
    return imx::check_launch("bn_stats");
}

extern "C" int imx_channel_sum_nhwc_bf16(const imx_bf16* x, float* out, int NP, int C, int Cp,
                                         imx_stream_t stream) {
    IMX_REQUIRE(x && out, "channel_sum: null pointer");
    IMX_REQUIRE(NP >= 0 && C >= 0 && Cp >= C, "channel_sum: bad size");
    IMX_REQUIRE(C % 8 == 0 && Cp % 8 == 0, "channel_sum: C and Cp must be multiples of 8");
    if ((long)NP * C == 0) return IMX_OK;
    const RowMap m = make_rowmap(C);
    dim3 grid(pick_blocks(NP, m.rows, 1, stats_ctas_per_sm((long)NP * C)), 1);
    stats_kernel<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(x), out, nullptr, NP, C, m.cg, m.tpr, m.rows, Cp / 8);
    return imx::check_launch("channel_sum");
}

extern "C" int imx_act_bwd_bias_nhwc_bf16(const imx_bf16* dy, const imx_bf16* y, imx_bf16* dx, float* bias_grad,
                                          int64_t NP, int C, int act, imx_stream_t stream) {
    IMX_REQUIRE(dy && y && dx, "act_bwd_bias: null pointer");
    IMX_REQUIRE(NP >= 0 && C >= 0 && C % 8 == 0, "act_bwd_bias: C (%d) must be a non-negative multiple of 8", C);
    IMX_REQUIRE(act == 1 || act == 2, "act_bwd_bias: act must be 1 (LeakyReLU 0.2) or 2 (ReLU)");
    if (NP * C == 0) return IMX_OK;
    const RowMap m = make_rowmap(C);
    dim3 grid(pick_blocks(NP, m.rows, 1, 2), 1);
    act_bwd_sum_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(y), reinterpret_cast<uint4*>(dx), bias_grad, NP,
        m.cg, m.tpr, m.rows, act == 1 ? 0.2f : 0.0f);
    return imx::check_launch("act_bwd_bias");
}

extern "C" int imx_bn_finalize(const float* sum, const float* sumsq, float count, float eps,
                               float* mean, float* rstd, float* running_mean, float* running_var,
                               float momentum, int G, int C, imx_stream_t stream) {
    IMX_REQUIRE(sum && sumsq && mean && rstd, "bn_finalize: null pointer");
    IMX_REQUIRE(G >= 0 && C >= 0, "bn_finalize: negative size");
    IMX_REQUIRE(count > 0.0f, "bn_finalize: count must be positive");
    IMX_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "bn_finalize: running_mean/var must come together");
    IMX_REQUIRE(running_mean == nullptr || G == 1, "bn_finalize: running stats need G == 1");
    const int n = G * C;
    if (n == 0) return IMX_OK;
    finalize_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(sum, sumsq, count, eps, mean, rstd,
                                                                      running_mean, running_var, momentum, n);
    return imx::check_launch("bn_finalize");
}

extern "C" int imx_spade_modulate_fwd(const imx_bf16* x, const float* mean, const float* rstd,
                                      const float* w, const float* b,
                                      const void* gamma, const void* beta, imx_bf16* y,
                                      int N, int H, int W, int C,
                                      int stats_per_instance, int gb_mode, int act, int up2x,
                                      imx_stream_t stream) {
    int rc = check_mod_args("spade_modulate_fwd", x, mean, rstd, gamma, beta, N, H, W, C, gb_mode, act, up2x);
    if (rc) return rc;
    IMX_REQUIRE(y, "spade_modulate_fwd: null output");
    if ((long)N * H * W * C == 0) return IMX_OK;
    const RowMap m = make_rowmap(C);
    ModArgs a{reinterpret_cast<const uint4*>(x), nullptr, mean, rstd, w, b, gamma, beta,
              N, H, W, C, stats_per_instance, gb_mode, up2x, act_slope(act), nullptr};
    dim3 grid(pick_blocks((long)H * W, m.rows, N, 6), N);
    auto kern = (w || b) ? modulate_kernel<0, true> : modulate_kernel<0, false>;
    kern<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        a, reinterpret_cast<uint4*>(y), nullptr, nullptr, nullptr, 0.0f, 1, m.cg, m.tpr, m.rows);
    return imx::check_launch("spade_modulate_fwd");
}

extern "C" int imx_spade_modulate_bwd_reduce(const imx_bf16* x, const imx_bf16* dy,
                                             const float* mean, const float* rstd,
                                             const float* w, const float* b,
                                             const void* gamma, const void* beta,
                                             void* dgamma, void* dbeta, float* red, float* gb_sums,
                                             int N, int H, int W, int C,
                                             int stats_per_instance, int gb_mode, int act,
                                             imx_stream_t stream) {
    int rc = check_mod_args("spade_modulate_bwd_reduce", x, mean, rstd, gamma, beta, N, H, W, C, gb_mode, act, 0);
    if (rc) return rc;
    IMX_REQUIRE(dy && red, "spade_modulate_bwd_reduce: null pointer");
    IMX_REQUIRE(gb_mode == 2 || (dgamma && dbeta), "spade_modulate_bwd_reduce: dgamma/dbeta missing");
    if ((long)N * H * W * C == 0) return IMX_OK;
    const RowMap m = make_rowmap(C);
    ModArgs a{reinterpret_cast<const uint4*>(x), reinterpret_cast<const uint4*>(dy), mean, rstd, w, b, gamma, beta,
              N, H, W, C, stats_per_instance, gb_mode, 0, act_slope(act), gb_mode == 0 ? gb_sums : nullptr};
    dim3 grid(pick_blocks((long)H * W, m.rows, N), N);
    auto kern = (w || b) ? modulate_kernel<1, true> : modulate_kernel<1, false>;
    kern<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        a, nullptr, dgamma, dbeta, red, 0.0f, 1, m.cg, m.tpr, m.rows);
    return imx::check_launch("spade_modulate_bwd_reduce");
}

extern "C" int imx_spade_modulate_bwd_dx(const imx_bf16* x, const imx_bf16* dy,
                                         const float* mean, const float* rstd,
                                         const float* w, const float* b,
                                         const void* gamma, const void* beta,
                                         const float* red, float count, imx_bf16* dx,
                                         int N, int H, int W, int C,
                                         int stats_per_instance, int gb_mode, int act,
                                         int use_batch_stats, imx_stream_t stream) {
    int rc = check_mod_args("spade_modulate_bwd_dx", x, mean, rstd, gamma, beta, N, H, W, C, gb_mode, act, 0);
    if (rc) return rc;
    IMX_REQUIRE(dy && dx, "spade_modulate_bwd_dx: null pointer");
    IMX_REQUIRE(!use_batch_stats || (red && count > 0.0f), "spade_modulate_bwd_dx: reductions/count missing");
    if ((long)N * H * W * C == 0) return IMX_OK;
    const RowMap m = make_rowmap(C);
    ModArgs a{reinterpret_cast<const uint4*>(x), reinterpret_cast<const uint4*>(dy), mean, rstd, w, b, gamma, beta,
              N, H, W, C, stats_per_instance, gb_mode, 0, act_slope(act), nullptr};
    dim3 grid(pick_blocks((long)H * W, m.rows, N, 6), N);
    auto kern = (w || b) ? modulate_kernel<2, true> : modulate_kernel<2, false>;
    kern<<<grid, kThreads, 0, (cudaStream_t)stream>>>(
        a, reinterpret_cast<uint4*>(dx), nullptr, nullptr, const_cast<float*>(red),
        use_batch_stats ? 1.0f / count : 0.0f, use_batch_stats, m.cg, m.tpr, m.rows);
    return imx::check_launch("spade_modulate_bwd_dx");
}
