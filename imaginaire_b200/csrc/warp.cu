// Flow-warp kernels.
//
//  * imx_resample2d_{fwd,bwd}_f32 : the arithmetic of the reference's
//    resample2d CUDA extension
//    (imaginaire/third_party/resample2d/src/resample2d_kernel.cu:15-75 forward,
//    :78-129 d/d image, :131-203 d/d flow), kernel_size == 1 only (every call
//    site uses 1; larger values index out of bounds in the reference).
//  * imx_gridsample_warp_{fwd,bwd}_f32 : the arithmetic of
//    imaginaire/model_utils/fs_vid2vid.py:14-39 resample() =
//    F.grid_sample(bilinear, border, align_corners=True) on linspace(-1,1)+flow,
//    with the grid never materialised.
//
// HBM-bound: one thread per output pixel; the flow-derived tap indices and
// weights stay in registers for the whole channel loop (the flow is read once
// per pixel, not once per channel as in the reference); consecutive lanes own
// consecutive x so flow loads / output stores are 128-byte coalesced, and the
// 2-D block (32 x 8 pixels) keeps the taps shared by neighbouring rows in L1.
#include <cstdlib>
#include <type_traits>
#include "imx_common.cuh"
#include "warp_window.cuh"
#include "../../include/imx.h"

namespace {

constexpr int TX = 32, TY = 8;

// Channels handled per block in z: split the channel loop when the image alone
// cannot fill the 148 SMs a few times over.
inline int pick_channels_per_block(int B, int C, int H, int W) {
    long blocks = (long)B * ((H + TY - 1) / TY) * ((W + TX - 1) / TX);
    int cpb = C;
    while (cpb > 4 && blocks * ((C + cpb - 1) / cpb) < 4L * imx::kNumSMs) cpb = (cpb + 1) / 2;
    return cpb;
}

// ---------------------------------------------------------------- resample2d

struct Taps {
    int xL, xR, yT, yB;
    float alpha, beta;
};

__device__ __forceinline__ Taps r2d_taps(float dx, float dy, int x, int y, int cw, int ch) {
    // resample2d_kernel.cu:44-55 (forward), :171-176 (d/d flow)
    Taps t;
    float xf = static_cast<float>(x) + dx;
    float yf = static_cast<float>(y) + dy;
    float fx = floorf(xf), fy = floorf(yf);
    t.alpha = xf - fx;
    t.beta = yf - fy;
    t.xL = max(min(int(fx), cw - 1), 0);
    t.xR = max(min(int(fx + 1.0f), cw - 1), 0);
    t.yT = max(min(int(fy), ch - 1), 0);
    t.yB = max(min(int(fy + 1.0f), ch - 1), 0);
    return t;
}

// Channel groups: the 4 * U gathers of U channels are issued before any of them is used.  A thread
// that walks the channels one at a time has four 4-byte gathers in flight, three of which usually
// hit L1 / L2, so only a quarter of its outstanding requests move new DRAM bytes and the kernel
// sits at ~0.4 of the HBM peak whatever the flow looks like (measured with a constant flow).
// The minimum-CTA launch bound matters as in spade_norm.cu: without it ptxas recycles the
// destination registers of the loads and serialises the channels again.
constexpr int kMinCtas = 4;
constexpr long kGroupMinElems = 8L << 20;   // output elements from which the grouped kernels pay off
constexpr int kMinCtasBwd = 3;   // 20 loads in flight per group plus both sets of taps: 64 registers spill

template <int U>
__device__ __forceinline__ void r2d_fwd_group(const float* __restrict__ ip, float* __restrict__ op, size_t hwi, size_t hw,
                                              int oTL, int oTR, int oBL, int oBR, double w00, double w01, double w10,
                                              float w11) {
    float tl[U], tr[U], bl[U], br[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const float* q = ip + (size_t)u * hwi;
        tl[u] = __ldg(q + oTL);
        tr[u] = __ldg(q + oTR);
        bl[u] = __ldg(q + oBL);
        br[u] = __ldg(q + oBR);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        float v = 0.0f;
        v += static_cast<float>(w00 * (double)tl[u]);
        v += static_cast<float>(w01 * (double)tr[u]);
        v += static_cast<float>(w10 * (double)bl[u]);
        v += __fmul_rn(w11, br[u]);
        __stcs(op + (size_t)u * hw, v);
    }
}

// kGroup false (small launches, a few waves of CTAs at most): one channel at a time and no
// minimum-CTA hint, which keeps more CTAs resident; there the latency of a thread's dependent
// chain (flow -> taps -> gathers -> store) is hidden by other warps rather than by its own loads.
template <bool kGroup>
__global__ void __launch_bounds__(TX* TY, kGroup ? kMinCtas : 0)
resample2d_fwd_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                      float* __restrict__ out, int C, int H, int W, int Hi, int Wi,
                      int cpb, int nchunk) {
    const int x = blockIdx.x * TX + threadIdx.x;
    const int y = blockIdx.y * TY + threadIdx.y;
    const int b = blockIdx.z / nchunk;
    const int c0 = (blockIdx.z % nchunk) * cpb;
    if (x >= W || y >= H) return;
    const size_t hw = (size_t)H * W, hwi = (size_t)Hi * Wi;
    const float* fl = flow + (size_t)b * 2 * hw + (size_t)y * W + x;
    const float dx = __ldg(fl), dy = __ldg(fl + hw);
    const Taps t = r2d_taps(dx, dy, x, y, W, H);  // clamps against the OUTPUT dims (:31-34)
    // The reference forms the three weights that contain the literal `1.` in double
    // ((1. - alpha) promotes) and rounds each product to float before the float
    // accumulation; `alpha * beta * I` has no double operand and stays in float (:59-62).
    const double a = t.alpha, be = t.beta;
    const double w00 = (1. - a) * (1. - be), w01 = a * (1. - be);
    const double w10 = (1. - a) * be;
    const float w11 = __fmul_rn(t.alpha, t.beta);
    const int oTL = t.yT * Wi + t.xL, oTR = t.yT * Wi + t.xR;
    const int oBL = t.yB * Wi + t.xL, oBR = t.yB * Wi + t.xR;
    const int c1 = min(C, c0 + cpb);
    const float* ip = img + ((size_t)b * C + c0) * hwi;
    float* op = out + ((size_t)b * C + c0) * hw + (size_t)y * W + x;
    int c = c0;
    if (kGroup) {
        for (; c + 4 <= c1; c += 4, ip += 4 * hwi, op += 4 * hw)
            r2d_fwd_group<4>(ip, op, hwi, hw, oTL, oTR, oBL, oBR, w00, w01, w10, w11);
        if (c1 - c == 3) r2d_fwd_group<3>(ip, op, hwi, hw, oTL, oTR, oBL, oBR, w00, w01, w10, w11);
        else if (c1 - c == 2) r2d_fwd_group<2>(ip, op, hwi, hw, oTL, oTR, oBL, oBR, w00, w01, w10, w11);
        else if (c1 - c == 1) r2d_fwd_group<1>(ip, op, hwi, hw, oTL, oTR, oBL, oBR, w00, w01, w10, w11);
    } else {
#pragma unroll 4
        for (; c < c1; ++c, ip += hwi, op += hw) {
            float v = 0.0f;
            v += static_cast<float>(w00 * (double)__ldg(ip + oTL));
            v += static_cast<float>(w01 * (double)__ldg(ip + oTR));
            v += static_cast<float>(w10 * (double)__ldg(ip + oBL));
            v += __fmul_rn(w11, __ldg(ip + oBR));
            __stcs(op, v);
        }
    }
}

// One pass over the channels produces both gradients: the 4-tap scatter into
// grad_img (atomics, like the reference) and the per-pixel flow gradient, which
// the reference computes in a second kernel that re-reads everything.
__global__ void __launch_bounds__(TX* TY, kMinCtasBwd)
resample2d_bwd_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                      const float* __restrict__ gout, float* __restrict__ gimg,
                      float* __restrict__ gflow, int C, int H, int W, int Hi, int Wi,
                      int trunc_alpha) {
    const int x = blockIdx.x * TX + threadIdx.x;
    const int y = blockIdx.y * TY + threadIdx.y;
    const int b = blockIdx.z;
    if (x >= W || y >= H) return;
    const size_t hw = (size_t)H * W, hwi = (size_t)Hi * Wi;
    const float* fl = flow + (size_t)b * 2 * hw + (size_t)y * W + x;
    const float dx = __ldg(fl), dy = __ldg(fl + hw);
    const float xf = static_cast<float>(x) + dx;
    const float yf = static_cast<float>(y) + dy;
    // d/d image (:106-118): weights from xf - int(xf) in the reference, taps clamped
    // against the INPUT image dims.
    const float sa = trunc_alpha ? xf - static_cast<float>(int(xf)) : xf - floorf(xf);
    const float sb = trunc_alpha ? yf - static_cast<float>(int(yf)) : yf - floorf(yf);
    const int sxL = max(min(int(floorf(xf)), Wi - 1), 0);
    const int sxR = max(min(int(floorf(xf) + 1.0f), Wi - 1), 0);
    const int syT = max(min(int(floorf(yf)), Hi - 1), 0);
    const int syB = max(min(int(floorf(yf) + 1.0f), Hi - 1), 0);
    const float s00 = (1 - sa) * (1 - sb), s01 = sa * (1 - sb), s10 = (1 - sa) * sb, s11 = sa * sb;
    // d/d flow (:171-199): taps clamped against the flow dims, gamma from floor.
    const Taps t = r2d_taps(dx, dy, x, y, W, H);
    const float gx = 1.0f - t.beta;   // channel 0 (dx): gamma = 1 - (yf - floor(yf))
    const float gy = 1.0f - t.alpha;  // channel 1 (dy): gamma = 1 - (xf - floor(xf))
    const int oTL = t.yT * Wi + t.xL, oTR = t.yT * Wi + t.xR;
    const int oBL = t.yB * Wi + t.xL, oBR = t.yB * Wi + t.xR;
    float accx = 0.0f, accy = 0.0f;
    const float* ip = img + (size_t)b * C * hwi;
    const float* gp = gout + (size_t)b * C * hw + (size_t)y * W + x;
    float* gi = gimg ? gimg + (size_t)b * C * hwi : nullptr;
    // Groups of up to four channels: all loads of the group first (see r2d_fwd_group), then the
    // per-channel work in channel order, so the flow gradient keeps the reference's summation order.
    auto group = [&](auto tag) {
        constexpr int U = decltype(tag)::value;
        float g[U], iTL[U], iTR[U], iBL[U], iBR[U];
#pragma unroll
        for (int u = 0; u < U; ++u) g[u] = __ldcs(gp + (size_t)u * hw);
        if (gflow) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const float* q = ip + (size_t)u * hwi;
                iTL[u] = __ldg(q + oTL); iTR[u] = __ldg(q + oTR);
                iBL[u] = __ldg(q + oBL); iBR[u] = __ldg(q + oBR);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (gi) {
                float* d = gi + (size_t)u * hwi;
                atomicAdd(d + syT * Wi + sxL, s00 * g[u]);
                atomicAdd(d + syT * Wi + sxR, s01 * g[u]);
                atomicAdd(d + syB * Wi + sxL, s10 * g[u]);
                atomicAdd(d + syB * Wi + sxR, s11 * g[u]);
            }
            if (gflow) {
                // Same operation order as the reference, no FMA contraction, so the flow
                // gradient is bit-exact against the CPU restatement.
                const float a0 = __fmul_rn(gx, g[u]), a1 = __fmul_rn(1.0f - gx, g[u]);
                accx = __fadd_rn(accx, __fmul_rn(a0, iTR[u]));
                accx = __fsub_rn(accx, __fmul_rn(a0, iTL[u]));
                accx = __fadd_rn(accx, __fmul_rn(a1, iBR[u]));
                accx = __fsub_rn(accx, __fmul_rn(a1, iBL[u]));
                const float b0 = __fmul_rn(gy, g[u]), b1 = __fmul_rn(1.0f - gy, g[u]);
                accy = __fadd_rn(accy, __fmul_rn(b0, iBL[u]));
                accy = __fsub_rn(accy, __fmul_rn(b0, iTL[u]));
                accy = __fadd_rn(accy, __fmul_rn(b1, iBR[u]));
                accy = __fsub_rn(accy, __fmul_rn(b1, iTR[u]));
            }
        }
        ip += (size_t)U * hwi;
        gp += (size_t)U * hw;
        if (gi) gi += (size_t)U * hwi;
    };
    int c = 0;
    for (; c + 4 <= C; c += 4) group(std::integral_constant<int, 4>{});
    if (C - c == 3) group(std::integral_constant<int, 3>{});
    else if (C - c == 2) group(std::integral_constant<int, 2>{});
    else if (C - c == 1) group(std::integral_constant<int, 1>{});
    if (gflow) {
        float* gf = gflow + (size_t)b * 2 * hw + (size_t)y * W + x;
        __stcs(gf, accx);
        __stcs(gf + hw, accy);
    }
}

// --------------------------------------------------- grid_sample-exact warp

struct GsTaps {
    int o_nw, o_ne, o_sw, o_se;   // element offsets inside one H x W plane
    bool e_ok, s_ok;
    float nw, ne, sw, se;         // bilinear weights
    float wx, wy;                 // fractional parts
    bool gx_live, gy_live;        // clip gradient masks
};

__device__ __forceinline__ float linspace_m1_p1(int i, int n) {
    // torch.linspace(-1, 1, n)[i] (fp32): symmetric two-sided evaluation with a fused
    // multiply-add, as ATen computes it.
    const float step = 2.0f / static_cast<float>(n - 1);
    return (i < n / 2) ? fmaf(step, static_cast<float>(i), -1.0f)
                       : fmaf(-step, static_cast<float>(n - 1 - i), 1.0f);
}

__device__ __forceinline__ GsTaps gs_taps(float fx, float fy, int x, int y, int H, int W) {
    GsTaps t;
    // fs_vid2vid.py:27-30: grid + flow / ((size - 1) / 2)
    const float cx = linspace_m1_p1(x, W) + __fdiv_rn(fx, (static_cast<float>(W) - 1.0f) / 2.0f);
    const float cy = linspace_m1_p1(y, H) + __fdiv_rn(fy, (static_cast<float>(H) - 1.0f) / 2.0f);
    // grid_sample, align_corners=True: ix = (c + 1) * ((size - 1) / 2), border clip
    float ix = __fmul_rn(cx + 1.0f, static_cast<float>(W - 1) / 2.0f);
    float iy = __fmul_rn(cy + 1.0f, static_cast<float>(H - 1) / 2.0f);
    t.gx_live = (ix > 0.0f) && (ix < static_cast<float>(W - 1));
    t.gy_live = (iy > 0.0f) && (iy < static_cast<float>(H - 1));
    ix = fminf(static_cast<float>(W - 1), fmaxf(ix, 0.0f));
    iy = fminf(static_cast<float>(H - 1), fmaxf(iy, 0.0f));
    const float xw = floorf(ix), yn = floorf(iy);
    const float w = ix - xw, e = 1.0f - w, n = iy - yn, s = 1.0f - n;
    t.wx = w; t.wy = n;
    t.nw = __fmul_rn(s, e); t.ne = __fmul_rn(s, w);
    t.sw = __fmul_rn(n, e); t.se = __fmul_rn(n, w);
    const int ixw = int(xw), iyn = int(yn);
    t.e_ok = (ixw + 1) < W;
    t.s_ok = (iyn + 1) < H;
    const int ixe = t.e_ok ? ixw + 1 : ixw, iys = t.s_ok ? iyn + 1 : iyn;
    t.o_nw = iyn * W + ixw; t.o_ne = iyn * W + ixe;
    t.o_sw = iys * W + ixw; t.o_se = iys * W + ixe;
    return t;
}

template <int U>
__device__ __forceinline__ void gs_fwd_group(const float* __restrict__ ip, float* __restrict__ op, size_t hw, const GsTaps& t) {
    // The tap offsets are always in range (gs_taps folds a missing east / south neighbour onto the
    // pixel itself), so every load is unconditional and the validity test is a select afterwards.
    float vnw[U], vne[U], vsw[U], vse[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const float* q = ip + (size_t)u * hw;
        vnw[u] = __ldg(q + t.o_nw);
        vne[u] = __ldg(q + t.o_ne);
        vsw[u] = __ldg(q + t.o_sw);
        vse[u] = __ldg(q + t.o_se);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        float o = __fmul_rn(vnw[u], t.nw);
        o = fmaf(t.e_ok ? vne[u] : 0.0f, t.ne, o);
        o = fmaf(t.s_ok ? vsw[u] : 0.0f, t.sw, o);
        o = fmaf((t.e_ok && t.s_ok) ? vse[u] : 0.0f, t.se, o);
        __stcs(op + (size_t)u * hw, o);
    }
}

// kGroup false: one channel at a time and no minimum-CTA hint (32 registers, 8 CTAs per SM).  With only 3-4 channels a
// thread lives too briefly for its own loads to overlap much, and occupancy wins (measured at
// B=8, C=3: 58 us one-by-one, 66 us grouped; at C=64 grouped is 0.47 vs 0.37 of the HBM peak).
template <bool kGroup>
__global__ void __launch_bounds__(TX* TY, kGroup ? kMinCtas : 0)
gridsample_warp_fwd_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                           float* __restrict__ out, int C, int H, int W, int cpb, int nchunk) {
    const int x = blockIdx.x * TX + threadIdx.x;
    const int y = blockIdx.y * TY + threadIdx.y;
    const int b = blockIdx.z / nchunk;
    const int c0 = (blockIdx.z % nchunk) * cpb;
    if (x >= W || y >= H) return;
    const size_t hw = (size_t)H * W;
    const float* fl = flow + (size_t)b * 2 * hw + (size_t)y * W + x;
    const GsTaps t = gs_taps(__ldg(fl), __ldg(fl + hw), x, y, H, W);
    const int c1 = min(C, c0 + cpb);
    const float* ip = img + ((size_t)b * C + c0) * hw;
    float* op = out + ((size_t)b * C + c0) * hw + (size_t)y * W + x;
    int c = c0;
    if (kGroup) {
        for (; c + 4 <= c1; c += 4, ip += 4 * hw, op += 4 * hw) gs_fwd_group<4>(ip, op, hw, t);
        if (c1 - c == 3) gs_fwd_group<3>(ip, op, hw, t);
        else if (c1 - c == 2) gs_fwd_group<2>(ip, op, hw, t);
        else if (c1 - c == 1) gs_fwd_group<1>(ip, op, hw, t);
    } else {
#pragma unroll 4
        for (; c < c1; ++c, ip += hw, op += hw) {
            const float vnw = __ldg(ip + t.o_nw);
            const float vne = t.e_ok ? __ldg(ip + t.o_ne) : 0.0f;
            const float vsw = t.s_ok ? __ldg(ip + t.o_sw) : 0.0f;
            const float vse = (t.e_ok && t.s_ok) ? __ldg(ip + t.o_se) : 0.0f;
            float o = __fmul_rn(vnw, t.nw);
            o = fmaf(vne, t.ne, o);
            o = fmaf(vsw, t.sw, o);
            o = fmaf(vse, t.se, o);
            __stcs(op, o);
        }
    }
}

template <bool kGroup>
__global__ void __launch_bounds__(TX* TY, kGroup ? kMinCtasBwd : 0)
gridsample_warp_bwd_kernel(const float* __restrict__ img, const float* __restrict__ flow,
                           const float* __restrict__ gout, float* __restrict__ gimg,
                           float* __restrict__ gflow, int C, int H, int W) {
    const int x = blockIdx.x * TX + threadIdx.x;
    const int y = blockIdx.y * TY + threadIdx.y;
    const int b = blockIdx.z;
    if (x >= W || y >= H) return;
    const size_t hw = (size_t)H * W;
    const float* fl = flow + (size_t)b * 2 * hw + (size_t)y * W + x;
    const GsTaps t = gs_taps(__ldg(fl), __ldg(fl + hw), x, y, H, W);
    const float e = 1.0f - t.wx, s = 1.0f - t.wy;
    float gix = 0.0f, giy = 0.0f;
    const float* ip = img + (size_t)b * C * hw;
    const float* gp = gout + (size_t)b * C * hw + (size_t)y * W + x;
    float* gi = gimg ? gimg + (size_t)b * C * hw : nullptr;
    const bool se_ok = t.e_ok && t.s_ok;
    auto group = [&](auto tag) {   // all loads of up to four channels first, see r2d_fwd_group
        constexpr int U = decltype(tag)::value;
        float g[U], vnw[U], vne[U], vsw[U], vse[U];
#pragma unroll
        for (int u = 0; u < U; ++u) g[u] = __ldcs(gp + (size_t)u * hw);
        if (gflow) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const float* q = ip + (size_t)u * hw;
                vnw[u] = __ldg(q + t.o_nw); vne[u] = __ldg(q + t.o_ne);
                vsw[u] = __ldg(q + t.o_sw); vse[u] = __ldg(q + t.o_se);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (gi) {
                float* d = gi + (size_t)u * hw;
                atomicAdd(d + t.o_nw, t.nw * g[u]);
                if (t.e_ok) atomicAdd(d + t.o_ne, t.ne * g[u]);
                if (t.s_ok) atomicAdd(d + t.o_sw, t.sw * g[u]);
                if (se_ok) atomicAdd(d + t.o_se, t.se * g[u]);
            }
            if (gflow) {
                const float ne = t.e_ok ? vne[u] : 0.0f, sw = t.s_ok ? vsw[u] : 0.0f, se = se_ok ? vse[u] : 0.0f;
                gix += g[u] * ((ne - vnw[u]) * s + (se - sw) * t.wy);
                giy += g[u] * ((sw - vnw[u]) * e + (se - ne) * t.wx);
            }
        }
        ip += (size_t)U * hw;
        gp += (size_t)U * hw;
        if (gi) gi += (size_t)U * hw;
    };
    int c = 0;
    if (kGroup) {
        for (; c + 4 <= C; c += 4) group(std::integral_constant<int, 4>{});
        if (C - c == 3) group(std::integral_constant<int, 3>{});
        else if (C - c == 2) group(std::integral_constant<int, 2>{});
        else if (C - c == 1) group(std::integral_constant<int, 1>{});
    } else {
        for (; c < C; ++c) group(std::integral_constant<int, 1>{});
    }
    if (gflow) {
        // d(ix)/d(flow_x) = ((W-1)/2) / ((W-1)/2) = 1 inside the border, 0 where clipped.
        float* gf = gflow + (size_t)b * 2 * hw + (size_t)y * W + x;
        __stcs(gf, t.gx_live ? gix : 0.0f);
        __stcs(gf + hw, t.gy_live ? giy : 0.0f);
    }
}

}  // namespace

extern "C" int imx_resample2d_fwd_f32(const float* img, const float* flow, float* out,
                                      int B, int C, int H, int W, int Hi, int Wi,
                                      int kernel_size, imx_stream_t stream) {
    if (B == 0 || C == 0 || H == 0 || W == 0) return IMX_OK;
    IMX_REQUIRE(img && flow && out, "resample2d_fwd: null pointer");
    IMX_REQUIRE(kernel_size == 1, "resample2d_fwd: kernel_size must be 1 (got %d)", kernel_size);
    IMX_REQUIRE(B >= 0 && C >= 0 && H >= 0 && W >= 0, "resample2d_fwd: negative size");
    IMX_REQUIRE(Hi >= H && Wi >= W, "resample2d_fwd: image (%dx%d) smaller than flow (%dx%d)", Hi, Wi, H, W);
    if ((long)B * C * H * W == 0) return IMX_OK;
    {
        // the shared-memory window does not pay for this arithmetic (three double-precision products per element,
        // see warp_window.cu); kept behind a switch for measurements
        const char* on = getenv("IMX_WARPWIN_R2D");
        if (on && on[0] == '1' && imx::warpwin::applicable(img, B, C, H, W, Hi, Wi))
            return imx::warpwin::resample2d_fwd(img, flow, out, B, C, H, W, Hi, Wi, (cudaStream_t)stream);
    }
    const int cpb = pick_channels_per_block(B, C, H, W);
    const int nchunk = (C + cpb - 1) / cpb;
    IMX_REQUIRE((long)B * nchunk <= 65535, "resample2d_fwd: batch*channel-chunks too large");
    dim3 grid((W + TX - 1) / TX, (H + TY - 1) / TY, B * nchunk), block(TX, TY);
    if ((long)B * C * H * W >= kGroupMinElems)
        resample2d_fwd_kernel<true><<<grid, block, 0, (cudaStream_t)stream>>>(img, flow, out, C, H, W, Hi, Wi, cpb, nchunk);
    else
        resample2d_fwd_kernel<false><<<grid, block, 0, (cudaStream_t)stream>>>(img, flow, out, C, H, W, Hi, Wi, cpb, nchunk);
    return imx::check_launch("resample2d_fwd");
}

extern "C" int imx_resample2d_bwd_f32(const float* img, const float* flow, const float* grad_out,
                                      float* grad_img, float* grad_flow,
                                      int B, int C, int H, int W, int Hi, int Wi,
                                      int kernel_size, int trunc_alpha, imx_stream_t stream) {
    if (B == 0 || H == 0 || W == 0) return IMX_OK;
    IMX_REQUIRE(img && flow && grad_out, "resample2d_bwd: null pointer");
    IMX_REQUIRE(kernel_size == 1, "resample2d_bwd: kernel_size must be 1 (got %d)", kernel_size);
    IMX_REQUIRE(B >= 0 && C >= 0 && H >= 0 && W >= 0, "resample2d_bwd: negative size");
    IMX_REQUIRE(Hi >= H && Wi >= W, "resample2d_bwd: image smaller than flow");
    IMX_REQUIRE(B <= 65535, "resample2d_bwd: batch too large");
    if ((long)B * H * W == 0) return IMX_OK;
    dim3 grid((W + TX - 1) / TX, (H + TY - 1) / TY, B), block(TX, TY);
    resample2d_bwd_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(img, flow, grad_out, grad_img, grad_flow,
                                                                    C, H, W, Hi, Wi, trunc_alpha);
    return imx::check_launch("resample2d_bwd");
}

extern "C" int imx_gridsample_warp_fwd_f32(const float* img, const float* flow, float* out,
                                           int B, int C, int H, int W, imx_stream_t stream) {
    if (B == 0 || C == 0 || H == 0 || W == 0) return IMX_OK;
    IMX_REQUIRE(img && flow && out, "gridsample_warp_fwd: null pointer");
    IMX_REQUIRE(B >= 0 && C >= 0 && H >= 0 && W >= 0, "gridsample_warp_fwd: negative size");
    if ((long)B * C * H * W == 0) return IMX_OK;
    IMX_REQUIRE(H >= 2 && W >= 2, "gridsample_warp_fwd: H and W must be >= 2");
    if (imx::warpwin::applicable(img, B, C, H, W, H, W))
        return imx::warpwin::gridsample_fwd(img, flow, out, B, C, H, W, (cudaStream_t)stream);
    const int cpb = pick_channels_per_block(B, C, H, W);
    const int nchunk = (C + cpb - 1) / cpb;
    IMX_REQUIRE((long)B * nchunk <= 65535, "gridsample_warp_fwd: batch*channel-chunks too large");
    dim3 grid((W + TX - 1) / TX, (H + TY - 1) / TY, B * nchunk), block(TX, TY);
    if (cpb >= 8) gridsample_warp_fwd_kernel<true><<<grid, block, 0, (cudaStream_t)stream>>>(img, flow, out, C, H, W, cpb, nchunk);
    else gridsample_warp_fwd_kernel<false><<<grid, block, 0, (cudaStream_t)stream>>>(img, flow, out, C, H, W, cpb, nchunk);
    return imx::check_launch("gridsample_warp_fwd");
}

extern "C" int imx_gridsample_warp_bwd_f32(const float* img, const float* flow, const float* grad_out,
                                           float* grad_img, float* grad_flow,
                                           int B, int C, int H, int W, imx_stream_t stream) {
    if (B == 0 || H == 0 || W == 0) return IMX_OK;
    IMX_REQUIRE(img && flow && grad_out, "gridsample_warp_bwd: null pointer");
    IMX_REQUIRE(B >= 0 && C >= 0 && H >= 0 && W >= 0, "gridsample_warp_bwd: negative size");
    IMX_REQUIRE(B <= 65535, "gridsample_warp_bwd: batch too large");
    if ((long)B * H * W == 0) return IMX_OK;
    IMX_REQUIRE(H >= 2 && W >= 2, "gridsample_warp_bwd: H and W must be >= 2");
    dim3 grid((W + TX - 1) / TX, (H + TY - 1) / TY, B), block(TX, TY);
    if (C >= 8) gridsample_warp_bwd_kernel<true><<<grid, block, 0, (cudaStream_t)stream>>>(img, flow, grad_out, grad_img, grad_flow, C, H, W);
    else gridsample_warp_bwd_kernel<false><<<grid, block, 0, (cudaStream_t)stream>>>(img, flow, grad_out, grad_img, grad_flow, C, H, W);
    return imx::check_launch("gridsample_warp_bwd");
}
