// Convolutions with 8 or fewer channels on one side (the 5x5
// to-RGB layer 256 -> 3, its gradients, and the first layers of the image networks).
//
// ref: torch.nn.Conv2d inside Conv2dBlock, imaginaire/layers/conv.py:93-104;
//      `conv_img256` / `conv_img512`, imaginaire/generators/spade.py:318-323
//
// The implicit-GEMM kernel needs 64 input channels per K step and at least 32 output channels
// per tile; a layer with 3 (padded 8) channels on one side wastes 8-10x of both the tensor
// pipe and, worse, of the operand traffic (the 256-channel activation is re-read once per
// filter tap).  Both cases become a 1x1 GEMM on the tensor cores plus one memory-bound pass:
//
//   few INPUT channels  (Cin_p = 8):  col[p][(tap, ci)] = x[p*stride + tap - pad][ci]  (this
//       file, imx_im2col_c8_nhwc_bf16), then y = col . W^T as a 1x1 convolution with
//       K = taps * 8.
//   few OUTPUT channels (Cout <= 8):  P[p][(tap, co)] = x[p] . W[co, :, tap]  as a 1x1
//       convolution with N = taps * Cout columns (fp32 output), then
//       y[q][co] = bias[co] + sum_tap P[q + tap - pad][(tap, co)]  (imx_shift_add_taps_f32).
//       The activation is read once instead of once per tap.
#include "imx_common.cuh"
#include "imx_tma_host.cuh"
#include "../../include/imx.h"

namespace {

// one thread per (output pixel, tap): a 16-byte read of the 8-channel pixel, a 16-byte write
__global__ void im2col_c8_kernel(const uint4* __restrict__ x, uint4* __restrict__ col, int N, int H, int W,
                                 int Ho, int Wo, int KH, int KW, int stride, int pad) {
    const int taps = KH * KW;
    const long total = (long)N * Ho * Wo * taps;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int tap = (int)(i % taps);
        long p = i / taps;
        const int ox = (int)(p % Wo); p /= Wo;
        const int oy = (int)(p % Ho);
        const int n = (int)(p / Ho);
        const int kh = tap / KW, kw = tap - kh * KW;
        const int iy = oy * stride + kh - pad, ix = ox * stride + kw - pad;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (iy >= 0 && iy < H && ix >= 0 && ix < W) v = __ldg(x + ((size_t)n * H + iy) * W + ix);
        imx::stg_stream(col + i, v);
    }
}

constexpr int TX = 32, TY = 16;          // output tile of one block
constexpr int kShiftThreads = 256;

// Block = one TX x TY output tile of one image.  Its warps read the P rows of the tile plus halo
// (whole rows: coalesced) and scatter every (tap, co) entry into the shared-memory accumulator
// of the output pixel it belongs to; then the tile is written out as 8-channel bf16 pixels.
__global__ void __launch_bounds__(kShiftThreads)
shift_add_kernel(const float* __restrict__ P, const float* __restrict__ bias, uint4* __restrict__ y,
                 int H, int W, int ncols, int KH, int KW, int pad, int Cout, int tiles_x) {
    __shared__ float acc[TY * TX * 8];
    const int n = blockIdx.y;
    const int ty0 = (blockIdx.x / tiles_x) * TY, tx0 = (blockIdx.x % tiles_x) * TX;
    for (int i = threadIdx.x; i < TY * TX * 8; i += kShiftThreads) acc[i] = 0.0f;
    __syncthreads();
    // input pixels that can contribute: q + tap - pad for q in the tile
    const int y_lo = ty0 - pad, x_lo = tx0 - pad;
    const int ry = TY + KH - 1, rx = TX + KW - 1;
    const int used = KH * KW * Cout;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = warp; r < ry * rx; r += kShiftThreads / 32) {
        const int py = y_lo + r / rx, px = x_lo + r % rx;
        if (py < 0 || py >= H || px < 0 || px >= W) continue;      // warp-uniform
        const float* row = P + (((size_t)n * H + py) * W + px) * ncols;
        for (int c = lane; c < used; c += 32) {
            const int tap = c / Cout, co = c - tap * Cout;
            const int kh = tap / KW, kw = tap - kh * KW;
            const int qy = py - kh + pad - ty0, qx = px - kw + pad - tx0;   // output pixel inside the tile
            if (qy >= 0 && qy < TY && qx >= 0 && qx < TX) atomicAdd(&acc[(qy * TX + qx) * 8 + co], __ldg(row + c));
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < TY * TX; i += kShiftThreads) {
        const int qy = ty0 + i / TX, qx = tx0 + i % TX;
        if (qy >= H || qx >= W) continue;
        float f[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = k < Cout ? acc[i * 8 + k] + (bias ? __ldg(bias + k) : 0.0f) : 0.0f;
        y[((size_t)n * H + qy) * W + qx] = make_uint4(imx::pack_bf16(f[0], f[1]), imx::pack_bf16(f[2], f[3]),
                                                      imx::pack_bf16(f[4], f[5]), imx::pack_bf16(f[6], f[7]));
    }
}

}  // namespace

extern "C" int imx_im2col_c8_nhwc_bf16(const imx_bf16* x, imx_bf16* col, int N, int H, int W,
                                       int KH, int KW, int stride, int pad, imx_stream_t stream) {
    IMX_REQUIRE(x && col, "im2col_c8: null pointer");
    IMX_REQUIRE(N >= 0 && H > 0 && W > 0, "im2col_c8: bad size");
    IMX_REQUIRE(KH > 0 && KW > 0 && KH <= 16 && KW <= 16 && stride >= 1 && pad >= 0, "im2col_c8: bad filter geometry");
    IMX_REQUIRE(imx::aligned16(x) && imx::aligned16(col), "im2col_c8: pointers must be 16-byte aligned");
    const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
    IMX_REQUIRE(Ho > 0 && Wo > 0, "im2col_c8: empty output");
    const long total = (long)N * Ho * Wo * KH * KW;
    if (total == 0) return IMX_OK;
    long blocks = (total + 255) / 256;
    if (blocks > 32L * imx::kNumSMs) blocks = 32L * imx::kNumSMs;
    im2col_c8_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(col), N, H, W, Ho, Wo, KH, KW, stride, pad);
    return imx::check_launch("im2col_c8");
}

extern "C" int imx_shift_add_taps_f32(const float* P, const float* bias, imx_bf16* y, int N, int H, int W,
                                      int ncols, int KH, int KW, int pad, int Cout, imx_stream_t stream) {
    IMX_REQUIRE(P && y, "shift_add_taps: null pointer");
    IMX_REQUIRE(N >= 0 && N <= 65535 && H > 0 && W > 0, "shift_add_taps: bad size");
    IMX_REQUIRE(KH > 0 && KW > 0 && KH <= 16 && KW <= 16 && pad >= 0, "shift_add_taps: bad filter geometry");
    IMX_REQUIRE(2 * pad == KH - 1 && 2 * pad == KW - 1, "shift_add_taps: needs a 'same' convolution (pad = (K-1)/2)");
    IMX_REQUIRE(Cout >= 1 && Cout <= 8, "shift_add_taps: Cout (%d) must be in [1, 8]", Cout);
    IMX_REQUIRE(ncols >= KH * KW * Cout, "shift_add_taps: P has %d columns, needs %d", ncols, KH * KW * Cout);
    IMX_REQUIRE(imx::aligned16(y), "shift_add_taps: y must be 16-byte aligned");
    if (N == 0) return IMX_OK;
    const int tiles_x = (W + TX - 1) / TX, tiles_y = (H + TY - 1) / TY;
    dim3 grid(tiles_x * tiles_y, N);
    shift_add_kernel<<<grid, kShiftThreads, 0, (cudaStream_t)stream>>>(P, bias, reinterpret_cast<uint4*>(y), H, W, ncols,
                                                                       KH, KW, pad, Cout, tiles_x);
    return imx::check_launch("shift_add_taps");
}
