// The two remaining native operators of the reference, used by
// FlowNet2 when the ground-truth flow of the vid2vid trainers is computed on the fly.
//
//  * imx_channelnorm_{fwd,bwd}_f32 : L2 norm over channels and its gradient
//    ref: imaginaire/third_party/channelnorm/src/channelnorm_kernel.cu:18-61, :64-96
//  * imx_correlation_{fwd,bwd}_f32 : FlowNetC cost volume
//    ref: imaginaire/third_party/correlation/src/correlation_cuda_kernel.cu:46-69 (padded
//    channels-last copies), :72-149 (forward), :152-241 / :243-336 (gradients);
//    sizes from correlation_cuda.cc:25-35.
//
// Both are small memory/latency-bound operators.  The correlation keeps the reference's
// assignment of work to the 32 lanes of a warp (lane = channel mod 32 in the forward, lane =
// displacement mod 32 in the backward) and its reduction order, so the results are bit-identical
// to the C restatement in oracle/flownet_ops_oracle.c; what changes is the grouping: several
// warps per CTA share the padded rows through L1, the padded copies are made by a tiled
// transpose, and nothing is launched per batch item.
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

// ------------------------------------------------------------------ channelnorm

__global__ void channelnorm_fwd_kernel(const float* __restrict__ in, float* __restrict__ out, int C, long hw, long total) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long b = i / hw, p = i - b * hw;
        const float* src = in + b * C * hw + p;
        float r = 0.0f;
        for (int c = 0; c < C; ++c) {
            const float v = __ldg(src + (long)c * hw);
            r = fmaf(v, v, r);
        }
        out[i] = sqrtf(r);
    }
}

__global__ void channelnorm_bwd_kernel(const float* __restrict__ in, const float* __restrict__ out,
                                       const float* __restrict__ gout, float* __restrict__ gin, int C, long hw, long total) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long b = i / (C * hw), p = i % hw;
        const long o = b * hw + p;
        const float num = __fmul_rn(__ldg(gout + o), __ldg(in + i));
        gin[i] = (float)((double)num / ((double)__ldg(out + o) + 1e-9));
    }
}

// ------------------------------------------------------------------ correlation

struct CorrGeom {
    int B, C, H, W, pad, ksize, max_disp, stride1, stride2;
    int ph, pw, krad, drad, dsize, oc, oh, ow;
};

// [B][C][H][W] -> zero-padded [B][pH][pW][C]; 32 x 32 (channel x column) tiles through shared memory
__global__ void corr_pad_nhwc_kernel(const float* __restrict__ in, float* __restrict__ r, CorrGeom g) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z / g.H, y = blockIdx.z % g.H;
    const int x0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const int c = c0 + k, x = x0 + threadIdx.x;
        tile[k][threadIdx.x] = (c < g.C && x < g.W) ? __ldg(in + (((size_t)n * g.C + c) * g.H + y) * g.W + x) : 0.0f;
    }
    __syncthreads();
    for (int k = threadIdx.y; k < 32; k += blockDim.y) {
        const int x = x0 + k, c = c0 + threadIdx.x;
        if (x < g.W && c < g.C) r[(((size_t)n * g.ph + y + g.pad) * g.pw + x + g.pad) * g.C + c] = tile[threadIdx.x][k];
    }
}

constexpr int kCorrWarps = 4;

// one warp per output pixel, lane = channel mod 32
__global__ void __launch_bounds__(kCorrWarps * 32)
corr_fwd_kernel(const float* __restrict__ r1, const float* __restrict__ r2, float* __restrict__ out, CorrGeom g) {
    const int lane = threadIdx.x & 31;
    const long pix = (long)blockIdx.x * kCorrWarps + (threadIdx.x >> 5);
    const long npix = (long)g.B * g.oh * g.ow;
    if (pix >= npix) return;
    const int bx = (int)(pix % g.ow), by = (int)((pix / g.ow) % g.oh), n = (int)(pix / ((long)g.ow * g.oh));
    const int y1 = by * g.stride1 + g.max_disp, x1 = bx * g.stride1 + g.max_disp;
    const size_t pdimxc = (size_t)g.pw * g.C, pdimyxc = (size_t)g.ph * pdimxc;
    const float nelems = (float)(g.ksize * g.ksize * g.C);
    const float* b1 = r1 + n * pdimyxc;
    const float* b2 = r2 + n * pdimyxc;
    for (int tj = -g.drad; tj <= g.drad; ++tj)
        for (int ti = -g.drad; ti <= g.drad; ++ti) {
            const int x2 = x1 + ti * g.stride2, y2 = y1 + tj * g.stride2;
            float acc = 0.0f;
            for (int j = -g.krad; j <= g.krad; ++j)
                for (int i = -g.krad; i <= g.krad; ++i) {
                    const float* p1 = b1 + (size_t)(y1 + j) * pdimxc + (size_t)(x1 + i) * g.C;
                    const float* p2 = b2 + (size_t)(y2 + j) * pdimxc + (size_t)(x2 + i) * g.C;
                    for (int ch = lane; ch < g.C; ch += 32) acc = fmaf(__ldg(p1 + ch), __ldg(p2 + ch), acc);
                }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) acc = __fadd_rn(acc, __shfl_down_sync(0xffffffffu, acc, off));
            if (lane == 0) {
                const int tc = (tj + g.drad) * g.dsize + (ti + g.drad);
                out[(((size_t)n * g.oc + tc) * g.oh + by) * g.ow + bx] = __fdiv_rn(acc, nelems);
            }
        }
}

constexpr int kCorrBwdWarps = 8;

// one warp per (pixel, channel), lane = displacement mod 32; kWhich 1: d/d input1 (reads padded
// input2), 2: d/d input2 (reads padded input1).  stride1 == 1 (enforced by the host code).
template <int kWhich>
__global__ void __launch_bounds__(kCorrBwdWarps * 32)
corr_bwd_kernel(const float* __restrict__ r, const float* __restrict__ gout, float* __restrict__ gin, CorrGeom g) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long pix = blockIdx.x;
    const int bx = (int)(pix % g.W), by = (int)((pix / g.W) % g.H), n = (int)(pix / ((long)g.W * g.H));
    const int y = by + g.pad, x = bx + g.pad;
    const size_t pdimxc = (size_t)g.pw * g.C, pdimyxc = (size_t)g.ph * pdimxc;
    const float nelems = (float)(g.ksize * g.ksize * g.C);
    int xmin0 = 0, xmax0 = -1, ymin0 = 0, ymax0 = -1;
    if (kWhich == 1) {
        xmin0 = x - g.krad - g.max_disp; ymin0 = y - g.krad - g.max_disp;
        xmax0 = x + g.krad - g.max_disp; ymax0 = y + g.krad - g.max_disp;
        if (xmax0 < 0 || ymax0 < 0 || xmin0 >= g.ow || ymin0 >= g.oh) return;      // gin is zero-filled by the host code
        xmin0 = max(0, xmin0); xmax0 = min(g.ow - 1, xmax0);
        ymin0 = max(0, ymin0); ymax0 = min(g.oh - 1, ymax0);
    }
    for (int c = warp; c < g.C; c += kCorrBwdWarps) {
        float ps = 0.0f;
        for (int tc = lane; tc < g.oc; tc += 32) {
            const int i2 = (tc % g.dsize - g.drad) * g.stride2, j2 = (tc / g.dsize - g.drad) * g.stride2;
            int xmin = xmin0, xmax = xmax0, ymin = ymin0, ymax = ymax0;
            float val;
            if (kWhich == 1) {
                val = __ldg(r + n * pdimyxc + (size_t)(y + j2) * pdimxc + (size_t)(x + i2) * g.C + c);
            } else {
                xmin = x - g.krad - g.max_disp - i2; ymin = y - g.krad - g.max_disp - j2;
                xmax = x + g.krad - g.max_disp - i2; ymax = y + g.krad - g.max_disp - j2;
                if (xmax < 0 || ymax < 0 || xmin >= g.ow || ymin >= g.oh) continue;
                xmin = max(0, xmin); xmax = min(g.ow - 1, xmax);
                ymin = max(0, ymin); ymax = min(g.oh - 1, ymax);
                val = __ldg(r + n * pdimyxc + (size_t)(y - j2) * pdimxc + (size_t)(x - i2) * g.C + c);
            }
            const float* gp = gout + ((size_t)n * g.oc + tc) * g.oh * g.ow;
            for (int j = ymin; j <= ymax; ++j)
                for (int i = xmin; i <= xmax; ++i) ps = fmaf(__ldg(gp + (size_t)j * g.ow + i), val, ps);
        }
        // the reference adds the 32 partial sums in lane order
        float sum = 0.0f;
#pragma unroll
        for (int t = 0; t < 32; ++t) sum = __fadd_rn(sum, __shfl_sync(0xffffffffu, ps, t));
        if (lane == 0) gin[(((size_t)n * g.C + c) * g.H + by) * g.W + bx] = __fdiv_rn(sum, nelems);
    }
}

int make_geom(CorrGeom& g, const char* who, int B, int C, int H, int W, int pad, int ksize, int max_disp, int stride1,
              int stride2) {
    IMX_REQUIRE(B >= 0 && C > 0 && H > 0 && W > 0, "%s: bad input size", who);
    IMX_REQUIRE(pad >= 0 && ksize >= 1 && (ksize & 1) && max_disp >= 0 && stride1 >= 1 && stride2 >= 1,
                "%s: bad geometry (odd kernel_size, strides >= 1)", who);
    g.B = B; g.C = C; g.H = H; g.W = W; g.pad = pad; g.ksize = ksize; g.max_disp = max_disp;
    g.stride1 = stride1; g.stride2 = stride2;
    g.ph = H + 2 * pad; g.pw = W + 2 * pad;
    g.krad = (ksize - 1) / 2; g.drad = max_disp / stride2; g.dsize = 2 * g.drad + 1;
    const int border = g.krad + max_disp;
    g.oc = g.dsize * g.dsize;
    g.oh = (int)ceilf((float)(g.ph - 2 * border) / (float)stride1);
    g.ow = (int)ceilf((float)(g.pw - 2 * border) / (float)stride1);
    IMX_REQUIRE(g.oh > 0 && g.ow > 0, "%s: empty output (padded input smaller than the displacement border)", who);
    // The reference does not check its geometry (kernel_size > 1 reads before the padded plane:
    // y1 starts at max_displacement, not at the border radius).  Refuse what would read out of bounds.
    const int lo = max_disp - g.drad * stride2 - g.krad;
    const int hi_y = (g.oh - 1) * stride1 + max_disp + g.drad * stride2 + g.krad;
    const int hi_x = (g.ow - 1) * stride1 + max_disp + g.drad * stride2 + g.krad;
    IMX_REQUIRE(lo >= 0 && hi_y < g.ph && hi_x < g.pw, "%s: this geometry reads outside the padded input "
                "(the reference would too); use kernel_size 1 or a larger pad_size", who);
    IMX_REQUIRE((long)B * H <= 65535 * 1L && (C + 31) / 32 <= 65535, "%s: batch * height too large", who);
    return IMX_OK;
}

int pad_copy(const float* in, float* r, const CorrGeom& g, cudaStream_t s) {
    const size_t bytes = (size_t)g.B * g.ph * g.pw * g.C * sizeof(float);
    cudaError_t e = cudaMemsetAsync(r, 0, bytes, s);
    if (e != cudaSuccess) return imx::set_error((int)e, "correlation: workspace memset: %s", cudaGetErrorString(e));
    dim3 grid((g.W + 31) / 32, (g.C + 31) / 32, g.B * g.H), block(32, 8);
    corr_pad_nhwc_kernel<<<grid, block, 0, s>>>(in, r, g);
    return imx::check_launch("correlation_pad");
}

}  // namespace

extern "C" int imx_channelnorm_fwd_f32(const float* input, float* output, int B, int C, int H, int W, int norm_deg,
                                       imx_stream_t stream) {
    IMX_REQUIRE(B >= 0 && C >= 0 && H >= 0 && W >= 0, "channelnorm_fwd: negative size");
    IMX_REQUIRE(norm_deg == 2, "channelnorm_fwd: norm_deg must be 2 (the reference ignores it and always computes the L2 norm)");
    const long total = (long)B * H * W;
    if (total == 0) return IMX_OK;
    IMX_REQUIRE(input && output, "channelnorm_fwd: null pointer");
    long blocks = (total + 255) / 256;
    if (blocks > 16L * imx::kNumSMs) blocks = 16L * imx::kNumSMs;
    channelnorm_fwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(input, output, C, (long)H * W, total);
    return imx::check_launch("channelnorm_fwd");
}

extern "C" int imx_channelnorm_bwd_f32(const float* input, const float* output, const float* grad_output,
                                       float* grad_input, int B, int C, int H, int W, int norm_deg, imx_stream_t stream) {
    IMX_REQUIRE(B >= 0 && C >= 0 && H >= 0 && W >= 0, "channelnorm_bwd: negative size");
    IMX_REQUIRE(norm_deg == 2, "channelnorm_bwd: norm_deg must be 2");
    const long total = (long)B * C * H * W;
    if (total == 0) return IMX_OK;
    IMX_REQUIRE(input && output && grad_output && grad_input, "channelnorm_bwd: null pointer");
    long blocks = (total + 255) / 256;
    if (blocks > 16L * imx::kNumSMs) blocks = 16L * imx::kNumSMs;
    channelnorm_bwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(input, output, grad_output, grad_input, C,
                                                                         (long)H * W, total);
    return imx::check_launch("channelnorm_bwd");
}

extern "C" int imx_correlation_output_size(int C, int H, int W, int pad_size, int kernel_size, int max_displacement,
                                           int stride1, int stride2, int* out_channels, int* out_h, int* out_w) {
    CorrGeom g;
    int rc = make_geom(g, "correlation_output_size", 1, C, H, W, pad_size, kernel_size, max_displacement, stride1, stride2);
    if (rc) return rc;
    IMX_REQUIRE(out_channels && out_h && out_w, "correlation_output_size: null pointer");
    *out_channels = g.oc; *out_h = g.oh; *out_w = g.ow;
    return IMX_OK;
}

extern "C" int imx_correlation_fwd_f32(const float* input1, const float* input2, float* rbot1, float* rbot2, float* output,
                                       int B, int C, int H, int W, int pad_size, int kernel_size, int max_displacement,
                                       int stride1, int stride2, imx_stream_t stream) {
    CorrGeom g;
    int rc = make_geom(g, "correlation_fwd", B, C, H, W, pad_size, kernel_size, max_displacement, stride1, stride2);
    if (rc) return rc;
    if (B == 0) return IMX_OK;
    IMX_REQUIRE(input1 && input2 && rbot1 && rbot2 && output, "correlation_fwd: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    if ((rc = pad_copy(input1, rbot1, g, s))) return rc;
    if ((rc = pad_copy(input2, rbot2, g, s))) return rc;
    const long npix = (long)B * g.oh * g.ow;
    corr_fwd_kernel<<<(unsigned)((npix + kCorrWarps - 1) / kCorrWarps), kCorrWarps * 32, 0, s>>>(rbot1, rbot2, output, g);
    return imx::check_launch("correlation_fwd");
}

extern "C" int imx_correlation_bwd_f32(const float* input1, const float* input2, const float* grad_output,
                                       float* rbot1, float* rbot2, float* grad_input1, float* grad_input2,
                                       int B, int C, int H, int W, int pad_size, int kernel_size, int max_displacement,
                                       int stride1, int stride2, imx_stream_t stream) {
    CorrGeom g;
    int rc = make_geom(g, "correlation_bwd", B, C, H, W, pad_size, kernel_size, max_displacement, stride1, stride2);
    if (rc) return rc;
    IMX_REQUIRE(stride1 == 1, "correlation_bwd: stride1 must be 1 (the reference's gradient kernels index out of bounds "
                              "for any other value; every FlowNet2 call site uses 1)");
    IMX_REQUIRE(pad_size >= g.drad * stride2, "correlation_bwd: pad_size must cover the displacement range "
                                              "(the reference reads out of bounds otherwise)");
    if (B == 0) return IMX_OK;
    IMX_REQUIRE(input1 && input2 && grad_output && rbot1 && rbot2 && grad_input1 && grad_input2, "correlation_bwd: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    if ((rc = pad_copy(input1, rbot1, g, s))) return rc;
    if ((rc = pad_copy(input2, rbot2, g, s))) return rc;
    const size_t bytes = (size_t)B * C * H * W * sizeof(float);
    cudaError_t e = cudaMemsetAsync(grad_input1, 0, bytes, s);
    if (e == cudaSuccess) e = cudaMemsetAsync(grad_input2, 0, bytes, s);
    if (e != cudaSuccess) return imx::set_error((int)e, "correlation_bwd: memset: %s", cudaGetErrorString(e));
    const unsigned blocks = (unsigned)((long)B * H * W);
    corr_bwd_kernel<1><<<blocks, kCorrBwdWarps * 32, 0, s>>>(rbot2, grad_output, grad_input1, g);
    corr_bwd_kernel<2><<<blocks, kCorrBwdWarps * 32, 0, s>>>(rbot1, grad_output, grad_input2, g);
    return imx::check_launch("correlation_bwd");
}
