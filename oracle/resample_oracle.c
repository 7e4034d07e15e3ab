/*
 * resample_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C restatement of the flow-warp arithmetic of StephChabr/imaginaire, used by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg as the checker / CPU
 * baseline for the CUDA kernels in imaginaire_b200/csrc/warp.cu.
 *
 *   r2d_*  follows imaginaire/third_party/resample2d/src/resample2d_kernel.cu
 *          :15-75  (forward), :78-129 (d/d image), :131-203 (d/d flow), kernel_size == 1.
 *          Pinned: the reference's own kernel bodies are compiled for host cores
 *          (oracle/ref_cuda/, output in oracle/_ref/) and this restatement equals them bit for
 *          bit -- forward and both gradients -- on tests/golden/ref_resample2d_*.npz
 *          (oracle/make_golden_native_ops.py) and on fresh inputs where oracle/_ref is present
 *          (tests/test_native_ops_reference_pin.py).  That is the -ffp-contract=off build; the
 *          build that fuses multiply-adds the way nvcc -fmad=true does differs from it by a
 *          rounding in the fourth tap and in the flow gradient's terms (held to 1e-5).
 *   gsw_*  follows imaginaire/model_utils/fs_vid2vid.py:14-39 resample() =
 *          F.grid_sample(bilinear, border, align_corners=True) over
 *          linspace(-1,1)+flow/((size-1)/2).  Pinned bit-exactly against the reference's
 *          own resample() run on CPU torch (tests/golden/warp_*.npz, oracle/make_golden.py).
 *
 * Build: gcc -O2 -fPIC -shared -ffp-contract=off -fopenmp (see oracle/build_oracle.py).
 * -ffp-contract=off matters: the fused multiply-adds below are explicit fmaf() calls and
 * nothing else may be contracted.
 */
#include <math.h>
#include <stddef.h>
#include <string.h>

static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* resample2d_kernel.cu:15-75.  img: B x C x Hi x Wi, flow: B x 2 x H x W, out: B x C x H x W */
void r2d_forward(const float* img, const float* flow, float* out, int B, int C, int H, int W, int Hi, int Wi) {
    const size_t hw = (size_t)H * W, hwi = (size_t)Hi * Wi;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                const float dx = flow[((size_t)b * 2 + 0) * hw + (size_t)y * W + x];
                const float dy = flow[((size_t)b * 2 + 1) * hw + (size_t)y * W + x];
                const float xf = (float)x + dx, yf = (float)y + dy;
                const float alpha = xf - floorf(xf), beta = yf - floorf(yf);
                /* clamped against the OUTPUT dims, :52-55 */
                const int xL = clampi((int)floorf(xf), 0, W - 1), xR = clampi((int)(floorf(xf) + 1), 0, W - 1);
                const int yT = clampi((int)floorf(yf), 0, H - 1), yB = clampi((int)(floorf(yf) + 1), 0, H - 1);
                for (int c = 0; c < C; ++c) {
                    const float* ip = img + ((size_t)b * C + c) * hwi;
                    float val = 0.0f;
                    /* (1. - alpha) promotes to double; each product is rounded to float, :59-62 */
                    val += (float)((1. - alpha) * (1. - beta) * ip[(size_t)yT * Wi + xL]);
                    val += (float)((alpha) * (1. - beta) * ip[(size_t)yT * Wi + xR]);
                    val += (float)((1. - alpha) * (beta)*ip[(size_t)yB * Wi + xL]);
                    val += (float)((alpha) * (beta)*ip[(size_t)yB * Wi + xR]);
                    out[((size_t)b * C + c) * hw + (size_t)y * W + x] = val;
                }
            }
}

/* resample2d_kernel.cu:78-129 (scatter; sequential here, so deterministic) and :131-203.
 * gimg (B x C x Hi x Wi) is accumulated into (caller zero-fills), gflow is overwritten.
 * trunc_alpha != 0: the reference's `xf - int(xf)`; 0: floor. */
void r2d_backward(const float* img, const float* flow, const float* gout, float* gimg, float* gflow, int B, int C,
                  int H, int W, int Hi, int Wi, int trunc_alpha) {
    const size_t hw = (size_t)H * W, hwi = (size_t)Hi * Wi;
    for (int b = 0; b < B; ++b)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                const float dx = flow[((size_t)b * 2 + 0) * hw + (size_t)y * W + x];
                const float dy = flow[((size_t)b * 2 + 1) * hw + (size_t)y * W + x];
                const float xf = (float)x + dx, yf = (float)y + dy;
                if (gimg) {
                    const float alpha = trunc_alpha ? xf - (float)(int)xf : xf - floorf(xf);
                    const float beta = trunc_alpha ? yf - (float)(int)yf : yf - floorf(yf);
                    const int xL = clampi((int)floorf(xf), 0, Wi - 1), xR = clampi((int)(floorf(xf) + 1), 0, Wi - 1);
                    const int yT = clampi((int)floorf(yf), 0, Hi - 1), yB = clampi((int)(floorf(yf) + 1), 0, Hi - 1);
                    for (int c = 0; c < C; ++c) {
                        const float g = gout[((size_t)b * C + c) * hw + (size_t)y * W + x];
                        float* gp = gimg + ((size_t)b * C + c) * hwi;
                        gp[(size_t)yT * Wi + xL] += (1 - alpha) * (1 - beta) * g;
                        gp[(size_t)yT * Wi + xR] += (alpha) * (1 - beta) * g;
                        gp[(size_t)yB * Wi + xL] += (1 - alpha) * (beta)*g;
                        gp[(size_t)yB * Wi + xR] += (alpha) * (beta)*g;
                    }
                }
                if (gflow) {
                    const int xL = clampi((int)floorf(xf), 0, W - 1), xR = clampi((int)(floorf(xf) + 1), 0, W - 1);
                    const int yT = clampi((int)floorf(yf), 0, H - 1), yB = clampi((int)(floorf(yf) + 1), 0, H - 1);
                    /* channel 0 (c % 2 == 0 branch, :187-199): gamma = 1 - (yf - floor(yf)) */
                    {
                        const float gamma = 1 - (yf - floorf(yf));
                        float output = 0.0f;
                        for (int ch = 0; ch < C; ++ch) {
                            const float g = gout[((size_t)b * C + ch) * hw + (size_t)y * W + x];
                            const float* ip = img + ((size_t)b * C + ch) * hwi;
                            output += (gamma)*g * ip[(size_t)yT * Wi + xR];
                            output -= (gamma)*g * ip[(size_t)yT * Wi + xL];
                            output += (1 - gamma) * g * ip[(size_t)yB * Wi + xR];
                            output -= (1 - gamma) * g * ip[(size_t)yB * Wi + xL];
                        }
                        gflow[((size_t)b * 2 + 0) * hw + (size_t)y * W + x] = output;
                    }
                    /* channel 1 (c % 2 == 1 branch, :173-185): gamma = 1 - (xf - floor(xf)) */
                    {
                        const float gamma = 1 - (xf - floorf(xf));
                        float output = 0.0f;
                        for (int ch = 0; ch < C; ++ch) {
                            const float g = gout[((size_t)b * C + ch) * hw + (size_t)y * W + x];
                            const float* ip = img + ((size_t)b * C + ch) * hwi;
                            output += (gamma)*g * ip[(size_t)yB * Wi + xL];
                            output -= (gamma)*g * ip[(size_t)yT * Wi + xL];
                            output += (1 - gamma) * g * ip[(size_t)yB * Wi + xR];
                            output -= (1 - gamma) * g * ip[(size_t)yT * Wi + xR];
                        }
                        gflow[((size_t)b * 2 + 1) * hw + (size_t)y * W + x] = output;
                    }
                }
            }
}

/* torch.linspace(-1, 1, n)[i] in fp32 (two-sided, fused multiply-add), fs_vid2vid.py:57-62 */
static float linspace_m1_p1(int i, int n) {
    const float step = 2.0f / (float)(n - 1);
    return (i < n / 2) ? fmaf(step, (float)i, -1.0f) : fmaf(-step, (float)(n - 1 - i), 1.0f);
}

/* fs_vid2vid.py:14-39 with F.grid_sample(bilinear, border, align_corners=True) written out.
 * img, out: B x C x H x W; flow: B x 2 x H x W. */
void gsw_forward(const float* img, const float* flow, float* out, int B, int C, int H, int W) {
    const size_t hw = (size_t)H * W;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                const float fx = flow[((size_t)b * 2 + 0) * hw + (size_t)y * W + x];
                const float fy = flow[((size_t)b * 2 + 1) * hw + (size_t)y * W + x];
                const float cx = linspace_m1_p1(x, W) + fx / (((float)W - 1.0f) / 2.0f);
                const float cy = linspace_m1_p1(y, H) + fy / (((float)H - 1.0f) / 2.0f);
                float ix = (cx + 1.0f) * ((float)(W - 1) / 2.0f);
                float iy = (cy + 1.0f) * ((float)(H - 1) / 2.0f);
                ix = fminf((float)(W - 1), fmaxf(ix, 0.0f));
                iy = fminf((float)(H - 1), fmaxf(iy, 0.0f));
                const float xw = floorf(ix), yn = floorf(iy);
                const float w = ix - xw, e = 1.0f - w, n = iy - yn, s = 1.0f - n;
                const float nw = s * e, ne = s * w, sw = n * e, se = n * w;
                const int ixw = (int)xw, iyn = (int)yn;
                const int e_ok = (ixw + 1) < W, s_ok = (iyn + 1) < H;
                for (int c = 0; c < C; ++c) {
                    const float* ip = img + ((size_t)b * C + c) * hw;
                    const float vnw = ip[(size_t)iyn * W + ixw];
                    const float vne = e_ok ? ip[(size_t)iyn * W + ixw + 1] : 0.0f;
                    const float vsw = s_ok ? ip[(size_t)(iyn + 1) * W + ixw] : 0.0f;
                    const float vse = (e_ok && s_ok) ? ip[(size_t)(iyn + 1) * W + ixw + 1] : 0.0f;
                    float o = vnw * nw;
                    o = fmaf(vne, ne, o);
                    o = fmaf(vsw, sw, o);
                    o = fmaf(vse, se, o);
                    out[((size_t)b * C + c) * hw + (size_t)y * W + x] = o;
                }
            }
}
