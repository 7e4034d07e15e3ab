/*
 * flownet_ops_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C restatement of the two remaining native ops of StephChabr/imaginaire (SURVEY.md
 * section 8(f) rank 4), the checker for imaginaire_b200/csrc/flownet_ops.cu:
 *
 *   cn_*    imaginaire/third_party/channelnorm/src/channelnorm_kernel.cu
 *           :18-61 (forward: L2 norm over channels), :64-96 (backward)
 *   corr_*  imaginaire/third_party/correlation/src/correlation_cuda_kernel.cu
 *           :46-69 (padded channels-last copy), :72-149 (forward), :152-241 (d/d input1),
 *           :243-336 (d/d input2); output size from correlation_cuda.cc:25-35.
 *
 * Pinned: the reference's own kernel bodies are compiled for host cores (oracle/ref_cuda/,
 * output in oracle/_ref/, 32-lane blocks run as fibers with emulated shuffles and barriers) and
 * this restatement equals their multiply-add-fused build bit for bit on
 * tests/golden/ref_{channelnorm,correlation}_*.npz and on fresh inputs where oracle/_ref is
 * present (tests/test_native_ops_reference_pin.py); the -ffp-contract=off build is within 1e-6.
 * The restatement keeps the reference's ORDER of floating-point operations, including the
 * per-lane partial sums of the 32-thread blocks and their reduction trees, so that the CUDA
 * kernels (which keep the same order) can be compared bit for bit.  `acc += a * b` in the
 * reference is contracted to a fused multiply-add by nvcc's default -fmad=true; it is written
 * as an explicit fmaf() here.  Build with -ffp-contract=off (oracle/build_oracle.py).
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#define LANES 32

/* channelnorm_kernel.cu:18-61.  in: B x C x H x W, out: B x 1 x H x W */
void cn_forward(const float* in, float* out, int B, int C, int H, int W) {
    const size_t hw = (size_t)H * W;
    for (int b = 0; b < B; ++b)
        for (size_t p = 0; p < hw; ++p) {
            float result = 0.0f;
            for (int c = 0; c < C; ++c) {
                const float v = in[((size_t)b * C + c) * hw + p];
                result = fmaf(v, v, result);
            }
            out[(size_t)b * hw + p] = sqrtf(result);
        }
}

/* channelnorm_kernel.cu:64-96: gradInput = gradOutput * input / (output + 1e-9); the literal is a
 * double, so the quotient is formed in double and rounded once. */
void cn_backward(const float* in, const float* out, const float* gout, float* gin, int B, int C, int H, int W) {
    const size_t hw = (size_t)H * W;
    for (int b = 0; b < B; ++b)
        for (int c = 0; c < C; ++c)
            for (size_t p = 0; p < hw; ++p) {
                const size_t i = ((size_t)b * C + c) * hw + p, o = (size_t)b * hw + p;
                const float num = gout[o] * in[i];
                gin[i] = (float)((double)num / ((double)out[o] + 1e-9));
            }
}

/* correlation_cuda.cc:25-35 */
void corr_output_size(int C, int H, int W, int pad, int ksize, int max_disp, int stride1, int stride2,
                      int* out_c, int* out_h, int* out_w) {
    const int krad = (ksize - 1) / 2, border = krad + max_disp;
    const int ph = H + 2 * pad, pw = W + 2 * pad;
    const int d = (max_disp / stride2) * 2 + 1;
    (void)C;
    *out_c = d * d;
    *out_h = (int)ceilf((float)(ph - 2 * border) / (float)stride1);
    *out_w = (int)ceilf((float)(pw - 2 * border) / (float)stride1);
}

/* correlation_cuda_kernel.cu:46-69: zero-padded channels-last copy [B][pH][pW][C] */
static float* padded_nhwc(const float* in, int B, int C, int H, int W, int pad) {
    const int ph = H + 2 * pad, pw = W + 2 * pad;
    float* r = (float*)calloc((size_t)B * ph * pw * C, sizeof(float));
    for (int b = 0; b < B; ++b)
        for (int c = 0; c < C; ++c)
            for (int y = 0; y < H; ++y)
                for (int x = 0; x < W; ++x)
                    r[(((size_t)b * ph + y + pad) * pw + x + pad) * C + c] = in[(((size_t)b * C + c) * H + y) * W + x];
    return r;
}

/* correlation_cuda_kernel.cu:72-149.  One 32-lane block per output pixel; lane l owns channels
 * l, l+32, ...; lane 0 of the shfl_down tree holds the sum. */
void corr_forward(const float* in1, const float* in2, float* out, int B, int C, int H, int W,
                  int pad, int ksize, int max_disp, int stride1, int stride2) {
    int oc, oh, ow;
    corr_output_size(C, H, W, pad, ksize, max_disp, stride1, stride2, &oc, &oh, &ow);
    const int ph = H + 2 * pad, pw = W + 2 * pad;
    (void)ph;
    const int krad = (ksize - 1) / 2, drad = max_disp / stride2, dsize = 2 * drad + 1;
    const int nelems = ksize * ksize * C;
    float* r1 = padded_nhwc(in1, B, C, H, W, pad);
    float* r2 = padded_nhwc(in2, B, C, H, W, pad);
    const size_t pdimxc = (size_t)pw * C, pdimyxc = (size_t)(H + 2 * pad) * pdimxc;
    for (int n = 0; n < B; ++n)
        for (int by = 0; by < oh; ++by)
            for (int bx = 0; bx < ow; ++bx) {
                const int y1 = by * stride1 + max_disp, x1 = bx * stride1 + max_disp;
                for (int tj = -drad; tj <= drad; ++tj)
                    for (int ti = -drad; ti <= drad; ++ti) {
                        const int x2 = x1 + ti * stride2, y2 = y1 + tj * stride2;
                        float acc[LANES];
                        for (int l = 0; l < LANES; ++l) {
                            float a = 0.0f;
                            for (int j = -krad; j <= krad; ++j)
                                for (int i = -krad; i <= krad; ++i)
                                    for (int ch = l; ch < C; ch += LANES)
                                        a = fmaf(r1[n * pdimyxc + (y1 + j) * pdimxc + (size_t)(x1 + i) * C + ch],
                                                 r2[n * pdimyxc + (y2 + j) * pdimxc + (size_t)(x2 + i) * C + ch], a);
                            acc[l] = a;
                        }
                        for (int off = 16; off > 0; off /= 2)
                            for (int l = 0; l < off; ++l) acc[l] = acc[l] + acc[l + off];
                        const int tc = (tj + drad) * dsize + (ti + drad);
                        out[(((size_t)n * oc + tc) * oh + by) * ow + bx] = acc[0] / (float)nelems;
                    }
            }
    free(r1);
    free(r2);
}

/* correlation_cuda_kernel.cu:152-241 (which = 1: d/d input1, reads the padded input2) and
 * :243-336 (which = 2: d/d input2, reads the padded input1).  One 32-lane block per (y, x, c);
 * lane t owns output channels t, t+32, ...; the 32 partial sums are added sequentially. */
static void corr_backward_one(int which, const float* other, const float* gout, float* gin, int B, int C, int H, int W,
                              int pad, int ksize, int max_disp, int stride1, int stride2) {
    int oc, oh, ow;
    corr_output_size(C, H, W, pad, ksize, max_disp, stride1, stride2, &oc, &oh, &ow);
    const int pw = W + 2 * pad;
    const int krad = (ksize - 1) / 2, drad = max_disp / stride2, dsize = 2 * drad + 1;
    const float nelems = (float)(ksize * ksize * C);
    float* r = padded_nhwc(other, B, C, H, W, pad);
    const size_t pdimxc = (size_t)pw * C, pdimyxc = (size_t)(H + 2 * pad) * pdimxc;
    memset(gin, 0, (size_t)B * C * H * W * sizeof(float));
    for (int n = 0; n < B; ++n)
        for (int by = 0; by < H; ++by)
            for (int bx = 0; bx < W; ++bx)
                for (int c = 0; c < C; ++c) {
                    const int y = by * stride1 + pad, x = bx * stride1 + pad;
                    int xmin0 = 0, xmax0 = 0, ymin0 = 0, ymax0 = 0;
                    if (which == 1) {
                        xmin0 = (x - krad - max_disp) / stride1; ymin0 = (y - krad - max_disp) / stride1;
                        xmax0 = (x + krad - max_disp) / stride1; ymax0 = (y + krad - max_disp) / stride1;
                        if (xmax0 < 0 || ymax0 < 0 || xmin0 >= ow || ymin0 >= oh) continue;
                        if (xmin0 > xmax0 || ymin0 > ymax0) continue;
                        if (xmin0 < 0) xmin0 = 0;
                        if (xmax0 > ow - 1) xmax0 = ow - 1;
                        if (ymin0 < 0) ymin0 = 0;
                        if (ymax0 > oh - 1) ymax0 = oh - 1;
                    }
                    float prod[LANES];
                    for (int t = 0; t < LANES; ++t) {
                        float ps = 0.0f;
                        for (int tc = t; tc < oc; tc += LANES) {
                            const int i2 = (tc % dsize - drad) * stride2, j2 = (tc / dsize - drad) * stride2;
                            int xmin = xmin0, xmax = xmax0, ymin = ymin0, ymax = ymax0;
                            float val;
                            if (which == 1) {
                                val = r[n * pdimyxc + (size_t)(y + j2) * pdimxc + (size_t)(x + i2) * C + c];
                            } else {
                                xmin = (x - krad - max_disp - i2) / stride1; ymin = (y - krad - max_disp - j2) / stride1;
                                xmax = (x + krad - max_disp - i2) / stride1; ymax = (y + krad - max_disp - j2) / stride1;
                                if (xmax < 0 || ymax < 0 || xmin >= ow || ymin >= oh) continue;
                                if (xmin > xmax || ymin > ymax) continue;
                                if (xmin < 0) xmin = 0;
                                if (xmax > ow - 1) xmax = ow - 1;
                                if (ymin < 0) ymin = 0;
                                if (ymax > oh - 1) ymax = oh - 1;
                                val = r[n * pdimyxc + (size_t)(y - j2) * pdimxc + (size_t)(x - i2) * C + c];
                            }
                            for (int j = ymin; j <= ymax; ++j)
                                for (int i = xmin; i <= xmax; ++i)
                                    ps = fmaf(gout[(((size_t)n * oc + tc) * oh + j) * ow + i], val, ps);
                        }
                        prod[t] = ps;
                    }
                    float sum = 0.0f;
                    for (int t = 0; t < LANES; ++t) sum += prod[t];
                    gin[(((size_t)n * C + c) * H + (y - pad)) * W + (x - pad)] = sum / nelems;
                }
    free(r);
}

void corr_backward(const float* in1, const float* in2, const float* gout, float* gin1, float* gin2, int B, int C, int H,
                   int W, int pad, int ksize, int max_disp, int stride1, int stride2) {
    corr_backward_one(1, in2, gout, gin1, B, C, H, W, pad, ksize, max_disp, stride1, stride2);
    corr_backward_one(2, in1, gout, gin2, B, C, H, W, pad, ksize, max_disp, stride1, stride2);
}
