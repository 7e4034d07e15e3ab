// Flow warps, forward, through a shared-memory window.
//
// The per-pixel kernels of warp.cu gather four taps per output element straight from global memory.  On the
// SURVEY flow (|d flow| ~ 0.44 px per px) the 32 lanes of one gather span two to three 128-byte lines, and the
// L1 replays one wavefront per line: measured 1.45 output elements per clock and SM (0.47 of the HBM peak) with
// the L1 pipe 89 % busy while DRAM moved only the algorithmic bytes.
//
// Here a CTA owns a 64 x 16 tile of output pixels of one image.  It
//   1. reads its flow tile once and derives every pixel's taps and weights (kept in registers for the channel loop),
//   2. reduces the taps' bounding box over the tile and picks the smallest of three TMA boxes that holds it,
//   3. streams that source WINDOW of channel after channel into a 4-stage shared-memory ring with one
//      cp.async.bulk.tensor per channel (a dedicated producer warp; the window never passes through the LSU),
//   4. reads the four taps of a pixel with LDS (conflict degree 1-2 instead of 2-3 line replays) and writes the
//      outputs as full 128-byte rows.
// A tile whose bounding box exceeds the largest box (a flow discontinuity, far out-of-range flows) takes the
// per-pixel gathers for that tile only.  Arithmetic per output element is exactly that of warp.cu (bit-exact with
// the oracles): the window only changes where a tap is read from.
//
// Measured (B=8, C=64, 512x1024, fraction of the 6544 GB/s HBM peak; profiles/r02_warp_window.md):
//   grid_sample form   smooth flows (8 px on an 8x16 .. 1x1 grid) 0.67-0.76 against 0.57-0.65 per pixel; on the SURVEY
//                      flow (8 px on a 32x64 grid, |d flow| ~0.5-0.7 px per px with as much shear) half the tiles fit
//                      no box and the result is the per-pixel kernel's 0.47
//   resample2d form    slower than the per-pixel kernel on every flow (0.40-0.55 against 0.48-0.61): its three
//                      double-precision products per element cost six F2F conversions, and with four pixels per thread
//                      they wait on the shared-memory reads in a block (96 registers, 2 CTAs per SM).  The entry point
//                      in warp.cu therefore keeps the per-pixel kernel for resample2d; the instantiation stays for
//                      IMX_WARPWIN_R2D=1 experiments.
//
// Tensor maps cover this GPU's image tensor only.
#include <cuda.h>
#include <climits>
#include <cstdlib>
#include <type_traits>
#include "imx_common.cuh"
#include "imx_ptx.cuh"
#include "imx_tma_host.cuh"
#include "warp_window.cuh"

namespace imx {
namespace warpwin {

namespace ptx = imx::ptx;

constexpr int TW = 64, TH = 16;          // output tile
constexpr int PPT = 4;                   // pixels per thread (rows ty, ty + 4, ty + 8, ty + 12)
constexpr int CONSUMERS = 256;           // 8 warps
constexpr int THREADS = CONSUMERS + 32;  // + the producer warp
constexpr int STAGES = 4;
// Box (i, j) = width i x height j: widths hold a stretch of up to ~1.05 / 1.3 / 1.55 px per px in x (plus the +1 tap and
// the 16-byte alignment of the box start), heights ~1.2 / 1.45 / 1.7 in y.  Width and height are chosen independently.
constexpr int NBW = 3, NBH = 3, NBOX = NBW * NBH;
__host__ __device__ constexpr int box_w(int i) { return i == 0 ? 72 : (i == 1 ? 88 : 104); }
__host__ __device__ constexpr int box_h(int j) { return j == 0 ? 20 : (j == 1 ? 24 : 28); }
constexpr int STAGE_FLOATS = box_w(NBW - 1) * box_h(NBH - 1);   // 2912 floats = 11.4 KB
constexpr int SMEM_BYTES = STAGES * STAGE_FLOATS * 4;

struct Maps {
    CUtensorMap m[NBOX];
};

// ------------------------------------------------------------------ the two samplers
// A sampler turns (flow, pixel) into: the top-left tap (absolute image coordinates), whether the right / bottom
// taps are one step away or folded onto the same column / row, and its weights; combine() is the arithmetic of the
// corresponding per-pixel kernel in warp.cu.

struct R2DSampler {   // resample2d_kernel.cu:15-75
    struct Pix {
        double w00, w01, w10;
        float w11;
    };
    __device__ static void taps(float dx, float dy, int x, int y, int H, int W, int& xL, int& yT, int& dxr, int& dyb,
                                Pix& p) {
        const float xf = static_cast<float>(x) + dx, yf = static_cast<float>(y) + dy;
        const float fx = floorf(xf), fy = floorf(yf);
        const float alpha = xf - fx, beta = yf - fy;
        xL = max(min(int(fx), W - 1), 0);
        yT = max(min(int(fy), H - 1), 0);
        dxr = max(min(int(fx + 1.0f), W - 1), 0) - xL;
        dyb = max(min(int(fy + 1.0f), H - 1), 0) - yT;
        const double a = alpha, be = beta;
        p.w00 = (1. - a) * (1. - be);
        p.w01 = a * (1. - be);
        p.w10 = (1. - a) * be;
        p.w11 = __fmul_rn(alpha, beta);
    }
    __device__ static float combine(const Pix& p, float tl, float tr, float bl, float br, int, int) {
        float v = 0.0f;
        v += static_cast<float>(p.w00 * (double)tl);
        v += static_cast<float>(p.w01 * (double)tr);
        v += static_cast<float>(p.w10 * (double)bl);
        v += __fmul_rn(p.w11, br);
        return v;
    }
};

struct GSSampler {    // fs_vid2vid.py:14-39 = grid_sample(bilinear, border, align_corners=True) on linspace + flow
    struct Pix {
        float nw, ne, sw, se;
    };
    __device__ static float linspace_m1_p1(int i, int n) {
        const float step = 2.0f / static_cast<float>(n - 1);
        return (i < n / 2) ? fmaf(step, static_cast<float>(i), -1.0f) : fmaf(-step, static_cast<float>(n - 1 - i), 1.0f);
    }
    __device__ static void taps(float fx, float fy, int x, int y, int H, int W, int& xL, int& yT, int& dxr, int& dyb,
                                Pix& p) {
        const float cx = linspace_m1_p1(x, W) + __fdiv_rn(fx, (static_cast<float>(W) - 1.0f) / 2.0f);
        const float cy = linspace_m1_p1(y, H) + __fdiv_rn(fy, (static_cast<float>(H) - 1.0f) / 2.0f);
        float ix = __fmul_rn(cx + 1.0f, static_cast<float>(W - 1) / 2.0f);
        float iy = __fmul_rn(cy + 1.0f, static_cast<float>(H - 1) / 2.0f);
        ix = fminf(static_cast<float>(W - 1), fmaxf(ix, 0.0f));
        iy = fminf(static_cast<float>(H - 1), fmaxf(iy, 0.0f));
        const float xw = floorf(ix), yn = floorf(iy);
        const float w = ix - xw, e = 1.0f - w, n = iy - yn, s = 1.0f - n;
        xL = int(xw);
        yT = int(yn);
        const bool e_ok = (xL + 1) < W, s_ok = (yT + 1) < H;
        dxr = e_ok ? 1 : 0;
        dyb = s_ok ? 1 : 0;
        p.nw = __fmul_rn(s, e);
        p.ne = __fmul_rn(s, w);
        p.sw = __fmul_rn(n, e);
        p.se = __fmul_rn(n, w);
    }
    __device__ static float combine(const Pix& p, float nw, float ne, float sw, float se, int e_ok, int s_ok) {
        // a missing east / south neighbour enters as the VALUE zero, like grid_sample does it (its tap was folded
        // onto the pixel itself, so the load is always in range)
        float o = __fmul_rn(nw, p.nw);
        o = fmaf(e_ok ? ne : 0.0f, p.ne, o);
        o = fmaf(s_ok ? sw : 0.0f, p.sw, o);
        o = fmaf((e_ok && s_ok) ? se : 0.0f, p.se, o);
        return o;
    }
};

// ------------------------------------------------------------------ kernel

template <class S>
__global__ void __launch_bounds__(THREADS, std::is_same<S, R2DSampler>::value ? 2 : 3)
warp_fwd_window_kernel(const __grid_constant__ Maps maps, const float* __restrict__ img, const float* __restrict__ flow,
                       float* __restrict__ out, int C, int H, int W, int Hi, int Wi, int cpb, int nchunk) {
    // No integer round-trip on this pointer: the tap reads must stay LDS (shared state space), not generic loads.
    extern __shared__ __align__(128) float ring[];
    __shared__ uint64_t full[STAGES], empty[STAGES];
    __shared__ int red[8][4];
    __shared__ int win[4];   // x0, y0, box index (-1: no box fits), unused

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int b = blockIdx.z / nchunk;
    const int c0 = (blockIdx.z % nchunk) * cpb;
    const int c1 = min(C, c0 + cpb);
    const int x = blockIdx.x * TW + (tid & (TW - 1));
    const int ty = (tid >> 6) & 3;
    const size_t hw = (size_t)H * W, hwi = (size_t)Hi * Wi;

    if (tid == 0) {
        if (ptx::smem_u32(ring) & 127u) __trap();      // TMA destinations sit on 128-byte boundaries
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full[s], 1);
            ptx::mbar_init(&empty[s], CONSUMERS / 32);
        }
        ptx::fence_mbar_init();
    }

    // 1. taps of this thread's pixels
    int oTL[PPT], dxr[PPT], dyb[PPT];      // oTL: absolute (xL, yT) first, window-relative offset later
    int yTa[PPT];
    bool live[PPT];
    typename S::Pix pix[PPT];
    int xmin = INT_MAX, xmax = INT_MIN, ymin = INT_MAX, ymax = INT_MIN;
    if (warp < CONSUMERS / 32) {
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
            const int y = blockIdx.y * TH + ty + 4 * k;
            live[k] = (x < W) && (y < H);
            oTL[k] = yTa[k] = dxr[k] = dyb[k] = 0;
            if (live[k]) {
                const float* fl = flow + (size_t)b * 2 * hw + (size_t)y * W + x;
                S::taps(__ldg(fl), __ldg(fl + hw), x, y, H, W, oTL[k], yTa[k], dxr[k], dyb[k], pix[k]);
                xmin = min(xmin, oTL[k]);
                xmax = max(xmax, oTL[k] + dxr[k]);
                ymin = min(ymin, yTa[k]);
                ymax = max(ymax, yTa[k] + dyb[k]);
            }
        }
        xmin = __reduce_min_sync(0xffffffffu, xmin);
        xmax = __reduce_max_sync(0xffffffffu, xmax);
        ymin = __reduce_min_sync(0xffffffffu, ymin);
        ymax = __reduce_max_sync(0xffffffffu, ymax);
        if (lane == 0) {
            red[warp][0] = xmin; red[warp][1] = xmax; red[warp][2] = ymin; red[warp][3] = ymax;
        }
    }
    __syncthreads();
    if (tid == 0) {
        int x0 = INT_MAX, x1 = INT_MIN, y0 = INT_MAX, y1 = INT_MIN;
#pragma unroll
        for (int w = 0; w < CONSUMERS / 32; ++w) {
            x0 = min(x0, red[w][0]); x1 = max(x1, red[w][1]);
            y0 = min(y0, red[w][2]); y1 = max(y1, red[w][3]);
        }
        int bi = -1, bj = -1;
        if (x0 <= x1) {
            x0 &= ~3;       // the box must start on a 16-byte boundary of the row (TMA's global address granularity)
#pragma unroll
            for (int q = NBW - 1; q >= 0; --q)
                if (x1 - x0 < box_w(q)) bi = q;
#pragma unroll
            for (int q = NBH - 1; q >= 0; --q)
                if (y1 - y0 < box_h(q)) bj = q;
        }
        win[0] = x0; win[1] = y0; win[2] = (bi < 0 || bj < 0) ? -1 : bi * NBH + bj;
    }
    __syncthreads();
    const int wx0 = win[0], wy0 = win[1], kbox = win[2];

    if (kbox < 0) {
        // ---- the tile's taps do not fit a window: per-pixel gathers, as in warp.cu
        if (warp >= CONSUMERS / 32) return;
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
            if (!live[k]) continue;
            const int y = blockIdx.y * TH + ty + 4 * k;
            const int o = yTa[k] * Wi + oTL[k];
            const float* ip = img + ((size_t)b * C + c0) * hwi;
            float* op = out + ((size_t)b * C + c0) * hw + (size_t)y * W + x;
            for (int c = c0; c < c1; ++c, ip += hwi, op += hw) {
                const float tl = __ldg(ip + o), tr = __ldg(ip + o + dxr[k]);
                const float bl = __ldg(ip + o + dyb[k] * Wi), br = __ldg(ip + o + dyb[k] * Wi + dxr[k]);
                __stcs(op, S::combine(pix[k], tl, tr, bl, br, dxr[k], dyb[k]));
            }
        }
        return;
    }
    const int pitch = box_w(kbox / NBH);
    const uint32_t stage_bytes = (uint32_t)(pitch * box_h(kbox % NBH) * 4);

    if (warp == CONSUMERS / 32) {
        // ---- producer: one bulk tensor copy per channel into the ring
        if (lane == 0) {
            // constant indices only: a dynamically indexed kernel parameter is copied to local memory, and a tensor
            // map must be read from parameter / constant / global space
            const CUtensorMap* m = &maps.m[0];
#pragma unroll
            for (int q = 1; q < NBOX; ++q)
                if (kbox == q) m = &maps.m[q];
            ptx::prefetch_tmap(m);
            for (int c = c0, it = 0; c < c1; ++c, ++it) {
                const int s = it % STAGES;
                if (it >= STAGES) ptx::mbar_wait(&empty[s], ((it / STAGES) - 1) & 1);
                ptx::mbar_arrive_expect_tx(&full[s], stage_bytes);
                ptx::tma_load_3d(ring + (size_t)s * STAGE_FLOATS, m, &full[s], wx0, wy0, b * C + c);
            }
        }
        return;
    }

    // ---- consumers
#pragma unroll
    for (int k = 0; k < PPT; ++k) oTL[k] = live[k] ? (yTa[k] - wy0) * pitch + (oTL[k] - wx0) : 0;
    float* op = out + ((size_t)b * C + c0) * hw + (size_t)(blockIdx.y * TH + ty) * W + x;
    for (int c = c0, it = 0; c < c1; ++c, ++it, op += hw) {
        const int s = it % STAGES;
        ptx::mbar_wait(&full[s], (it / STAGES) & 1);
        const float* wp = ring + (size_t)s * STAGE_FLOATS;
        float tl[PPT], tr[PPT], bl[PPT], br[PPT];
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
            const float* q = wp + oTL[k];
            tl[k] = q[0];
            tr[k] = q[dxr[k]];
            bl[k] = q[dyb[k] * pitch];
            br[k] = q[dyb[k] * pitch + dxr[k]];
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&empty[s]);     // the window's values are in registers
#pragma unroll
        for (int k = 0; k < PPT; ++k)
            if (live[k]) __stcs(op + (size_t)(4 * k) * W, S::combine(pix[k], tl[k], tr[k], bl[k], br[k], dxr[k], dyb[k]));
    }
}

// ------------------------------------------------------------------ host side

static int encode_f32_box(CUtensorMap* m, const float* base, int Wi, int Hi, long planes, int bw, int bh) {
    auto fn = get_encode_fn();
    if (!fn) return set_error(IMX_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not available");
    const cuuint64_t dims[3] = {(cuuint64_t)Wi, (cuuint64_t)Hi, (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)Wi * 4, (cuuint64_t)Wi * Hi * 4};
    const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(IMX_ERR_DRIVER, "warp window: cuTensorMapEncodeTiled failed (CUresult %d)", (int)r);
    return IMX_OK;
}

bool applicable(const float* img, int B, int C, int H, int W, int Hi, int Wi) {
    // TMA needs 16-byte aligned rows; small problems keep the per-pixel kernels (a 64 x 16 tile per CTA would leave
    // most SMs idle, and the window pays off over many channels only)
    const char* off = getenv("IMX_WARPWIN");        // IMX_WARPWIN=0: per-pixel kernels only (A/B measurements)
    if (off && off[0] == '0') return false;
    if ((Wi & 3) || !aligned16(img) || C < 8) return false;
    if ((long)B * C * H * W < (8L << 20)) return false;
    return true;
}

template <class S>
static int launch(const float* img, const float* flow, float* out, int B, int C, int H, int W, int Hi, int Wi,
                  cudaStream_t stream) {
    Maps maps;
    for (int k = 0; k < NBOX; ++k) {
        int rc = encode_f32_box(&maps.m[k], img, Wi, Hi, (long)B * C, box_w(k / NBH), box_h(k % NBH));
        if (rc != IMX_OK) return rc;
    }
    static bool attr_set[kMaxDevices] = {};
    auto kern = warp_fwd_window_kernel<S>;
    cudaError_t e = ensure_dyn_smem(kern, SMEM_BYTES, attr_set);
    if (e != cudaSuccess) return set_error((int)e, "warp window: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const long tiles = (long)B * ((H + TH - 1) / TH) * ((W + TW - 1) / TW);
    // Channels per CTA.  Few: the CTAs of one wave then work on the same channels of neighbouring tiles at the same
    // time, so the overlap of their windows (2-3x the tile) is served by L2; with all 64 channels in one CTA the
    // CTAs drift apart after the first wave and DRAM reads grew to 1.5-2.4x the algorithmic bytes (ncu).  Many: the
    // per-CTA set-up (flow tile, taps, bounding box, pipeline fill) is amortised.
    const char* cpb_env = getenv("IMX_WARPWIN_CPB");
    int cpb = cpb_env ? atoi(cpb_env) : 32;
    if (cpb < 1 || cpb > C) cpb = C;
    const int nchunk = (C + cpb - 1) / cpb;
    if ((long)B * nchunk > 65535) return set_error(IMX_ERR_TOO_LARGE, "warp window: batch*channel-chunks too large");
    dim3 grid((W + TW - 1) / TW, (H + TH - 1) / TH, B * nchunk);
    kern<<<grid, THREADS, SMEM_BYTES, stream>>>(maps, img, flow, out, C, H, W, Hi, Wi, cpb, nchunk);
    return check_launch("warp_fwd_window");
}

int resample2d_fwd(const float* img, const float* flow, float* out, int B, int C, int H, int W, int Hi, int Wi,
                   cudaStream_t stream) {
    return launch<R2DSampler>(img, flow, out, B, C, H, W, Hi, Wi, stream);
}

int gridsample_fwd(const float* img, const float* flow, float* out, int B, int C, int H, int W, cudaStream_t stream) {
    return launch<GSSampler>(img, flow, out, B, C, H, W, H, W, stream);
}

}  // namespace warpwin
}  // namespace imx
