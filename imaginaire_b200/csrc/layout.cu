// Layout / resize helpers around the SPADE block:
// label-map nearest resize into the NHWC bf16 conv operand layout, nearest x2
// feature upsample, NCHW fp32 <-> NHWC bf16 boundary conversion and the conv-weight
// repack (W / sigma, bf16, GEMM layout).  All plain HBM-bound copies with 128-bit
// accesses on the NHWC side.
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

using imx::pack_bf16;

// dst[n,y,x,0:Cp] from src[n,0:L,sy,sx].  One thread per 8 output channels; the
// NCHW source is read with stride H*W between channels (label maps are small and
// L2-resident: the resize runs once per resolution per forward).
__global__ void nearest_resize_kernel(const float* __restrict__ src, uint4* __restrict__ dst,
                                      int N, int L, int H, int W, int h, int w, int Cp) {
    const int cg = Cp / 8;
    const long total = (long)N * h * w * cg;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cg);
        long p = i / cg;
        const int x = (int)(p % w); p /= w;
        const int y = (int)(p % h);
        const int n = (int)(p / h);
        // torch 'nearest': src index = floor(dst_index * (in / out)), scale in fp32
        const float sy_f = (float)H / (float)h, sx_f = (float)W / (float)w;
        const int sy = min((int)floorf((float)y * sy_f), H - 1);
        const int sx = min((int)floorf((float)x * sx_f), W - 1);
        const float* s = src + ((size_t)n * L * H + sy) * W + sx;
        float f[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int c = c8 * 8 + k;
            f[k] = c < L ? __ldg(s + (size_t)c * H * W) : 0.0f;
        }
        dst[i] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
    }
}

__global__ void upsample2x_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst,
                                  int N, int H, int W, int cg) {
    // one thread per source 16-byte chunk, written to the 2x2 destination block
    const long total = (long)N * H * W * cg;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cg);
        long p = i / cg;
        const int x = (int)(p % W); p /= W;
        const int y = (int)(p % H);
        const int n = (int)(p / H);
        const uint4 v = imx::ldg_stream(src + i);
        uint4* d = dst + (((size_t)n * 2 * H + 2 * y) * 2 * W + 2 * x) * cg + c8;
        imx::stg_stream(d, v);
        imx::stg_stream(d + cg, v);
        imx::stg_stream(d + (size_t)2 * W * cg, v);
        imx::stg_stream(d + (size_t)2 * W * cg + cg, v);
    }
}

__global__ void upsample2x_bwd_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst,
                                      int N, int H, int W, int cg) {
    const long total = (long)N * H * W * cg;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cg);
        long p = i / cg;
        const int x = (int)(p % W); p /= W;
        const int y = (int)(p % H);
        const int n = (int)(p / H);
        const uint4* s = src + (((size_t)n * 2 * H + 2 * y) * 2 * W + 2 * x) * cg + c8;
        const uint4 a = imx::ldg_stream(s), b = imx::ldg_stream(s + cg);
        const uint4 c = imx::ldg_stream(s + (size_t)2 * W * cg), d = imx::ldg_stream(s + (size_t)2 * W * cg + cg);
        const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
        const uint32_t cv[4] = {c.x, c.y, c.z, c.w}, dv[4] = {d.x, d.y, d.z, d.w};
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float lo = imx::bf16_lo(av[k]) + imx::bf16_lo(bv[k]) + imx::bf16_lo(cv[k]) + imx::bf16_lo(dv[k]);
            const float hi = imx::bf16_hi(av[k]) + imx::bf16_hi(bv[k]) + imx::bf16_hi(cv[k]) + imx::bf16_hi(dv[k]);
            o[k] = pack_bf16(lo, hi);
        }
        imx::stg_stream(dst + i, make_uint4(o[0], o[1], o[2], o[3]));
    }
}

// NCHW fp32 -> NHWC bf16 through a 32(pixels) x 32(channels) shared-memory transpose so
// both sides are coalesced.
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                    int C, long HW, int Cp) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const long p0 = (long)blockIdx.x * 32;
    const int c0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int c = c0 + j;
        const long p = p0 + threadIdx.x;
        tile[j][threadIdx.x] = (c < C && p < HW) ? __ldg(src + ((size_t)n * C + c) * HW + p) : 0.0f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const long p = p0 + j;
        const int c = c0 + threadIdx.x;
        if (p < HW && c < Cp) dst[((size_t)n * HW + p) * Cp + c] = __float2bfloat16_rn(tile[threadIdx.x][j]);
    }
}

// The same for few channels (Cp <= 64: label maps, images, the discriminator's label + image input).  The 32 x 32
// tiles above write 64-byte pieces of 80-byte pixels (Cp = 40) and run their second channel tile three quarters
// empty.  Here a CTA owns 128 consecutive pixels and ALL channels: every channel plane is read in 512-byte rows, ten
// loads in flight per thread, and the CTA's output is one contiguous piece of 128 * Cp * 2 bytes written in 16-byte
// chunks.  8 x 38 x 256 x 512 (the discriminator input): 150 -> ~50 us.
constexpr int kStripPx = 128, kStripBatch = 10;

__global__ void __launch_bounds__(256)
nchw_to_nhwc_strip_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int C, long HW, int Cp) {
    extern __shared__ float strip[];            // [kStripPx][Cp + 1]
    const int n = blockIdx.y;
    const long p0 = (long)blockIdx.x * kStripPx;
    const int npx = (int)min((long)kStripPx, HW - p0);
    const int pitch = Cp + 1;
    const int px = threadIdx.x & (kStripPx - 1), half = threadIdx.x >> 7;
    float* row = strip + px * pitch;
    if (px < npx) {
        const float* s0 = src + (size_t)n * C * HW + p0 + px;
        for (int c0 = half; c0 < C; c0 += 2 * kStripBatch) {
            float v[kStripBatch];
#pragma unroll
            for (int u = 0; u < kStripBatch; ++u) {
                const int c = c0 + 2 * u;
                v[u] = c < C ? __ldcs(s0 + (size_t)c * HW) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < kStripBatch; ++u) {
                const int c = c0 + 2 * u;
                if (c < C) row[c] = v[u];
            }
        }
        for (int c = C + half; c < Cp; c += 2) row[c] = 0.0f;
    }
    __syncthreads();
    const int cpp = Cp >> 3;                    // 16-byte chunks per pixel
    const unsigned m = 0xFFFFFFFFu / (unsigned)cpp + 1u;
    uint4* out = reinterpret_cast<uint4*>(dst + ((size_t)n * HW + p0) * Cp);
    for (int q = threadIdx.x; q < npx * cpp; q += 256) {
        const int p = cpp > 1 ? (int)__umulhi((unsigned)q, m) : q;
        const float* r = strip + p * pitch + (q - p * cpp) * 8;
        uint4 o;
        o.x = imx::pack_bf16(r[0], r[1]); o.y = imx::pack_bf16(r[2], r[3]);
        o.z = imx::pack_bf16(r[4], r[5]); o.w = imx::pack_bf16(r[6], r[7]);
        out[q] = o;
    }
}

__global__ void nhwc_to_nchw_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst,
                                    int C, long HW, int Cp, int apply_tanh) {
    __shared__ float tile[32][33];
    const int n = blockIdx.z;
    const long p0 = (long)blockIdx.x * 32;
    const int c0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const long p = p0 + j;
        const int c = c0 + threadIdx.x;
        tile[j][threadIdx.x] = (p < HW && c < C) ? __bfloat162float(src[((size_t)n * HW + p) * Cp + c]) : 0.0f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int c = c0 + j;
        const long p = p0 + threadIdx.x;
        if (c < C && p < HW) {
            float v = tile[threadIdx.x][j];
            if (apply_tanh) v = tanhf(v);
            dst[((size_t)n * C + c) * HW + p] = v;
        }
    }
}

// packed[co][tap][ci_p] = w[co][ci][tap] * inv_sigma                       (kMode == kPackPlain)
// packed[ci][T-1-tap][co_p] = w[co][ci][tap] * inv_sigma                   (kMode == kPackFlip)
// kMode == kPackGB: the plain pack of the interleaved gamma/beta filter of pack_gb_weight_kernel (packed row
// R -> channel (R / 128) * 64 + R % 64 of gamma (R % 128 < 64) or beta), each with its own scale.
//
// A [inner x taps] -> [taps x inner] transpose per row of the packed matrix through shared memory.  A CTA owns
// 128 inner indices of `rpb` consecutive packed rows (8 for 3x3: 36 KB of fp32 per CTA, 5 CTAs per SM):
// the training step packs ~0.9 G filter elements per iteration and the first version (one row, 4.6 KB per CTA,
// 4 loads per thread, a division per element) ran at a quarter of the HBM peak.  Plain: every row is one contiguous
// run of 128 * taps floats, copied by the warps of that row, kPackBatch loads in flight per thread.  Flip: the rows are
// consecutive input channels, so for every output channel the CTA reads ONE run of rpb * taps floats
// (w[co][r0 .. r0 + rpb)[:]) instead of rpb runs of `taps` floats.  Index arithmetic by multiplication
// (__umulhi with 2^32 / d, exact for the index range here).  The bf16 rows leave as 4-byte pairs, 256 B per row.
constexpr int kPackTile = 128;
constexpr int kPackMaxTaps = 49;   // up to 7x7; larger kernels take the scalar path
constexpr int kPackPlain = 0, kPackFlip = 1, kPackGB = 2;
constexpr int kPackBatch = 18;     // loads in flight per thread: 8 left the kernel at half of the HBM peak (~50 KB per SM)

__device__ __forceinline__ unsigned magic_of(int d) { return d > 1 ? 0xFFFFFFFFu / (unsigned)d + 1u : 0u; }
__device__ __forceinline__ int div_magic(int e, int d, unsigned m) { return d > 1 ? (int)__umulhi((unsigned)e, m) : e; }

// rows per CTA: the largest of 8 / 4 / 2 / 1 whose staging tile fits 48 KB
__host__ __device__ constexpr int pack_rows_per_cta(int taps) {
    return 8 * (kPackTile * (taps | 1) + taps) * 4 <= 48 * 1024 ? 8
         : 4 * (kPackTile * (taps | 1) + taps) * 4 <= 48 * 1024 ? 4
         : 2 * (kPackTile * (taps | 1) + taps) * 4 <= 48 * 1024 ? 2 : 1;
}

// One tile with run-time tap count and ragged edges (any tile of a filter with an unusual tap count; the edge tiles of
// the others).
template <int kMode, int rpb>
__device__ __forceinline__ void pack_tile_generic(const float* __restrict__ w, const float* __restrict__ w2,
                                                  const float* __restrict__ inv_sigma, const float* __restrict__ inv_sigma2,
                                                  __nv_bfloat16* __restrict__ packed, int Cout, int Cin, int taps, int Cp,
                                                  int r0, int i0, float* tile) {
    constexpr bool kFlip = kMode == kPackFlip;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rows_total = kMode == kPackGB ? 2 * ((Cout + 63) / 64 * 64) : (kFlip ? Cin : Cout);
    const int inner = kFlip ? Cout : Cin;
    const int nrows = min(rpb, rows_total - r0);
    const int n_in = max(0, min(kPackTile, inner - i0));   // valid inner entries (0 for a pure padding tile)
    const int pitch = taps | 1;                            // odd: consecutive j fall into different banks
    const int rowstride = kPackTile * pitch + taps;        // = taps (mod 32): a flip run (rr, t) walks the banks
    const unsigned m_taps = magic_of(taps);
    if (!kFlip) {
        constexpr int nsub = 8 / rpb;                      // warps per row
        const int row = warp % rpb, sub = warp / rpb;
        int c = r0 + row;                                  // source output channel
        const float* src_base = w;
        float s = 1.0f;
        bool valid = row < nrows;
        if (kMode == kPackGB) {
            const int within = c & 127;
            const bool is_beta = within >= 64;
            c = (c >> 7) * 64 + (within & 63);
            valid = valid && c < Cout;
            src_base = is_beta ? w2 : w;
            const float* sc = is_beta ? inv_sigma2 : inv_sigma;
            s = sc ? __ldg(sc) : 1.0f;
        } else {
            s = inv_sigma ? __ldg(inv_sigma) : 1.0f;
        }
        const int n = valid ? n_in * taps : 0;
        const float* src = src_base + ((size_t)c * Cin + i0) * taps;
        float* dst = tile + row * rowstride;
        constexpr int step = 32 * nsub;
        for (int e0 = sub * 32 + lane; e0 < n; e0 += kPackBatch * step) {
            float v[kPackBatch];
#pragma unroll
            for (int u = 0; u < kPackBatch; ++u) {
                const int e = e0 + u * step;
                v[u] = e < n ? __ldg(src + e) : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < kPackBatch; ++u) {
                const int e = e0 + u * step;
                if (e < n) dst[pitch == taps ? e : e + div_magic(e, taps, m_taps)] = v[u] * s;
            }
        }
        if (kMode == kPackGB && row < nrows && !valid) {   // rows past C of the last 64-channel block: zeros
            for (int e = sub * 32 + lane; e < n_in * pitch; e += step) dst[e] = 0.0f;
        }
    } else {
        const float s = inv_sigma ? __ldg(inv_sigma) : 1.0f;
        // One run per output channel j: w[i0 + j][r0 .. r0 + nrows)[:], L = nrows * taps <= 96 floats (the 48 KB
        // staging limit), read by ONE warp with lane-fixed positions lane, lane + 32, lane + 64 -- their (row, tap)
        // and shared-memory offsets are the same for every run, so the loop body is loads and stores only
        // (flat indexing cost ~20 instructions per element and left the kernel issue-bound at 0.40 of the HBM peak).
        const int L = nrows * taps;
        int sidx[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int rem = lane + 32 * k;
            const int rr = div_magic(rem, taps, m_taps);
            sidx[k] = rem < L ? rr * rowstride + (rem - rr * taps) : -1;
        }
        const size_t CT = (size_t)Cin * taps;
        const float* run0 = w + ((size_t)i0 * Cin + r0) * taps + lane;
        constexpr int kRuns = kPackBatch / 3;              // runs in flight per warp
        for (int j0 = warp; j0 < n_in; j0 += 8 * kRuns) {
            float v[kRuns][3];
#pragma unroll
            for (int q = 0; q < kRuns; ++q) {
                const int j = j0 + 8 * q;
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    v[q][k] = (j < n_in && sidx[k] >= 0) ? __ldg(run0 + (size_t)j * CT + 32 * k) : 0.0f;
            }
#pragma unroll
            for (int q = 0; q < kRuns; ++q) {
                const int j = j0 + 8 * q;
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    if (j < n_in && sidx[k] >= 0) tile[sidx[k] + j * pitch] = v[q][k] * s;
            }
        }
    }
    __syncthreads();
    const int npair = min(kPackTile, Cp - i0) >> 1;        // Cp is a multiple of 8
    const int nrt = nrows * taps;
    for (int rt = warp; rt < nrt; rt += 8) {
        const int row = div_magic(rt, taps, m_taps), t = rt - row * taps;
        const float* srow = tile + row * rowstride + t;
        __nv_bfloat162* dst = reinterpret_cast<__nv_bfloat162*>(
            packed + ((size_t)(r0 + row) * taps + (kFlip ? taps - 1 - t : t)) * Cp + i0);
        for (int jp = lane; jp < npair; jp += 32) {
            const int j = 2 * jp;
            const float a = j < n_in ? srow[j * pitch] : 0.0f;
            const float b = j + 1 < n_in ? srow[(j + 1) * pitch] : 0.0f;
            dst[jp] = __floats2bfloat162_rn(a, b);
        }
    }
}

// grid: plain / gamma-beta (inner tiles, row groups); flip (row groups, inner tiles) -- consecutive CTAs of a flip pack
// take consecutive row groups of the same 128 output channels, i.e. adjacent runs of the same filter rows.
template <int kMode, int rpb>
__global__ void __launch_bounds__(256, 5)
pack_weight_rows_kernel(const float* __restrict__ w, const float* __restrict__ w2, const float* __restrict__ inv_sigma,
                        const float* __restrict__ inv_sigma2, __nv_bfloat16* __restrict__ packed, int Cout, int Cin,
                        int taps, int Cp) {
    extern __shared__ float tile[];   // [rpb] x (128 * pitch + taps): element (row, j, t) at row * rowstride + j * pitch + t
    constexpr bool kFlip = kMode == kPackFlip;
    pack_tile_generic<kMode, rpb>(w, w2, inv_sigma, inv_sigma2, packed, Cout, Cin, taps, Cp,
                                  (kFlip ? blockIdx.x : blockIdx.y) * rpb, (kFlip ? blockIdx.y : blockIdx.x) * kPackTile, tile);
}

// The same for the tap counts the networks use (3x3, 4x4, 5x5), everything but the pointers a compile-time constant:
// a full tile costs ~5 instructions per element (load, scale, shared store; two shared loads, convert, store per
// pair) where the run-time-indexed loops above issue ~40 and are bound by instruction issue, not by memory
// (profiles/r02_pack_kernel.md).  Ragged tiles fall through to pack_tile_generic.
template <int kMode, int kTaps>
__global__ void __launch_bounds__(256, 5)
pack_weight_taps_kernel(const float* __restrict__ w, const float* __restrict__ w2, const float* __restrict__ inv_sigma,
                        const float* __restrict__ inv_sigma2, __nv_bfloat16* __restrict__ packed, int Cout, int Cin, int Cp) {
    extern __shared__ float tile[];
    constexpr bool kFlip = kMode == kPackFlip;
    constexpr int rpb = pack_rows_per_cta(kTaps);
    constexpr int pitch = kTaps | 1, rowstride = kPackTile * pitch + kTaps;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rows_total = kMode == kPackGB ? 2 * ((Cout + 63) / 64 * 64) : (kFlip ? Cin : Cout);
    const int inner = kFlip ? Cout : Cin;
    const int r0 = (kFlip ? blockIdx.x : blockIdx.y) * rpb;
    const int i0 = (kFlip ? blockIdx.y : blockIdx.x) * kPackTile;
    bool full = rows_total - r0 >= rpb && inner - i0 >= kPackTile && Cp - i0 >= kPackTile;
    if (kMode == kPackGB) {            // the rows of a group lie in one 64-channel block: the last one decides
        const int rl = r0 + rpb - 1;
        full = full && ((rl >> 7) * 64 + (rl & 63)) < Cout;
    }
    if (!full) {
        pack_tile_generic<kMode, rpb>(w, w2, inv_sigma, inv_sigma2, packed, Cout, Cin, kTaps, Cp, r0, i0, tile);
        return;
    }
    if (!kFlip) {
        constexpr int nsub = 8 / rpb, step = 32 * nsub, T = 4 * kTaps / nsub;   // T loads per thread
        static_assert((4 * kTaps) % nsub == 0, "a row must divide evenly over its warps");
        const int row = warp % rpb, sub = warp / rpb;
        int c = r0 + row;
        const float* base = w;
        const float* sc = inv_sigma;
        if (kMode == kPackGB) {
            const bool is_beta = (c & 127) >= 64;
            c = (c >> 7) * 64 + (c & 63);
            base = is_beta ? w2 : w;
            sc = is_beta ? inv_sigma2 : inv_sigma;
        }
        const float s = sc ? __ldg(sc) : 1.0f;
        const int e_first = sub * 32 + lane;
        const float* src = base + ((size_t)c * Cin + i0) * kTaps + e_first;
        float* dst = tile + row * rowstride;
#pragma unroll
        for (int b = 0; b < T; b += kPackBatch) {
            float v[kPackBatch];
#pragma unroll
            for (int u = 0; u < kPackBatch; ++u)
                if (b + u < T) v[u] = __ldg(src + (b + u) * step);
#pragma unroll
            for (int u = 0; u < kPackBatch; ++u)
                if (b + u < T) {
                    const int e = e_first + (b + u) * step;
                    dst[pitch == kTaps ? e : e + e / kTaps] = v[u] * s;
                }
        }
    } else {
        constexpr int L = rpb * kTaps, K = (L + 31) / 32, kRuns = kPackBatch / K, kRunsPerWarp = kPackTile / 8;
        static_assert(L <= 96, "a run is read by one warp in at most three loads");
        const float s = inv_sigma ? __ldg(inv_sigma) : 1.0f;
        const bool last_ok = lane + 32 * (K - 1) < L;
        int sidx[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int rem = lane + 32 * k, rr = rem / kTaps;
            sidx[k] = rr * rowstride + (rem - rr * kTaps);
        }
        const size_t CT = (size_t)Cin * kTaps;
        const float* run0 = w + ((size_t)i0 * Cin + r0) * kTaps + lane;
#pragma unroll
        for (int q0 = 0; q0 < kRunsPerWarp; q0 += kRuns) {
            float v[kRuns][K];
#pragma unroll
            for (int q = 0; q < kRuns; ++q)
                if (q0 + q < kRunsPerWarp) {
                    const float* pr = run0 + (size_t)(warp + 8 * (q0 + q)) * CT;
#pragma unroll
                    for (int k = 0; k < K; ++k)
                        if (k < K - 1 || last_ok) v[q][k] = __ldg(pr + 32 * k);
                }
#pragma unroll
            for (int q = 0; q < kRuns; ++q)
                if (q0 + q < kRunsPerWarp) {
                    float* tj = tile + (warp + 8 * (q0 + q)) * pitch;
#pragma unroll
                    for (int k = 0; k < K; ++k)
                        if (k < K - 1 || last_ok) tj[sidx[k]] = v[q][k] * s;
                }
        }
    }
    __syncthreads();
    constexpr int nrt = rpb * kTaps;
    const float* sl = tile + 2 * lane * pitch;
#pragma unroll 3
    for (int rt = warp; rt < nrt; rt += 8) {
        const int row = rt / kTaps, t = rt - row * kTaps;
        const float* sr = sl + row * rowstride + t;
        __nv_bfloat162* dst = reinterpret_cast<__nv_bfloat162*>(
            packed + ((size_t)(r0 + row) * kTaps + (kFlip ? kTaps - 1 - t : t)) * Cp + i0) + lane;
        const float a0 = sr[0], b0 = sr[pitch], a1 = sr[64 * pitch], b1 = sr[65 * pitch];
        dst[0] = __floats2bfloat162_rn(a0, b0);
        dst[32] = __floats2bfloat162_rn(a1, b1);
    }
}

template <int kMode>
void launch_pack_rows(dim3 grid, cudaStream_t st, const float* w, const float* w2, const float* s1, const float* s2,
                      __nv_bfloat16* packed, int Cout, int Cin, int taps, int Cp) {
    const int rpb = pack_rows_per_cta(taps);
    const size_t smem = (size_t)rpb * (kPackTile * (taps | 1) + taps) * sizeof(float);
    if (taps == 9) { pack_weight_taps_kernel<kMode, 9><<<grid, 256, smem, st>>>(w, w2, s1, s2, packed, Cout, Cin, Cp); return; }
    if (taps == 16) { pack_weight_taps_kernel<kMode, 16><<<grid, 256, smem, st>>>(w, w2, s1, s2, packed, Cout, Cin, Cp); return; }
    if (taps == 25) { pack_weight_taps_kernel<kMode, 25><<<grid, 256, smem, st>>>(w, w2, s1, s2, packed, Cout, Cin, Cp); return; }
    switch (rpb) {
        case 8: pack_weight_rows_kernel<kMode, 8><<<grid, 256, smem, st>>>(w, w2, s1, s2, packed, Cout, Cin, taps, Cp); break;
        case 4: pack_weight_rows_kernel<kMode, 4><<<grid, 256, smem, st>>>(w, w2, s1, s2, packed, Cout, Cin, taps, Cp); break;
        case 2: pack_weight_rows_kernel<kMode, 2><<<grid, 256, smem, st>>>(w, w2, s1, s2, packed, Cout, Cin, taps, Cp); break;
        default: pack_weight_rows_kernel<kMode, 1><<<grid, 256, smem, st>>>(w, w2, s1, s2, packed, Cout, Cin, taps, Cp); break;
    }
}

// Splits the transposed, flipped filter of a stride-2 convolution, full[ci][KH-1-kh][KW-1-kw][co_p] (the dgrad pack),
// into the four parity-class sub-filters of imx_conv2d_igemm_fprop_window_bf16's stride-2 data gradient, stored one
// after the other in class order (a, b) = (0,0), (0,1), (1,0), (1,1):
//   cls[a][b][ci][ty][tx][:] = full[ci][KH-1-kh][KW-1-kw][:],  kh = k0(a) + 2 (T(a) - 1 - ty),  k0(a) = (a + pad) mod 2
// One thread per 16-byte chunk of `full` (coalesced both ways: whole co_p runs move).
struct ParitySplit { long long base[2][2]; int ty[2], tx[2]; };   // class offsets in 16-byte chunks, taps per class

__global__ void split_dgrad_s2_kernel(const uint4* __restrict__ full, uint4* __restrict__ out, ParitySplit ps, long total,
                                      int KH, int KW, int c8n, int pad) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % c8n);
        long t = i / c8n;
        const int fx = (int)(t % KW); t /= KW;
        const int fy = (int)(t % KH);
        const long ci = t / KH;
        const int kh = KH - 1 - fy, kw = KW - 1 - fx;
        const int a = (kh + pad) & 1, b = (kw + pad) & 1;          // parity of the input pixels this tap reaches
        const int ty = ps.ty[a] - 1 - (kh >> 1), tx = ps.tx[b] - 1 - (kw >> 1);   // (kh - k0) / 2 == kh >> 1 as k0 = kh & 1
        out[ps.base[a][b] + ((ci * ps.ty[a] + ty) * ps.tx[b] + tx) * c8n + c8] = __ldg(full + i);
    }
}

// Scalar fallback for very large kernels.
__global__ void pack_weight_kernel(const float* __restrict__ w, const float* __restrict__ inv_sigma,
                                   __nv_bfloat16* __restrict__ packed, int Cout, int Cin, int KH, int KW,
                                   int Cp, int transpose_flip) {
    const float s = inv_sigma ? __ldg(inv_sigma) : 1.0f;
    const int rows = transpose_flip ? Cin : Cout;    // rows of the packed matrix
    const int inner = transpose_flip ? Cout : Cin;   // valid extent of the innermost packed dim
    const long total = (long)rows * KH * KW * Cp;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int ci = (int)(i % Cp);
        long t = i / Cp;
        const int kw = (int)(t % KW); t /= KW;
        const int kh = (int)(t % KH);
        const int r = (int)(t / KH);
        float v = 0.0f;
        if (ci < inner) {
            if (!transpose_flip) v = __ldg(w + (((size_t)r * Cin + ci) * KH + kh) * KW + kw) * s;
            else v = __ldg(w + (((size_t)ci * Cin + r) * KH + (KH - 1 - kh)) * KW + (KW - 1 - kw)) * s;
        }
        packed[i] = __float2bfloat16_rn(v);
    }
}

// Interleaved gamma/beta filter for the fused projection+modulation conv: packed row
// (c / 64) * 128 + (c % 64) holds gamma's filter for channel c, row + 64 holds beta's.
__global__ void pack_gb_weight_kernel(const float* __restrict__ wg, const float* __restrict__ wb,
                                      const float* __restrict__ isg, const float* __restrict__ isb,
                                      __nv_bfloat16* __restrict__ packed, int C, int Ch, int KH, int KW, int Cp) {
    const float sg = isg ? __ldg(isg) : 1.0f, sb = isb ? __ldg(isb) : 1.0f;
    const int Cr = (C + 63) / 64 * 64;
    const long total = (long)2 * Cr * KH * KW * Cp;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int ci = (int)(i % Cp);
        long t = i / Cp;
        const int kw = (int)(t % KW); t /= KW;
        const int kh = (int)(t % KH);
        const int row = (int)(t / KH);
        const int blk = row / 128, within = row % 128;
        const bool is_beta = within >= 64;
        const int c = blk * 64 + (within % 64);
        float v = 0.0f;
        if (ci < Ch && c < C) {
            const float* src = is_beta ? wb : wg;
            v = __ldg(src + (((size_t)c * Ch + ci) * KH + kh) * KW + kw) * (is_beta ? sb : sg);
        }
        packed[i] = __float2bfloat16_rn(v);
    }
}

inline int grid_for(long total, int threads) {
    long b = (total + threads - 1) / threads;
    long cap = 16L * imx::kNumSMs;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

// ------------------------------------------------------------------ 2x2 / stride-2 max pooling
// ref: the perceptual loss' VGG-19 (torchvision layout: nn.MaxPool2d(2, 2) between the conv
// stages, imaginaire/losses/perceptual.py:175-192 (torchvision vgg19 .features: MaxPool2d(2, 2) at indices 4, 9, 18, 27)).  NHWC bf16, one thread per output pixel and
// 8-channel group.  ATen's kernel also writes int64 argmax indices (8 B per output element); here
// the backward recomputes the argmax from the saved input: the first of (0,0), (0,1), (1,0), (1,1)
// that equals the window maximum takes the gradient, as in ATen.  Floor mode: an odd last row /
// column is not pooled and gets a zero gradient.
__device__ __forceinline__ uint32_t bf16x2_max_nan(uint32_t a, uint32_t b) {
    const __nv_bfloat162 r = __hmax2_nan(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
    return *reinterpret_cast<const uint32_t*>(&r);
}
__device__ __forceinline__ uint4 max4_u4(const uint4& a, const uint4& b, const uint4& c, const uint4& d) {
    uint4 r;
    r.x = bf16x2_max_nan(bf16x2_max_nan(a.x, b.x), bf16x2_max_nan(c.x, d.x));
    r.y = bf16x2_max_nan(bf16x2_max_nan(a.y, b.y), bf16x2_max_nan(c.y, d.y));
    r.z = bf16x2_max_nan(bf16x2_max_nan(a.z, b.z), bf16x2_max_nan(c.z, d.z));
    r.w = bf16x2_max_nan(bf16x2_max_nan(a.w, b.w), bf16x2_max_nan(c.w, d.w));
    return r;
}

__global__ void __launch_bounds__(256, 4)
maxpool2x2_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int N, int H, int W, int cg) {
    const int Ho = H / 2, Wo = W / 2;
    const long total = (long)N * Ho * Wo * cg;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cg);
        long p = i / cg;
        const int x = (int)(p % Wo); p /= Wo;
        const int y = (int)(p % Ho);
        const int n = (int)(p / Ho);
        const uint4* q = src + (((size_t)n * H + 2 * y) * W + 2 * x) * cg + c8;
        const uint4 a = imx::ldg_stream(q), b = imx::ldg_stream(q + cg);
        const uint4 c = imx::ldg_stream(q + (size_t)W * cg), d = imx::ldg_stream(q + (size_t)W * cg + cg);
        imx::stg_stream(dst + i, max4_u4(a, b, c, d));
    }
}

// 16-bit lanes of `v` equal to the lanes of `m` and not yet taken -> gradient lanes of g, else 0;
// `taken` accumulates the lanes already routed (0xffff per lane).
__device__ __forceinline__ uint32_t route2(uint32_t v, uint32_t m, uint32_t g, uint32_t& taken) {
    // compare as floats so that -0 == +0, like ATen's value comparison
    const bool lo = imx::bf16_lo(v) == imx::bf16_lo(m), hi = imx::bf16_hi(v) == imx::bf16_hi(m);
    uint32_t mask = (lo ? 0x0000ffffu : 0u) | (hi ? 0xffff0000u : 0u);
    mask &= ~taken;
    taken |= mask;
    return g & mask;
}
__device__ __forceinline__ uint4 route8(const uint4& v, const uint4& m, const uint4& g, uint4& taken) {
    return make_uint4(route2(v.x, m.x, g.x, taken.x), route2(v.y, m.y, g.y, taken.y),
                      route2(v.z, m.z, g.z, taken.z), route2(v.w, m.w, g.w, taken.w));
}

__global__ void __launch_bounds__(256, 4)
maxpool2x2_bwd_kernel(const uint4* __restrict__ src, const uint4* __restrict__ gout, uint4* __restrict__ gsrc,
                      int N, int H, int W, int cg) {
    const int Ho = H / 2, Wo = W / 2;
    const long total = (long)N * Ho * Wo * cg;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cg);
        long p = i / cg;
        const int x = (int)(p % Wo); p /= Wo;
        const int y = (int)(p % Ho);
        const int n = (int)(p / Ho);
        const size_t o = (((size_t)n * H + 2 * y) * W + 2 * x) * cg + c8;
        const size_t row = (size_t)W * cg;
        const uint4 a = imx::ldg_stream(src + o), b = imx::ldg_stream(src + o + cg);
        const uint4 c = imx::ldg_stream(src + o + row), d = imx::ldg_stream(src + o + row + cg);
        const uint4 g = imx::ldg_stream(gout + i);
        const uint4 m = max4_u4(a, b, c, d);
        uint4 taken = make_uint4(0, 0, 0, 0);
        imx::stg_stream(gsrc + o, route8(a, m, g, taken));
        imx::stg_stream(gsrc + o + cg, route8(b, m, g, taken));
        imx::stg_stream(gsrc + o + row, route8(c, m, g, taken));
        imx::stg_stream(gsrc + o + row + cg, route8(d, m, g, taken));
    }
}

// ------------------------------------------------------------------ bilinear resize, NCHW fp32
// ref: the multi-resolution patch discriminator halves its (label, image) input between the
// discriminators with F.interpolate(scale_factor=0.5, mode='bilinear', align_corners=True),
// imaginaire/discriminators/spade.py:86-88 (multires_patch.py:169-171).  Arithmetic of ATen's
// upsample_bilinear2d (align_corners=True): source index = dst * (in - 1) / (out - 1) in fp32, the
// integer part picks the taps, the fraction the weights.  One thread per output element, x
// fastest: ATen's kernel walks the (n, c) planes inside each thread and takes 0.35 ms for the
// 8 x 38 x 256 x 512 discriminator input; this one is bound by the read of the input.
__global__ void __launch_bounds__(256)
bilinear_resize_ac_kernel(const float* __restrict__ src, float* __restrict__ dst, long planes, int H, int W, int Ho, int Wo,
                          float rh, float rw) {
    const long total = planes * Ho * Wo;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int x = (int)(i % Wo);
        long p = i / Wo;
        const int y = (int)(p % Ho);
        const long pl = p / Ho;
        const float h1r = rh * (float)y, w1r = rw * (float)x;
        const int h1 = (int)h1r, w1 = (int)w1r;
        const int h1p = h1 < H - 1 ? 1 : 0, w1p = w1 < W - 1 ? 1 : 0;
        const float h1l = h1r - (float)h1, h0l = 1.0f - h1l;
        const float w1l = w1r - (float)w1, w0l = 1.0f - w1l;
        const float* q = src + (pl * H + h1) * (long)W + w1;
        const float v00 = __ldg(q), v01 = __ldg(q + w1p), v10 = __ldg(q + (long)h1p * W), v11 = __ldg(q + (long)h1p * W + w1p);
        dst[i] = h0l * (w0l * v00 + w1l * v01) + h1l * (w0l * v10 + w1l * v11);
    }
}

}  // namespace

extern "C" int imx_nearest_resize_nchw_f32_to_nhwc_bf16(const float* src, imx_bf16* dst,
                                                        int N, int L, int H, int W,
                                                        int h, int w, int Cp, imx_stream_t stream) {
    IMX_REQUIRE(src && dst, "nearest_resize: null pointer");
    IMX_REQUIRE(N >= 0 && L > 0 && H > 0 && W > 0 && h > 0 && w > 0, "nearest_resize: bad size");
    IMX_REQUIRE(Cp >= L && Cp % 8 == 0, "nearest_resize: Cp (%d) must be >= L (%d) and a multiple of 8", Cp, L);
    const long total = (long)N * h * w * (Cp / 8);
    if (total == 0) return IMX_OK;
    nearest_resize_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        src, reinterpret_cast<uint4*>(dst), N, L, H, W, h, w, Cp);
    return imx::check_launch("nearest_resize");
}

extern "C" int imx_upsample2x_nhwc_bf16(const imx_bf16* src, imx_bf16* dst,
                                        int N, int H, int W, int C, imx_stream_t stream) {
    IMX_REQUIRE(src && dst, "upsample2x: null pointer");
    IMX_REQUIRE(N >= 0 && H >= 0 && W >= 0 && C >= 0 && C % 8 == 0, "upsample2x: bad size (C %% 8 must be 0)");
    const long total = (long)N * H * W * (C / 8);
    if (total == 0) return IMX_OK;
    upsample2x_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), N, H, W, C / 8);
    return imx::check_launch("upsample2x");
}

extern "C" int imx_upsample2x_bwd_nhwc_bf16(const imx_bf16* src, imx_bf16* dst,
                                            int N, int H, int W, int C, imx_stream_t stream) {
    IMX_REQUIRE(src && dst, "upsample2x_bwd: null pointer");
    IMX_REQUIRE(N >= 0 && H >= 0 && W >= 0 && C >= 0 && C % 8 == 0, "upsample2x_bwd: bad size (C %% 8 must be 0)");
    const long total = (long)N * H * W * (C / 8);
    if (total == 0) return IMX_OK;
    upsample2x_bwd_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), N, H, W, C / 8);
    return imx::check_launch("upsample2x_bwd");
}

// ---------------------------------------------------------------------------------------------
// Bilinear x2 upsampling (align_corners = False) fused with the lateral addition of a feature
// pyramid: out = up2x(src) + lat.  ref: FPSE discriminator top-down path,
// imaginaire/discriminators/fpse.py:119-122 (nn.Upsample(scale_factor=2, mode='bilinear') + lateral).
// For scale 2 the taps are fixed: an even output row 2k reads rows k-1, k with weights .25, .75
// (row 0 reads row 0 only), an odd row 2k+1 reads rows k, k+1 with weights .75, .25 (the last
// row reads row h-1 only); columns alike.
namespace {

__device__ __forceinline__ void up2x_taps(int o, int n, int& i0, int& i1, float& w0, float& w1) {
    const int k = o >> 1;
    if (o & 1) { i0 = k; i1 = min(k + 1, n - 1); w0 = 0.75f; w1 = 0.25f; }
    else { i0 = max(k - 1, 0); i1 = k; w0 = 0.25f; w1 = 0.75f; }
}

__device__ __forceinline__ void unpack8u(const uint4& v, float (&f)[8]) {
    f[0] = imx::bf16_lo(v.x); f[1] = imx::bf16_hi(v.x); f[2] = imx::bf16_lo(v.y); f[3] = imx::bf16_hi(v.y);
    f[4] = imx::bf16_lo(v.z); f[5] = imx::bf16_hi(v.z); f[6] = imx::bf16_lo(v.w); f[7] = imx::bf16_hi(v.w);
}

__global__ void bilinear_up2x_add_kernel(const uint4* __restrict__ src, const uint4* __restrict__ lat,
                                         uint4* __restrict__ out, int N, int h, int w, int cg) {
    const int H = 2 * h, W = 2 * w;
    const long total = (long)N * H * W * cg;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cg);
        long p = i / cg;
        const int ox = (int)(p % W); p /= W;
        const int oy = (int)(p % H);
        const int n = (int)(p / H);
        int y0, y1, x0, x1;
        float wy0, wy1, wx0, wx1;
        up2x_taps(oy, h, y0, y1, wy0, wy1);
        up2x_taps(ox, w, x0, x1, wx0, wx1);
        const uint4* b = src + (size_t)n * h * w * cg + c8;
        float a[8], t[8], acc[8];
        unpack8u(__ldg(b + ((size_t)y0 * w + x0) * cg), a);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = wy0 * wx0 * a[k];
        unpack8u(__ldg(b + ((size_t)y0 * w + x1) * cg), t);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(wy0 * wx1, t[k], acc[k]);
        unpack8u(__ldg(b + ((size_t)y1 * w + x0) * cg), t);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(wy1 * wx0, t[k], acc[k]);
        unpack8u(__ldg(b + ((size_t)y1 * w + x1) * cg), t);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(wy1 * wx1, t[k], acc[k]);
        if (lat) {
            unpack8u(imx::ldg_stream(lat + i), t);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[k] += t[k];
        }
        imx::stg_stream(out + i, make_uint4(imx::pack_bf16(acc[0], acc[1]), imx::pack_bf16(acc[2], acc[3]),
                                            imx::pack_bf16(acc[4], acc[5]), imx::pack_bf16(acc[6], acc[7])));
    }
}

// Gradient with respect to src: source pixel k receives from output rows 2k-1 (.25), 2k (.75, or 1
// for k = 0), 2k+1 (.75, or 1 for k = h-1), 2k+2 (.25); columns alike.
__device__ __forceinline__ void up2x_bwd_taps(int k, int n, int (&o)[4], float (&wt)[4]) {
    o[0] = 2 * k - 1; wt[0] = k >= 1 ? 0.25f : 0.0f;
    o[1] = 2 * k;     wt[1] = k == 0 ? 1.0f : 0.75f;
    o[2] = 2 * k + 1; wt[2] = k == n - 1 ? 1.0f : 0.75f;
    o[3] = 2 * k + 2; wt[3] = k <= n - 2 ? 0.25f : 0.0f;
}

__global__ void bilinear_up2x_bwd_kernel(const uint4* __restrict__ dout, uint4* __restrict__ dsrc, int N, int h, int w, int cg) {
    const int H = 2 * h, W = 2 * w;
    const long total = (long)N * h * w * cg;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % cg);
        long p = i / cg;
        const int x = (int)(p % w); p /= w;
        const int y = (int)(p % h);
        const int n = (int)(p / h);
        int oy[4], ox[4];
        float wy[4], wx[4];
        up2x_bwd_taps(y, h, oy, wy);
        up2x_bwd_taps(x, w, ox, wx);
        const uint4* b = dout + (size_t)n * H * W * cg + c8;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            if (wy[a] == 0.0f) continue;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (wx[c] == 0.0f) continue;
                float t[8];
                unpack8u(__ldg(b + ((size_t)oy[a] * W + ox[c]) * cg), t);
                const float wgt = wy[a] * wx[c];
#pragma unroll
                for (int k = 0; k < 8; ++k) acc[k] = fmaf(wgt, t[k], acc[k]);
            }
        }
        imx::stg_stream(dsrc + i, make_uint4(imx::pack_bf16(acc[0], acc[1]), imx::pack_bf16(acc[2], acc[3]),
                                             imx::pack_bf16(acc[4], acc[5]), imx::pack_bf16(acc[6], acc[7])));
    }
}

}  // namespace

extern "C" int imx_bilinear_up2x_add_nhwc_bf16(const imx_bf16* src, const imx_bf16* lateral, imx_bf16* out,
                                               int N, int h, int w, int C, imx_stream_t stream) {
    IMX_REQUIRE(src && out, "bilinear_up2x_add: null pointer");
    IMX_REQUIRE(N >= 0 && h >= 0 && w >= 0 && C >= 0 && C % 8 == 0, "bilinear_up2x_add: bad size (C %% 8 must be 0)");
    const long total = (long)N * 4 * h * w * (C / 8);
    if (total == 0) return IMX_OK;
    bilinear_up2x_add_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(src), reinterpret_cast<const uint4*>(lateral), reinterpret_cast<uint4*>(out), N, h, w, C / 8);
    return imx::check_launch("bilinear_up2x_add");
}

extern "C" int imx_bilinear_up2x_bwd_nhwc_bf16(const imx_bf16* grad_out, imx_bf16* grad_src,
                                               int N, int h, int w, int C, imx_stream_t stream) {
    IMX_REQUIRE(grad_out && grad_src, "bilinear_up2x_bwd: null pointer");
    IMX_REQUIRE(N >= 0 && h >= 0 && w >= 0 && C >= 0 && C % 8 == 0, "bilinear_up2x_bwd: bad size (C %% 8 must be 0)");
    const long total = (long)N * h * w * (C / 8);
    if (total == 0) return IMX_OK;
    bilinear_up2x_bwd_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(grad_out), reinterpret_cast<uint4*>(grad_src), N, h, w, C / 8);
    return imx::check_launch("bilinear_up2x_bwd");
}

extern "C" int imx_bilinear_resize_ac_nchw_f32(const float* src, float* dst, int64_t planes, int H, int W, int Ho, int Wo,
                                               imx_stream_t stream) {
    IMX_REQUIRE(planes >= 0 && H >= 0 && W >= 0 && Ho >= 0 && Wo >= 0, "bilinear_resize_ac: negative size");
    const long total = (long)planes * Ho * Wo;
    if (total == 0) return IMX_OK;
    IMX_REQUIRE(src && dst, "bilinear_resize_ac: null pointer");
    IMX_REQUIRE(H >= 1 && W >= 1, "bilinear_resize_ac: empty input");
    const float rh = Ho > 1 ? (float)(H - 1) / (float)(Ho - 1) : 0.0f;
    const float rw = Wo > 1 ? (float)(W - 1) / (float)(Wo - 1) : 0.0f;
    bilinear_resize_ac_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(src, dst, (long)planes, H, W, Ho, Wo, rh, rw);
    return imx::check_launch("bilinear_resize_ac");
}

extern "C" int imx_maxpool2x2_nhwc_bf16(const imx_bf16* src, imx_bf16* dst, int N, int H, int W, int C, imx_stream_t stream) {
    IMX_REQUIRE(src && dst, "maxpool2x2: null pointer");
    IMX_REQUIRE(N >= 0 && H >= 0 && W >= 0 && C >= 0 && C % 8 == 0, "maxpool2x2: bad size (C %% 8 must be 0)");
    const long total = (long)N * (H / 2) * (W / 2) * (C / 8);
    if (total == 0) return IMX_OK;
    maxpool2x2_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), N, H, W, C / 8);
    return imx::check_launch("maxpool2x2");
}

extern "C" int imx_maxpool2x2_bwd_nhwc_bf16(const imx_bf16* src, const imx_bf16* grad_out, imx_bf16* grad_src,
                                            int N, int H, int W, int C, imx_stream_t stream) {
    IMX_REQUIRE(src && grad_out && grad_src, "maxpool2x2_bwd: null pointer");
    IMX_REQUIRE(N >= 0 && H >= 0 && W >= 0 && C >= 0 && C % 8 == 0, "maxpool2x2_bwd: bad size (C %% 8 must be 0)");
    const long total = (long)N * (H / 2) * (W / 2) * (C / 8);
    if (total == 0) return IMX_OK;
    maxpool2x2_bwd_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint4*>(src), reinterpret_cast<const uint4*>(grad_out), reinterpret_cast<uint4*>(grad_src),
        N, H, W, C / 8);
    return imx::check_launch("maxpool2x2_bwd");
}

extern "C" int imx_nchw_f32_to_nhwc_bf16(const float* src, imx_bf16* dst, int N, int C, int H, int W,
                                         int Cp, imx_stream_t stream) {
    IMX_REQUIRE(src && dst, "nchw_to_nhwc: null pointer");
    IMX_REQUIRE(N >= 0 && C >= 0 && H >= 0 && W >= 0 && Cp >= C, "nchw_to_nhwc: bad size");
    IMX_REQUIRE(N <= 65535, "nchw_to_nhwc: batch too large");
    const long HW = (long)H * W;
    if ((long)N * HW * Cp == 0) return IMX_OK;
    if (Cp <= 64 && Cp % 8 == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
        dim3 sgrid((unsigned)((HW + kStripPx - 1) / kStripPx), N);
        nchw_to_nhwc_strip_kernel<<<sgrid, 256, (size_t)kStripPx * (Cp + 1) * sizeof(float), (cudaStream_t)stream>>>(
            src, reinterpret_cast<__nv_bfloat16*>(dst), C, HW, Cp);
        return imx::check_launch("nchw_to_nhwc");
    }
    dim3 grid((unsigned)((HW + 31) / 32), (Cp + 31) / 32, N), block(32, 8);
    IMX_REQUIRE(grid.y <= 65535, "nchw_to_nhwc: too many channels");
    nchw_to_nhwc_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(
        src, reinterpret_cast<__nv_bfloat16*>(dst), C, HW, Cp);
    return imx::check_launch("nchw_to_nhwc");
}

extern "C" int imx_nhwc_bf16_to_nchw_f32(const imx_bf16* src, float* dst, int N, int C, int H, int W,
                                         int Cp, int apply_tanh, imx_stream_t stream) {
    IMX_REQUIRE(src && dst, "nhwc_to_nchw: null pointer");
    IMX_REQUIRE(N >= 0 && C >= 0 && H >= 0 && W >= 0 && Cp >= C, "nhwc_to_nchw: bad size");
    IMX_REQUIRE(N <= 65535, "nhwc_to_nchw: batch too large");
    const long HW = (long)H * W;
    if ((long)N * HW * C == 0) return IMX_OK;
    dim3 grid((unsigned)((HW + 31) / 32), (C + 31) / 32, N), block(32, 8);
    IMX_REQUIRE(grid.y <= 65535, "nhwc_to_nchw: too many channels");
    nhwc_to_nchw_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const __nv_bfloat16*>(src), dst, C, HW, Cp, apply_tanh);
    return imx::check_launch("nhwc_to_nchw");
}

extern "C" int imx_pack_conv_weight_bf16(const float* w, const float* inv_sigma, imx_bf16* packed,
                                         int Cout, int Cin, int KH, int KW, int Cin_p,
                                         int transpose_flip, imx_stream_t stream) {
    IMX_REQUIRE(w && packed, "pack_conv_weight: null pointer");
    IMX_REQUIRE(Cout > 0 && Cin > 0 && KH > 0 && KW > 0, "pack_conv_weight: bad size");
    IMX_REQUIRE(Cin_p % 8 == 0 && Cin_p >= (transpose_flip ? Cout : Cin),
                "pack_conv_weight: padded inner dim (%d) too small or not a multiple of 8", Cin_p);
    const int taps = KH * KW;
    const int rows = transpose_flip ? Cin : Cout;
    if (taps <= kPackMaxTaps && rows <= 65535) {
        const int rpb = pack_rows_per_cta(taps);
        dim3 grid((Cin_p + kPackTile - 1) / kPackTile, (rows + rpb - 1) / rpb);
        if (transpose_flip) grid = dim3(grid.y, grid.x);
        if (transpose_flip)
            launch_pack_rows<kPackFlip>(grid, (cudaStream_t)stream, w, nullptr, inv_sigma, nullptr,
                                        reinterpret_cast<__nv_bfloat16*>(packed), Cout, Cin, taps, Cin_p);
        else
            launch_pack_rows<kPackPlain>(grid, (cudaStream_t)stream, w, nullptr, inv_sigma, nullptr,
                                         reinterpret_cast<__nv_bfloat16*>(packed), Cout, Cin, taps, Cin_p);
        return imx::check_launch("pack_conv_weight");
    }
    const long total = (long)rows * taps * Cin_p;
    pack_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        w, inv_sigma, reinterpret_cast<__nv_bfloat16*>(packed), Cout, Cin, KH, KW, Cin_p, transpose_flip);
    return imx::check_launch("pack_conv_weight");
}

extern "C" int imx_split_dgrad_s2_filter_bf16(const imx_bf16* full, imx_bf16* out, int Cin, int KH, int KW, int Cout_p,
                                             int pad, int64_t* class_offsets, imx_stream_t stream) {
    IMX_REQUIRE(full && out, "split_dgrad_s2_filter: null pointer");
    IMX_REQUIRE(Cin > 0 && KH >= 2 && KW >= 2 && KH <= 16 && KW <= 16 && pad >= 0, "split_dgrad_s2_filter: bad filter geometry");
    IMX_REQUIRE(Cout_p > 0 && Cout_p % 8 == 0, "split_dgrad_s2_filter: Cout_p must be a positive multiple of 8");
    IMX_REQUIRE(((reinterpret_cast<uintptr_t>(full) | reinterpret_cast<uintptr_t>(out)) & 15) == 0, "split_dgrad_s2_filter: pointers must be 16-byte aligned");
    ParitySplit ps;
    for (int a = 0; a < 2; ++a) {
        const int k0 = (a + pad) & 1;
        ps.ty[a] = (KH - k0 + 1) / 2;
        ps.tx[a] = (KW - k0 + 1) / 2;
    }
    const int c8n = Cout_p / 8;
    long long off = 0;
    for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b) {
            ps.base[a][b] = off;
            if (class_offsets) class_offsets[a * 2 + b] = off * 8;      // in elements
            off += (long long)Cin * ps.ty[a] * ps.tx[b] * c8n;
        }
    const long total = (long)Cin * KH * KW * c8n;
    long blocks = (total + 255) / 256;
    if (blocks > 8L * imx::kNumSMs) blocks = 8L * imx::kNumSMs;
    split_dgrad_s2_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4*>(full),
                                                                         reinterpret_cast<uint4*>(out), ps, total, KH, KW, c8n, pad);
    return imx::check_launch("split_dgrad_s2_filter");
}

extern "C" int imx_pack_spade_gb_weight_bf16(const float* w_gamma, const float* w_beta,
                                             const float* inv_sigma_gamma, const float* inv_sigma_beta,
                                             imx_bf16* packed, int C, int Ch, int KH, int KW, int Ch_p,
                                             imx_stream_t stream) {
    IMX_REQUIRE(w_gamma && w_beta && packed, "pack_spade_gb_weight: null pointer");
    IMX_REQUIRE(C > 0 && Ch > 0 && KH > 0 && KW > 0, "pack_spade_gb_weight: bad size");
    IMX_REQUIRE(Ch_p % 8 == 0 && Ch_p >= Ch, "pack_spade_gb_weight: bad Ch_p");
    const int taps_gb = KH * KW, rows_gb = 2 * ((C + 63) / 64 * 64);
    if (taps_gb <= kPackMaxTaps && rows_gb <= 65535 * 8) {
        const int rpb = pack_rows_per_cta(taps_gb);
        dim3 grid((Ch_p + kPackTile - 1) / kPackTile, (rows_gb + rpb - 1) / rpb);
        if (grid.y <= 65535) {
            launch_pack_rows<kPackGB>(grid, (cudaStream_t)stream, w_gamma, w_beta, inv_sigma_gamma, inv_sigma_beta,
                                      reinterpret_cast<__nv_bfloat16*>(packed), C, Ch, taps_gb, Ch_p);
            return imx::check_launch("pack_spade_gb_weight");
        }
    }
    const long total = (long)2 * ((C + 63) / 64 * 64) * KH * KW * Ch_p;
    pack_gb_weight_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(
        w_gamma, w_beta, inv_sigma_gamma, inv_sigma_beta, reinterpret_cast<__nv_bfloat16*>(packed),
        C, Ch, KH, KW, Ch_p);
    return imx::check_launch("pack_spade_gb_weight");
}
