// Convolution weight gradient as an implicit GEMM on tcgen05 (SURVEY.md section 8 rows
// a4, a6, backward): for every filter tap
//     dW[co, tap, ci] = sum_{pixels} dY[pixel, co] * X[pixel @ tap, ci]
// i.e. M = Cout (128-row tiles), N = Cin (<= 256), K = N*Ho*Wo output pixels.
//
// ref: autograd of torch.nn.Conv2d inside Conv2dBlock, imaginaire/layers/conv.py:93-104
//
// Both operands are "MN-major" for the tensor core (the GEMM's M/N index - the channel -
// is the contiguous one in NHWC memory), which tcgen05 reads directly from the
// 128-byte-swizzled TMA boxes: a K step is a 64-pixel spatial patch; dY arrives as two
// {64 co, tw, th} boxes, X as BN/64 {64 ci, tw, th} boxes at the tap's spatial offset
// (zero padding = TMA out-of-bounds fill, stride 2 = tensor-map element strides).
//
// The pixel reduction is split over CTAs (split-K) so that small-channel / large-image
// layers still fill 148 SMs; each work unit (co tile, ci tile, tap, split) writes an fp32
// partial tile to the caller's workspace and a second kernel sums the splits into the
// PyTorch-layout gradient [Cout][Cin][KH][KW].
#include "imx_common.cuh"
#include "imx_ptx.cuh"
#include "imx_tma_host.cuh"
#include "../../include/imx.h"

namespace {

using namespace imx::ptx;
using imx::aligned16;
using imx::encode_tmap;
using imx::num_sms;

constexpr int BM = 128;      // output channels per tile
constexpr int BKP = 64;      // pixels per K step
constexpr int UMMA_K = 16;
constexpr int kThreads = 256;
constexpr int kEpiWarp0 = 4;
constexpr int BOX_BYTES = BKP * 64 * 2;   // one {64 ch x 64 px} box
constexpr int A_BYTES = 2 * BOX_BYTES;

// T = filter taps accumulated per work unit.  The dY tile (A operand) of a pixel step is the
// same for every tap, so a unit that owns T taps loads it once and issues T MMAs into T TMEM
// accumulators: with narrow Cin tiles (BN <= 128) a single tap would need more shared-memory
// fill bandwidth (A + B bytes per MMA cycle) than L2 can deliver.  T * BN = 256 columns per
// accumulator stage, two stages.
template <int BN> struct WCfg {
    static constexpr int T = 256 / BN;
    static constexpr int TAP_BYTES = (BN / 64) * BOX_BYTES;      // X boxes of one tap
    static constexpr int B_BYTES = T * TAP_BYTES;                  // = 32 KB for every BN
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = 4;
    static constexpr int TMEM_COLS = 2 * T * BN;                   // = 512
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
};

struct WgradArgs {
    int N, Ho, Wo;
    int KH, KW, stride, pad;
    int tw_log2, tw, th, tiles_x, tiles_y, ptiles;   // pixel tiles (K steps) in total
    int co_tiles, ci_tiles, splits, total_units;
    int co_r, ci_r;                                   // tile-rounded extents of the workspace
    float* ws;
};

struct Unit { int co_t, ci_t, tap0, ntaps, split; };

template <int T>
__device__ __forceinline__ Unit decode_unit(int u, const WgradArgs& p) {
    Unit r;
    const int taps = p.KH * p.KW;
    const int groups = (taps + T - 1) / T;
    // Tap group fastest, pixel split slowest: the CTAs of one wave then work on the same pixel
    // range (same dY boxes, overlapping X boxes), so those are fetched from HBM once and hit L2
    // for every other unit of the wave.
    const int g = u % groups; u /= groups;
    r.tap0 = g * T;
    r.ntaps = taps - r.tap0 < T ? taps - r.tap0 : T;
    r.ci_t = u % p.ci_tiles; u /= p.ci_tiles;
    r.co_t = u % p.co_tiles;
    r.split = u / p.co_tiles;
    return r;
}

template <int BN>
__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                  const WgradArgs p) {
    using C = WCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + C::STAGES * A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + C::STAGES;
    uint64_t* tmem_full = bars + 2 * C::STAGES;
    uint64_t* tmem_empty = bars + 2 * C::STAGES + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * C::STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_dy);
        prefetch_tmap(&tmap_x);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < C::STAGES; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 4 * 32);
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc(tmem_ptr, C::TMEM_COLS);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_ptr;

    const int tiles_per_img = p.tiles_x * p.tiles_y;
    // pixel tiles [k_begin, k_end) of split s
    auto k_range = [&](int split, int& kb, int& ke) {
        const int per = (p.ptiles + p.splits - 1) / p.splits;
        kb = split * per;
        ke = kb + per < p.ptiles ? kb + per : p.ptiles;
        if (kb > ke) kb = ke;
    };

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
                const Unit un = decode_unit<C::T>(u, p);
                int kb, ke;
                k_range(un.split, kb, ke);
                // Per-unit constants hoisted out of the pixel loop (this single thread's
                // instruction latency is on the critical path: ~6 TMA issues per 512 MMA cycles).
                int dx[C::T], dy[C::T];
#pragma unroll
                for (int t = 0; t < C::T; ++t) {
                    const int tap = un.tap0 + t;
                    const int kh = tap / p.KW;
                    dx[t] = tap - kh * p.KW - p.pad;
                    dy[t] = kh - p.pad;
                }
                const int co0 = un.co_t * BM, ci0 = un.ci_t * BN;
                int img = kb / tiles_per_img;
                int rem = kb - img * tiles_per_img;
                int ty = rem / p.tiles_x, tx = rem - ty * p.tiles_x;
                for (int k = kb; k < ke; ++k) {
                    const int oy0 = ty * p.th, ox0 = tx * p.tw;
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_arrive_expect_tx(&full_bar[stage], A_BYTES + un.ntaps * C::TAP_BYTES);
                    uint8_t* sa = smem_a + stage * A_BYTES;
                    tma_load_4d(sa, &tmap_dy, &full_bar[stage], co0, ox0, oy0, img);
                    tma_load_4d(sa + BOX_BYTES, &tmap_dy, &full_bar[stage], co0 + 64, ox0, oy0, img);
                    uint8_t* sb = smem_b + stage * C::B_BYTES;
                    const int ix0 = ox0 * p.stride, iy0 = oy0 * p.stride;
#pragma unroll
                    for (int t = 0; t < C::T; ++t) {
                        if (t < un.ntaps) {
#pragma unroll
                            for (int j = 0; j < BN / 64; ++j)
                                tma_load_4d(sb + t * C::TAP_BYTES + j * BOX_BYTES, &tmap_x, &full_bar[stage],
                                            ci0 + j * 64, ix0 + dx[t], iy0 + dy[t], img);
                        }
                    }
                    if (++tx == p.tiles_x) {
                        tx = 0;
                        if (++ty == p.tiles_y) { ty = 0; ++img; }
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            int iter = 0;
            for (int u = blockIdx.x; u < p.total_units; u += gridDim.x, ++iter) {
                const Unit un = decode_unit<C::T>(u, p);
                int kb, ke;
                k_range(un.split, kb, ke);
                const int as = iter & 1;
                const uint32_t aphase = (iter >> 1) & 1;
                mbar_wait(&tmem_empty[as], aphase ^ 1);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + as * (C::T * BN);
                // The X boxes of the unit's taps lie back to back in the stage (64-channel groups
                // BOX_BYTES apart = the descriptor's LBO), so ONE MMA with N = ntaps * BN covers all
                // taps: accumulator columns [t * BN, (t + 1) * BN) belong to tap t.  Both operands
                // are MN-major.
                const uint32_t idesc = make_idesc_bf16_f32(BM, un.ntaps * BN, 1, 1);
                for (int k = kb; k < ke; ++k) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after_sync();
                    const uint32_t a0 = smem_u32(smem_a + stage * A_BYTES);
                    const uint32_t b0 = smem_u32(smem_b + stage * C::B_BYTES);
#pragma unroll
                    for (int j = 0; j < BKP / UMMA_K; ++j) {
                        // 16 pixels = 16 rows of 128 B inside every 64-channel box
                        const uint64_t da = make_mnmajor_sw128_desc(a0 + j * UMMA_K * 128, 1024, BOX_BYTES);
                        const uint64_t db = make_mnmajor_sw128_desc(b0 + j * UMMA_K * 128, 1024, BOX_BYTES);
                        umma_bf16_ss(d_tmem, da, db, idesc, (k > kb) || (j > 0));
                    }
                    umma_commit(&empty_bar[stage]);
                    if (k == ke - 1) umma_commit(&tmem_full[as]);
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                if (kb == ke) {
                    // empty split (cannot happen with the host's split choice; keep the
                    // pipeline consistent anyway): nothing was accumulated
                    umma_commit(&tmem_full[as]);
                }
            }
        }
    } else if (warp >= kEpiWarp0) {
        const int q = warp & 3;
        const int row = q * 32 + lane;   // output channel inside the tile
        const int taps = p.KH * p.KW;
        int iter = 0;
        for (int u = blockIdx.x; u < p.total_units; u += gridDim.x, ++iter) {
            const Unit un = decode_unit<C::T>(u, p);
            int kb, ke;
            k_range(un.split, kb, ke);
            const int as = iter & 1;
            const uint32_t aphase = (iter >> 1) & 1;
            mbar_wait(&tmem_full[as], aphase);
            tc_fence_after_sync();
            const uint32_t t_acc = tmem_base + as * (C::T * BN) + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
            for (int cc = 0; cc < un.ntaps * (BN / 32); ++cc) {
                const int t = cc / (BN / 32), ch = cc - t * (BN / 32);
                float* dst = p.ws + ((((size_t)un.split * taps + un.tap0 + t) * p.co_r + un.co_t * BM + row) * p.ci_r) +
                             un.ci_t * BN;
                uint32_t v[32];
                tmem_ld_32x32(t_acc + t * BN + ch * 32, v);
                tmem_ld_wait();
                if (cc == un.ntaps * (BN / 32) - 1) {
                    tc_fence_before_sync();
                    mbar_arrive(&tmem_empty[as]);
                }
                float4* o = reinterpret_cast<float4*>(dst + ch * 32);
#pragma unroll
                for (int g = 0; g < 8; ++g) {
                    float4 f;
                    if (kb == ke) f = make_float4(0.f, 0.f, 0.f, 0.f);
                    else f = make_float4(__uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]),
                                         __uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3]));
                    o[g] = f;
                }
            }
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

// dw[co][ci][kh][kw] = sum_split ws[split][tap][co][ci]
__global__ void wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int Cout, int Cin,
                                    int taps, int splits, int co_r, int ci_r) {
    const long total = (long)Cout * Cin * taps;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        // consecutive threads -> consecutive ci (coalesced workspace reads)
        const int ci = (int)(i % Cin);
        long t = i / Cin;
        const int tap = (int)(t % taps);
        const int co = (int)(t / taps);
        float acc = 0.0f;
        for (int s = 0; s < splits; ++s)
            acc += __ldg(ws + (((size_t)s * taps + tap) * co_r + co) * ci_r + ci);
        dw[((size_t)co * Cin + ci) * taps + tap] = acc;
    }
}

// The same reduction with coalesced stores, and optionally <dw, W> for the spectral-norm gradient on the way.
// A CTA owns `kRedRows` output channels x a tile of 128 input channels: each thread sums the splits of one (co, ci)
// for every tap (consecutive threads -> consecutive ci: 512-byte workspace rows), parks the taps in shared memory and
// the CTA writes its run of 128 * taps floats of dw[co][ci0..][:] contiguously.  wgrad_reduce_kernel's stores are
// `taps` floats apart (one 4-byte word per 36-byte stride for a 3x3 filter), which held it at ~1.4 TB/s.
// dot (fp32, zeroed by the caller) += sum dw * W over the CTA's elements: W has dw's layout, so it is read in the
// same contiguous runs; one atomic per CTA.
constexpr int kRedRows = 2, kRedTile = 128;

__global__ void __launch_bounds__(kRedRows * kRedTile)
wgrad_reduce_tiled_kernel(const float* __restrict__ ws, float* __restrict__ dw, const float* __restrict__ Wt,
                          float* __restrict__ dot, int Cout, int Cin, int taps, int splits, int co_r, int ci_r) {
    extern __shared__ float rtile[];            // [kRedRows][kRedTile][taps + 1]
    __shared__ float rsum[kRedRows * kRedTile / 32];
    const int sub = threadIdx.x / kRedTile, j = threadIdx.x % kRedTile;
    const int co = blockIdx.y * kRedRows + sub;
    const int ci0 = blockIdx.x * kRedTile;
    const int n_ci = min(kRedTile, Cin - ci0);
    const int pitch = taps + 1;
    float* tile = rtile + (size_t)sub * kRedTile * pitch;
    if (co < Cout && j < n_ci) {
        const float* src = ws + (size_t)co * ci_r + ci0 + j;
        const size_t tap_stride = (size_t)co_r * ci_r, split_stride = (size_t)taps * tap_stride;
        // four taps at a time: four independent loads per split in flight (one by one the loop is a chain of
        // dependent 4-byte loads and the kernel ran at ~1.2 TB/s; 16-byte loads with a warp per output channel were
        // tried and lost: the 5x5 layers then fit only two rows of shared memory per CTA)
        for (int t0 = 0; t0 < taps; t0 += 4) {
            float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            for (int sp = 0; sp < splits; ++sp) {
                float v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    v[u] = (t0 + u < taps) ? __ldcs(src + sp * split_stride + (size_t)(t0 + u) * tap_stride) : 0.0f;
#pragma unroll
                for (int u = 0; u < 4; ++u) acc[u] += v[u];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (t0 + u < taps) tile[j * pitch + t0 + u] = acc[u];
        }
    }
    __syncthreads();
    float part = 0.0f;
    if (co < Cout) {
        const size_t base = ((size_t)co * Cin + ci0) * taps;
        for (int e = j; e < n_ci * taps; e += kRedTile) {
            const float v = tile[(e / taps) * pitch + e % taps];
            dw[base + e] = v;
            if (Wt) part = fmaf(v, __ldg(Wt + base + e), part);
        }
    }
    if (Wt) {
        part = imx::warp_sum(part);
        if ((threadIdx.x & 31) == 0) rsum[threadIdx.x >> 5] = part;
        __syncthreads();
        if (threadIdx.x == 0) {
            float tot = 0.0f;
            for (int i = 0; i < kRedRows * kRedTile / 32; ++i) tot += rsum[i];
            atomicAdd(dot, tot);
        }
    }
}

// The same with the tap count a template constant (1x1, 3x3, 4x4, 5x5): all taps of a split are loaded before any
// is used (kTaps independent 512-byte-per-warp loads in flight per thread instead of four), and the second phase has
// no division.  The run-time version above issued ~40 instructions per element and ran at half of the HBM peak
// (2048 x 2048 x 3x3, one split: 139 us for 453 MB).
template <int kTaps>
__global__ void __launch_bounds__(kRedRows * kRedTile, kTaps <= 9 ? 4 : 3)
wgrad_reduce_taps_kernel(const float* __restrict__ ws, float* __restrict__ dw, const float* __restrict__ Wt,
                         float* __restrict__ dot, int Cout, int Cin, int splits, int co_r, int ci_r) {
    constexpr int pitch = kTaps | 1;            // odd: consecutive ci in different banks
    extern __shared__ float rtile[];            // [kRedRows][kRedTile * pitch]
    __shared__ float rsum[kRedRows * kRedTile / 32];
    const int sub = threadIdx.x / kRedTile, j = threadIdx.x % kRedTile;
    const int co = blockIdx.y * kRedRows + sub;
    const int ci0 = blockIdx.x * kRedTile;
    const int n_ci = min(kRedTile, Cin - ci0);
    float* tile = rtile + sub * (kRedTile * pitch);
    if (co < Cout && j < n_ci) {
        const float* src = ws + (size_t)co * ci_r + ci0 + j;
        const size_t tap_stride = (size_t)co_r * ci_r, split_stride = (size_t)kTaps * tap_stride;
        float acc[kTaps];
#pragma unroll
        for (int t = 0; t < kTaps; ++t) acc[t] = __ldcs(src + t * tap_stride);
        for (int sp = 1; sp < splits; ++sp) {
            const float* s2 = src + sp * split_stride;
            float v[kTaps];
#pragma unroll
            for (int t = 0; t < kTaps; ++t) v[t] = __ldcs(s2 + t * tap_stride);
#pragma unroll
            for (int t = 0; t < kTaps; ++t) acc[t] += v[t];
        }
#pragma unroll
        for (int t = 0; t < kTaps; ++t) tile[j * pitch + t] = acc[t];
    }
    __syncthreads();
    float part = 0.0f;
    if (co < Cout) {
        const size_t base = ((size_t)co * Cin + ci0) * kTaps;
        const int n = n_ci * kTaps;
        float* out = dw + base + j;
        const float* wsrc = Wt ? Wt + base + j : nullptr;
        if (n == kRedTile * kTaps) {                       // full tile: no predicate per element
#pragma unroll
            for (int k = 0; k < kTaps; ++k) {
                const int e = j + k * kRedTile;
                const float v = tile[pitch == kTaps ? e : e + e / kTaps];
                out[k * kRedTile] = v;
                if (Wt) part = fmaf(v, __ldg(wsrc + k * kRedTile), part);
            }
        } else {
#pragma unroll
            for (int k = 0; k < kTaps; ++k) {
                const int e = j + k * kRedTile;
                if (e < n) {
                    const float v = tile[pitch == kTaps ? e : e + e / kTaps];
                    out[k * kRedTile] = v;
                    if (Wt) part = fmaf(v, __ldg(wsrc + k * kRedTile), part);
                }
            }
        }
    }
    if (Wt) {
        part = imx::warp_sum(part);
        if ((threadIdx.x & 31) == 0) rsum[threadIdx.x >> 5] = part;
        __syncthreads();
        if (threadIdx.x == 0) {
            float tot = 0.0f;
            for (int i = 0; i < kRedRows * kRedTile / 32; ++i) tot += rsum[i];
            atomicAdd(dot, tot);
        }
    }
}

template <int kTaps>
void launch_reduce_taps(dim3 grid, cudaStream_t s, const float* ws, float* dw, const float* Wt, float* dot, int Cout, int Cin,
                        int splits, int co_r, int ci_r) {
    const size_t smem = (size_t)kRedRows * kRedTile * (kTaps | 1) * sizeof(float);
    wgrad_reduce_taps_kernel<kTaps><<<grid, kRedRows * kRedTile, smem, s>>>(ws, dw, Wt, dot, Cout, Cin, splits, co_r, ci_r);
}

int launch_reduce(const float* ws, float* dw, const float* Wt, float* dot, int Cout, int Cin, int taps, int splits,
                  int co_r, int ci_r, cudaStream_t s) {
    const size_t smem = (size_t)kRedRows * kRedTile * (taps + 1) * sizeof(float);
    const int gy = (Cout + kRedRows - 1) / kRedRows;
    if (smem <= 48 * 1024 && gy <= 65535) {
        dim3 grid((Cin + kRedTile - 1) / kRedTile, gy);
        switch (taps) {
            case 1: launch_reduce_taps<1>(grid, s, ws, dw, Wt, dot, Cout, Cin, splits, co_r, ci_r); break;
            case 9: launch_reduce_taps<9>(grid, s, ws, dw, Wt, dot, Cout, Cin, splits, co_r, ci_r); break;
            case 16: launch_reduce_taps<16>(grid, s, ws, dw, Wt, dot, Cout, Cin, splits, co_r, ci_r); break;
            case 25: launch_reduce_taps<25>(grid, s, ws, dw, Wt, dot, Cout, Cin, splits, co_r, ci_r); break;
            default:
                wgrad_reduce_tiled_kernel<<<grid, kRedRows * kRedTile, smem, s>>>(ws, dw, Wt, dot, Cout, Cin, taps, splits,
                                                                                co_r, ci_r);
        }
        return imx::check_launch("conv2d_wgrad_reduce");
    }
    if (Wt) return imx::set_error(IMX_ERR_UNSUPPORTED, "conv2d_wgrad: fused <dw, W> needs taps <= 47 and Cout <= 131070");
    const long total = (long)Cout * Cin * taps;
    long blocks = (total + 255) / 256;
    if (blocks > 16L * imx::kNumSMs) blocks = 16L * imx::kNumSMs;
    wgrad_reduce_kernel<<<(int)blocks, 256, 0, s>>>(ws, dw, Cout, Cin, taps, splits, co_r, ci_r);
    return imx::check_launch("conv2d_wgrad_reduce");
}

constexpr int kMaxPeers = 8;
struct PeerPtrs { const float* p[kMaxPeers]; };

// Data-parallel weight gradient without a separate collective: every rank's split-K partials live
// in peer-mapped (NVLink) memory; this rank sums, over ALL ranks and splits, the output channels
// [co_begin, co_end) it owns and writes scale * sum into its own gradient tensor (PyTorch layout).
// The other ranks' channel ranges are fetched later by gather_rows_peers_kernel.
__global__ void wgrad_reduce_peers_kernel(PeerPtrs ws, int world, float* __restrict__ dw, int co_begin, int co_end,
                                          int Cin, int taps, int splits, int co_r, int ci_r, float scale, int accumulate) {
    const long total = (long)(co_end - co_begin) * Cin * taps;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int ci = (int)(i % Cin);
        long t = i / Cin;
        const int tap = (int)(t % taps);
        const int co = co_begin + (int)(t / taps);
        float acc = 0.0f;
        for (int q = 0; q < world; ++q) {
            const float* w = ws.p[q];
            for (int s = 0; s < splits; ++s) acc += __ldg(w + (((size_t)s * taps + tap) * co_r + co) * ci_r + ci);
        }
        float* dst = dw + ((size_t)co * Cin + ci) * taps + tap;
        *dst = accumulate ? fmaf(scale, acc, *dst) : scale * acc;
    }
}

struct GatherDesc { long long offset, rows, row_elems; };   // mirrored by imaginaire_b200/grad_sync.py (3 x int64)

// All-gather of the reduced gradients: tensor d (rows x row_elems floats at `offset` of the
// symmetric arena) takes rows [q*rows/world, (q+1)*rows/world) from peer q's arena, for every q
// other than this rank.  grid = (blocks, n_tensors).
__global__ void gather_rows_peers_kernel(PeerPtrs arena, float* __restrict__ mine, const GatherDesc* __restrict__ descs,
                                         int world, int rank) {
    const GatherDesc d = descs[blockIdx.y];
    for (int q = 0; q < world; ++q) {
        if (q == rank) continue;
        const long long lo = d.rows * q / world * d.row_elems, hi = d.rows * (q + 1) / world * d.row_elems;
        const float* src = arena.p[q] + d.offset;
        float* dst = mine + d.offset;
        const long long n = hi - lo;
        if (n <= 0) continue;
        const bool vec = (((uintptr_t)(src + lo) | (uintptr_t)(dst + lo)) & 15) == 0;
        const long long n4 = vec ? n / 4 : 0;
        const long long stride = (long long)gridDim.x * blockDim.x;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride)
            reinterpret_cast<float4*>(dst + lo)[i] = __ldcs(reinterpret_cast<const float4*>(src + lo) + i);
        for (long long i = n4 * 4 + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride)
            dst[lo + i] = __ldcs(src + lo + i);
    }
}

struct Plan {
    int BN, Ho, Wo, tw, th, tw_log2, tiles_x, tiles_y, ptiles, co_tiles, ci_tiles, splits;
    long ws_floats;
};

int make_plan(Plan& pl, int N, int H, int W, int Cin_p, int Cout_p, int KH, int KW, int stride, int pad) {
    pl.Ho = (H + 2 * pad - KH) / stride + 1;
    pl.Wo = (W + 2 * pad - KW) / stride + 1;
    if (pl.Ho <= 0 || pl.Wo <= 0) return imx::set_error(IMX_ERR_INVALID_ARG, "conv2d_wgrad: empty output");
    pl.BN = Cin_p > 128 ? 256 : (Cin_p > 64 ? 128 : 64);
    int tw = 64;
    while (tw > 8 && tw > pl.Wo) tw >>= 1;
    pl.tw = tw; pl.th = BKP / tw;
    pl.tw_log2 = 0;
    while ((1 << pl.tw_log2) < tw) ++pl.tw_log2;
    pl.tiles_x = (pl.Wo + pl.tw - 1) / pl.tw;
    pl.tiles_y = (pl.Ho + pl.th - 1) / pl.th;
    const long ptiles = (long)N * pl.tiles_x * pl.tiles_y;
    if (ptiles > 0x7fffffffL) return imx::set_error(IMX_ERR_TOO_LARGE, "conv2d_wgrad: too many pixel tiles");
    pl.ptiles = (int)ptiles;
    pl.co_tiles = (Cout_p + BM - 1) / BM;
    pl.ci_tiles = (Cin_p + pl.BN - 1) / pl.BN;
    const int T = 256 / pl.BN;
    const long base_units = (long)pl.co_tiles * pl.ci_tiles * ((KH * KW + T - 1) / T);
    // Split the pixel reduction so that the units fill the 148 persistent CTAs.  Cost model (us):
    // rounds x (K steps per unit x 0.3 + 1) for the GEMM, plus the partial-sum traffic
    // (workspace written, then read by the reduction kernel) at ~3 TB/s.
    long max_splits = ptiles / 4 > 1 ? ptiles / 4 : 1;
    if (max_splits > 64) max_splits = 64;
    long splits = 1;
    double best = 1e30;
    const double out_bytes = 4.0 * KH * KW * (double)pl.co_tiles * BM * (double)pl.ci_tiles * pl.BN;
    for (long sp = 1; sp <= max_splits; ++sp) {
        if (sp > 1 && (sp - 1) * ((ptiles + sp - 1) / sp) >= ptiles) continue;   // would leave an empty split
        const long units = base_units * sp;
        const long rounds = (units + imx::kNumSMs - 1) / imx::kNumSMs;
        const double ksteps = (double)((ptiles + sp - 1) / sp);
        const double cost = rounds * (ksteps * 0.3 + 1.0) + (double)sp * out_bytes / 3.0e6;
        if (cost < best - 1e-9) { best = cost; splits = sp; }
    }
    pl.splits = (int)splits;
    pl.ws_floats = splits * KH * KW * (long)pl.co_tiles * BM * (long)pl.ci_tiles * pl.BN;
    return IMX_OK;
}

template <int BN>
int launch_wgrad(const CUtensorMap& tdy, const CUtensorMap& tx, const WgradArgs& a, cudaStream_t stream) {
    using C = WCfg<BN>;
    auto kern = conv_wgrad_kernel<BN>;
    static bool attr_set[imx::kMaxDevices] = {};
    {
        cudaError_t e = imx::ensure_dyn_smem(kern, C::SMEM_BYTES, attr_set);
        if (e != cudaSuccess) return imx::set_error((int)e, "conv2d_wgrad: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    const int grid = a.total_units < num_sms() ? a.total_units : num_sms();
    kern<<<grid, kThreads, C::SMEM_BYTES, stream>>>(tdy, tx, a);
    return imx::check_launch("conv2d_wgrad");
}

int check_wgrad_args(int N, int H, int W, int Cin, int Cin_p, int Cout, int Cout_p, int KH, int KW, int stride, int pad) {
    IMX_REQUIRE(N >= 0 && H > 0 && W > 0, "conv2d_wgrad: bad input size");
    IMX_REQUIRE(Cin > 0 && Cin_p >= Cin && Cin_p % 8 == 0, "conv2d_wgrad: Cin_p must be >= Cin and a multiple of 8");
    IMX_REQUIRE(Cout > 0 && Cout_p >= Cout && Cout_p % 8 == 0, "conv2d_wgrad: Cout_p must be >= Cout and a multiple of 8");
    IMX_REQUIRE(KH > 0 && KW > 0 && KH <= 16 && KW <= 16, "conv2d_wgrad: bad kernel size");
    IMX_REQUIRE(stride >= 1 && stride <= 2 && pad >= 0, "conv2d_wgrad: stride must be 1 or 2, pad >= 0");
    return IMX_OK;
}

}  // namespace

extern "C" int64_t imx_conv2d_wgrad_workspace_bytes(int N, int H, int W, int Cin, int Cin_p,
                                                    int Cout, int Cout_p, int KH, int KW,
                                                    int stride, int pad) {
    if (check_wgrad_args(N, H, W, Cin, Cin_p, Cout, Cout_p, KH, KW, stride, pad)) return -1;
    if (N == 0) return 0;
    Plan pl;
    if (make_plan(pl, N, H, W, Cin_p, Cout_p, KH, KW, stride, pad)) return -1;
    return pl.ws_floats * 4;
}

namespace {

// The split-K GEMM: fills workspace[split][tap][co_r][ci_r] (fp32) and reports the plan.
int wgrad_gemm(const imx_bf16* x, const imx_bf16* dy, void* workspace, int64_t workspace_bytes, int N, int H, int W,
               int Cin, int Cin_p, int Cout, int Cout_p, int KH, int KW, int stride, int pad, cudaStream_t s,
               WgradArgs& a, Plan& pl) {
    IMX_REQUIRE(x && dy, "conv2d_wgrad: null pointer");
    int rc = check_wgrad_args(N, H, W, Cin, Cin_p, Cout, Cout_p, KH, KW, stride, pad);
    if (rc) return rc;
    rc = make_plan(pl, N, H, W, Cin_p, Cout_p, KH, KW, stride, pad);
    if (rc) return rc;
    IMX_REQUIRE(workspace && workspace_bytes >= pl.ws_floats * 4,
                "conv2d_wgrad: workspace too small (%lld bytes given, %lld needed)", (long long)workspace_bytes,
                (long long)pl.ws_floats * 4);
    IMX_REQUIRE(aligned16(x) && aligned16(dy) && aligned16(workspace), "conv2d_wgrad: pointers must be 16-byte aligned");

    a = WgradArgs{};
    a.N = N; a.Ho = pl.Ho; a.Wo = pl.Wo;
    a.KH = KH; a.KW = KW; a.stride = stride; a.pad = pad;
    a.tw = pl.tw; a.th = pl.th; a.tw_log2 = pl.tw_log2;
    a.tiles_x = pl.tiles_x; a.tiles_y = pl.tiles_y; a.ptiles = pl.ptiles;
    a.co_tiles = pl.co_tiles; a.ci_tiles = pl.ci_tiles; a.splits = pl.splits;
    const int T = 256 / pl.BN;
    a.total_units = pl.co_tiles * pl.ci_tiles * ((KH * KW + T - 1) / T) * pl.splits;
    a.co_r = pl.co_tiles * BM; a.ci_r = pl.ci_tiles * pl.BN;
    a.ws = reinterpret_cast<float*>(workspace);

    CUtensorMap tdy, tx;
    const uint64_t dims_dy[4] = {(uint64_t)Cout_p, (uint64_t)pl.Wo, (uint64_t)pl.Ho, (uint64_t)N};
    const uint64_t str_dy[3] = {(uint64_t)Cout_p * 2, (uint64_t)pl.Wo * Cout_p * 2, (uint64_t)pl.Ho * pl.Wo * Cout_p * 2};
    const uint32_t box_dy[4] = {64u, (uint32_t)pl.tw, (uint32_t)pl.th, 1u};
    const uint32_t es1[4] = {1u, 1u, 1u, 1u};
    rc = encode_tmap(&tdy, dy, 4, dims_dy, str_dy, box_dy, es1);
    if (rc) return rc;
    const uint64_t dims_x[4] = {(uint64_t)Cin_p, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    const uint64_t str_x[3] = {(uint64_t)Cin_p * 2, (uint64_t)W * Cin_p * 2, (uint64_t)H * W * Cin_p * 2};
    const uint32_t box_x[4] = {64u, (uint32_t)((pl.tw - 1) * stride + 1), (uint32_t)((pl.th - 1) * stride + 1), 1u};
    const uint32_t es_x[4] = {1u, (uint32_t)stride, (uint32_t)stride, 1u};
    rc = encode_tmap(&tx, x, 4, dims_x, str_x, box_x, es_x);
    if (rc) return rc;

    switch (pl.BN) {
        case 256: return launch_wgrad<256>(tdy, tx, a, s);
        case 128: return launch_wgrad<128>(tdy, tx, a, s);
        default: return launch_wgrad<64>(tdy, tx, a, s);
    }
}

}  // namespace

extern "C" int imx_conv2d_igemm_wgrad_bf16(const imx_bf16* x, const imx_bf16* dy, float* dw,
                                           void* workspace, int64_t workspace_bytes,
                                           int N, int H, int W, int Cin, int Cin_p,
                                           int Cout, int Cout_p, int KH, int KW,
                                           int stride, int pad, imx_stream_t stream) {
    IMX_REQUIRE(dw, "conv2d_wgrad: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    if (N == 0) {
        int rc = check_wgrad_args(N, H, W, Cin, Cin_p, Cout, Cout_p, KH, KW, stride, pad);
        if (rc) return rc;
        cudaError_t e = cudaMemsetAsync(dw, 0, (size_t)Cout * Cin * KH * KW * 4, s);
        return e == cudaSuccess ? IMX_OK : imx::set_error((int)e, "conv2d_wgrad: memset failed");
    }
    WgradArgs a;
    Plan pl;
    int rc = wgrad_gemm(x, dy, workspace, workspace_bytes, N, H, W, Cin, Cin_p, Cout, Cout_p, KH, KW, stride, pad, s, a, pl);
    if (rc) return rc;
    return launch_reduce(a.ws, dw, nullptr, nullptr, Cout, Cin, KH * KW, pl.splits, a.co_r, a.ci_r, s);
}

// The weight gradient plus dot += <dw, w_orig> (the inner product the spectral-norm gradient needs), formed while the
// split-K partial sums are reduced.  dot: one fp32, zeroed by the caller.
extern "C" int imx_conv2d_igemm_wgrad_dot_bf16(const imx_bf16* x, const imx_bf16* dy, float* dw, const float* w_orig,
                                               float* dot, void* workspace, int64_t workspace_bytes,
                                               int N, int H, int W, int Cin, int Cin_p,
                                               int Cout, int Cout_p, int KH, int KW,
                                               int stride, int pad, imx_stream_t stream) {
    IMX_REQUIRE(dw && w_orig && dot, "conv2d_wgrad_dot: null pointer");
    IMX_REQUIRE(N > 0, "conv2d_wgrad_dot: empty batch");
    cudaStream_t s = (cudaStream_t)stream;
    WgradArgs a;
    Plan pl;
    int rc = wgrad_gemm(x, dy, workspace, workspace_bytes, N, H, W, Cin, Cin_p, Cout, Cout_p, KH, KW, stride, pad, s, a, pl);
    if (rc) return rc;
    return launch_reduce(a.ws, dw, w_orig, dot, Cout, Cin, KH * KW, pl.splits, a.co_r, a.ci_r, s);
}

// Data-parallel variant, step 1: only the split-K GEMM.  plan_out = {splits, co_r, ci_r}.
extern "C" int imx_conv2d_igemm_wgrad_partials_bf16(const imx_bf16* x, const imx_bf16* dy,
                                                    void* workspace, int64_t workspace_bytes, int* plan_out,
                                                    int N, int H, int W, int Cin, int Cin_p,
                                                    int Cout, int Cout_p, int KH, int KW,
                                                    int stride, int pad, imx_stream_t stream) {
    IMX_REQUIRE(plan_out, "conv2d_wgrad_partials: null pointer");
    IMX_REQUIRE(N > 0, "conv2d_wgrad_partials: empty batch");
    WgradArgs a;
    Plan pl;
    int rc = wgrad_gemm(x, dy, workspace, workspace_bytes, N, H, W, Cin, Cin_p, Cout, Cout_p, KH, KW, stride, pad,
                        (cudaStream_t)stream, a, pl);
    if (rc) return rc;
    plan_out[0] = pl.splits; plan_out[1] = a.co_r; plan_out[2] = a.ci_r;
    return IMX_OK;
}

// Step 2: see wgrad_reduce_peers_kernel.  ws_peers: HOST array of `world` device pointers (this
// rank's own workspace included, any order as long as every rank passes all of them).
extern "C" int imx_wgrad_reduce_peers_f32(const void* const* ws_peers, int world, float* dw, int co_begin, int co_end,
                                          int Cout, int Cin, int taps, int splits, int co_r, int ci_r,
                                          float scale, int accumulate, imx_stream_t stream) {
    IMX_REQUIRE(ws_peers && dw, "wgrad_reduce_peers: null pointer");
    IMX_REQUIRE(world >= 1 && world <= kMaxPeers, "wgrad_reduce_peers: world size %d outside [1, %d]", world, kMaxPeers);
    IMX_REQUIRE(0 <= co_begin && co_begin <= co_end && co_end <= Cout && Cout <= co_r && Cin <= ci_r && taps > 0 && splits > 0,
                "wgrad_reduce_peers: bad geometry");
    PeerPtrs pp{};
    for (int q = 0; q < world; ++q) {
        IMX_REQUIRE(ws_peers[q], "wgrad_reduce_peers: null peer pointer %d", q);
        pp.p[q] = reinterpret_cast<const float*>(ws_peers[q]);
    }
    const long total = (long)(co_end - co_begin) * Cin * taps;
    if (total == 0) return IMX_OK;
    long blocks = (total + 255) / 256;
    if (blocks > 16L * imx::kNumSMs) blocks = 16L * imx::kNumSMs;
    wgrad_reduce_peers_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(pp, world, dw, co_begin, co_end, Cin, taps, splits,
                                                                            co_r, ci_r, scale, accumulate);
    return imx::check_launch("wgrad_reduce_peers");
}

// Step 3 (once per backward): see gather_rows_peers_kernel.  arena_peers: HOST array of `world`
// device pointers to every rank's gradient arena (same layout), descs: DEVICE array of n_tensors
// {offset, rows, row_elems} records (int64).
extern "C" int imx_gather_rows_peers_f32(const void* const* arena_peers, int world, int rank, const void* descs,
                                         int n_tensors, int64_t max_elems, imx_stream_t stream) {
    IMX_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "gather_rows_peers: bad world / rank");
    IMX_REQUIRE(n_tensors >= 0 && n_tensors <= 65535, "gather_rows_peers: bad tensor count");
    if (n_tensors == 0 || world == 1) return IMX_OK;
    IMX_REQUIRE(arena_peers && descs && max_elems > 0, "gather_rows_peers: null pointer");
    PeerPtrs pp{};
    for (int q = 0; q < world; ++q) {
        IMX_REQUIRE(arena_peers[q], "gather_rows_peers: null peer pointer %d", q);
        pp.p[q] = reinterpret_cast<const float*>(arena_peers[q]);
    }
    long blocks = (max_elems / world / 4 + 255) / 256;
    if (blocks > 8L * imx::kNumSMs) blocks = 8L * imx::kNumSMs;
    if (blocks < 1) blocks = 1;
    gather_rows_peers_kernel<<<dim3((unsigned)blocks, n_tensors), 256, 0, (cudaStream_t)stream>>>(
        pp, const_cast<float*>(pp.p[rank]), reinterpret_cast<const GatherDesc*>(descs), world, rank);
    return imx::check_launch("gather_rows_peers");
}

// ---------------------------------------------------------------------------------------------
// Weight gradient for callers that must not allocate.
//
// imx_conv2d_igemm_wgrad_bf16 needs a split-K workspace that the caller allocates.  Some hosts
// cannot do that at the point where they need the gradient: a training server that reserves all
// of its device memory once at start-up and then forbids every later cudaMalloc (a late
// allocation can fail, or stall the whole device while the driver maps memory), or a plugin
// that runs on a thread which simply has no allocator to call.  For them the workspace lives in
// this translation unit as a fixed-size device array that is part of the loaded library image:
// every GPU that loads the library carries IMX_WGRAD_NOALLOC_SCRATCH_BYTES of it until the
// process exits, and no call ever allocates or frees anything.
//
// `g_wgrad_noalloc_scratch` is a device symbol.  In host code its name is only a stand-in, not a
// device address, so the host asks the runtime for the real address once per device
// (cudaGetSymbolAddress, with that device current) and hands that pointer to the kernels.
static __device__ __align__(256) unsigned char g_wgrad_noalloc_scratch[IMX_WGRAD_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kMaxDevices] = {nullptr};

// Device address of the scratch array on the CURRENT device (looked up on first use per device).
int noalloc_scratch_for_current_device(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return imx::set_error((int)e, "conv2d_wgrad_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices)
        return imx::set_error(IMX_ERR_UNSUPPORTED, "conv2d_wgrad_noalloc: device index %d outside [0, %d)", dev, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_wgrad_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr)
            return imx::set_error((int)e, "conv2d_wgrad_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                                  cudaGetErrorString(e));
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return IMX_OK;
}

}  // namespace

extern "C" int imx_conv2d_igemm_wgrad_bf16_noalloc(const imx_bf16* x, const imx_bf16* dy, float* dw,
                                                   int N, int H, int W, int Cin, int Cin_p,
                                                   int Cout, int Cout_p, int KH, int KW,
                                                   int stride, int pad, imx_stream_t stream) {
    // Size first: a layer whose split-K partials do not fit is refused before anything runs.
    const int64_t need = imx_conv2d_wgrad_workspace_bytes(N, H, W, Cin, Cin_p, Cout, Cout_p, KH, KW, stride, pad);
    if (need < 0) return IMX_ERR_INVALID_ARG;   // message already set by the size query
    if (need > (int64_t)IMX_WGRAD_NOALLOC_SCRATCH_BYTES)
        return imx::set_error(IMX_ERR_TOO_LARGE,
                              "conv2d_wgrad_noalloc: this layer needs %lld bytes of split-K scratch but the built-in "
                              "array holds %lld; use imx_conv2d_igemm_wgrad_bf16 with a caller-owned workspace",
                              (long long)need, (long long)IMX_WGRAD_NOALLOC_SCRATCH_BYTES);
    void* scratch = nullptr;
    int rc = noalloc_scratch_for_current_device(&scratch);
    if (rc) return rc;
    // Nothing in the scratch has to start at zero: every (split, tap, Cout, Cin) slot that the
    // reduction kernel reads is overwritten by this call's GEMM units first, so whatever the
    // previous call left behind is never observed.
    return imx_conv2d_igemm_wgrad_bf16(x, dy, dw, scratch, (int64_t)IMX_WGRAD_NOALLOC_SCRATCH_BYTES, N, H, W, Cin, Cin_p,
                                       Cout, Cout_p, KH, KW, stride, pad, stream);
}
