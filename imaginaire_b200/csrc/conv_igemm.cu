// Implicit-GEMM convolution on the 5th-generation tensor cores (SURVEY.md section 8
// rows a4, a6): tcgen05.mma with fp32 accumulators in TMEM, operands staged by TMA.
//
// ref: torch.nn.Conv2d inside Conv2dBlock  imaginaire/layers/conv.py:93-104, 262-327
//      SPADE gamma/beta projection + modulation  imaginaire/layers/activation_norm.py:226-233
//
// GEMM view:  D[pixel, co] = sum_{kh,kw,ci} A[pixel @ tap(kh,kw), ci] * B[co, (kh,kw,ci)]
//   M tile = 128 output pixels = a (th x tw) spatial patch of one image (tw * th = 128),
//   N tile = BN output channels (32..256),  K step = one filter tap x 64 input channels.
// The A tile of a tap is ONE 4-D TMA box {64 ch, tw, th, 1} of the NHWC activation at the
// tap's spatial offset: the zero padding (and any ragged edge) is TMA out-of-bounds fill,
// so no im2col buffer and no halo logic exist.  Stride-2 convolutions use the tensor map's
// element strides.  The B tile is a {64, 1, BN} box of the packed filter
// [Cout][KH*KW][Cin_p].  Both land in shared memory in the 128-byte-swizzled K-major
// layout that tcgen05.mma consumes directly.
//
// Warp roles (256 threads, persistent CTAs, one per SM):
//   warp 0   TMA producer (one lane)            smem ring: full[]/empty[] mbarriers
//   warp 1   MMA issuer (one lane)              tcgen05.commit -> empty[] / tmem_full[]
//   warp 2   TMEM allocator
//   warps 4-7 epilogue: tcgen05.ld -> bias/act/residual/(SPADE modulation) -> global
// TMEM holds two accumulator stages (2 x BN columns) so the epilogue of tile i overlaps
// the main loop of tile i+1.
#include "imx_common.cuh"
#include "imx_ptx.cuh"
#include "../../include/imx.h"

#include "imx_tma_host.cuh"

namespace {

using namespace imx::ptx;
using imx::encode_tmap;
using imx::num_sms;

constexpr int BM = 128;          // pixels per tile
constexpr int BK = 64;           // channels per K step (128 bytes of bf16: one swizzle row)
constexpr int UMMA_K = 16;
constexpr int kThreads = 256;
constexpr int kEpiWarp0 = 4;     // first epilogue warp (warp % 4 == TMEM lane quarter)
constexpr int A_TILE_BYTES = BM * BK * 2;

// MT = 128-pixel sub-tiles per CTA tile.  With a narrow channel tile (BN <= 128) one MMA per
// K step would need more shared-memory fill (A + B bytes per MMA cycle) than L2 sustains, so
// such layers use MT = 2: one 256-pixel TMA box, two MMAs that share the B tile, two TMEM
// accumulators per stage.
template <int BN, int MT> struct Cfg {
    static constexpr int A_BYTES = MT * A_TILE_BYTES;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (196608 / STAGE_BYTES) > 8 ? 8 : (196608 / STAGE_BYTES);
    static constexpr int TMEM_COLS = (2 * MT * BN < 32) ? 32 : 2 * MT * BN;
    static constexpr int STATS_SMEM_CH = 2048;   // per-CTA statistics accumulators for up to this many channels
    static constexpr int EPI_SMEM_BYTES = 6144;  // per-tile epilogue coefficients, two stages
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/ +
                                      2 * STATS_SMEM_CH * 4 + EPI_SMEM_BYTES;
};

struct ConvArgs {
    // geometry
    int N, H, W, Ho, Wo;
    int Cout, Cout_p;
    int KH, KW, stride, pad_y, pad_x;
    // where output pixel (oy, ox) goes: row oy * oys + oy0, column ox * oxs + ox0 of an out_H x out_W tensor (the whole
    // tensor unless the launch fills one parity class of a stride-2 data gradient)
    int out_H, out_W, oys, oxs, oy0, ox0;
    int tw_log2, tw, th;
    int tiles_x, tiles_y, tiles_n, total_tiles;
    int kc_blocks;
    int splits, k_per_split;     // split-K: work item = (tile, split); each split owns k_per_split K steps
    float* ws;                   // split-K: fp32 [pixels][Cout_p] partial sums (zero-filled by the host code)
    int flags;
    // plain epilogue
    const float* scale;          // optional: one fp32 on the device, the accumulator is multiplied by it before the
                                 // bias (1 / sigma of a spectrally normalised layer whose bf16 filter is packed unscaled)
    const float* bias;
    const uint16_t* residual;
    void* out;
    float* stats;
    // SPADE-fused epilogue (mode 1)
    const float* scale_beta;     // mode 1: `scale` multiplies the gamma accumulators, `scale_beta` the beta ones
    const float* bias_beta;
    const uint16_t* x;
    const float* mean;
    const float* rstd;
    const float* nw;
    const float* nb;
    int C, per_instance, up2x;
    float act_slope;             // activation as its negative-side slope: LeakyReLU 0.2, ReLU 0, none 1
};

// named barrier of the four epilogue warps
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

__device__ __forceinline__ float act_apply(float v, float slope) {
    return fmaxf(v, 0.0f) + slope * fminf(v, 0.0f);   // branch-free for all three activations
}

// kMode 0: plain conv epilogue.   kMode 1: SPADE gamma/beta projection + modulation.
template <int BN, int kMode, int MT>
__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const ConvArgs p) {
    using C = Cfg<BN, MT>;
    constexpr int A_BYTES = C::A_BYTES;
    extern __shared__ uint8_t smem_raw[];
    // 128-byte swizzle atoms must sit on 1024-byte boundaries.
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + C::STAGES * A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + C::STAGES;
    uint64_t* tmem_full = bars + 2 * C::STAGES;
    uint64_t* tmem_empty = bars + 2 * C::STAGES + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * C::STAGES + 4);
    // Batch statistics are accumulated per CTA in shared memory over all of its tiles and flushed
    // once at the end: 2*Cout global atomics per CTA instead of per tile (the per-tile version
    // serialises on the few L2 slices that hold the 2*Cout accumulators).
    float* s_stats = reinterpret_cast<float*>(smem + C::STAGES * C::STAGE_BYTES + 256);
    float* s_epi = s_stats + 2 * C::STATS_SMEM_CH;
    const bool stats_in_smem = kMode == 0 && (p.flags & IMX_EPI_STATS) && !(p.flags & IMX_EPI_STATS_INSTANCE) &&
                               p.Cout <= C::STATS_SMEM_CH;
    if (stats_in_smem)
        for (int i = threadIdx.x; i < 2 * p.Cout; i += kThreads) s_stats[i] = 0.0f;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmap_a);
        prefetch_tmap(&tmap_b);
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

    const int taps = p.KH * p.KW;
    const int k_iters = taps * p.kc_blocks;
    const int tiles_per_img = p.tiles_x * p.tiles_y;

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int stage = 0;
            uint32_t phase = 0;
            for (int item = blockIdx.x; item < p.total_tiles * p.splits; item += gridDim.x) {
                const int tile = item / p.splits, split = item - tile * p.splits;
                const int n_tile = tile % p.tiles_n;
                const int m_tile = tile / p.tiles_n;
                const int img = m_tile / tiles_per_img;
                const int rem = m_tile - img * tiles_per_img;
                const int ty = rem / p.tiles_x, tx = rem - ty * p.tiles_x;
                const int iy0 = ty * p.th * p.stride - p.pad_y;
                const int ix0 = tx * p.tw * p.stride - p.pad_x;
                const int n0 = n_tile * BN;
                const int kb = split * p.k_per_split;
                const int ke = kb + p.k_per_split < k_iters ? kb + p.k_per_split : k_iters;
                int tap = kb / p.kc_blocks;
                int kc = kb - tap * p.kc_blocks;
                int kh = tap / p.KW, kw = tap - kh * p.KW;
                for (int k = kb; k < ke; ++k) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_arrive_expect_tx(&full_bar[stage], C::STAGE_BYTES);
                    tma_load_4d(smem_a + stage * A_BYTES, &tmap_a, &full_bar[stage], kc * BK, ix0 + kw, iy0 + kh, img);
                    tma_load_3d(smem_b + stage * C::B_BYTES, &tmap_b, &full_bar[stage], kc * BK, tap, n0);
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                    if (++kc == p.kc_blocks) {
                        kc = 0; ++tap;
                        if (++kw == p.KW) { kw = 0; ++kh; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            constexpr uint32_t idesc = make_idesc_bf16_f32(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int iter = 0;
            for (int item = blockIdx.x; item < p.total_tiles * p.splits; item += gridDim.x, ++iter) {
                const int split = item % p.splits;
                const int kb = split * p.k_per_split;
                const int ke = kb + p.k_per_split < k_iters ? kb + p.k_per_split : k_iters;
                const int as = iter & 1;
                const uint32_t aphase = (iter >> 1) & 1;
                mbar_wait(&tmem_empty[as], aphase ^ 1);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + as * (MT * BN);
                for (int k = kb; k < ke; ++k) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after_sync();
                    const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + stage * C::B_BYTES));
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) {
                        const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + stage * A_BYTES + mt * A_TILE_BYTES));
#pragma unroll
                        for (int j = 0; j < BK / UMMA_K; ++j) {
                            // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in (addr >> 4)
                            umma_bf16_ss(d_tmem + mt * BN, da + 2 * j, db + 2 * j, idesc, (k > kb) || (j != 0));
                        }
                    }
                    umma_commit(&empty_bar[stage]);
                    if (k == ke - 1) umma_commit(&tmem_full[as]);
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ===== epilogue =====
        const int q = warp & 3;                 // TMEM lane quarter owned by this warp
        const int row = q * 32 + lane;          // row of a 128-pixel sub-tile
        int iter = 0;
        for (int item = blockIdx.x; item < p.total_tiles * p.splits; item += gridDim.x, ++iter) {
            const int tile = item / p.splits;
            const int as = iter & 1;
            const uint32_t aphase = (iter >> 1) & 1;
            const int n_tile = tile % p.tiles_n;
            const int m_tile = tile / p.tiles_n;
            const int img = m_tile / tiles_per_img;
            const int rem = m_tile - img * tiles_per_img;
            const int ty = rem / p.tiles_x, tx = rem - ty * p.tiles_x;
            const int n0 = n_tile * BN;

            if (kMode == 0 && p.splits > 1) {
                // split-K: add the raw fp32 partial sums into the workspace; bias / activation /
                // residual are applied by the finishing kernel once every split has landed.
                mbar_wait(&tmem_full[as], aphase);
                tc_fence_after_sync();
#pragma unroll 1
                for (int mt = 0; mt < MT; ++mt) {
                    const int prow = mt * BM + row;
                    const int pr = prow >> p.tw_log2, pc = prow & (p.tw - 1);
                    const int oy = ty * p.th + pr, ox = tx * p.tw + pc;
                    const bool valid = (oy < p.Ho) && (ox < p.Wo);
                    const size_t pix = ((size_t)img * p.out_H + oy * p.oys + p.oy0) * p.out_W + ox * p.oxs + p.ox0;
                    const uint32_t t_acc = tmem_base + as * (MT * BN) + mt * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
                    for (int ch = 0; ch < BN / 32; ++ch) {
                        uint32_t v[32];
                        tmem_ld_32x32(t_acc + ch * 32, v);
                        tmem_ld_wait();
                        if (ch == BN / 32 - 1 && mt == MT - 1) {
                            tc_fence_before_sync();
                            mbar_arrive(&tmem_empty[as]);
                        }
                        const int nb = n0 + ch * 32;
                        if (!valid || nb >= p.Cout_p) continue;
                        float* dst = p.ws + pix * p.Cout_p + nb;
#pragma unroll
                        for (int g = 0; g < 8; ++g)
                            if (nb + g * 4 < p.Cout_p)
                                imx::red_add_v4(dst + g * 4, __uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]),
                                                __uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3]));
                    }
                }
                continue;
            }
            // Per-channel epilogue coefficients of this tile go to shared memory once (double
            // buffered by accumulator stage), so the per-element work reads them as broadcast
            // LDS instead of 32 dependent global loads per chunk.
            if constexpr (kMode == 0) {
                if (p.flags & IMX_EPI_BIAS) {
                    for (int i = threadIdx.x - kEpiWarp0 * 32; i < BN; i += 128)
                        s_epi[as * BN + i] = (n0 + i < p.Cout) ? __ldg(p.bias + n0 + i) : 0.0f;
                    epi_bar_sync();
                }
            } else {
                // channels c0 + [0, 64) of every 128-column group: {mean, rstd*w, b, 1 + bias_gamma}, bias_beta
                float4* cf = reinterpret_cast<float4*>(s_epi) + as * (BN / 2);
                float* cbb = s_epi + 4 * BN + as * (BN / 2);
                for (int i = threadIdx.x - kEpiWarp0 * 32; i < BN / 2; i += 128) {
                    const int c = (n0 / 128 + (i >> 6)) * 64 + (i & 63);
                    float4 v = make_float4(0.f, 0.f, 0.f, 1.f);
                    float bb = 0.0f;
                    if (c < p.C) {
                        const int so = (p.per_instance ? img * p.C : 0) + c;
                        v.x = __ldg(p.mean + so);
                        v.y = __ldg(p.rstd + so) * (p.nw ? __ldg(p.nw + c) : 1.0f);
                        v.z = p.nb ? __ldg(p.nb + c) : 0.0f;
                        v.w = 1.0f + (p.bias ? __ldg(p.bias + c) : 0.0f);
                        bb = p.bias_beta ? __ldg(p.bias_beta + c) : 0.0f;
                    }
                    cf[i] = v;
                    cbb[i] = bb;
                }
                epi_bar_sync();
            }

            bool waited = false;
#pragma unroll 1
            for (int mt = 0; mt < MT; ++mt) {
            const int prow = mt * BM + row;         // pixel index inside the CTA tile
            const int pr = prow >> p.tw_log2, pc = prow & (p.tw - 1);
            const int oy = ty * p.th + pr, ox = tx * p.tw + pc;
            const bool valid = (oy < p.Ho) && (ox < p.Wo);
            const size_t pix = kMode == 0 ? ((size_t)img * p.out_H + oy * p.oys + p.oy0) * p.out_W + ox * p.oxs + p.ox0
                                         : ((size_t)img * p.Ho + oy) * p.Wo + ox;
            const uint32_t t_acc = tmem_base + as * (MT * BN) + mt * BN + (static_cast<uint32_t>(q * 32) << 16);

            if constexpr (kMode == 0) {
                constexpr int NCH = BN / 32;
                const float acc_scale = p.scale ? __ldg(p.scale) : 1.0f;
                const bool has_res = (p.flags & IMX_EPI_RESIDUAL) != 0;
                const uint4* rbase = reinterpret_cast<const uint4*>(p.residual + pix * p.Cout_p + n0);
                // residual of chunk ch: four 16-byte pieces of this pixel's row (issued two chunks ahead)
                auto load_res = [&](int ch, uint4 (&r)[4]) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        r[g] = make_uint4(0u, 0u, 0u, 0u);
                        if (has_res && valid && ch < NCH && n0 + ch * 32 + g * 8 < p.Cout_p) r[g] = __ldg(rbase + ch * 4 + g);
                    }
                };
                auto process = [&](int ch, const uint32_t (&v)[32], const uint4 (&r)[4]) {
                    const int nb = n0 + ch * 32;
                    if (nb >= p.Cout_p) return;   // warp-uniform
                    float f[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * acc_scale;
                    if (p.flags & IMX_EPI_BIAS) {
                        const float4* sb = reinterpret_cast<const float4*>(s_epi + as * BN + ch * 32);
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            const float4 b4 = sb[g];
                            f[4 * g] += b4.x; f[4 * g + 1] += b4.y; f[4 * g + 2] += b4.z; f[4 * g + 3] += b4.w;
                        }
                    }
                    if (has_res) {
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            f[g * 8 + 0] += imx::bf16_lo(r[g].x); f[g * 8 + 1] += imx::bf16_hi(r[g].x);
                            f[g * 8 + 2] += imx::bf16_lo(r[g].y); f[g * 8 + 3] += imx::bf16_hi(r[g].y);
                            f[g * 8 + 4] += imx::bf16_lo(r[g].z); f[g * 8 + 5] += imx::bf16_hi(r[g].z);
                            f[g * 8 + 6] += imx::bf16_lo(r[g].w); f[g * 8 + 7] += imx::bf16_hi(r[g].w);
                        }
                    }
                    if (p.flags & IMX_EPI_LRELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = f[j] > 0.0f ? f[j] : 0.2f * f[j];
                    }
                    if (p.flags & IMX_EPI_RELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
                    }
                    if (p.flags & IMX_EPI_TANH) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = tanhf(f[j]);
                    }
                    if (p.flags & IMX_EPI_OUT_F32) {
                        if (valid) {
                            float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + pix * p.Cout_p + nb);
#pragma unroll
                            for (int g = 0; g < 8; ++g)
                                if (nb + g * 4 < p.Cout_p) o[g] = make_float4(f[g * 4], f[g * 4 + 1], f[g * 4 + 2], f[g * 4 + 3]);
                        }
                    } else {
                        uint32_t o[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) o[j] = imx::pack_bf16(f[2 * j], f[2 * j + 1]);
                        if (valid) {
                            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out) + pix * p.Cout_p + nb);
#pragma unroll
                            for (int g = 0; g < 4; ++g)
                                if (nb + g * 8 < p.Cout_p) dst[g] = make_uint4(o[4 * g], o[4 * g + 1], o[4 * g + 2], o[4 * g + 3]);
                        }
                        if (p.flags & IMX_EPI_STATS) {
                            // statistics of the values as stored (bf16-rounded), so a later
                            // normalisation of the stored tensor is self-consistent
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                f[2 * j] = valid ? imx::bf16_lo(o[j]) : 0.0f;
                                f[2 * j + 1] = valid ? imx::bf16_hi(o[j]) : 0.0f;
                            }
                        }
                    }
                    if (p.flags & IMX_EPI_STATS) {
                        if (p.flags & IMX_EPI_OUT_F32) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) f[j] = valid ? f[j] : 0.0f;
                        }
                        // Column sums over the warp's 32 rows: lane L ends up with channel nb + L.
                        float q2[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) q2[j] = f[j] * f[j];
                        const float s_mine = imx::warp_transpose_sum32(f, lane);
                        const float q_mine = imx::warp_transpose_sum32(q2, lane);
                        if (stats_in_smem) {
                            if (nb + lane < p.Cout) {
                                atomicAdd(s_stats + nb + lane, s_mine);
                                atomicAdd(s_stats + p.Cout + nb + lane, q_mine);
                            }
                        } else if (nb + lane < p.Cout) {
                            // layout (2, G, Cout): G = 1 (batch statistics) or N (instance statistics)
                            const bool inst = (p.flags & IMX_EPI_STATS_INSTANCE) != 0;
                            float* sdst = p.stats + (inst ? (size_t)img * p.Cout : 0) + nb + lane;
                            atomicAdd(sdst, s_mine);
                            atomicAdd(sdst + (size_t)(inst ? p.N : 1) * p.Cout, q_mine);
                        }
                    }
                };
                // Software pipeline over the 32-column chunks: the TMEM load of chunk ch+1 and the
                // residual loads of chunk ch+2 are in flight while chunk ch is processed.
                uint32_t va[32], vb[32];
                uint4 ra[4], rb[4];
                load_res(0, ra);
                load_res(1, rb);
                if (!waited) {
                    mbar_wait(&tmem_full[as], aphase);
                    tc_fence_after_sync();
                    waited = true;
                }
                tmem_ld_32x32(t_acc, va);
#pragma unroll 1
                for (int ch = 0; ch < NCH; ch += 2) {
                    tmem_ld_wait();
                    if (ch + 1 < NCH) tmem_ld_32x32(t_acc + (ch + 1) * 32, vb);
                    else if (mt == MT - 1) { tc_fence_before_sync(); mbar_arrive(&tmem_empty[as]); }
                    process(ch, va, ra);
                    load_res(ch + 2, ra);
                    if (ch + 1 < NCH) {
                        tmem_ld_wait();
                        if (ch + 2 < NCH) tmem_ld_32x32(t_acc + (ch + 2) * 32, va);
                        else if (mt == MT - 1) { tc_fence_before_sync(); mbar_arrive(&tmem_empty[as]); }
                        process(ch + 1, vb, rb);
                        load_res(ch + 3, rb);
                    }
                }
            } else {
                // ---- SPADE: columns [g*128, g*128+64) = gamma, [g*128+64, g*128+128) = beta of
                //      channels c0 + [0, 64);  y = act(((x-mean)*rstd*w+b) * (1+gamma) + beta)
                const int oyx = p.up2x ? (oy >> 1) : oy, oxx = p.up2x ? (ox >> 1) : ox;
                const int Hx = p.up2x ? (p.Ho >> 1) : p.Ho, Wx = p.up2x ? (p.Wo >> 1) : p.Wo;
                const size_t xpix = ((size_t)img * Hx + oyx) * Wx + oxx;
                const float4* cf = reinterpret_cast<const float4*>(s_epi) + as * (BN / 2);
                const float* cbb = s_epi + 4 * BN + as * (BN / 2);
                const float sc_g = p.scale ? __ldg(p.scale) : 1.0f, sc_b = p.scale_beta ? __ldg(p.scale_beta) : 1.0f;
                constexpr int NCH = BN / 64;     // chunks of 32 channels (32 gamma + 32 beta columns)
                auto chan0 = [&](int cc) { return (n0 / 128 + (cc >> 1)) * 64 + (cc & 1) * 32; };
                auto load_x = [&](int cc, uint4 (&xr)[4]) {
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        xr[t] = make_uint4(0u, 0u, 0u, 0u);
                        if (valid && cc < NCH && chan0(cc) < p.C)
                            xr[t] = __ldg(reinterpret_cast<const uint4*>(p.x + xpix * p.C + chan0(cc)) + t);
                    }
                };
                uint4 xa[4], xb[4];
                load_x(0, xa);
                load_x(1, xb);
                if (!waited) {
                    mbar_wait(&tmem_full[as], aphase);
                    tc_fence_after_sync();
                    waited = true;
                }
                auto process = [&](int cc, const uint4 (&xr4)[4]) {
                    const int g = cc >> 1, h = cc & 1;
                    uint32_t vg[32], vb[32];
                    tmem_ld_32x32(t_acc + g * 128 + h * 32, vg);
                    tmem_ld_32x32(t_acc + g * 128 + 64 + h * 32, vb);
                    tmem_ld_wait();
                    if (cc == NCH - 1 && mt == MT - 1) {
                        tc_fence_before_sync();
                        mbar_arrive(&tmem_empty[as]);
                    }
                    const int cb = chan0(cc);
                    if (cb >= p.C) return;   // warp-uniform (C % 32 == 0 enforced on the host)
                    const uint32_t xr[16] = {xr4[0].x, xr4[0].y, xr4[0].z, xr4[0].w, xr4[1].x, xr4[1].y, xr4[1].z, xr4[1].w,
                                             xr4[2].x, xr4[2].y, xr4[2].z, xr4[2].w, xr4[3].x, xr4[3].y, xr4[3].z, xr4[3].w};
                    uint32_t o[16];
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        float y2[2];
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int c = j + e;
                            const float xv = e == 0 ? imx::bf16_lo(xr[j >> 1]) : imx::bf16_hi(xr[j >> 1]);
                            const float4 k4 = cf[g * 64 + h * 32 + c];
                            const float xh = fmaf(xv - k4.x, k4.y, k4.z);
                            const float gam1 = fmaf(__uint_as_float(vg[c]), sc_g, k4.w);
                            const float bet = fmaf(__uint_as_float(vb[c]), sc_b, cbb[g * 64 + h * 32 + c]);
                            y2[e] = act_apply(fmaf(xh, gam1, bet), p.act_slope);
                        }
                        o[j >> 1] = imx::pack_bf16(y2[0], y2[1]);
                    }
                    if (valid) {
                        uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out) + pix * p.C + cb);
#pragma unroll
                        for (int t = 0; t < 4; ++t) dst[t] = make_uint4(o[4 * t], o[4 * t + 1], o[4 * t + 2], o[4 * t + 3]);
                    }
                };
#pragma unroll 1
                for (int cc = 0; cc < NCH; cc += 2) {
                    process(cc, xa);
                    load_x(cc + 2, xa);
                    process(cc + 1, xb);
                    load_x(cc + 3, xb);
                }
            }
            }   // mt
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (stats_in_smem) {
        if ((p.Cout & 3) == 0) {
            for (int i = threadIdx.x * 4; i < 2 * p.Cout; i += kThreads * 4)
                imx::red_add_v4(p.stats + i, s_stats[i], s_stats[i + 1], s_stats[i + 2], s_stats[i + 3]);
        } else {
            for (int i = threadIdx.x; i < 2 * p.Cout; i += kThreads) atomicAdd(p.stats + i, s_stats[i]);
        }
    }
    if (warp == 2) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

// ------------------------------------------------------------------ host side

template <int BN, int kMode, int MT = 1>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const ConvArgs& a, cudaStream_t stream) {
    using C = Cfg<BN, MT>;
    static bool attr_set[imx::kMaxDevices] = {};   // per (BN, mode, MT) instantiation and device
    auto kern = conv_igemm_kernel<BN, kMode, MT>;
    {
        cudaError_t e = imx::ensure_dyn_smem(kern, C::SMEM_BYTES, attr_set);
        if (e != cudaSuccess) return imx::set_error((int)e, "conv_igemm: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    }
    const long items = (long)a.total_tiles * a.splits;
    const int grid = items < num_sms() ? (int)items : num_sms();
    kern<<<grid, kThreads, C::SMEM_BYTES, stream>>>(ta, tb, a);
    return imx::check_launch("conv_igemm");
}

// Common geometry + tensor maps.  `rows_b` = number of rows of the packed filter matrix.
struct OutWindow { int pad_y, pad_x, Ho, Wo, out_H, out_W, oys, oxs, oy0, ox0; };

int setup(ConvArgs& a, CUtensorMap& ta, CUtensorMap& tb, const void* x, const void* wpk, int N, int H, int W,
          int Cin_p, int rows_b, int KH, int KW, int stride, int pad, int BN, int MT = 1, const OutWindow* win = nullptr) {
    a.N = N; a.H = H; a.W = W;
    a.KH = KH; a.KW = KW; a.stride = stride; a.pad_y = pad; a.pad_x = pad;
    a.Ho = (H + 2 * pad - KH) / stride + 1;
    a.Wo = (W + 2 * pad - KW) / stride + 1;
    if (win) {           // explicit output size and per-axis left padding (rows / columns past the input read zeros)
        a.pad_y = win->pad_y; a.pad_x = win->pad_x;
        a.Ho = win->Ho; a.Wo = win->Wo;
    }
    a.out_H = win ? win->out_H : a.Ho; a.out_W = win ? win->out_W : a.Wo;
    a.oys = win ? win->oys : 1; a.oxs = win ? win->oxs : 1;
    a.oy0 = win ? win->oy0 : 0; a.ox0 = win ? win->ox0 : 0;
    if (a.Ho <= 0 || a.Wo <= 0) return imx::set_error(IMX_ERR_INVALID_ARG, "conv_igemm: empty output (%d x %d)", a.Ho, a.Wo);
    // pixel tile: tw = largest power of two <= min(Wo, 128) (at least 8 so th <= 16), th = 128 / tw
    int tw = 128;
    while (tw > 8 && tw > a.Wo) tw >>= 1;
    a.tw = tw; a.th = MT * BM / tw;
    a.tw_log2 = 0;
    while ((1 << a.tw_log2) < tw) ++a.tw_log2;
    a.tiles_x = (a.Wo + a.tw - 1) / a.tw;
    a.tiles_y = (a.Ho + a.th - 1) / a.th;
    a.tiles_n = (rows_b + BN - 1) / BN;
    const long total = (long)N * a.tiles_x * a.tiles_y * a.tiles_n;
    if (total > 0x7fffffffL) return imx::set_error(IMX_ERR_TOO_LARGE, "conv_igemm: too many tiles");
    a.total_tiles = (int)total;
    a.kc_blocks = (Cin_p + BK - 1) / BK;
    a.splits = 1;
    a.k_per_split = KH * KW * a.kc_blocks;
    a.ws = nullptr;

    const uint64_t dims_a[4] = {(uint64_t)Cin_p, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    const uint64_t str_a[3] = {(uint64_t)Cin_p * 2, (uint64_t)W * Cin_p * 2, (uint64_t)H * W * Cin_p * 2};
    const uint32_t box_a[4] = {(uint32_t)BK, (uint32_t)((a.tw - 1) * stride + 1), (uint32_t)((a.th - 1) * stride + 1), 1u};
    const uint32_t es_a[4] = {1u, (uint32_t)stride, (uint32_t)stride, 1u};
    if (box_a[1] > 256 || box_a[2] > 256) return imx::set_error(IMX_ERR_UNSUPPORTED, "conv_igemm: TMA box too large for stride %d", stride);
    int rc = encode_tmap(&ta, x, 4, dims_a, str_a, box_a, es_a);
    if (rc) return rc;
    const uint64_t dims_b[3] = {(uint64_t)Cin_p, (uint64_t)(KH * KW), (uint64_t)rows_b};
    const uint64_t str_b[2] = {(uint64_t)Cin_p * 2, (uint64_t)KH * KW * Cin_p * 2};
    const uint32_t box_b[3] = {(uint32_t)BK, 1u, (uint32_t)BN};
    const uint32_t es_b[3] = {1u, 1u, 1u};
    return encode_tmap(&tb, wpk, 3, dims_b, str_b, box_b, es_b);
}

using imx::aligned16;

}  // namespace

namespace {

// Applies bias / residual / activation to the split-K sums and stores the output tensor.
__global__ void splitk_finish_kernel(const float* __restrict__ ws, const float* __restrict__ scale,
                                     const float* __restrict__ bias,
                                     const uint16_t* __restrict__ residual, void* __restrict__ out,
                                     long pixels, int Cout, int Cout_p, int flags) {
    const int c4n = Cout_p / 4;
    const long total = pixels * c4n;
    const float sc = scale ? __ldg(scale) : 1.0f;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % c4n) * 4;
        const long pix = i / c4n;
        const float4 v4 = __ldg(reinterpret_cast<const float4*>(ws) + i);
        float f[4] = {v4.x * sc, v4.y * sc, v4.z * sc, v4.w * sc};
        if (flags & IMX_EPI_BIAS) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (c + j < Cout) f[j] += __ldg(bias + c + j);
        }
        if (flags & IMX_EPI_RESIDUAL) {
            const uint2 r = __ldg(reinterpret_cast<const uint2*>(residual + pix * Cout_p + c));
            f[0] += imx::bf16_lo(r.x); f[1] += imx::bf16_hi(r.x);
            f[2] += imx::bf16_lo(r.y); f[3] += imx::bf16_hi(r.y);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (flags & IMX_EPI_LRELU) f[j] = f[j] > 0.0f ? f[j] : 0.2f * f[j];
            if (flags & IMX_EPI_RELU) f[j] = fmaxf(f[j], 0.0f);
            if (flags & IMX_EPI_TANH) f[j] = tanhf(f[j]);
        }
        if (flags & IMX_EPI_OUT_F32)
            reinterpret_cast<float4*>(out)[i] = make_float4(f[0], f[1], f[2], f[3]);
        else
            reinterpret_cast<uint2*>(out)[i] = make_uint2(imx::pack_bf16(f[0], f[1]), imx::pack_bf16(f[2], f[3]));
    }
}

struct FpropPlan { int BN, MT, splits, k_per_split; long ws_bytes; };

// Tile shape and split-K factor from a small cost model, in units of one K step of a full-rate
// tile (~0.3 us).  Narrow channel tiles (BN <= 128) with 128-pixel tiles are operand-bandwidth
// bound (1.7x slower per pixel), 256-pixel tiles are not.  Layers with fewer tiles than SMs split
// the K loop (taps x channel blocks) over several CTAs, which add fp32 partial sums into a
// workspace; that costs a memset and a finishing kernel (~15 units).
void plan_fprop(FpropPlan& pl, int N, int H, int W, int Cin_p, int Cout, int Cout_p, int KH, int KW, int stride, int pad,
                int flags, bool allow_split, const OutWindow* win = nullptr) {
    pl.BN = Cout_p > 128 ? 256 : (Cout_p > 64 ? 128 : (Cout_p > 32 ? 64 : 32));
    const int Ho = win ? win->Ho : (H + 2 * pad - KH) / stride + 1, Wo = win ? win->Wo : (W + 2 * pad - KW) / stride + 1;
    if (win) allow_split = false;      // the finishing kernel addresses a dense output
    const long k_iters = (long)KH * KW * ((Cin_p + BK - 1) / BK);
    const int tiles_n = (Cout + pl.BN - 1) / pl.BN;
    const bool mt2_ok = pl.BN <= 128 &&
                        (stride == 1 || Wo >= 128 || ((256 / (Wo < 128 ? Wo : 128)) - 1) * stride + 1 <= 256);
    if (flags & IMX_EPI_STATS) allow_split = false;
    double best = 1e30;
    pl.MT = 1; pl.splits = 1;
    for (int mt = 1; mt <= (mt2_ok ? 2 : 1); ++mt) {
        int tw = 128;
        while (tw > 8 && tw > Wo) tw >>= 1;
        const int th = mt * BM / tw;
        const long tiles = (long)N * ((Wo + tw - 1) / tw) * ((Ho + th - 1) / th) * tiles_n;
        // K-step time relative to a 128 x 256 tile: the larger of the MMA time and the operand
        // fill time (A: mt x 16 KB, B: BN / 8 KB against 48 KB); measured 0.85 for BN = 128, MT = 1.
        const double t_mma = mt * pl.BN / 256.0, t_fill = (mt * 16.0 + pl.BN / 8.0) / 48.0 * 1.27;
        const double t_step = t_mma > t_fill ? t_mma : t_fill;
        const double ws_units = (double)N * Ho * Wo * Cout_p * 4.0 / 2.0e6 / 0.3;   // one pass over the workspace
        for (int sp = 1; sp <= (allow_split && tiles < imx::kNumSMs ? 16 : 1); ++sp) {
            const long per = (k_iters + sp - 1) / sp;
            if (sp > 1 && (per < 8 || (long)(sp - 1) * per >= k_iters)) continue;
            const long rounds = (tiles * sp + imx::kNumSMs - 1) / imx::kNumSMs;
            const double cost = rounds * (per * t_step + 2.0) + (sp > 1 ? 15.0 + ws_units * (sp + 2) : 0.0);
            if (cost < best - 1e-9) { best = cost; pl.MT = mt; pl.splits = sp; }
        }
    }
    pl.k_per_split = (int)((k_iters + pl.splits - 1) / pl.splits);
    pl.ws_bytes = pl.splits > 1 ? (long)N * Ho * Wo * Cout_p * 4 : 0;
}

int check_fprop_args(int N, int H, int W, int Cin_p, int Cout, int Cout_p, int KH, int KW, int stride, int pad) {
    IMX_REQUIRE(N >= 0 && H > 0 && W > 0, "conv2d_fprop: bad input size");
    IMX_REQUIRE(Cin_p > 0 && Cin_p % 8 == 0, "conv2d_fprop: Cin_p (%d) must be a positive multiple of 8", Cin_p);
    IMX_REQUIRE(Cout > 0 && Cout_p >= Cout && Cout_p % 8 == 0, "conv2d_fprop: Cout_p (%d) must be >= Cout (%d) and a multiple of 8", Cout_p, Cout);
    IMX_REQUIRE(KH > 0 && KW > 0 && KH <= 16 && KW <= 16, "conv2d_fprop: bad kernel size");
    IMX_REQUIRE(stride >= 1 && stride <= 2 && pad >= 0, "conv2d_fprop: stride must be 1 or 2, pad >= 0");
    return IMX_OK;
}

int fprop_impl(const imx_bf16* x, const imx_bf16* wpk, const float* scale, const float* bias, const imx_bf16* residual,
               void* out, float* stats, void* workspace, int64_t workspace_bytes, int N, int H, int W, int Cin_p, int Cout,
               int Cout_p, int KH, int KW, int stride, int pad, int flags, imx_stream_t stream,
               const OutWindow* win = nullptr) {
    IMX_REQUIRE(x && wpk && out, "conv2d_fprop: null pointer");
    int rc = check_fprop_args(N, H, W, Cin_p, Cout, Cout_p, KH, KW, stride, pad);
    if (rc) return rc;
    IMX_REQUIRE(!(flags & IMX_EPI_BIAS) || bias, "conv2d_fprop: bias flag without bias");
    IMX_REQUIRE(!(flags & IMX_EPI_RESIDUAL) || residual, "conv2d_fprop: residual flag without residual");
    IMX_REQUIRE(!(flags & IMX_EPI_STATS) || stats, "conv2d_fprop: stats flag without stats buffer");
    IMX_REQUIRE(aligned16(x) && aligned16(wpk) && aligned16(out) && aligned16(residual), "conv2d_fprop: pointers must be 16-byte aligned");
    if (N == 0) return IMX_OK;
    FpropPlan pl;
    plan_fprop(pl, N, H, W, Cin_p, Cout, Cout_p, KH, KW, stride, pad, flags, workspace != nullptr, win);
    if (pl.splits > 1) {
        IMX_REQUIRE(workspace_bytes >= pl.ws_bytes && aligned16(workspace),
                    "conv2d_fprop: split-K workspace too small (%lld bytes given, %lld needed) or misaligned",
                    (long long)workspace_bytes, (long long)pl.ws_bytes);
    }
    const int BN = pl.BN, MT = pl.MT;
    ConvArgs a{};
    CUtensorMap ta, tb;
    rc = setup(a, ta, tb, x, wpk, N, H, W, Cin_p, Cout, KH, KW, stride, pad, BN, MT, win);
    if (rc) return rc;
    a.Cout = Cout; a.Cout_p = Cout_p; a.flags = flags;
    a.scale = scale; a.bias = bias; a.residual = residual; a.out = out; a.stats = stats;
    a.splits = pl.splits; a.k_per_split = pl.k_per_split; a.ws = reinterpret_cast<float*>(workspace);
    cudaStream_t s = (cudaStream_t)stream;
    if (pl.splits > 1) {
        cudaError_t e = cudaMemsetAsync(workspace, 0, (size_t)pl.ws_bytes, s);
        if (e != cudaSuccess) return imx::set_error((int)e, "conv2d_fprop: workspace memset: %s", cudaGetErrorString(e));
    }
    if (MT == 2) {
        switch (BN) {
            case 128: rc = launch<128, 0, 2>(ta, tb, a, s); break;
            case 64: rc = launch<64, 0, 2>(ta, tb, a, s); break;
            default: rc = launch<32, 0, 2>(ta, tb, a, s); break;
        }
    } else {
        switch (BN) {
            case 256: rc = launch<256, 0>(ta, tb, a, s); break;
            case 128: rc = launch<128, 0>(ta, tb, a, s); break;
            case 64: rc = launch<64, 0>(ta, tb, a, s); break;
            default: rc = launch<32, 0>(ta, tb, a, s); break;
        }
    }
    if (rc || pl.splits == 1) return rc;
    const long pixels = (long)N * a.Ho * a.Wo;
    long blocks = (pixels * (Cout_p / 4) + 255) / 256;
    if (blocks > 8L * imx::kNumSMs) blocks = 8L * imx::kNumSMs;
    splitk_finish_kernel<<<(int)blocks, 256, 0, s>>>(a.ws, scale, bias, reinterpret_cast<const uint16_t*>(residual), out, pixels,
                                                     Cout, Cout_p, flags);
    return imx::check_launch("conv2d_fprop_splitk_finish");
}

}  // namespace

extern "C" int imx_conv2d_igemm_fprop_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* bias,
                                           const imx_bf16* residual, void* out, float* stats,
                                           int N, int H, int W, int Cin_p,
                                           int Cout, int Cout_p, int KH, int KW,
                                           int stride, int pad, int flags, imx_stream_t stream) {
    return fprop_impl(x, wpk, nullptr, bias, residual, out, stats, nullptr, 0, N, H, W, Cin_p, Cout, Cout_p, KH, KW, stride,
                      pad, flags, stream);
}

// y = act(scale * conv(x, wpk) + bias) (+ residual): `scale` is ONE fp32 in device memory (e.g. 1 / sigma of a
// spectrally normalised layer), so that the bf16 filter can be packed once per optimizer step instead of once per
// application with its current sigma folded in.  workspace may be null (no split-K).
extern "C" int imx_conv2d_igemm_fprop_scaled_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* scale,
                                                  const float* bias, const imx_bf16* residual, void* out, float* stats,
                                                  void* workspace, int64_t workspace_bytes,
                                                  int N, int H, int W, int Cin_p,
                                                  int Cout, int Cout_p, int KH, int KW,
                                                  int stride, int pad, int flags, imx_stream_t stream) {
    return fprop_impl(x, wpk, scale, bias, residual, out, stats, workspace, workspace_bytes, N, H, W, Cin_p, Cout, Cout_p,
                      KH, KW, stride, pad, flags, stream);
}

// One piece of a larger output: y[n][oy * oys + oy0][ox * oxs + ox0][:] = act(scale * sum_taps x[n][oy + kh - pad_y]
// [ox + kw - pad_x][:] . w[:][kh][kw][:] + bias) for oy < Ho, ox < Wo, stride 1, rows / columns outside the input read
// as zeros on every side.  The data gradient of a stride-2 convolution is four such launches, one per parity class of
// the input pixel, each with the taps of that parity (imaginaire_b200/functional.py: conv2d_dgrad) -- a quarter of the
// multiply-adds of the zero-insertion form and no zero-filled copy of dy.
extern "C" int imx_conv2d_igemm_fprop_window_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* scale,
                                                  const float* bias, const imx_bf16* residual, void* out,
                                                  int N, int H, int W, int Cin_p, int Cout, int Cout_p, int KH, int KW,
                                                  int pad_y, int pad_x, int Ho, int Wo, int out_H, int out_W,
                                                  int out_stride_y, int out_stride_x, int out_y0, int out_x0,
                                                  int flags, imx_stream_t stream) {
    IMX_REQUIRE(pad_y >= 0 && pad_x >= 0 && Ho > 0 && Wo > 0, "conv2d_fprop_window: bad padding or output size");
    IMX_REQUIRE(out_stride_y >= 1 && out_stride_x >= 1 && out_y0 >= 0 && out_x0 >= 0 &&
                (Ho - 1) * out_stride_y + out_y0 < out_H && (Wo - 1) * out_stride_x + out_x0 < out_W,
                "conv2d_fprop_window: the window leaves the %d x %d output", out_H, out_W);
    IMX_REQUIRE(!(flags & IMX_EPI_STATS), "conv2d_fprop_window: no statistics epilogue");
    const OutWindow win{pad_y, pad_x, Ho, Wo, out_H, out_W, out_stride_y, out_stride_x, out_y0, out_x0};
    return fprop_impl(x, wpk, scale, bias, residual, out, nullptr, nullptr, 0, N, H, W, Cin_p, Cout, Cout_p, KH, KW, 1,
                      pad_y > pad_x ? pad_y : pad_x, flags, stream, &win);
}

extern "C" int64_t imx_conv2d_fprop_workspace_bytes(int N, int H, int W, int Cin_p, int Cout, int Cout_p,
                                                    int KH, int KW, int stride, int pad, int flags) {
    if (check_fprop_args(N, H, W, Cin_p, Cout, Cout_p, KH, KW, stride, pad)) return -1;
    if (N == 0) return 0;
    FpropPlan pl;
    plan_fprop(pl, N, H, W, Cin_p, Cout, Cout_p, KH, KW, stride, pad, flags, true);
    return pl.ws_bytes;
}

extern "C" int imx_conv2d_igemm_fprop_splitk_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* bias,
                                                  const imx_bf16* residual, void* out, float* stats,
                                                  void* workspace, int64_t workspace_bytes,
                                                  int N, int H, int W, int Cin_p,
                                                  int Cout, int Cout_p, int KH, int KW,
                                                  int stride, int pad, int flags, imx_stream_t stream) {
    return fprop_impl(x, wpk, nullptr, bias, residual, out, stats, workspace, workspace_bytes, N, H, W, Cin_p, Cout, Cout_p,
                      KH, KW, stride, pad, flags, stream);
}

static int spade_gb_impl(const imx_bf16* hidden, const imx_bf16* wgb, const float* scale_gamma, const float* scale_beta,
                         const float* bias_gamma, const float* bias_beta,
                         const imx_bf16* x, const float* mean, const float* rstd,
                         const float* w, const float* b, imx_bf16* y,
                         int N, int H, int W, int Ch_p, int C, int KH, int KW, int pad,
                         int stats_per_instance, int act, int up2x,
                         imx_stream_t stream) {
    IMX_REQUIRE(hidden && wgb && x && mean && rstd && y, "spade_gb_conv_modulate: null pointer");
    IMX_REQUIRE(N >= 0 && H > 0 && W > 0, "spade_gb_conv_modulate: bad size");
    IMX_REQUIRE(Ch_p > 0 && Ch_p % 8 == 0, "spade_gb_conv_modulate: Ch_p must be a positive multiple of 8");
    IMX_REQUIRE(C > 0 && C % 32 == 0, "spade_gb_conv_modulate: C (%d) must be a multiple of 32", C);
    IMX_REQUIRE(KH > 0 && KW > 0 && KH <= 16 && KW <= 16 && pad >= 0, "spade_gb_conv_modulate: bad kernel size");
    IMX_REQUIRE(2 * pad == KH - 1 && 2 * pad == KW - 1, "spade_gb_conv_modulate: needs a 'same' convolution (pad = (K-1)/2)");
    IMX_REQUIRE(act >= 0 && act <= 2, "spade_gb_conv_modulate: bad act");
    IMX_REQUIRE(!up2x || (H % 2 == 0 && W % 2 == 0), "spade_gb_conv_modulate: up2x needs even H, W");
    IMX_REQUIRE((w != nullptr) || (b == nullptr), "spade_gb_conv_modulate: norm bias without norm weight");
    IMX_REQUIRE(aligned16(hidden) && aligned16(wgb) && aligned16(x) && aligned16(y), "spade_gb_conv_modulate: pointers must be 16-byte aligned");
    if (N == 0) return IMX_OK;
    const int rows_b = 2 * ((C + 63) / 64 * 64);
    const int BN = rows_b >= 256 ? 256 : 128;
    ConvArgs a{};
    CUtensorMap ta, tb;
    int rc = setup(a, ta, tb, hidden, wgb, N, H, W, Ch_p, rows_b, KH, KW, 1, pad, BN);
    if (rc) return rc;
    a.out = y; a.bias = bias_gamma; a.bias_beta = bias_beta;
    a.scale = scale_gamma; a.scale_beta = scale_beta;
    a.x = x; a.mean = mean; a.rstd = rstd; a.nw = w; a.nb = b;
    a.C = C; a.per_instance = stats_per_instance; a.up2x = up2x;
    a.act_slope = act == 1 ? 0.2f : (act == 2 ? 0.0f : 1.0f);
    cudaStream_t s = (cudaStream_t)stream;
    return BN == 256 ? launch<256, 1>(ta, tb, a, s) : launch<128, 1>(ta, tb, a, s);
}

extern "C" int imx_spade_gb_conv_modulate_bf16(const imx_bf16* hidden, const imx_bf16* wgb,
                                               const float* bias_gamma, const float* bias_beta,
                                               const imx_bf16* x, const float* mean, const float* rstd,
                                               const float* w, const float* b, imx_bf16* y,
                                               int N, int H, int W, int Ch_p, int C, int KH, int KW, int pad,
                                               int stats_per_instance, int act, int up2x,
                                               imx_stream_t stream) {
    return spade_gb_impl(hidden, wgb, nullptr, nullptr, bias_gamma, bias_beta, x, mean, rstd, w, b, y, N, H, W, Ch_p, C, KH,
                         KW, pad, stats_per_instance, act, up2x, stream);
}

// The same with the gamma accumulators multiplied by *scale_gamma and the beta accumulators by *scale_beta (one fp32
// each in device memory, either may be null) before their biases: 1 / sigma of the two spectrally normalised
// projections when the interleaved filter is packed unscaled, once per optimizer step.
extern "C" int imx_spade_gb_conv_modulate_scaled_bf16(const imx_bf16* hidden, const imx_bf16* wgb,
                                                      const float* scale_gamma, const float* scale_beta,
                                                      const float* bias_gamma, const float* bias_beta,
                                                      const imx_bf16* x, const float* mean, const float* rstd,
                                                      const float* w, const float* b, imx_bf16* y,
                                                      int N, int H, int W, int Ch_p, int C, int KH, int KW, int pad,
                                                      int stats_per_instance, int act, int up2x,
                                                      imx_stream_t stream) {
    return spade_gb_impl(hidden, wgb, scale_gamma, scale_beta, bias_gamma, bias_beta, x, mean, rstd, w, b, y, N, H, W, Ch_p, C,
                         KH, KW, pad, stats_per_instance, act, up2x, stream);
}
