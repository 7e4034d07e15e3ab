// Spectral-norm weight preparation.
//
// ref: torch.nn.utils.spectral_norm as configured by imaginaire/layers/weight_norm.py:85
//      (dim 0, one power iteration per training-mode forward, eps 1e-12) and the EMA-side
//      recompute imaginaire/utils/model_average.py:183-197.
//
// The reference's hook runs ~16 small ATen / cuBLAS launches per layer per forward (two
// GEMVs, two normalisations, clones, a third GEMV + dot for sigma, a full W / sigma
// materialisation) and autograd adds ~8 more in the backward.  Here one forward is four
// launches that read W twice, and the backward is two:
//   t = W^T u (partial sums over row slices) -> v = t / max(|t|, eps)
//   s = W v                                   -> u = s / max(|s|, eps), sigma = u . s
// 1/sigma is then folded into the bf16 repack (imx_pack_conv_weight_bf16) and
//   dL/dW = G / sigma - (<G, W> / sigma^2) u v^T          (G = dL/d(W / sigma))
// is applied in place on the fp32 weight gradient.
// HBM-bound: 2 reads of W (4 B / element each) forward; read G, W + read/write G backward.
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

constexpr int kT = 256;

// partial[s][c] = sum_{r in slice s} W[r][c] * u[r];  grid (ceil(C / (kT*4)), S)
__global__ void __launch_bounds__(kT)
sn_wt_u_kernel(const float* __restrict__ W, const float* __restrict__ u, float* __restrict__ partial,
               int R, int C, int rows_per_slice) {
    const int c0 = (blockIdx.x * kT + threadIdx.x) * 4;
    const int r0 = blockIdx.y * rows_per_slice;
    const int r1 = min(R, r0 + rows_per_slice);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (c0 + 3 < C && (C & 3) == 0) {
        const float* p = W + (size_t)r0 * C + c0;
#pragma unroll 4
        for (int r = r0; r < r1; ++r, p += C) {
            const float4 w = __ldg(reinterpret_cast<const float4*>(p));
            const float ur = __ldg(u + r);
            a0 = fmaf(w.x, ur, a0); a1 = fmaf(w.y, ur, a1); a2 = fmaf(w.z, ur, a2); a3 = fmaf(w.w, ur, a3);
        }
    } else if (c0 < C) {
        for (int r = r0; r < r1; ++r) {
            const float ur = __ldg(u + r);
            const float* p = W + (size_t)r * C + c0;
            a0 = fmaf(__ldg(p), ur, a0);
            if (c0 + 1 < C) a1 = fmaf(__ldg(p + 1), ur, a1);
            if (c0 + 2 < C) a2 = fmaf(__ldg(p + 2), ur, a2);
            if (c0 + 3 < C) a3 = fmaf(__ldg(p + 3), ur, a3);
        }
    }
    float* o = partial + (size_t)blockIdx.y * C + c0;
    if (c0 < C) o[0] = a0;
    if (c0 + 1 < C) o[1] = a1;
    if (c0 + 2 < C) o[2] = a2;
    if (c0 + 3 < C) o[3] = a3;
}

__device__ __forceinline__ float block_sum(float v, float* sh) {
    v = imx::warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) sh[w] = v;
    __syncthreads();
    float t = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.0f;
    if (w == 0) t = imx::warp_sum(t);
    if (threadIdx.x == 0) sh[0] = t;
    __syncthreads();
    t = sh[0];
    __syncthreads();
    return t;
}

// One block: t = sum of the slices, v = t / max(|t|, eps)  (written to v and v_copy)
__global__ void __launch_bounds__(1024)
sn_finish_v_kernel(const float* __restrict__ partial, float* __restrict__ v, float* __restrict__ v_copy,
                   int C, int S, float eps) {
    __shared__ float sh[32];
    float sq = 0.0f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float t = 0.0f;
        for (int s = 0; s < S; ++s) t += partial[(size_t)s * C + c];
        v[c] = t;
        sq = fmaf(t, t, sq);
    }
    const float nrm = sqrtf(block_sum(sq, sh));
    const float inv = 1.0f / fmaxf(nrm, eps);
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float t = v[c] * inv;
        v[c] = t;
        if (v_copy) v_copy[c] = t;
    }
}

// s[r] = W[r,:] . v; one warp per row; grid ceil(R / 8) blocks of 8 warps
__global__ void __launch_bounds__(kT)
sn_w_v_kernel(const float* __restrict__ W, const float* __restrict__ v, float* __restrict__ s, int R, int C) {
    const int r = blockIdx.x * (kT / 32) + (threadIdx.x >> 5);
    if (r >= R) return;
    const int lane = threadIdx.x & 31;
    const float* p = W + (size_t)r * C;
    float acc = 0.0f;
    if ((C & 3) == 0) {
        for (int c = lane * 4; c < C; c += 128) {
            const float4 w = __ldg(reinterpret_cast<const float4*>(p + c));
            const float4 x = __ldg(reinterpret_cast<const float4*>(v + c));
            acc = fmaf(w.x, x.x, acc); acc = fmaf(w.y, x.y, acc); acc = fmaf(w.z, x.z, acc); acc = fmaf(w.w, x.w, acc);
        }
    } else {
        for (int c = lane; c < C; c += 32) acc = fmaf(__ldg(p + c), __ldg(v + c), acc);
    }
    acc = imx::warp_sum(acc);
    if (lane == 0) s[r] = acc;
}

// One block.  update_u: u = s / max(|s|, eps).  sigma = u . s; writes 1/sigma, u_copy (and v_copy
// when the power iteration did not run).
__global__ void __launch_bounds__(1024)
sn_finish_u_kernel(const float* __restrict__ s, float* __restrict__ u, float* __restrict__ u_copy,
                   const float* __restrict__ v, float* __restrict__ v_copy, float* __restrict__ inv_sigma,
                   int R, int C, float eps, int update_u) {
    __shared__ float sh[32];
    if (update_u) {
        float sq = 0.0f;
        for (int r = threadIdx.x; r < R; r += blockDim.x) sq = fmaf(s[r], s[r], sq);
        const float inv = 1.0f / fmaxf(sqrtf(block_sum(sq, sh)), eps);
        for (int r = threadIdx.x; r < R; r += blockDim.x) u[r] = s[r] * inv;
        __syncthreads();
    }
    float d = 0.0f;
    for (int r = threadIdx.x; r < R; r += blockDim.x) {
        const float ur = u[r];
        d = fmaf(ur, s[r], d);
        if (u_copy) u_copy[r] = ur;
    }
    const float sigma = block_sum(d, sh);
    if (threadIdx.x == 0) *inv_sigma = 1.0f / sigma;
    if (!update_u && v_copy)
        for (int c = threadIdx.x; c < C; c += blockDim.x) v_copy[c] = v[c];
}

// dot += sum G * W   (fp32 atomics; caller zero-fills)
__global__ void __launch_bounds__(kT)
sn_dot_kernel(const float* __restrict__ G, const float* __restrict__ W, float* __restrict__ dot, long n) {
    __shared__ float sh[32];
    float acc = 0.0f;
    const long stride = (long)gridDim.x * kT * 4;
    for (long i = ((long)blockIdx.x * kT + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n && (n & 3) == 0) {
            const float4 g = __ldg(reinterpret_cast<const float4*>(G + i));
            const float4 w = __ldg(reinterpret_cast<const float4*>(W + i));
            acc = fmaf(g.x, w.x, acc); acc = fmaf(g.y, w.y, acc); acc = fmaf(g.z, w.z, acc); acc = fmaf(g.w, w.w, acc);
        } else {
            for (long j = i; j < n && j < i + 4; ++j) acc = fmaf(G[j], W[j], acc);
        }
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) atomicAdd(dot, acc);
}

// G[r][c] = G[r][c] * inv_sigma - (dot * inv_sigma^2) * u[r] * v[c]
__global__ void __launch_bounds__(kT)
sn_grad_kernel(float* __restrict__ G, const float* __restrict__ u, const float* __restrict__ v,
               const float* __restrict__ inv_sigma, const float* __restrict__ dot, int R, int C) {
    const float is = __ldg(inv_sigma);
    const float coef = __ldg(dot) * is * is;
    const long n = (long)R * C;
    for (long i = (long)blockIdx.x * kT + threadIdx.x; i < n; i += (long)gridDim.x * kT) {
        const int r = (int)(i / C), c = (int)(i - (long)r * C);
        G[i] = fmaf(G[i], is, -coef * __ldg(u + r) * __ldg(v + c));
    }
}

// The same for C % 4 == 0 (every filter but the few with 3 input channels): a thread keeps its four columns of v and
// walks down the rows, four 16-byte loads of G in flight.  sn_grad_kernel spends a 64-bit division per element and
// moves 4 bytes per instruction: 1.8 ms per iteration for 5.5 GB, a third of what the memory system allows.
__global__ void __launch_bounds__(kT)
sn_grad_rows_kernel(float* __restrict__ G, const float* __restrict__ u, const float* __restrict__ v,
                    const float* __restrict__ inv_sigma, const float* __restrict__ dot, int R, int C4) {
    const int c4 = blockIdx.x * kT + threadIdx.x;
    if (c4 >= C4) return;
    const float is = __ldg(inv_sigma);
    const float coef = -__ldg(dot) * is * is;
    // v is often a slice of the batched power-iteration output at an odd offset: four scalar loads, once per thread
    const float4 v4 = make_float4(__ldg(v + 4 * c4), __ldg(v + 4 * c4 + 1), __ldg(v + 4 * c4 + 2), __ldg(v + 4 * c4 + 3));
    float4* g = reinterpret_cast<float4*>(G) + c4;
    const int step = gridDim.y;
    int r = blockIdx.y;
    for (; r + 3 * step < R; r += 4 * step) {
        float4 a[4];
        float ur[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            a[k] = g[(size_t)(r + k * step) * C4];
            ur[k] = coef * __ldg(u + r + k * step);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            a[k].x = fmaf(a[k].x, is, ur[k] * v4.x); a[k].y = fmaf(a[k].y, is, ur[k] * v4.y);
            a[k].z = fmaf(a[k].z, is, ur[k] * v4.z); a[k].w = fmaf(a[k].w, is, ur[k] * v4.w);
            g[(size_t)(r + k * step) * C4] = a[k];
        }
    }
    for (; r < R; r += step) {
        float4 a = g[(size_t)r * C4];
        const float ur = coef * __ldg(u + r);
        a.x = fmaf(a.x, is, ur * v4.x); a.y = fmaf(a.y, is, ur * v4.y);
        a.z = fmaf(a.z, is, ur * v4.z); a.w = fmaf(a.w, is, ur * v4.w);
        g[(size_t)r * C4] = a;
    }
}

int launch_sn_grad(float* G, const float* u, const float* v, const float* inv_sigma, const float* dot, int R, int C,
                   cudaStream_t st) {
    if ((C & 3) == 0 && (reinterpret_cast<uintptr_t>(G) & 15) == 0) {
        const int C4 = C / 4;
        const int gx = (C4 + kT - 1) / kT;
        int gy = (8 * imx::kNumSMs + gx - 1) / gx;
        if (gy > R) gy = R;
        if (gy > 65535) gy = 65535;
        sn_grad_rows_kernel<<<dim3(gx, gy), kT, 0, st>>>(G, u, v, inv_sigma, dot, R, C4);
        return imx::check_launch("spectral_norm_weight_grad");
    }
    const long n = (long)R * C;
    long b2 = (n + kT - 1) / kT;
    if (b2 > 8L * imx::kNumSMs) b2 = 8L * imx::kNumSMs;
    sn_grad_kernel<<<(int)b2, kT, 0, st>>>(G, u, v, inv_sigma, dot, R, C);
    return imx::check_launch("spectral_norm_weight_grad");
}

inline int slices_for(int R, int C) {
    const int col_blocks = (C + kT * 4 - 1) / (kT * 4);
    int s = (2 * imx::kNumSMs + col_blocks - 1) / col_blocks;
    if (s > R / 8) s = R / 8;
    if (s < 1) s = 1;
    if (s > 64) s = 64;
    return s;
}

}  // namespace

extern "C" int64_t imx_spectral_norm_scratch_bytes(int R, int C) {
    if (R <= 0 || C <= 0) return -1;
    // partial sums of W^T u over the row slices, then s = W v
    return ((int64_t)slices_for(R, C) * C + R + 64) * 4;
}

extern "C" int imx_spectral_norm_sigma_f32(const float* W, float* u, float* v, float* u_copy, float* v_copy,
                                           float* inv_sigma, void* scratch, int64_t scratch_bytes,
                                           int R, int C, int n_power_iterations, float eps,
                                           imx_stream_t stream) {
    IMX_REQUIRE(R > 0 && C > 0, "spectral_norm_sigma: bad size");
    IMX_REQUIRE(W && u && v && inv_sigma && scratch, "spectral_norm_sigma: null pointer");
    IMX_REQUIRE(n_power_iterations >= 0, "spectral_norm_sigma: negative iteration count");
    IMX_REQUIRE(scratch_bytes >= imx_spectral_norm_scratch_bytes(R, C), "spectral_norm_sigma: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int S = slices_for(R, C);
    float* partial = reinterpret_cast<float*>(scratch);
    float* s = partial + (size_t)S * C;
    const int rows_per_slice = (R + S - 1) / S;
    const dim3 g1((C + kT * 4 - 1) / (kT * 4), S);
    for (int it = 0; it < n_power_iterations; ++it) {
        sn_wt_u_kernel<<<g1, kT, 0, st>>>(W, u, partial, R, C, rows_per_slice);
        sn_finish_v_kernel<<<1, 1024, 0, st>>>(partial, v, v_copy, C, S, eps);
        sn_w_v_kernel<<<(R + kT / 32 - 1) / (kT / 32), kT, 0, st>>>(W, v, s, R, C);
        const bool last = it == n_power_iterations - 1;
        sn_finish_u_kernel<<<1, 1024, 0, st>>>(s, u, last ? u_copy : nullptr, v, nullptr, inv_sigma, R, C, eps, 1);
    }
    if (n_power_iterations == 0) {
        sn_w_v_kernel<<<(R + kT / 32 - 1) / (kT / 32), kT, 0, st>>>(W, v, s, R, C);
        sn_finish_u_kernel<<<1, 1024, 0, st>>>(s, u, u_copy, v, v_copy, inv_sigma, R, C, eps, 0);
    }
    return imx::check_launch("spectral_norm_sigma");
}

extern "C" int imx_spectral_norm_weight_grad_f32(float* G, const float* W, const float* u, const float* v,
                                                 const float* inv_sigma, float* dot_scratch, int R, int C,
                                                 imx_stream_t stream) {
    IMX_REQUIRE(R > 0 && C > 0, "spectral_norm_weight_grad: bad size");
    IMX_REQUIRE(G && W && u && v && inv_sigma && dot_scratch, "spectral_norm_weight_grad: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(dot_scratch, 0, sizeof(float), st);
    if (e != cudaSuccess) return imx::set_error((int)e, "spectral_norm_weight_grad: memset: %s", cudaGetErrorString(e));
    const long n = (long)R * C;
    long blocks = (n / 4 + kT - 1) / kT;
    if (blocks > 4L * imx::kNumSMs) blocks = 4L * imx::kNumSMs;
    if (blocks < 1) blocks = 1;
    sn_dot_kernel<<<(int)blocks, kT, 0, st>>>(G, W, dot_scratch, n);
    return launch_sn_grad(G, u, v, inv_sigma, dot_scratch, R, C, st);
}

// The correction alone, for a caller that already holds dot = <G, W> (imx_conv2d_igemm_wgrad_dot_bf16 forms it while
// it reduces the split-K partial sums).
extern "C" int imx_spectral_norm_weight_grad_apply_f32(float* G, const float* u, const float* v, const float* inv_sigma,
                                                       const float* dot, int R, int C, imx_stream_t stream) {
    IMX_REQUIRE(R > 0 && C > 0, "spectral_norm_weight_grad_apply: bad size");
    IMX_REQUIRE(G && u && v && inv_sigma && dot, "spectral_norm_weight_grad_apply: null pointer");
    return launch_sn_grad(G, u, v, inv_sigma, dot, R, C, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// All spectrally normalised layers of a network in one go.  The per-layer version above is a
// chain of four small launches; a generator forward has ~90 such layers whose iterations depend
// on the filters only, so they are batched: grid.y = layer, descriptors in device memory, five
// launches per network forward in total.
namespace {

struct SnDesc {          // mirrored by imaginaire_b200/layers/conv.py (8 x int64)
    const float* W;
    float* u;
    float* v;
    long long t_off;     // offsets (floats) into the shared scratch: t[C] (zeroed by the caller), s[R]
    long long s_off;
    long long out_off;   // offset (floats) into the output buffer: [1/sigma | u copy (R) | v copy (C)]
    long long R;
    long long C;
};

__global__ void __launch_bounds__(kT)
snb_wt_u_kernel(const SnDesc* __restrict__ descs, float* __restrict__ scratch, int rows_per_slice) {
    const SnDesc d = descs[blockIdx.y];
    const int R = (int)d.R, C = (int)d.C;
    const int col_blocks = (C + kT * 4 - 1) / (kT * 4);
    const int slices = (R + rows_per_slice - 1) / rows_per_slice;
    if ((int)blockIdx.x >= col_blocks * slices) return;
    const int cb = blockIdx.x % col_blocks, sl = blockIdx.x / col_blocks;
    const int c0 = (cb * kT + threadIdx.x) * 4;
    const int r0 = sl * rows_per_slice, r1 = min(R, r0 + rows_per_slice);
    if (c0 >= C) return;
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    if (c0 + 3 < C && (C & 3) == 0) {
        const float* p = d.W + (size_t)r0 * C + c0;
#pragma unroll 4
        for (int r = r0; r < r1; ++r, p += C) {
            const float4 w = __ldg(reinterpret_cast<const float4*>(p));
            const float ur = __ldg(d.u + r);
            a[0] = fmaf(w.x, ur, a[0]); a[1] = fmaf(w.y, ur, a[1]); a[2] = fmaf(w.z, ur, a[2]); a[3] = fmaf(w.w, ur, a[3]);
        }
    } else {
        for (int r = r0; r < r1; ++r) {
            const float ur = __ldg(d.u + r);
            for (int k = 0; k < 4 && c0 + k < C; ++k) a[k] = fmaf(__ldg(d.W + (size_t)r * C + c0 + k), ur, a[k]);
        }
    }
    float* t = scratch + d.t_off + c0;
    for (int k = 0; k < 4 && c0 + k < C; ++k) atomicAdd(t + k, a[k]);
}

__global__ void __launch_bounds__(1024)
snb_finish_v_kernel(const SnDesc* __restrict__ descs, float* __restrict__ scratch, float* __restrict__ out, float eps) {
    __shared__ float sh[32];
    const SnDesc d = descs[blockIdx.x];
    const int R = (int)d.R, C = (int)d.C;
    const float* t = scratch + d.t_off;
    float sq = 0.0f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) sq = fmaf(t[c], t[c], sq);
    const float inv = 1.0f / fmaxf(sqrtf(block_sum(sq, sh)), eps);
    float* vc = out + d.out_off + 1 + R;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float x = t[c] * inv;
        d.v[c] = x;
        vc[c] = x;
    }
}

__global__ void __launch_bounds__(kT)
snb_w_v_kernel(const SnDesc* __restrict__ descs, float* __restrict__ scratch) {
    const SnDesc d = descs[blockIdx.y];
    const int R = (int)d.R, C = (int)d.C;
    const int r = blockIdx.x * (kT / 32) + (threadIdx.x >> 5);
    if (r >= R) return;
    const int lane = threadIdx.x & 31;
    const float* p = d.W + (size_t)r * C;
    float acc = 0.0f;
    if ((C & 3) == 0) {
        for (int c = lane * 4; c < C; c += 128) {
            const float4 w = __ldg(reinterpret_cast<const float4*>(p + c));
            const float4 x = *reinterpret_cast<const float4*>(d.v + c);
            acc = fmaf(w.x, x.x, acc); acc = fmaf(w.y, x.y, acc); acc = fmaf(w.z, x.z, acc); acc = fmaf(w.w, x.w, acc);
        }
    } else {
        for (int c = lane; c < C; c += 32) acc = fmaf(__ldg(p + c), d.v[c], acc);
    }
    acc = imx::warp_sum(acc);
    if (lane == 0) scratch[d.s_off + r] = acc;
}

__global__ void __launch_bounds__(1024)
snb_finish_u_kernel(const SnDesc* __restrict__ descs, const float* __restrict__ scratch, float* __restrict__ out,
                    float eps, int update_u) {
    __shared__ float sh[32];
    const SnDesc d = descs[blockIdx.x];
    const int R = (int)d.R, C = (int)d.C;
    const float* s = scratch + d.s_off;
    float* o = out + d.out_off;
    if (update_u) {
        float sq = 0.0f;
        for (int r = threadIdx.x; r < R; r += blockDim.x) sq = fmaf(s[r], s[r], sq);
        const float inv = 1.0f / fmaxf(sqrtf(block_sum(sq, sh)), eps);
        for (int r = threadIdx.x; r < R; r += blockDim.x) d.u[r] = s[r] * inv;
        __syncthreads();
    }
    float dd = 0.0f;
    for (int r = threadIdx.x; r < R; r += blockDim.x) {
        const float ur = d.u[r];
        dd = fmaf(ur, s[r], dd);
        o[1 + r] = ur;
    }
    const float sigma = block_sum(dd, sh);
    if (threadIdx.x == 0) o[0] = 1.0f / sigma;
    if (!update_u)
        for (int c = threadIdx.x; c < C; c += blockDim.x) o[1 + R + c] = d.v[c];
}

}  // namespace

namespace {

// target = beta * target + (1 - beta) * W / sigma for every described layer (grid.y = layer).
__global__ void __launch_bounds__(256)
snb_ema_kernel(const SnDesc* __restrict__ descs, const long long* __restrict__ targets,
               const float* __restrict__ out, float beta) {
    const SnDesc d = descs[blockIdx.y];
    const long n = (long)d.R * d.C;
    float* t = reinterpret_cast<float*>(targets[blockIdx.y]);
    const float k = (1.0f - beta) * out[d.out_off];      // (1 - beta) / sigma
    const long n4 = ((reinterpret_cast<uintptr_t>(d.W) | reinterpret_cast<uintptr_t>(t)) & 15) == 0 ? n / 4 : 0;
    const long stride = (long)gridDim.x * blockDim.x;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 w = __ldg(reinterpret_cast<const float4*>(d.W) + i);
        float4 a = reinterpret_cast<float4*>(t)[i];
        a.x = fmaf(beta, a.x, k * w.x); a.y = fmaf(beta, a.y, k * w.y);
        a.z = fmaf(beta, a.z, k * w.z); a.w = fmaf(beta, a.w, k * w.w);
        reinterpret_cast<float4*>(t)[i] = a;
    }
    for (long i = n4 * 4 + blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += stride)
        t[i] = fmaf(beta, t[i], k * __ldg(d.W + i));
}

}  // namespace

extern "C" int imx_spectral_norm_ema_update_f32(const void* descs, const void* targets, int n_layers, int64_t max_elems,
                                                const float* sigma_out, float beta, imx_stream_t stream) {
    IMX_REQUIRE(n_layers >= 0 && n_layers <= 65535, "spectral_norm_ema_update: bad layer count");
    if (n_layers == 0) return IMX_OK;
    IMX_REQUIRE(descs && targets && sigma_out, "spectral_norm_ema_update: null pointer");
    IMX_REQUIRE(max_elems > 0, "spectral_norm_ema_update: bad size");
    long blocks = (max_elems / 4 + 255) / 256;
    if (blocks > 2 * imx::kNumSMs) blocks = 2 * imx::kNumSMs;
    if (blocks < 1) blocks = 1;
    snb_ema_kernel<<<dim3((unsigned)blocks, n_layers), 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const SnDesc*>(descs), reinterpret_cast<const long long*>(targets), sigma_out, beta);
    return imx::check_launch("spectral_norm_ema_update");
}

extern "C" int imx_spectral_norm_sigma_batched_f32(const void* descs, int n_layers, int max_R, int max_C,
                                                   void* scratch, float* out, int n_power_iterations, float eps,
                                                   imx_stream_t stream) {
    IMX_REQUIRE(n_layers >= 0 && n_layers <= 65535, "spectral_norm_sigma_batched: bad layer count");
    if (n_layers == 0) return IMX_OK;
    IMX_REQUIRE(descs && scratch && out, "spectral_norm_sigma_batched: null pointer");
    IMX_REQUIRE(max_R > 0 && max_C > 0, "spectral_norm_sigma_batched: bad size");
    IMX_REQUIRE(n_power_iterations == 0 || n_power_iterations == 1,
                "spectral_norm_sigma_batched: 0 or 1 power iteration (the scratch is zeroed once by the caller)");
    cudaStream_t st = (cudaStream_t)stream;
    const SnDesc* d = reinterpret_cast<const SnDesc*>(descs);
    float* sc = reinterpret_cast<float*>(scratch);
    if (n_power_iterations == 1) {
        const int rows_per_slice = 64;
        const int col_blocks = (max_C + kT * 4 - 1) / (kT * 4);
        const int slices = (max_R + rows_per_slice - 1) / rows_per_slice;
        snb_wt_u_kernel<<<dim3(col_blocks * slices, n_layers), kT, 0, st>>>(d, sc, rows_per_slice);
        snb_finish_v_kernel<<<n_layers, 1024, 0, st>>>(d, sc, out, eps);
    }
    snb_w_v_kernel<<<dim3((max_R + kT / 32 - 1) / (kT / 32), n_layers), kT, 0, st>>>(d, sc);
    snb_finish_u_kernel<<<n_layers, 1024, 0, st>>>(d, sc, out, eps, n_power_iterations);
    return imx::check_launch("spectral_norm_sigma_batched");
}
