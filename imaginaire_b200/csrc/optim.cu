// Multi-tensor Adam step (the optimizer pass over the ~456 M
// generator + discriminator parameters is the largest remaining memory-bound item of a training
// iteration).
//
// ref: get_optimizer_for_params 'adam' -> apex FusedAdam / torch.optim.Adam,
//      imaginaire/utils/trainer.py:261-306;  arithmetic of torch.optim.Adam (no amsgrad):
//        g  = grad (+ weight_decay * p)
//        m  = m + (1 - beta1) * (g - m)
//        v  = beta2 * v + (1 - beta2) * g * g
//        p -= (lr / (1 - beta1^t)) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps)
//
// One pass: 4 reads + 3 writes of fp32 per parameter (28 B) - the HBM floor of the update.  The
// tensors of a parameter group are handed over as host arrays of device pointers; they travel to
// the device in the kernel-argument space (up to kMaxTensors per launch, a table of
// (tensor, chunk) per CTA), so nothing is allocated or copied and ~500 tensors take ~12 launches.
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

// The table lives in the kernel-argument space (CUDA >= 12.1: up to 32 KB of parameters).
constexpr int kMaxTensors = 96;
constexpr int kMaxBlocks = 1184;   // 8 CTAs per SM
constexpr int kChunk = 32768;      // elements per CTA
constexpr int kThreadsOpt = 256;

struct AdamBatch {
    float* p[kMaxTensors];
    const float* g[kMaxTensors];
    float* m[kMaxTensors];
    float* v[kMaxTensors];
    long long n[kMaxTensors];
    int blk_chunk[kMaxBlocks];
    unsigned char blk_tensor[kMaxBlocks];
};

struct AdamHyper {
    float lr_over_bc1;       // lr / (1 - beta1^t)
    float inv_sqrt_bc2;      // 1 / sqrt(1 - beta2^t)
    float omb1, beta2, omb2; // 1 - beta1, beta2, 1 - beta2 (differences formed in double on the host)
    float eps, weight_decay;
};

// Step counters on the device (imx_adam_multi_tensor_f32_capturable): the bias corrections are formed in the kernel
// from the counter of the tensor, so that one captured launch stays valid as the counters advance.
struct AdamDevSteps {
    const float* step[kMaxTensors];   // null table entry: unused
    double lr, beta1, beta2;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamHyper& h) {
    if (h.weight_decay != 0.0f) g = fmaf(h.weight_decay, p, g);
    m = h.omb1 == 1.0f ? g : fmaf(h.omb1, g - m, m);     // lerp(m, g, 1 - beta1); exact for beta1 = 0
    v = fmaf(h.beta2, v, h.omb2 * g * g);
    const float denom = fmaf(sqrtf(v), h.inv_sqrt_bc2, h.eps);
    p -= h.lr_over_bc1 * (m / denom);
}

template <bool kDevStep>
__global__ void __launch_bounds__(kThreadsOpt)
adam_multi_tensor_kernel(const __grid_constant__ AdamBatch b, AdamHyper h, const __grid_constant__ AdamDevSteps ds) {
    const int t = b.blk_tensor[blockIdx.x];
    if (kDevStep) {
        // same double-precision formulas as the host path (one evaluation per thread: a few hundred cycles against
        // the 32768-element chunk the CTA streams)
        const double step = (double)__ldg(ds.step[t]);
        h.lr_over_bc1 = (float)(ds.lr / (1.0 - pow(ds.beta1, step)));
        h.inv_sqrt_bc2 = (float)(1.0 / sqrt(1.0 - pow(ds.beta2, step)));
    }
    const long long n = b.n[t];
    const long long base = (long long)b.blk_chunk[blockIdx.x] * kChunk;
    const long long end = base + kChunk < n ? base + kChunk : n;
    float* __restrict__ p = b.p[t];
    const float* __restrict__ g = b.g[t];
    float* __restrict__ m = b.m[t];
    float* __restrict__ v = b.v[t];
    const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                       reinterpret_cast<uintptr_t>(v)) & 15) == 0;
    if (vec) {
        // chunk starts are multiples of 4 elements, so 16-byte alignment carries over
        const long long end4 = base + ((end - base) & ~3LL);
        for (long long i = base + threadIdx.x * 4LL; i < end4; i += kThreadsOpt * 4LL) {
            float4 pp = *reinterpret_cast<float4*>(p + i);
            const float4 gg = __ldcs(reinterpret_cast<const float4*>(g + i));
            float4 mm = *reinterpret_cast<float4*>(m + i);
            float4 vv = *reinterpret_cast<float4*>(v + i);
            adam_one(pp.x, gg.x, mm.x, vv.x, h);
            adam_one(pp.y, gg.y, mm.y, vv.y, h);
            adam_one(pp.z, gg.z, mm.z, vv.z, h);
            adam_one(pp.w, gg.w, mm.w, vv.w, h);
            *reinterpret_cast<float4*>(p + i) = pp;
            *reinterpret_cast<float4*>(m + i) = mm;
            *reinterpret_cast<float4*>(v + i) = vv;
        }
        for (long long i = end4 + threadIdx.x; i < end; i += kThreadsOpt) adam_one(p[i], g[i], m[i], v[i], h);
    } else {
        for (long long i = base + threadIdx.x; i < end; i += kThreadsOpt) adam_one(p[i], g[i], m[i], v[i], h);
    }
}

}  // namespace

// steps == nullptr: `step` is the host-side step number; else steps[i] is tensor i's device counter.
static int adam_launch(const void* const* params, const void* const* grads, const void* const* exp_avg,
                       const void* const* exp_avg_sq, const void* const* steps, const int64_t* numel, int n_tensors,
                       double lr, double beta1, double beta2, double eps, double weight_decay, int64_t step,
                       imx_stream_t stream) {
    IMX_REQUIRE(n_tensors >= 0, "adam_multi_tensor: negative tensor count");
    if (n_tensors == 0) return IMX_OK;
    IMX_REQUIRE(params && grads && exp_avg && exp_avg_sq && numel, "adam_multi_tensor: null pointer table");
    IMX_REQUIRE(steps || step >= 1, "adam_multi_tensor: step counts from 1 (got %lld)", (long long)step);
    IMX_REQUIRE(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0 && eps >= 0.0 && lr >= 0.0,
                "adam_multi_tensor: bad hyper-parameters");
    AdamHyper h;
    const double bc1 = steps ? 1.0 : 1.0 - pow(beta1, (double)step);
    const double bc2 = steps ? 1.0 : 1.0 - pow(beta2, (double)step);
    h.lr_over_bc1 = (float)(lr / bc1);
    h.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    h.omb1 = (float)(1.0 - beta1); h.beta2 = (float)beta2; h.omb2 = (float)(1.0 - beta2);
    h.eps = (float)eps; h.weight_decay = (float)weight_decay;
    cudaStream_t s = (cudaStream_t)stream;

    AdamBatch b;
    AdamDevSteps ds;
    ds.lr = lr; ds.beta1 = beta1; ds.beta2 = beta2;
    int nt = 0, nb = 0;
    auto flush = [&]() -> int {
        if (nb == 0) return IMX_OK;
        if (steps) adam_multi_tensor_kernel<true><<<nb, kThreadsOpt, 0, s>>>(b, h, ds);
        else adam_multi_tensor_kernel<false><<<nb, kThreadsOpt, 0, s>>>(b, h, ds);
        nt = 0; nb = 0;
        return imx::check_launch("adam_multi_tensor");
    };
    for (int i = 0; i < n_tensors; ++i) {
        const long long n = numel[i];
        IMX_REQUIRE(n >= 0, "adam_multi_tensor: negative numel for tensor %d", i);
        if (n == 0) continue;
        IMX_REQUIRE(params[i] && grads[i] && exp_avg[i] && exp_avg_sq[i], "adam_multi_tensor: null tensor %d", i);
        const long long chunks = (n + kChunk - 1) / kChunk;
        long long c = 0;
        while (c < chunks) {
            if (nt == kMaxTensors || nb == kMaxBlocks) {
                int rc = flush();
                if (rc) return rc;
            }
            const int slot = nt++;
            b.p[slot] = (float*)params[i]; b.g[slot] = (const float*)grads[i];
            b.m[slot] = (float*)exp_avg[i]; b.v[slot] = (float*)exp_avg_sq[i];
            b.n[slot] = n;
            ds.step[slot] = steps ? (const float*)steps[i] : nullptr;
            if (steps) IMX_REQUIRE(steps[i] != nullptr, "adam_multi_tensor: null step counter for tensor %d", i);
            while (c < chunks && nb < kMaxBlocks) {
                b.blk_tensor[nb] = (unsigned char)slot;
                b.blk_chunk[nb] = (int)c;
                ++nb; ++c;
            }
        }
    }
    return flush();
}

extern "C" int imx_adam_multi_tensor_f32(const void* const* params, const void* const* grads,
                                         const void* const* exp_avg, const void* const* exp_avg_sq,
                                         const int64_t* numel, int n_tensors,
                                         double lr, double beta1, double beta2, double eps, double weight_decay,
                                         int64_t step, imx_stream_t stream) {
    return adam_launch(params, grads, exp_avg, exp_avg_sq, nullptr, numel, n_tensors, lr, beta1, beta2, eps, weight_decay,
                       step, stream);
}

extern "C" int imx_adam_multi_tensor_f32_capturable(const void* const* params, const void* const* grads,
                                                    const void* const* exp_avg, const void* const* exp_avg_sq,
                                                    const void* const* steps, const int64_t* numel, int n_tensors,
                                                    double lr, double beta1, double beta2, double eps,
                                                    double weight_decay, imx_stream_t stream) {
    IMX_REQUIRE(steps, "adam_multi_tensor_capturable: null step-counter table");
    return adam_launch(params, grads, exp_avg, exp_avg_sq, steps, numel, n_tensors, lr, beta1, beta2, eps, weight_decay, 0,
                       stream);
}
