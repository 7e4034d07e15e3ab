// Small-vector all-reduce over NVLink peer memory, fused with the batch-norm finalize (SURVEY.md section 8(e):
// SyncBatchNorm's statistics exchange; ref torch nn/modules/_functions.py:49-165, which all-gathers
// [mean, invstd, count] forward and all-reduces [sum dy, sum dy * xhat] backward with NCCL, once per layer and pass).
//
// A C3 training iteration runs 69 such exchanges of at most 4096 floats.  Through NCCL each is a clone, a
// latency-bound collective kernel and a finalize launch on the compute stream.  Here each is ONE single-CTA kernel:
//
//   every rank owns a region of symmetric (peer-mapped) memory:
//       counter | flags[2][8] | slots[2][8][kMaxFloats]
//   call number seq (a counter in the rank's own region, advanced by the kernel itself, so a launch captured in a
//   CUDA graph stays valid) uses slot set k = seq & 1:
//     1. push: the rank stores its vector into slots[k][rank] of EVERY rank (posted NVLink stores), fences, then
//        stores seq into flags[k][rank] of every rank;
//     2. wait: it spins on its OWN flags[k][*] (local memory) until all show seq;
//     3. sum: it adds the `world` vectors of its own slots[k] in rank order -- the same order on every rank, so all
//        ranks hold bit-identical results -- and either writes the sum or goes on to mean / rstd / running statistics.
//   Two slot sets suffice: a rank can only start call seq + 2 after finishing call seq + 1, which needed every peer's
//   flag for seq + 1, and a peer raises that flag only after it has finished reading the slots of call seq.
//
// All loads and stores of peer memory are plain st / ld (no TMA, no multicast).  A wait that lasts longer than ~10 s
// of SM clocks (a peer died) traps instead of hanging the GPU.
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

constexpr int kMaxPeers = 8;
constexpr int kMaxFloats = 8192;          // 2 x 4096 channels
constexpr int kThreadsX = 512;
// region layout (bytes): [0,256) counter; [256, 768) flags; slots from 1024
constexpr size_t kFlagsOff = 256, kSlotsOff = 1024;
constexpr size_t kRegionBytes = kSlotsOff + (size_t)2 * kMaxPeers * kMaxFloats * 4;

struct Peers {
    unsigned char* base[kMaxPeers];
};

__device__ __forceinline__ unsigned* flag_ptr(unsigned char* base, int k, int r) {
    return reinterpret_cast<unsigned*>(base + kFlagsOff) + k * kMaxPeers + r;
}
__device__ __forceinline__ float* slot_ptr(unsigned char* base, int k, int r) {
    return reinterpret_cast<float*>(base + kSlotsOff) + ((size_t)k * kMaxPeers + r) * kMaxFloats;
}

// Returns with slots[k][0..world) of this rank's region complete; `k` through the reference.
__device__ void exchange(const Peers& p, int world, int rank, const float* in, int n, int& k_out) {
    __shared__ unsigned s_seq;
    unsigned char* mine = p.base[rank];
    if (threadIdx.x == 0) {
        unsigned* ctr = reinterpret_cast<unsigned*>(mine);
        s_seq = *ctr + 1u;
        *ctr = s_seq;
    }
    __syncthreads();
    const unsigned seq = s_seq;
    const int k = (int)(seq & 1u);
    // 1. push my vector to every rank (my own copy included)
    const int n4 = n >> 2;
    for (int q = 0; q < world; ++q) {
        float* dst = slot_ptr(p.base[q], k, rank);
        for (int i = threadIdx.x; i < n4; i += blockDim.x)
            reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(in)[i];   // plain loads: `out` may alias `in`
        for (int i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) dst[i] = in[i];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) {
        volatile unsigned* f = flag_ptr(p.base[threadIdx.x], k, rank);
        *f = seq;
        // 2. wait for everybody's vector
        volatile unsigned* mf = flag_ptr(mine, k, threadIdx.x);
        const long long t0 = clock64();
        while (*mf != seq) {
            if (clock64() - t0 > 20000000000LL) {
                printf("imx: peer exchange timed out waiting for rank %d (seq %u, saw %u)\n", (int)threadIdx.x, seq, *mf);
                __trap();
            }
        }
    }
    __syncthreads();
    __threadfence_system();
    k_out = k;
}

__global__ void __launch_bounds__(kThreadsX)
peer_allreduce_kernel(Peers p, int world, int rank, const float* in, float* out, int n) {
    int k;
    exchange(p, world, rank, in, n, k);
    const float* base = slot_ptr(p.base[rank], k, 0);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float acc = 0.0f;
        for (int q = 0; q < world; ++q) acc += __ldcv(base + (size_t)q * kMaxFloats + i);
        out[i] = acc;
    }
}

// in: [sum (C) | sumsq (C)] of this rank; count: GLOBAL elements per channel.  Same arithmetic as finalize_kernel
// (spade_norm.cu) on the summed vectors.
__global__ void __launch_bounds__(kThreadsX)
peer_bn_finalize_kernel(Peers p, int world, int rank, const float* __restrict__ in, int C, float count, float eps,
                        float* __restrict__ mean, float* __restrict__ rstd, float* running_mean, float* running_var,
                        float momentum) {
    int k;
    exchange(p, world, rank, in, 2 * C, k);
    const float* base = slot_ptr(p.base[rank], k, 0);
    for (int i = threadIdx.x; i < C; i += blockDim.x) {
        float s = 0.0f, ss = 0.0f;
        for (int q = 0; q < world; ++q) {
            s += __ldcv(base + (size_t)q * kMaxFloats + i);
            ss += __ldcv(base + (size_t)q * kMaxFloats + C + i);
        }
        const float m = s / count;
        const float var = fmaxf(ss / count - m * m, 0.0f);
        mean[i] = m;
        rstd[i] = rsqrtf(var + eps);
        if (running_mean) {
            const float unbiased = count > 1.0f ? var * (count / (count - 1.0f)) : var;
            running_mean[i] = (1.0f - momentum) * running_mean[i] + momentum * m;
            running_var[i] = (1.0f - momentum) * running_var[i] + momentum * unbiased;
        }
    }
}

int check(const void* const* regions, int world, int rank, int n, Peers& p) {
    IMX_REQUIRE(regions, "peer exchange: null region table");
    IMX_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "peer exchange: bad world %d / rank %d", world, rank);
    IMX_REQUIRE(n >= 0 && n <= kMaxFloats, "peer exchange: %d floats exceed the slot size %d", n, kMaxFloats);
    for (int q = 0; q < world; ++q) {
        IMX_REQUIRE(regions[q] && (reinterpret_cast<uintptr_t>(regions[q]) & 15) == 0, "peer exchange: region %d null or unaligned", q);
        p.base[q] = (unsigned char*)regions[q];
    }
    return IMX_OK;
}

}  // namespace

extern "C" int64_t imx_peer_exchange_region_bytes(void) { return (int64_t)kRegionBytes; }

extern "C" int imx_peer_allreduce_sum_f32(const void* const* regions, int world, int rank, const float* in, float* out,
                                          int n, imx_stream_t stream) {
    Peers p;
    int rc = check(regions, world, rank, n, p);
    if (rc) return rc;
    IMX_REQUIRE(in && out, "peer_allreduce_sum: null pointer");
    if (n == 0) return IMX_OK;
    peer_allreduce_kernel<<<1, kThreadsX, 0, (cudaStream_t)stream>>>(p, world, rank, in, out, n);
    return imx::check_launch("peer_allreduce_sum");
}

extern "C" int imx_peer_bn_finalize_f32(const void* const* regions, int world, int rank, const float* sums, int C,
                                        float count, float eps, float* mean, float* rstd, float* running_mean,
                                        float* running_var, float momentum, imx_stream_t stream) {
    Peers p;
    int rc = check(regions, world, rank, 2 * C, p);
    if (rc) return rc;
    IMX_REQUIRE(sums && mean && rstd, "peer_bn_finalize: null pointer");
    IMX_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "peer_bn_finalize: running stats come in pairs");
    IMX_REQUIRE(count > 0.0f, "peer_bn_finalize: count must be positive");
    if (C == 0) return IMX_OK;
    peer_bn_finalize_kernel<<<1, kThreadsX, 0, (cudaStream_t)stream>>>(p, world, rank, sums, C, count, eps, mean, rstd,
                                                                      running_mean, running_var, momentum);
    return imx::check_launch("peer_bn_finalize");
}
