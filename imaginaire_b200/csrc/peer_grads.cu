// All-reduce (average) of gradient segments between the GPUs of one node through symmetric (peer-mapped) memory:
// a reduce-scatter and an all-gather over NVLink, written as two streaming kernels.
//
// ref: the gradient averaging of the reference's DistributedDataParallel wrapper
// (imaginaire/utils/trainer.py:193-216): bucketed NCCL all-reduce after / during the backward.
//
// Every rank has written its LOCAL gradients (PyTorch layout, fp32) into the same offsets of its own arena of
// symmetric memory.  A segment = one gradient tensor (offset, numel); rank q owns the q-th of `world` contiguous
// chunks of every segment (boundaries rounded to 4 floats).
//   reduce-scatter   mine[chunk(rank)] = scale * sum_q arena_q[chunk(rank)]      (float4 loads from every peer,
//                                                                                 summed in rank order)
//   all-gather       mine[chunk(q)]    = arena_q[chunk(q)]   for q != rank
// with a device-side barrier of all ranks before, between and after (the caller's: symmetric-memory barrier).
// Per rank and byte of gradient: (world-1)/world bytes read remotely in each phase -- the volume of a ring
// all-reduce -- as fully coalesced 16-byte loads, by kernels that use the whole GPU for a few milliseconds at the
// end of the backward instead of sharing the SMs with the persistent convolution kernels.
// All accesses to peer memory are plain ld / st (no TMA, no multicast).
#include "imx_common.cuh"
#include "../../include/imx.h"

namespace {

constexpr int kMaxPeers = 8;
struct PeerArena { float* p[kMaxPeers]; };
struct Seg { long long offset, numel; };     // mirrored by imaginaire_b200/grad_sync.py (2 x int64)

__device__ __forceinline__ long long chunk_begin(long long numel, int q, int world) {
    return q >= world ? numel : ((numel * q / world) & ~3LL);
}

__global__ void __launch_bounds__(256)
peer_reduce_scatter_kernel(PeerArena a, const Seg* __restrict__ segs, int world, int rank, float scale) {
    const Seg s = segs[blockIdx.y];
    const long long lo = chunk_begin(s.numel, rank, world), hi = chunk_begin(s.numel, rank + 1, world);
    const long long n = hi - lo;
    if (n <= 0) return;
    float* mine = a.p[rank] + s.offset + lo;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long n4 = n >> 2;                       // lo is a multiple of 4 and offsets are 256-byte aligned
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < kMaxPeers; ++q) {
            if (q < world) {
                const float4 v = __ldcs(reinterpret_cast<const float4*>(a.p[q] + s.offset + lo) + i);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
        }
        acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
        reinterpret_cast<float4*>(mine)[i] = acc;
    }
    for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
        float acc = 0.f;
        for (int q = 0; q < world; ++q) acc += __ldcs(a.p[q] + s.offset + lo + i);
        mine[i] = acc * scale;
    }
}

__global__ void __launch_bounds__(256)
peer_all_gather_kernel(PeerArena a, const Seg* __restrict__ segs, int world, int rank) {
    const Seg s = segs[blockIdx.y];
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (int d = 1; d < world; ++d) {
        const int q = (rank + d) % world;               // ranks start on different peers
        const long long lo = chunk_begin(s.numel, q, world), hi = chunk_begin(s.numel, q + 1, world);
        const long long n = hi - lo;
        if (n <= 0) continue;
        const float* src = a.p[q] + s.offset + lo;
        float* dst = a.p[rank] + s.offset + lo;
        const long long n4 = n >> 2;
        for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride)
            reinterpret_cast<float4*>(dst)[i] = __ldcs(reinterpret_cast<const float4*>(src) + i);
        for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride)
            dst[i] = __ldcs(src + i);
    }
}

int fill(const void* const* arena_peers, int world, int rank, const void* segs, int n_seg, PeerArena& a) {
    IMX_REQUIRE(arena_peers && segs, "peer gradient all-reduce: null pointer");
    IMX_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "peer gradient all-reduce: bad world %d / rank %d", world, rank);
    IMX_REQUIRE(n_seg >= 0 && n_seg <= 65535, "peer gradient all-reduce: bad segment count %d", n_seg);
    for (int q = 0; q < world; ++q) {
        IMX_REQUIRE(arena_peers[q] && (reinterpret_cast<uintptr_t>(arena_peers[q]) & 15) == 0,
                    "peer gradient all-reduce: arena %d null or unaligned", q);
        a.p[q] = (float*)arena_peers[q];
    }
    return IMX_OK;
}

int blocks_for(long long max_numel, int world) {
    long long b = (max_numel / world / 4 + 255) / 256;
    if (b > 4LL * imx::kNumSMs) b = 4LL * imx::kNumSMs;
    return b < 1 ? 1 : (int)b;
}

}  // namespace

extern "C" int imx_peer_reduce_scatter_f32(const void* const* arena_peers, int world, int rank, const void* segs,
                                           int n_seg, int64_t max_numel, float scale, imx_stream_t stream) {
    PeerArena a;
    int rc = fill(arena_peers, world, rank, segs, n_seg, a);
    if (rc) return rc;
    if (n_seg == 0) return IMX_OK;
    dim3 grid(blocks_for(max_numel, world), n_seg);
    peer_reduce_scatter_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a, (const Seg*)segs, world, rank, scale);
    return imx::check_launch("peer_reduce_scatter");
}

extern "C" int imx_peer_all_gather_f32(const void* const* arena_peers, int world, int rank, const void* segs, int n_seg,
                                       int64_t max_numel, imx_stream_t stream) {
    PeerArena a;
    int rc = fill(arena_peers, world, rank, segs, n_seg, a);
    if (rc) return rc;
    if (n_seg == 0 || world == 1) return IMX_OK;
    dim3 grid(blocks_for(max_numel, world), n_seg);
    peer_all_gather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a, (const Seg*)segs, world, rank);
    return imx::check_launch("peer_all_gather");
}
