// Shared helpers for the imx C-ABI kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>

#define IMX_OK 0
#define IMX_ERR_INVALID_ARG 10001
#define IMX_ERR_UNSUPPORTED 10002
#define IMX_ERR_DRIVER 10003
#define IMX_ERR_TOO_LARGE 10004

namespace imx {

// Thread-local last-error text, read through imx_last_error().
char* last_error_buf();
int set_error(int code, const char* fmt, ...);

inline int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error((int)e, "%s: %s", what, cudaGetErrorString(e));
    return IMX_OK;
}

#define IMX_REQUIRE(cond, ...)                                             \
    do {                                                                   \
        if (!(cond)) return imx::set_error(IMX_ERR_INVALID_ARG, __VA_ARGS__); \
    } while (0)

constexpr int kNumSMs = 148;

__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&p);
}

// 128-bit streaming global access (data touched once: keep it out of L1).
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream(void* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Vector reduction to global memory: one REDG.ADD.F32x4 instead of four scalar atomics (the
// accumulators of the statistics kernels are hammered by every CTA at the same time; the L2
// atomic units serialise per address, so fewer, wider operations matter).  p must be 16-byte aligned.
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Each lane holds v[0..31] (one value per column); returns in lane L the sum over all 32 lanes
// of column L.  Recursive halving: 31 shuffles instead of 32 x 5 for 32 separate butterflies.
// v is clobbered.  All indices are compile-time after unrolling, so v stays in registers.
__device__ __forceinline__ float warp_transpose_sum32(float (&v)[32], int lane) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const bool upper = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
            const float send = upper ? v[i] : v[i + off];
            const float keep = upper ? v[i + off] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    return v[0];
}

}  // namespace imx
