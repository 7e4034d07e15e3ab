// Host-side TMA descriptor (CUtensorMap) encoding through the driver entry point, so the
// library has no link-time dependency on libcuda.
#pragma once
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <mutex>
#include "imx_common.cuh"

namespace imx {

inline PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(f);
    });
    return fn;
}

inline int encode_tmap(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box, const uint32_t* estrides) {
    auto fn = get_encode_fn();
    if (!fn) return imx::set_error(IMX_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not available");
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims, strides_bytes,
                    box, estrides, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return imx::set_error(IMX_ERR_DRIVER, "cuTensorMapEncodeTiled failed (CUresult %d; rank %d dims %llu %llu %llu %llu box %u %u %u %u)",
                              (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1],
                              (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0),
                              box[0], box[1], rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return IMX_OK;
}

constexpr int kMaxDevices = 64;

inline int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < kMaxDevices) ? dev : 0;
}

// SM count of the CURRENT device (cached per device: one process may drive several GPUs).
inline int num_sms() {
    static int n[kMaxDevices] = {};
    const int dev = current_device();
    if (n[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = imx::kNumSMs;
        n[dev] = v;
    }
    return n[dev];
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-device setting: `done` is one flag array per kernel
// instantiation, indexed by the current device.
template <typename K>
inline cudaError_t ensure_dyn_smem(K kern, int bytes, bool (&done)[kMaxDevices]) {
    const int dev = current_device();
    if (done[dev]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done[dev] = true;
    return e;
}


inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace imx
