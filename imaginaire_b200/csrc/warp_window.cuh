// Forward flow warps through a shared-memory window (warp_window.cu); called by the entry points in warp.cu.
#pragma once
#include <cuda_runtime.h>

namespace imx {
namespace warpwin {

// True when the windowed kernels can and should take the launch (16-byte aligned rows, enough work and channels).
bool applicable(const float* img, int B, int C, int H, int W, int Hi, int Wi);

int resample2d_fwd(const float* img, const float* flow, float* out, int B, int C, int H, int W, int Hi, int Wi,
                   cudaStream_t stream);
int gridsample_fwd(const float* img, const float* flow, float* out, int B, int C, int H, int W, cudaStream_t stream);

}  // namespace warpwin
}  // namespace imx
