// TEST INFRASTRUCTURE ONLY -- the reference's majorant_kernel (ovr/devices/optix7/accel/sp_singlemc.cu:56-97) run ON THE GPU.
//
// Its body is cut out of the reference tree at build time (oracle/Makefile target `ref` -> oracle/_ref/majorant_kernel_extract.h,
// git-ignored, never committed) and compiled unchanged by nvcc for sm_100a; this file supplies only the environment the kernel text
// names: range1f, a DeviceTransferFunction with the four members it reads, clamp(), and the launch of
// SpacePartiton_SingleMC::compute_majorant (sp_singlemc.cu:123-138: one block of N threads per N cells, N = length of the opacity
// table, N floats of dynamic shared memory). It cannot be pinned on the host: `uint32_t i_lower = floorf(...) - 1` converts -1.f to
// an unsigned integer, which is undefined in C++ and resolved differently by x86 (wraps) and the GPU (cvt.rzi.u32.f32 saturates to 0);
// the device's behaviour is the reference's. Known defect of the kernel text: threads with i >= count return before
// they have loaded their entry of the opacity table, so cells of the LAST block see uninitialised table entries unless the cell
// count is a multiple of N -- the pin uses such grids only.
//
// Only tests/ load this library (tests/test_zz_ref_majorant_gpu.py); nothing in the product does.
#include <cuda_runtime.h>

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>

namespace ovr {
namespace optix7 {

struct range1f {
  float lower, upper;
};
struct vec2f_ {
  float x, y;
};
struct dims_ {
  long long n;
  __host__ __device__ long long long_product() const { return n; }
};
struct opacity_ {
  dims_ dims;
  void* rawptr;
};
struct DeviceTransferFunction {
  opacity_ opacity;
  vec2f_ value_range;
  float range_rcp_norm; // == 1.f / value_range.span()  (array.h:58, volume.cpp:153)
};

template<typename T>
__host__ __device__ inline T
clamp(T v, T lo, T hi) // gdt/math: min(max(v, lo), hi)
{
  return v < lo ? lo : (v > hi ? hi : v);
}
__host__ __device__ inline float
clamp(float v, float lo, float hi)
{
  return fminf(fmaxf(v, lo), hi);
}

__global__ void
#include "_ref/majorant_kernel_extract.h"

} // namespace optix7
} // namespace ovr

// ranges: n_cells (lower, upper) pairs, alphas: n_alpha floats -- DEVICE pointers; majorants: n_cells floats, device.
// Returns 0, or a CUDA error code. The launch configuration is compute_majorant's.
extern "C" int
ref_majorant_kernel_run(const float* ranges_dev, uint32_t n_cells, const float* alphas_dev, uint32_t n_alpha, float range_lo,
                        float range_hi, float* majorants_dev)
{
  using namespace ovr::optix7;
  if (n_alpha == 0 || n_alpha > 1024) return -1; // one thread per table entry: the reference cannot launch more either
  DeviceTransferFunction tfn;
  tfn.opacity.dims.n = n_alpha;
  tfn.opacity.rawptr = const_cast<float*>(alphas_dev);
  tfn.value_range.x = range_lo;
  tfn.value_range.y = range_hi;
  tfn.range_rcp_norm = 1.f / (range_hi - range_lo);
  const uint32_t N = n_alpha;
  majorant_kernel<<<(n_cells + N - 1) / N, N, N * sizeof(float)>>>(n_cells, tfn, (const range1f*)ranges_dev, majorants_dev);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  return (int)e;
}
