// Sparse (foveated) sampling: which pixels a frame renders. Replaces generate_sparse_sampling_mask_d
// (ovr/common/generate_mask.cu:55-124: Bernoulli mask with a Gaussian fall-off around the focus point, thresholded against a
// tiled spatio-temporal noise table, ovr/common/random/blue_noise.h:81-115) and the thrust::remove stream compaction that
// follows it (generate_mask.cu:89-97), consumed by compute_screen_position's "method 3" (shaders_common.h:437-449).
//
// B200 design: the surviving pixels are not listed in row-major order but in the order the render kernels walk the image
// (64x32-pixel super tiles, 8x4-pixel warp tiles inside them), so that the 32 consecutive list entries a warp takes are
// neighbouring pixels and their rays stay coherent in the texture cache. The SET of pixels is the reference's. Three small
// kernels (count per 1024-entry block, one-CTA scan, scatter with a ballot prefix) replace thrust; nothing is sorted.
#pragma once
#include "launch_params.cuh"

namespace b200k {

constexpr int SPARSE_BLOCK = 1024; // enumeration entries per CTA = 32 warp tiles

struct SparseParams {
  const float* noise;   // [tile_xy][tile_xy][tile_t] floats: (y % tile_xy, x % tile_xy, frame % tile_t), blue_noise.h:95-99
  int tile_xy, tile_t;
  int width, height;
  int super_w;          // image width in super tiles
  int n_entries;        // super tiles * 2048
  int frame_index;
  float mean_x, mean_y; // focus centre in [0, 1]^2
  float sigma_rcp2;     // 1 / focus_scale^2
  float base_noise;
};

// e-th entry of the tile-ordered enumeration of the image -> pixel; false for the padding beyond the image edge
__device__ __forceinline__ bool sparse_entry_pixel(const SparseParams& S, int e, int& x, int& y)
{
  const int s = e / (WARP_TILES_PER_SUPER * 32), r = e % (WARP_TILES_PER_SUPER * 32);
  const int t = r >> 5, lane = r & 31;
  const int b = t / TILE_CHUNK, i = t % TILE_CHUNK; // same walk as warp_tile_origin
  x = (s % S.super_w) * SUPER_TILE_W + ((b % SUPER_BLOCKS_X) * 4 + (i & 3)) * WARP_TILE_W + (lane & (WARP_TILE_W - 1));
  y = (s / S.super_w) * SUPER_TILE_H + ((b / SUPER_BLOCKS_X) * 4 + (i >> 2)) * WARP_TILE_H + lane / WARP_TILE_W;
  return x < S.width && y < S.height;
}

// generate_sparse_samples (generate_mask.cu:55-86), same arithmetic: IEEE divisions, __expf
__device__ __forceinline__ bool sparse_keep(const SparseParams& S, int x, int y)
{
  const float aspect = __fdiv_rn((float)S.width, (float)S.height);
  const float fx = __fdiv_rn((float)x, (float)S.width) - S.mean_x;
  const float fy = __fdiv_rn(__fdiv_rn((float)y, (float)S.height) - S.mean_y, aspect);
  const float p = (1.0f - S.base_noise) * __expf(-0.5f * (fx * fx + fy * fy) * S.sigma_rcp2) + S.base_noise;
  const float d = S.noise[((size_t)(y % S.tile_xy) * S.tile_xy + (size_t)(x % S.tile_xy)) * S.tile_t + (size_t)(S.frame_index % S.tile_t)];
  return d < p;
}

__global__ void __launch_bounds__(SPARSE_BLOCK) sparse_count_kernel(const SparseParams S, unsigned int* __restrict__ block_counts)
{
  const int e = blockIdx.x * SPARSE_BLOCK + threadIdx.x;
  int x, y;
  const bool keep = e < S.n_entries && sparse_entry_pixel(S, e, x, y) && sparse_keep(S, x, y);
  const int n = __syncthreads_count(keep);
  if (threadIdx.x == 0) block_counts[blockIdx.x] = (unsigned)n;
}

// exclusive scan of the block counts in place (one CTA, chunks of 1024 with a running carry); total -> *count
__global__ void __launch_bounds__(1024) sparse_scan_kernel(unsigned int* __restrict__ block_counts, int n_blocks, unsigned int* __restrict__ count)
{
  __shared__ unsigned int warp_sums[32];
  __shared__ unsigned int carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0u;
  __syncthreads();
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + threadIdx.x;
    const unsigned int v = i < n_blocks ? block_counts[i] : 0u;
    unsigned int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      unsigned int w = warp_sums[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int o = __shfl_up_sync(0xffffffffu, w, d);
        if (lane >= d) w += o;
      }
      warp_sums[lane] = w; // inclusive over warps
    }
    __syncthreads();
    const unsigned int carry = carry_s;
    const unsigned int before = carry + (warp > 0 ? warp_sums[warp - 1] : 0u) + inc - v;
    if (i < n_blocks) block_counts[i] = before;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + warp_sums[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = carry_s;
}

__global__ void __launch_bounds__(SPARSE_BLOCK) sparse_scatter_kernel(const SparseParams S, const unsigned int* __restrict__ block_offsets,
                                                                      int2* __restrict__ xy)
{
  __shared__ unsigned int warp_sums[SPARSE_BLOCK / 32];
  const int e = blockIdx.x * SPARSE_BLOCK + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int x = 0, y = 0;
  const bool keep = e < S.n_entries && sparse_entry_pixel(S, e, x, y) && sparse_keep(S, x, y);
  const unsigned int ballot = __ballot_sync(0xffffffffu, keep);
  if (lane == 0) warp_sums[warp] = (unsigned)__popc(ballot);
  __syncthreads();
  if (warp == 0) {
    const unsigned int v = warp_sums[lane];
    unsigned int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned int o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    warp_sums[lane] = inc - v; // exclusive
  }
  __syncthreads();
  if (keep) xy[block_offsets[blockIdx.x] + warp_sums[warp] + (unsigned)__popc(ballot & ((1u << lane) - 1u))] = make_int2(x, y);
}

// default noise table when the application supplies none (the reference embeds stbn_128x128x64.bin, which its repository
// does not ship; data/noise/blue_64x64x64.bin can be loaded with b200vr_set_sparse_noise): white noise on the reference's
// 8-bit value grid k / 255
__global__ void sparse_default_noise_kernel(float* __restrict__ noise, int n, unsigned int seed)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned int h = (unsigned)i * 0x9E3779B9u ^ seed;
  h ^= h >> 16, h *= 0x85EBCA6Bu, h ^= h >> 13, h *= 0xC2B2AE35u, h ^= h >> 16;
  noise[i] = (float)(h >> 24) / 255.f;
}

} // namespace b200k
