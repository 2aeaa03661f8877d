// Set-up and exchange kernels of the B200 volume device (sm_100a): synthetic volume, type conversion, macrocell value
// ranges + global range, majorants, sort-first tile pack/unpack, sort-last OVER compositing.
// Reference counterparts: value_range_kernel / majorant_kernel (ovr/devices/optix7/accel/sp_singlemc.cu:10-97),
// compute_scalar_range + convert_array1d (ovr/devices/optix7/array.cpp:27-90), alpha_blend (shaders_common.h:329-337).
#pragma once
#include "launch_params.cuh"

namespace b200k {

// ---- synthetic volume noise(seed): integer-exact fBm value noise -------------------------------------
__device__ __forceinline__ uint32_t fmix32(uint32_t h)
{
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}
__device__ __forceinline__ unsigned long long noise_lattice(uint32_t seed, uint32_t a, uint32_t b, uint32_t c, uint32_t o)
{
  return fmix32(seed ^ (a * 73856093u) ^ (b * 19349663u) ^ (c * 83492791u) ^ (o * 0x9E3779B9u)) >> 16;
}
__device__ __forceinline__ float noise_voxel(uint32_t seed, uint32_t x, uint32_t y, uint32_t z)
{
  unsigned long long sum = 0;
#pragma unroll
  for (uint32_t o = 0; o < 4; o++) {
    const uint32_t sh = 6 - o, P = 1u << sh;
    const uint32_t a = x >> sh, b = y >> sh, c = z >> sh;
    const unsigned long long fx = x & (P - 1), fy = y & (P - 1), fz = z & (P - 1);
    unsigned long long acc = 0;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) {
      const uint32_t dx = k & 1, dy = (k >> 1) & 1, dz = k >> 2;
      const unsigned long long w = (dx ? fx : P - fx) * (dy ? fy : P - fy) * (dz ? fz : P - fz);
      acc += w * noise_lattice(seed, a + dx, b + dy, c + dz, o);
    }
    sum += (unsigned long long)(8u >> o) * (acc << (3 * o));
  }
  return (float)((double)sum / (double)(15ull * 65536ull * 262144ull));
}

__global__ void noise_kernel(float* __restrict__ out, uint32_t seed, int ox, int oy, int oz, int nx, int ny, int nz)
{
  const size_t n = (size_t)nx * ny * nz;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % nx), y = (int)((i / nx) % ny), z = (int)(i / ((size_t)nx * ny));
    out[i] = noise_voxel(seed, (uint32_t)(ox + x), (uint32_t)(oy + y), (uint32_t)(oz + z));
  }
}

// ---- typed volume -> fp32 (array.cpp:68-108,300-306,335-346; array.h:64-106) ---------------------------------------
// 8/32-bit integers are read by the reference through cudaReadModeNormalizedFloat (value / type-max, signed clamped at
// -1); 16-bit integers and doubles are cast to float without normalisation.
template<typename T> __device__ __forceinline__ float to_f32(T v) { return (float)v; }
template<> __device__ __forceinline__ float to_f32<uint8_t>(uint8_t v) { return (float)v / 255.f; }
template<> __device__ __forceinline__ float to_f32<int8_t>(int8_t v) { return fmaxf((float)v / 127.f, -1.f); }
template<> __device__ __forceinline__ float to_f32<uint32_t>(uint32_t v) { return (float)v / 4294967295.f; }
template<> __device__ __forceinline__ float to_f32<int32_t>(int32_t v) { return fmaxf((float)v / 2147483647.f, -1.f); }

template<typename T>
__global__ void convert_kernel(float* __restrict__ out, const T* __restrict__ in, size_t n)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = to_f32<T>(in[i]);
}

// ---- macrocell value ranges: one CTA per macrocell, lanes along x, warps over (y,z) rows --------------------------
// Range over voxels 16c-2 .. 16c+17 (clamped): the tri-lerp footprint of every position in the cell plus slack
// (the reference kernel, sp_singlemc.cu:36-42, misses the upper apron). Also reduces the global
// [min,max] (compute_scalar_range, array.cpp:27-66) through ordered-int atomics.
__device__ __forceinline__ int float_to_ordered(float f)
{
  const int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__host__ __device__ __forceinline__ float ordered_to_float(int i)
{
  const int j = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
  return __int_as_float(j);
#else
  float f;
  memcpy(&f, &j, sizeof(f));
  return f;
#endif
}

constexpr int MC_APRON = 2;
constexpr int RANGE_THREADS = 256;

// One CTA per (cy, cz): it walks the 20 x 20 rows (16 + the apron in y and z) of that line of macrocells over their whole
// length in x, a warp per 32-voxel column chunk -- 128-byte reads, every sector used, the only re-reads are the y / z aprons
// ((20 / 16)^2 = 1.56 x the volume, mostly from L2) -- and folds each chunk into the cells it belongs to (own cell, and the
// neighbour whose apron it is in) with shared-memory atomics on order-preserving integers. The first version had one CTA per
// macrocell reading 400 rows of 20 floats: 0.65 TB/s on 512^3 (profiles/r1d_launches_C2.csv), 10 % of the copy bandwidth.
// Dynamic shared memory: 2 * mx ints.
__global__ void __launch_bounds__(RANGE_THREADS)
macrocell_range_kernel(const float* __restrict__ vol, int nx, int ny, int nz, int mx, int my, int mz,
                       float2* __restrict__ ranges, int* __restrict__ global_minmax)
{
  extern __shared__ int s_mm[]; // [0, mx): lower bounds, [mx, 2 mx): upper bounds, as ordered ints
  const int cy = blockIdx.x % my, cz = blockIdx.x / my;
  for (int i = threadIdx.x; i < mx; i += RANGE_THREADS) {
    s_mm[i] = float_to_ordered(INFINITY);
    s_mm[mx + i] = float_to_ordered(-INFINITY);
  }
  __syncthreads();
  const int by = max(cy * MC_SIZE - MC_APRON, 0), ey = min(cy * MC_SIZE + MC_SIZE + MC_APRON, ny);
  const int bz = max(cz * MC_SIZE - MC_APRON, 0), ez = min(cz * MC_SIZE + MC_SIZE + MC_APRON, nz);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = RANGE_THREADS / 32;
  const int rows_y = ey - by;
  const int nchunks = (nx + 31) / 32;
  for (int chunk = warp; chunk < nchunks; chunk += nwarps) {
    const int x = chunk * 32 + lane;
    float lo = INFINITY, hi = -INFINITY;
    if (x < nx) {
      const float* col = vol + ((size_t)bz * ny + by) * nx + x;
      for (int z = 0; z < ez - bz; z++) {
        const float* p = col + (size_t)z * ny * nx;
#pragma unroll 4
        for (int y = 0; y < rows_y; y++) {
          const float v = __ldg(p + (size_t)y * nx);
          lo = fminf(lo, v);
          hi = fmaxf(hi, v);
        }
      }
    }
    // voxel x belongs to cell x / 16, to the next cell when it is in that cell's lower apron, to the previous one when in its upper
    const int c = x / MC_SIZE, m = x % MC_SIZE;
    if (x < nx) {
      if (m >= MC_SIZE - MC_APRON && c + 1 < mx) {
        atomicMin(&s_mm[c + 1], float_to_ordered(lo));
        atomicMax(&s_mm[mx + c + 1], float_to_ordered(hi));
      }
      if (m < MC_APRON && c > 0) {
        atomicMin(&s_mm[c - 1], float_to_ordered(lo));
        atomicMax(&s_mm[mx + c - 1], float_to_ordered(hi));
      }
    }
    // own cell: the 16 lanes of a half warp share it
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((lane & 15) == 0 && x < nx) {
      atomicMin(&s_mm[c], float_to_ordered(lo));
      atomicMax(&s_mm[mx + c], float_to_ordered(hi));
    }
  }
  __syncthreads();
  int glo = float_to_ordered(INFINITY), ghi = float_to_ordered(-INFINITY);
  for (int i = threadIdx.x; i < mx; i += RANGE_THREADS) {
    const int a = s_mm[i], b = s_mm[mx + i];
    ranges[((size_t)cz * my + cy) * mx + i] = make_float2(ordered_to_float(a), ordered_to_float(b));
    glo = min(glo, a);
    ghi = max(ghi, b);
  }
  glo = __reduce_min_sync(0xffffffffu, glo);
  ghi = __reduce_max_sync(0xffffffffu, ghi);
  if ((threadIdx.x & 31) == 0 && threadIdx.x < mx) { // warps that held at least one cell
    atomicMin(global_minmax + 0, glo);
    atomicMax(global_minmax + 1, ghi);
  }
}

// ---- majorants: max of the opacity table over the (+-1 entry padded) index range of the cell's value range ---------
// majorant_kernel (sp_singlemc.cu:56-97) with its defects fixed: whole table staged by all threads
// before the barrier, index clamped before the float->int conversion, lookup window used for normalisation.
__global__ void majorant_kernel(const float2* __restrict__ ranges, int n_cells, const float* __restrict__ alphas, int n_alpha,
                                float win_lo, float win_hi, float* __restrict__ majorants)
{
  extern __shared__ float s_alpha[];
  for (int i = threadIdx.x; i < n_alpha; i += blockDim.x) s_alpha[i] = alphas[i];
  __syncthreads();
  const float rcp = 1.f / (win_hi - win_lo);
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < n_cells; c += gridDim.x * blockDim.x) {
    const float2 r = ranges[c];
    const float lo = __fmul_rn(__fsub_rn(clampf(r.x, win_lo, win_hi), win_lo), rcp);
    const float hi = __fmul_rn(__fsub_rn(clampf(r.y, win_lo, win_hi), win_lo), rcp);
    float fl = floorf(__fmaf_rn(lo, (float)(n_alpha - 1), 0.5f)) - 1.f;
    float fh = floorf(__fmaf_rn(hi, (float)(n_alpha - 1), 0.5f)) + 1.f;
    fl = clampf(fl, 0.f, (float)(n_alpha - 1));
    fh = clampf(fh, 0.f, (float)(n_alpha - 1));
    float m = 0.f;
    for (int k = (int)fl; k <= (int)fh; k++) m = fmaxf(m, s_alpha[k]);
    majorants[c] = m;
  }
}

// padded copy: cell (x, y, z) of the (mc+2)^3 grid = cell (clamp(x-1), clamp(y-1), clamp(z-1)) of the mc^3 grid
__global__ void majorant_pad_kernel(const float* __restrict__ maj, int mx, int my, int mz, float* __restrict__ pad)
{
  const int px = mx + 2, py = my + 2, pz = mz + 2;
  const int n = px * py * pz;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int x = i % px - 1, y = (i / px) % py - 1, z = i / (px * py) - 1;
    pad[i] = maj[min(max(x, 0), mx - 1) + mx * (min(max(y, 0), my - 1) + my * min(max(z, 0), mz - 1))];
  }
}

// occupancy mask of the padded grid: bit i of word w = (majorant of padded cell 32 w + i != 0)
__global__ void occupancy_bits_kernel(const float* __restrict__ maj_pad, int n_pad, unsigned int* __restrict__ bits, int n_words)
{
  for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += gridDim.x * blockDim.x) {
    unsigned int v = 0u;
    for (int b = 0; b < 32; b++) {
      const int i = w * 32 + b;
      if (i < n_pad && maj_pad[i] != 0.f) v |= 1u << b;
    }
    bits[w] = v;
  }
}

// ---- launch order of the super tiles: most expensive first ----------------------------------------------------------
// A warp tile is a serial chain (march, shade 32 samples, march ...) that can take a large part of the whole frame
// time; started late it becomes the tail of the launch. The owned super tiles are therefore launched in descending order of
// the warp time they took in the previous frame (frames of an interactive session are similar); without a history (all
// costs 0) the order is centre-out, the volume usually being in the middle of the image. Rank sort, one thread per tile.
__global__ void tile_order_kernel(const unsigned int* __restrict__ cost, int n_owned, int rank, int nranks, int super_w, int super_h,
                                  int* __restrict__ order, int mode)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_owned) return;
  if (mode == 0) { // row-major, for A/B runs
    order[i] = i;
    return;
  }
  auto key = [&](int m) { // larger = earlier
    const int s = rank + m * nranks;
    const float dx = (float)(s % super_w) + 0.5f - 0.5f * (float)super_w, dy = (float)(s / super_w) + 0.5f - 0.5f * (float)super_h;
    const unsigned int d = (unsigned int)fminf(4.f * (2.f * dx * 2.f * dx + dy * dy), 1048575.f); // tiles are 64 x 32 pixels
    if (mode == 3) { // Z-order over the super-tile grid: the tiles in flight form a compact block of the image
      unsigned int x = (unsigned int)(s % super_w), y = (unsigned int)(s / super_w), z = 0u;
      for (int b = 0; b < 10; b++) z |= ((x >> b) & 1u) << (2 * b) | ((y >> b) & 1u) << (2 * b + 1);
      return (unsigned long long)(1048575u - z);
    }
    if (mode == 4 || mode == 5) { // column-major (4: bottom-up, 5: top-down): tiles in flight form a vertical strip of the image
      const unsigned int x = (unsigned int)(s % super_w), y = (unsigned int)(s / super_w);
      return (unsigned long long)(1048575u - (x * (unsigned int)super_h + (mode == 4 ? y : (unsigned int)super_h - 1u - y)));
    }
    if (mode == 6) { // whole rows of super tiles, middle row first, then alternating outwards: row-major locality inside a row,
                     // the (usually heavy) middle of the image no longer starts halfway through the launch
      const int x = s % super_w, y = s / super_w, mid = super_h / 2;
      const int d2 = y >= mid ? 2 * (y - mid) : 2 * (mid - y) - 1;
      return (unsigned long long)(1048575u - (unsigned int)(d2 * super_w + x));
    }
    return ((unsigned long long)(mode == 2 ? cost[m] : 0u) << 20) | (unsigned long long)(1048575u - d);
  };
  const unsigned long long ki = key(i);
  int pos = 0;
  for (int j = 0; j < n_owned; j++) {
    const unsigned long long kj = key(j);
    pos += (kj > ki) || (kj == ki && j < i);
  }
  order[pos] = i;
}

// ---- sheet order of the warp tiles (shaded frames) -----------------------------------------------------------------
// All shadow rays are parallel (fixed directional light L, params.h:79). The primary rays of the pixels on one image line
// through the vanishing point of L lie in one plane through the eye that contains L, and so do all their shadow rays: the
// planes form a pencil around the axis (eye, L) and cut the volume into disjoint wedges. Warp tiles launched in the order of
// their plane angle phi (then along the plane) make the tiles in flight share a few wedges instead of spreading their
// shadow slabs over the whole volume: the voxels a wedge needs stay in L2 while its tiles run. Any order is a valid order
// (the frame does not depend on it); this one is chosen for the 1024^3 volume, where the rays of a row-major launch re-read
// the volume from HBM ~13 times (profiles/r1c_ncu_full_raymarch_q_C3.txt).
struct SheetParams {
  f3 cam_dir, cam_hor, cam_ver;
  f3 e1, e2, light;   // e1, e2: orthonormal basis of the plane perpendicular to the light direction; phi = atan2(d.e2, d.e1)
  float rw, rh;       // 1 / image size
  float inv_dphi;     // 1 / (plane angle per strip)
  int super_w, rank, nranks, n_warp_tiles;
};
__global__ void sheet_key_kernel(const SheetParams S, unsigned int* __restrict__ keys, unsigned int* __restrict__ vals)
{
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= S.n_warp_tiles) return;
  // k-th owned warp tile in the default (row-major) walk, as warp_tile_origin() with the identity order
  const int m = k / WARP_TILES_PER_SUPER, r = k % WARP_TILES_PER_SUPER;
  const int s = S.rank + m * S.nranks;
  const int b = r / TILE_CHUNK, i = r % TILE_CHUNK;
  const int px = (s % S.super_w) * SUPER_TILE_W + ((b % SUPER_BLOCKS_X) * 4 + (i & 3)) * WARP_TILE_W;
  const int py = (s / S.super_w) * SUPER_TILE_H + ((b / SUPER_BLOCKS_X) * 4 + (i >> 2)) * WARP_TILE_H;
  const float sx = ((float)px + 0.5f * WARP_TILE_W) * S.rw, sy = ((float)py + 0.5f * WARP_TILE_H) * S.rh;
  const f3 d = S.cam_dir + (sx - 0.5f) * S.cam_hor + (sy - 0.5f) * S.cam_ver; // centre ray of the tile (unnormalised)
  const float phi = atan2f(dot3(d, S.e2), dot3(d, S.e1));
  const int strip = min(max(__float2int_rd(phi * S.inv_dphi) + (1 << 18), 0), (1 << 19) - 1);
  const float along = dot3(d, S.light) * rsqrtf(dot3(d, d)); // cosine between the ray and the light: position along the strip
  const int a = min(max(__float2int_rn((along * 0.5f + 0.5f) * 4095.f), 0), 4095);
  keys[k] = ((unsigned int)strip << 12) | (unsigned int)a;
  vals[k] = (unsigned int)(px / WARP_TILE_W) | ((unsigned int)(py / WARP_TILE_H) << 16);
}

// ---- end of a ring-kernel sample: fixed-point sums -> frame (the tail of raymarch_q_kernel, per pixel of the owned super tiles) ----
__global__ void resolve_kernel(const LaunchParams P, int n_owned)
{
  const size_t n = (size_t)n_owned * SUPER_TILE_W * SUPER_TILE_H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int m = (int)(i / (SUPER_TILE_W * SUPER_TILE_H)), r = (int)(i % (SUPER_TILE_W * SUPER_TILE_H));
    const int s = P.rank + m * P.nranks;
    const int ix = (s % P.super_w) * SUPER_TILE_W + r % SUPER_TILE_W, iy = (s / P.super_w) * SUPER_TILE_H + r / SUPER_TILE_W;
    if (ix >= P.width || iy >= P.height) continue;
    const size_t pixel_index = (size_t)ix + (size_t)iy * (size_t)P.width;
    unsigned long long* const acc = P.acc + 3 * pixel_index;
    const unsigned long long a0 = acc[0], a1 = acc[1], a2 = acc[2];
    acc[0] = 0ull, acc[1] = 0ull, acc[2] = 0ull; // ready for the next sample / frame
    const float k = 1.f / 1073741824.f;
    f3 C = mk3((float)(unsigned int)a0 * k, (float)(unsigned int)(a0 >> 32) * k, (float)(unsigned int)a1 * k);
    f3 G = mk3((float)(unsigned int)(a1 >> 32) * k, (float)(unsigned int)a2 * k, (float)(unsigned int)(a2 >> 32) * k);
    const float A = P.ray_alpha[pixel_index];
    // render_raymarching (shaders_raymarching.cu:260-321): background is always 0 -> alpha_blend = C / A
    if (!P.premultiplied) {
      C = A > 0.f ? C / A : mk3(0.f);
      G = A > 0.f ? G / A : mk3(0.f);
    }
    float4 px4 = make_float4(C.x, C.y, C.z, A);
    float* g = P.grad + 3 * pixel_index;
    if (P.sidx > 0) { // running sums over the samples of this pixel
      const float4 prev = P.rgba[pixel_index];
      px4 = make_float4(prev.x + px4.x, prev.y + px4.y, prev.z + px4.z, prev.w + px4.w);
      G = G + mk3(g[0], g[1], g[2]);
    }
    if (P.sidx == P.spp - 1) {
      const float rspp = 1.f / (float)P.spp;
      px4 = make_float4(px4.x * rspp, px4.y * rspp, px4.z * rspp, px4.w * rspp);
      G = G * rspp;
      // frame write / progressive accumulation (shaders_raymarching.cu:389-406)
      if (P.accumulate) {
        if (P.frame_index == 1) {
          P.accum[pixel_index] = px4;
        }
        else {
          const float4 acc4 = P.accum[pixel_index];
          const float4 sum = make_float4(acc4.x + px4.x, acc4.y + px4.y, acc4.z + px4.z, acc4.w + px4.w);
          P.accum[pixel_index] = sum;
          const float fi = (float)P.frame_index;
          px4 = make_float4(sum.x / fi, sum.y / fi, sum.z / fi, sum.w / fi);
        }
      }
    }
    P.rgba[pixel_index] = px4;
    g[0] = G.x;
    g[1] = G.y;
    g[2] = G.z;
  }
}

__global__ void ring_init_kernel(unsigned int* seq, int cap, int shards, unsigned int* ctrl, int ctrl_words)
{
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < cap * shards; i += gridDim.x * blockDim.x) seq[i] = (unsigned int)(i % cap);
  if (blockIdx.x == 0)
    for (int i = threadIdx.x; i < ctrl_words; i += blockDim.x) ctrl[i] = 0u;
}
// start of a ring launch (one thread per ring = per CTA): draining warps of the previous launch may have drawn tickets beyond
// the last published block
__global__ void ring_start_kernel(unsigned int* ctrl, int stride, int rings, int warps_per_cta)
{
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < rings) {
    unsigned int* c = ctrl + (size_t)r * 3 * stride;
    c[stride] = c[0];
    c[2 * stride] = (unsigned int)warps_per_cta;
  }
  if (r == 0) {
    ctrl[(size_t)rings * 3 * stride] = (unsigned int)(rings * warps_per_cta);
    ctrl[(size_t)rings * 3 * stride + 1] = 0u; // march_rq.cuh: blocks published and not yet taken
    ctrl[(size_t)rings * 3 * stride + stride] = 0u;
  }
}

// ---- sort-first: dense slice <-> frame for the warp tiles owned by (rank, nranks) --------------------------------
// slice layout: [owned super tile m][y in 32][x in 64][rgba]; pixels outside the image are zero in the slice
template<bool PACK>
__global__ void tile_copy_kernel(float4* __restrict__ frame, float4* __restrict__ slice, int width, int height, int super_w,
                                 int rank, int nranks, int n_owned)
{
  const size_t n = (size_t)n_owned * SUPER_TILE_W * SUPER_TILE_H;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int m = (int)(i / (SUPER_TILE_W * SUPER_TILE_H)), r = (int)(i % (SUPER_TILE_W * SUPER_TILE_H));
    const int s = rank + m * nranks;
    const int x = (s % super_w) * SUPER_TILE_W + r % SUPER_TILE_W, y = (s / super_w) * SUPER_TILE_H + r / SUPER_TILE_W;
    const bool inb = x < width && y < height;
    if (PACK)
      slice[i] = inb ? frame[(size_t)y * width + x] : make_float4(0.f, 0.f, 0.f, 0.f);
    else if (inb)
      frame[(size_t)y * width + x] = slice[i];
  }
}

// ---- sort-last: front OVER back on premultiplied float4; `back` may live on a peer GPU (NVLink P2P ld.global) ------
__global__ void composite_over_kernel(const float4* __restrict__ front, const float4* __restrict__ back, float4* __restrict__ out,
                                      size_t n)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float4 f = front[i], b = back[i];
    const float tr = 1.f - f.w;
    out[i] = make_float4(f.x + tr * b.x, f.y + tr * b.y, f.z + tr * b.z, f.w + tr * b.w);
  }
}
// Direct-send compositing (sort-last): one pass blends the same pixel range of up to COMPOSITE_MAX_LAYERS premultiplied
// layers front to back (layer[0] nearest the eye) -- each layer read once, straight from the GPU that rendered it when the pointer
// is a peer mapping -- and writes the result once (possibly into the display GPU's frame). finalize: un-premultiply on the way out
// (alpha_blend, shaders_common.h:329-337). The blend is composite_over_kernel's, applied layer by layer in the same fp32 order.
constexpr int COMPOSITE_MAX_LAYERS = 16;
struct CompositeLayers {
  const float4* layer[COMPOSITE_MAX_LAYERS];
  int n;
};
__global__ void composite_layers_kernel(const __grid_constant__ CompositeLayers L, float4* out, size_t n, int finalize)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float4 v[COMPOSITE_MAX_LAYERS];
#pragma unroll
    for (int k = 0; k < COMPOSITE_MAX_LAYERS; k++)
      if (k < L.n) v[k] = __ldcs(L.layer[k] + i); // all loads in flight before the first blend (peer reads: ~2 us each)
    float4 a = v[0];
#pragma unroll
    for (int k = 1; k < COMPOSITE_MAX_LAYERS; k++)
      if (k < L.n) {
        const float tr = 1.f - a.w;
        a = make_float4(a.x + tr * v[k].x, a.y + tr * v[k].y, a.z + tr * v[k].z, a.w + tr * v[k].w);
      }
    if (finalize) a = a.w > 0.f ? make_float4(a.x / a.w, a.y / a.w, a.z / a.w, a.w) : make_float4(0.f, 0.f, 0.f, a.w);
    out[i] = a;
  }
}
__global__ void finalize_premultiplied_kernel(float4* __restrict__ px, size_t n)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float4 p = px[i];
    px[i] = p.w > 0.f ? make_float4(p.x / p.w, p.y / p.w, p.z / p.w, p.w) : make_float4(0.f, 0.f, 0.f, p.w);
  }
}

} // namespace b200k
