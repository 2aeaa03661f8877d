// Per-launch constant block of the B200 volume kernels. Plays the role of the reference's LaunchParams +
// DeviceStructuredRegularVolume + DeviceSpacePartiton_SingleMC (ovr/devices/optix7/params.h:55-99, volume.h:40-62,
// accel/spatial_partition.h:23-52) but is passed by value as a __grid_constant__ kernel argument: no SBT, no
// params upload, no pointer chase per sample.
#pragma once
#include "device_math.cuh"

namespace b200k {

constexpr int MC_SIZE = 16;        // macrocell edge in voxels (accel/spatial_partition.h:24)
constexpr int WARP_TILE_W = 8;     // pixels per warp tile: 8 x 4
constexpr int WARP_TILE_H = 4;
#ifndef B200VR_SUPER_W // tuning builds only (multiples of 32 x 16): distributed.py mirrors the two numbers
#define B200VR_SUPER_W 64
#define B200VR_SUPER_H 32
#endif
constexpr int SUPER_TILE_W = B200VR_SUPER_W;   // sort-first ownership granule: 64 x 32 pixels = 8 x 8 warp tiles
constexpr int SUPER_TILE_H = B200VR_SUPER_H;
constexpr int SUPER_BLOCKS_X = SUPER_TILE_W / (4 * WARP_TILE_W); // blocks of 4 x 4 warp tiles per row of a super tile
constexpr int WARP_TILES_PER_SUPER = (SUPER_TILE_W / WARP_TILE_W) * (SUPER_TILE_H / WARP_TILE_H);
constexpr int RENDER_THREADS = 128;
#ifndef B200VR_TILE_STATS
#define B200VR_TILE_STATS 0
#endif

struct Counters {
  unsigned long long primary_samples, shadow_samples, gradient_fetches, delta_samples, rays;
  unsigned long long tile_clocks_max, tile_clocks_sum; // longest warp tile / all warp tiles, in SM clocks (B200VR_DEBUG_TILES)
  unsigned long long drain_clocks_sum, warp_clocks_max, blocks_published, blocks_stolen; // ring kernel: after a warp's last tile
  unsigned long long dbg[10]; // -DB200VR_TILE_STATS=1, march_rq.cuh: the longest warp tile in detail
};

struct LaunchParams {
  // ---- volume ----
  cudaTextureObject_t tex;  // fp32 3-D array, linear filter, clamp, normalised coords (array.cpp:300-306)
  const float* lin;         // linear fp32 copy, x fastest (manual sampler, macrocell build)
  int nx, ny, nz;
  f3 rdim;                  // 1 / dims
  aff3 otw, wto;
  f3 clip_lo, clip_hi;      // object-space clip box (unit cube unless sort-last)
  int brick;                // sort-last: march the t lattice of the whole volume, own the samples whose midpoint is inside own_lo..own_hi
  f3 lat_lo, lat_hi;        // WORLD-space box of the whole volume (brick only)
  f3 own_lo, own_hi;        // WORLD-space box of this brick's own voxels (brick only)
  // ---- transfer function (global copies; staged into shared memory by every CTA) ----
  const float4* tf_color;
  const float* tf_alpha;
  int n_color, n_alpha;
  float tf_lower, tf_upper, tf_scale;
  // power-of-two table fast path of the texture-unit coordinate model (sampling.cuh): q = (int(v*mul + add_f) + add_i) >> shift
  // is the unit's 1/256-texel coordinate t; shift < 0 selects the generic 64-bit path
  float tfa_mul, tfa_addf, tfc_mul, tfc_addf;
  int tfa_addi, tfa_shift, tfc_addi, tfc_shift;
  // ---- macrocells ----
  const float* majorants;
  const float* maj_pad;     // the same majorants on a grid padded by one replicated cell on every side: (mc+2)^3, so the
  int mcp_sy, mcp_sz;       // cell of any position in [-eps, 1+eps] is addressable without per-axis clamps; strides in floats
  unsigned int mcp_bias;    // folds the float->int magic constant and the +1 pad offset of pad_index_magic()
  unsigned int mcp_last;    // number of padded cells - 1
  const unsigned int* occ_bits; // one bit per padded macrocell: majorant != 0 (build.cuh occupancy_bits_kernel)
  int occ_words;            // words in occ_bits
  i3 mc;
  f3 cells_per_unit;        // dims / 16: object-space position -> macrocell coordinate
  f3 spac_rcp;              // 1 / (16 / dims) computed like sp_singlemc.cu:99-111 (delta tracking DDA)
  // ---- sampling ----
  float step, base;
  // ---- camera (device form, device_impl.cpp:125-136) ----
  f3 cam_pos, cam_dir, cam_hor, cam_ver;
  lin3 wtc;                 // get_xfm_world_to_camera (shaders_common.h:276-289)
  f3 light_dir;             // normalize(light_directional_pos), params.h:79
  float light_ambient;
  int max_num_scatters;
  int pt_child_instance_xfm; // b200vr_params::pt_child_instance_xfm
  // ---- kernel-wide constants derived on the host so that they are plain constant-bank operands in the kernels ----
  f3 cam_oo;                // camera position in object space: xfmPoint(wto, cam_pos)
  f3 sh_d, sh_rcp;          // shadow-ray direction in object space wto.l * light_dir, and __frcp_rn of it
  int sh_deg_x, sh_deg_y, sh_deg_z; // |sh_d| < FLT_MIN on that axis (intersect_box, shaders_common.h:163-171)
  f3 sh_dcell;              // sh_d in macrocell units (sh_d * cells_per_unit)
  float sstep;              // shadow step (step * 10) * step  (shaders_raymarching.cu:220, A.7-q1)
  float log_ert;            // log2(1 - 0.9999f): early-ray-termination threshold on log2-transmittance
  lin3 wto_t;               // transpose(wto.l): xfmNormal(otw, n) = wto_t * n
  // ---- frame ----
  float4* rgba;
  float* grad;              // 3 floats per pixel (vec3f)
  float4* accum;
  int width, height;
  float rw, rh;             // 1 / size (size_rcp)
  int frame_index;
  int spp;
  int accumulate;
  int premultiplied;
  // ---- tiles ----
  int super_w, super_h;     // image size in super tiles
  int rank, nranks;         // sort-first ownership: super tile s belongs to rank s % nranks
  int n_warp_tiles;         // warp tiles this launch has to render
  unsigned int* tile_counter;
  const int* tile_order;    // launch order of the owned super tiles: tile_order[j] = slot m (super tile rank + m * nranks)
  const unsigned int* wt_order; // optional launch order of the owned WARP tiles (sheet order, build.cuh): entry k = (px / 8) | (py / 4) << 16
  unsigned int* tile_cost;  // per slot m: warp clocks / 64 spent on its warp tiles in THIS launch (orders the next one)
  Counters* counters;
  // ---- sparse (foveated) sampling, sparse.cuh: when set, warp tile k renders list entries 32k .. 32k+31 ----
  const int2* sparse_xy;    // compacted pixel coordinates (xs_and_ys, shaders_common.h:437-449); nullptr = whole image
  int sparse_count;
  // ---- deferred-shading ring (march_ring.cuh) ----
  unsigned int* ring_ctrl;  // [0] head, [1] tail, [2] warps still marching, [3] error
  unsigned int* ring_seq;   // per slot sequence number
  float4* ring_data;        // ring_cap blocks of 64 float4 (32 tasks)
  int ring_cap, ring_high;  // slots (power of two); fill level above which a producer shades a block before publishing
  int rq_self;              // march_rq.cuh: blocks of a tile its warp shades itself while the ring is above ring_high
  unsigned long long* acc;  // per pixel 3 words of fixed-point sums: (r|g), (b|Gx), (Gy|Gz)
  float* ray_alpha;         // per pixel opacity of the sample being rendered
  int sidx;                 // sample index of this launch (the spp loop runs on the host for the ring kernel)
};

// k-th warp tile of this launch -> pixel origin and the slot m of its super tile in the cost / order tables. The launch walks
// the owned super tiles in the order of P.tile_order (most expensive first, see tile_order_kernel); inside a super tile the
// 64 warp tiles are ordered so that 16 consecutive ones form a 4 x 4 block of tiles (32 x 16 pixels).
constexpr int TILE_CHUNK = 16;
__device__ __forceinline__ int warp_tile_origin(const LaunchParams& P, int k, int& px, int& py)
{
  const int m = P.tile_order[k / WARP_TILES_PER_SUPER], r = k % WARP_TILES_PER_SUPER;
  const int s = P.rank + m * P.nranks;
  const int sx = s % P.super_w, sy = s / P.super_w;
  const int b = r / TILE_CHUNK, i = r % TILE_CHUNK; // block (2 x 2 per super tile), tile inside the block (4 x 4)
  px = sx * SUPER_TILE_W + ((b % SUPER_BLOCKS_X) * 4 + (i & 3)) * WARP_TILE_W;
  py = sy * SUPER_TILE_H + ((b / SUPER_BLOCKS_X) * 4 + (i >> 2)) * WARP_TILE_H;
  return m;
}
// pixel of this lane in the k-th warp tile of the launch; returns the cost-table slot of the tile (-1: none). Dense frames:
// 8x4 pixels at warp_tile_origin; sparse sampling: the lane's entry of the compacted list (compute_screen_position method 3,
// shaders_common.h:437-449). Warp-uniform branch on a launch constant.
__device__ __forceinline__ int warp_tile_pixel(const LaunchParams& P, int k, int lane, int& ix, int& iy)
{
  if (P.sparse_xy != nullptr) {
    const int e = k * 32 + lane;
    const int2 c = e < P.sparse_count ? __ldg(P.sparse_xy + e) : make_int2(P.width, 0); // x = width: outside the image
    ix = c.x, iy = c.y;
    return -1;
  }
  int px, py, m;
  if (P.wt_order != nullptr) { // explicit order of the warp tiles (sheet order): the table holds the tile origin
    const unsigned int v = __ldg(P.wt_order + k);
    px = (int)(v & 0xffffu) * WARP_TILE_W, py = (int)(v >> 16) * WARP_TILE_H;
    m = ((px / SUPER_TILE_W) + (py / SUPER_TILE_H) * P.super_w - P.rank) / P.nranks; // cost-table slot of its super tile
  }
  else m = warp_tile_origin(P, k, px, py);
  ix = px + (lane & (WARP_TILE_W - 1)), iy = py + (lane / WARP_TILE_W);
  return m;
}
// a warp reports how long its tile took: next frame's launch order
__device__ __forceinline__ void tile_cost_add(const LaunchParams& P, int m, long long t_start)
{
  if (m < 0) return; // sparse sampling: no tile history
  const unsigned long long dt = (unsigned long long)(clock64() - t_start);
  atomicAdd(P.tile_cost + m, (unsigned int)(dt >> 6));
#if B200VR_TILE_STATS // two more atomics per tile on single words: measurably slows the launch, debugging only
  atomicMax(&P.counters->tile_clocks_max, dt);
  atomicAdd(&P.counters->tile_clocks_sum, dt);
#endif
}
__host__ __device__ __forceinline__ int owned_super_tiles(int rank, int nranks, int n_super)
{
  return rank < n_super ? (n_super - rank + nranks - 1) / nranks : 0;
}

} // namespace b200k
