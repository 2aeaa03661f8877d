// Ray marching WITHOUT gradient / shadow shading (SHADING_NONE): classification + front-to-back compositing only
// (raymarching, shaders_raymarching.cu:87-171 with the shading branch removed). This is the kernel of the sort-last bricks
// (shadow rays would leave the brick) and of the C2-noshade measurements.
//
// Every lane marches its own ray; there is nothing to defer, so the only thing between the kernel and the texture unit is
// latency. Inside a live macrocell the sample positions are known in advance (fixed t lattice): the kernel fetches the next
// NS_BATCH lattice samples together and then classifies / composites them in order, stopping at early termination (at most
// NS_BATCH - 1 fetches per ray are wasted). Empty macrocells are stepped over on the lattice exactly as in march_q.cuh.
#pragma once
#include "march_q.cuh"

namespace b200k {

#ifndef B200VR_NS_BATCH
#define B200VR_NS_BATCH 4
#endif
constexpr int NS_BATCH = B200VR_NS_BATCH;
#ifndef B200VR_NS_THREADS
#define B200VR_NS_THREADS 256
#endif
#ifndef B200VR_NS_MIN_CTAS
#define B200VR_NS_MIN_CTAS 4
#endif
constexpr int NS_THREADS = B200VR_NS_THREADS;
// Measured on C2-noshade (82 M samples): 128-thread CTAs 0.89 ms (nine resident copies of the tables take 99 KB away from
// L1), 256 x 4 / 512 x 2 / 1024 x 1 threads 0.51 ms; batches of 2 / 4 / 6 / 8 samples 0.92 / 0.51 / 0.53 / 0.54 ms (256 threads
// except batch 2); the per-lane state machine it replaces (march_sm.cuh) 0.90 ms.

// dynamic shared memory: [alpha pairs (n_alpha + 1) x float2][colour table (n_color + 1) x float4][occupancy bits]
__host__ __device__ __forceinline__ size_t ns_color_bytes(int n_color) { return sizeof(float4) * (size_t)(n_color + 1); }
__host__ __device__ __forceinline__ size_t ns_shared_bytes(int n_color, int n_alpha, int occ_words_in_smem)
{
  return q_alpha_bytes(n_alpha) + ns_color_bytes(n_color) + q_occ_bytes(occ_words_in_smem);
}

template<bool SKIP, bool POW2, bool OCCS>
__global__ void __launch_bounds__(NS_THREADS, B200VR_NS_MIN_CTAS)
raymarch_ns_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  float2* const s_apair = reinterpret_cast<float2*>(smem);
  float4* const s_color = reinterpret_cast<float4*>(smem + q_alpha_bytes(P.n_alpha));
  unsigned int* const s_occ = reinterpret_cast<unsigned int*>(smem + q_alpha_bytes(P.n_alpha) + ns_color_bytes(P.n_color));
  const unsigned int apair_base = pin_u32((unsigned int)__cvta_generic_to_shared(smem));
  const unsigned int color_base = pin_u32(apair_base + (unsigned int)q_alpha_bytes(P.n_alpha));
  const unsigned int occ_base = pin_u32(color_base + (unsigned int)ns_color_bytes(P.n_color));
  if (SKIP && OCCS)
    for (int i = threadIdx.x; i < P.occ_words; i += blockDim.x) s_occ[i] = P.occ_bits[i];
  for (int i = threadIdx.x; i <= P.n_alpha; i += blockDim.x) {
    const float v0 = P.tf_alpha[min(i, P.n_alpha - 1)], v1 = P.tf_alpha[min(i + 1, P.n_alpha - 1)];
    s_apair[i] = make_float2(v0, (v1 - v0) * (1.f / 256.f));
  }
  for (int i = threadIdx.x; i <= P.n_color; i += blockDim.x) s_color[i] = P.tf_color[min(i, P.n_color - 1)];
  __syncthreads();
  const f3 oo = P.cam_oo;
  const float step = P.step;
  unsigned int c_primary = 0, c_rays = 0; // per lane, over all tiles of this warp

  for (;;) {
    int k = 0;
    if (lane == 0) k = (int)atomicAdd(P.tile_counter, 1u);
    k = __shfl_sync(0xffffffffu, k, 0);
    if (k >= P.n_warp_tiles) break;
    int ix, iy;
    const int tile_slot = warp_tile_pixel(P, k, lane, ix, iy);
    const long long tile_t0 = clock64();
    const bool inside = ix < P.width && iy < P.height;
    const unsigned int pixel_index = (unsigned)ix + (unsigned)iy * (unsigned)P.width;

    for (int sidx = 0; sidx < P.spp; sidx++) {
      // ---- ray generation + clip (shaders_common.h:400-404, shaders_raymarching.cu:351-362, shaders_common.h:156-184) ----
      float sx = ((float)ix + .5f) * P.rw, sy = ((float)iy + .5f) * P.rh;
      if (P.spp > 1) {
        Tea rng((unsigned)P.frame_index, pixel_index); // RandomTEA(v0 = frame_index, v1 = pixel_index): draw number sidx + 1
        float2 j = rng.draw();
        for (int d = 0; d < sidx; d++) j = rng.draw();
        sx += (j.x - 0.5f) * P.rw;
        sy += (j.y - 0.5f) * P.rh;
      }
      const f3 dir = normalize3(P.cam_dir + (sx - 0.5f) * P.cam_hor + (sy - 0.5f) * P.cam_ver);
      const f3 od = xfm_vector(P.wto.l, dir);
      float tx = 0.f, t1 = FLT_MAX;
      bool alive;
      float own_exit = FLT_MAX, b0 = 0.f; // sort-last brick: samples with midpoint in [b0, own_exit) are this handle's
      if (!P.brick) alive = inside && intersect_box(tx, t1, oo, od, P.clip_lo, P.clip_hi);
      else {
        // both intervals from WORLD-space slab tests on the world ray: all ranks compute bit-identical parameters for a shared
        // cut plane, so the bricks partition the lattice samples exactly, and all ranks walk the same lattice (b200vr_set_brick)
        alive = inside && intersect_box(tx, t1, P.cam_pos, dir, P.lat_lo, P.lat_hi);
        alive = intersect_box(b0, own_exit, P.cam_pos, dir, P.own_lo, P.own_hi) && alive;
      }
      if (inside) c_rays++;
      float ty = fminf(t1, tx + step);
      alive = alive && ty > tx;
      if (P.brick && alive && 0.5f * (tx + ty) < b0) { // walk the lattice up to the brick: the same running sum as the march itself
        const float lim = fminf(b0, t1) - 5.f * step;
        while (ty < lim) {
          ty += step;
          ty += step;
          ty += step;
          ty += step;
        }
        do {
          tx = ty;
          ty = fminf(tx + step, t1);
        } while (ty > tx && 0.5f * (tx + ty) < b0);
        alive = ty > tx;
      }
      const f3 od_rcp = mk3(rcp_fast(od.x), rcp_fast(od.y), rcp_fast(od.z)); // cell exits only (see march_q.cuh)
      float A = 0.f;
      f3 C = mk3(0.f);
      float cell_exit = -1.f; // macrocell the ray is in: valid while the sample parameter stays below cell_exit

      while (alive) {
        const float tm = 0.5f * (tx + ty);
        if (!(tm < own_exit)) break; // the rest of the lattice belongs to the bricks behind this one
        if (SKIP && !(tm < cell_exit)) {
          const f3 p = mk3(__fmaf_rn(tm, od.x, oo.x), __fmaf_rn(tm, od.y, oo.y), __fmaf_rn(tm, od.z, oo.z));
          i3 c;
          const bool cell_empty = !occ_bit<OCCS>(P, occ_base, (unsigned)pad_cell(P, p, c));
          cell_exit = cell_exit_t(P, c, oo, od_rcp, od);
          if (cell_empty) {
            // every sample of this cell has alpha == 0 exactly and adds +0: advance t on the reference's lattice
            const float lim = fminf(cell_exit, t1) - 5.f * step;
            while (ty < lim) { // four whole steps stay inside the cell and the ray: min(., t1) is the identity
              ty += step;
              ty += step;
              ty += step;
              ty += step;
            }
            do {
              tx = ty;
              ty = fminf(tx + step, t1);
            } while (ty > tx && 0.5f * (tx + ty) < cell_exit);
            alive = ty > tx;
            cell_exit = -1.f; // the next sample is in another cell
            continue;
          }
        }
        // ---- the next NS_BATCH lattice samples of this (live) cell: positions first, fetches together ----
        const float lim = SKIP ? fminf(cell_exit, own_exit) : own_exit;
        float ta[NS_BATCH + 1], sv[NS_BATCH];
        bool ok[NS_BATCH];
        ta[0] = tx, ta[1] = ty;
#pragma unroll
        for (int i = 1; i < NS_BATCH; i++) ta[i + 1] = fminf(ta[i] + step, t1);
#pragma unroll
        for (int i = 0; i < NS_BATCH; i++) {
          const float tmi = 0.5f * (ta[i] + ta[i + 1]);
          ok[i] = i == 0 || (ok[i - 1] && ta[i + 1] > ta[i] && tmi < lim);
          sv[i] = 0.f;
          if (ok[i]) sv[i] = tex3D<float>(P.tex, __fmaf_rn(tmi, od.x, oo.x), __fmaf_rn(tmi, od.y, oo.y), __fmaf_rn(tmi, od.z, oo.z));
        }
        // ---- classify + composite in order (shaders_raymarching.cu:110-169 without shading) ----
#pragma unroll
        for (int i = 0; i < NS_BATCH; i++) {
          if (!ok[i] || !alive) break;
          c_primary++;
          float a = alpha_lookup<POW2>(P, apair_base, sv[i]);
          {
            const float adj = P.base * (ta[i + 1] - ta[i]); // opacity correction (shaders_raymarching.cu:118-122)
            if (!(fabsf(adj - 1.f) < 1e-7f)) a = clamp01(1.f - __powf(1.f - a, adj));
          }
          if (!SKIP || a != 0.f) {
            int ci, ca;
            tf_coord(tf_unit(P, sv[i]), P.n_color, P.tfc_mul, P.tfc_addf, P.tfc_addi, P.tfc_shift, ci, ca);
            const float4 c0 = lds_f32x4(color_base + 16u * (unsigned)ci), c1 = lds_f32x4(color_base + 16u * (unsigned)ci + 16u);
            const float cw = (float)ca * (1.f / 256.f);
            const f3 rgb = mk3(__fmaf_rn(cw, c1.x - c0.x, c0.x), __fmaf_rn(cw, c1.y - c0.y, c0.y), __fmaf_rn(cw, c1.z - c0.z, c0.z));
            const float w = (1.f - A) * a;
            C = C + clamp01(rgb) * w;
            A += w;
          }
          tx = ta[i + 1];
          ty = fminf(tx + step, t1);
          alive = ty > tx && A < 0.9999f;
        }
      }
      // render_raymarching (shaders_raymarching.cu:260-321): background is always 0 -> alpha_blend = C / A
      if (!P.premultiplied) C = A > 0.f ? C / A : mk3(0.f);
      if (inside) {
        float4 px4 = make_float4(C.x, C.y, C.z, A);
        float* g = P.grad + 3 * (size_t)pixel_index;
        if (sidx > 0) { // running sums over the samples of this pixel
          const float4 prev = P.rgba[pixel_index];
          px4 = make_float4(prev.x + px4.x, prev.y + px4.y, prev.z + px4.z, prev.w + px4.w);
        }
        if (sidx == P.spp - 1) {
          const float rspp = 1.f / (float)P.spp;
          px4 = make_float4(px4.x * rspp, px4.y * rspp, px4.z * rspp, px4.w * rspp);
          // frame write / progressive accumulation (shaders_raymarching.cu:389-406)
          if (P.accumulate) {
            if (P.frame_index == 1) {
              P.accum[pixel_index] = px4;
            }
            else {
              const float4 acc = P.accum[pixel_index];
              const float4 sum = make_float4(acc.x + px4.x, acc.y + px4.y, acc.z + px4.z, acc.w + px4.w);
              P.accum[pixel_index] = sum;
              const float fi = (float)P.frame_index;
              px4 = make_float4(sum.x / fi, sum.y / fi, sum.z / fi, sum.w / fi);
            }
          }
        }
        P.rgba[pixel_index] = px4;
        g[0] = 0.f; // no gradient without shading
        g[1] = 0.f;
        g[2] = 0.f;
      }
    }
    if (lane == 0) tile_cost_add(P, tile_slot, tile_t0);
  }

  const unsigned int n_primary = __reduce_add_sync(0xffffffffu, c_primary), n_rays = __reduce_add_sync(0xffffffffu, c_rays);
  if (lane == 0) { // one atomic set per warp
    atomicAdd(&P.counters->primary_samples, (unsigned long long)n_primary);
    atomicAdd(&P.counters->rays, (unsigned long long)n_rays);
  }
}

} // namespace b200k
