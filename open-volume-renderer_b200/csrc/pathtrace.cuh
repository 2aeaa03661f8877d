// Delta-tracking (Woodcock) path-tracing kernel for sm_100a with per-macrocell majorants walked by a resumable 3-D DDA.
// Replaces the reference's K6-K8 (ovr/devices/optix7/shaders_pathtracing.cu:269-289,385-404,477-542,603-721) and
// DeltaTrackingIter<true> (accel/spatial_partition.h:56-96) over DDAIter (accel/dda.h:30-132). The reference recurses
// through optixTrace once per bounce; here the walk is a loop with a small albedo stack in registers/local memory so the
// throughput product keeps the reference's inside-out multiplication order. Per-pixel RNG is RandomTEA seeded
// (frame_index, pixel_index); progressive accumulation as in the ray-march kernel.
#pragma once
#include "sampling.cuh"

namespace b200k {

struct DdaIter {
  f3 t_next;
  i3 cell;
  float next_cell_begin;

  // DDAIter::init (accel/dda.h:36-59); org/dir in macrocell units, t is the ray's own parameter
  __device__ __forceinline__ void init(f3 org, f3 dir, float t_min, i3 grid)
  {
    const f3 o = org + t_min * dir;
    const f3 fc = mk3(fmaxf(0.f, fminf((float)grid.x - 1.f, floorf(o.x))), fmaxf(0.f, fminf((float)grid.y - 1.f, floorf(o.y))),
                      fmaxf(0.f, fminf((float)grid.z - 1.f, floorf(o.z))));
    const f3 fe = mk3(dir.x > 0.f ? fc.x + 1.f : fc.x, dir.y > 0.f ? fc.y + 1.f : fc.y, dir.z > 0.f ? fc.z + 1.f : fc.z);
    const f3 ts = mk3(fabsf(1.f / dir.x), fabsf(1.f / dir.y), fabsf(1.f / dir.z));
    t_next = mk3(dir.x == 0.f ? FLT_MAX : __fmul_rn(fabsf(fe.x - o.x), ts.x), dir.y == 0.f ? FLT_MAX : __fmul_rn(fabsf(fe.y - o.y), ts.y),
                 dir.z == 0.f ? FLT_MAX : __fmul_rn(fabsf(fe.z - o.z), ts.z));
    cell.x = (int)fc.x;
    cell.y = (int)fc.y;
    cell.z = (int)fc.z;
    next_cell_begin = 0.f;
  }
};

// DeltaTrackingIter<true>::hashit (accel/spatial_partition.h:70-95) with DDAIter::next (accel/dda.h:61-132) inlined.
// As in the reference, `t` is not advanced across cells whose majorant is <= FLT_EPSILON.
__device__ __forceinline__ bool dda_hashit(const LaunchParams& P, DdaIter& it, f3 m_dir, float tnear, float tfar, float& rayt,
                                           float& majorant, float& tau)
{
  const i3 grid = P.mc;
  const i3 stop = { m_dir.x > 0.f ? grid.x : -1, m_dir.y > 0.f ? grid.y : -1, m_dir.z > 0.f ? grid.z : -1 };
  const f3 ts = mk3(fabsf(1.f / m_dir.x), fabsf(1.f / m_dir.y), fabsf(1.f / m_dir.z));
  const i3 delta = { m_dir.x > 0.f ? 1 : -1, m_dir.y > 0.f ? 1 : -1, m_dir.z > 0.f ? 1 : -1 };
  float t = it.next_cell_begin + tnear;
  for (;;) {
    if (it.cell.x == stop.x || it.cell.y == stop.y || it.cell.z == stop.z) return false;
    const float t_closest = fminf(fminf(it.t_next.x, it.t_next.y), it.t_next.z);
    const float cell_t0 = fmaxf(tnear + it.next_cell_begin, tnear);
    const float cell_t1 = fminf(tnear + t_closest, tfar);
    if (cell_t0 >= cell_t1) return false;
    // ---- lambda body ----
    bool go_on = true, found = false;
    majorant = __ldg(P.majorants + (it.cell.x + grid.x * (it.cell.y + grid.y * it.cell.z)));
    if (!(fabsf(majorant) <= FLT_EPSILON)) {
      tau = __fsub_rn(tau, __fmul_rn(cell_t1 - t, majorant));
      t = cell_t1;
      if (!(tau > 0.f)) {
        t = t + tau / majorant;
        found = true;
        it.next_cell_begin = t - tnear;
        rayt = t;
        go_on = false;
      }
    }
    // ---- advance ----
    if (go_on || fmaxf(tnear + it.next_cell_begin, tnear) >= cell_t1) {
      if (it.t_next.x == t_closest) { it.t_next.x += ts.x; it.cell.x += delta.x; if (it.cell.x == stop.x) return found; }
      if (it.t_next.y == t_closest) { it.t_next.y += ts.y; it.cell.y += delta.y; if (it.cell.y == stop.y) return found; }
      if (it.t_next.z == t_closest) { it.t_next.z += ts.z; it.cell.z += delta.z; if (it.cell.z == stop.z) return found; }
      it.next_cell_begin = t_closest;
    }
    if (!go_on) return found;
  }
}

// delta_tracking, use_dda == 1 (shaders_pathtracing.cu:385-404); ray already in object space
template<int SAMPLER>
__device__ __forceinline__ bool delta_tracking(const LaunchParams& P, const TfShared& tf, Tea& rng, f3 org, f3 dir, float t_min,
                                               float t_max, float& t_hit, f3& albedo, unsigned int& n_samples)
{
  const f3 m_org = org * P.spac_rcp, m_dir = dir * P.spac_rcp;
  DdaIter it;
  it.init(m_org, m_dir, t_min, P.mc);
  float tau = -logf(1.f - rng.draw().x);
  float t = t_min, majorant = 0.f;
  while (dda_hashit(P, it, m_dir, t_min, t_max, t, majorant, tau)) {
    const f3 p = org + t * dir;
    const float s = sample_volume<SAMPLER>(P, p);
    n_samples++;
    const float a = classify_alpha(P, tf, s); // no opacity correction: extinction = alpha per unit length
    if (rng.draw().x * majorant < a) {
      albedo = classify_color(P, tf, s);
      t_hit = t;
      return true;
    }
    tau = -logf(1.f - rng.draw().x);
  }
  return false;
}

// uniform_sample_sphere (shaders_common.h:339-354); the reference evaluates 2*M_PI*s.x in double
__device__ __forceinline__ f3 uniform_sphere(float2 s)
{
  const float phi = (float)(2.0 * 3.14159265358979323846 * (double)s.x);
  const float cos_t = 1.f - 2.f * s.y;
  const float sin_t = 2.f * sqrtf(s.y * (1.f - s.y));
  float sp, cp;
  sincosf(phi, &sp, &cp);
  return mk3(cp * sin_t, sp * sin_t, cos_t);
}

constexpr int PT_MAX_STACK = 16; // max_num_scatters = 24 allows 12 stored bounces (scatter_index advances by 2)

template<int SAMPLER>
__global__ void __launch_bounds__(RENDER_THREADS)
pathtrace_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const TfShared tf = stage_transfer_function(P, smem);
  const int lane = threadIdx.x & 31;
  const f3 oo = xfm_point(P.wto, P.cam_pos);
  unsigned long long n_delta = 0, n_rays = 0;

  for (;;) {
    int k = 0;
    if (lane == 0) k = (int)atomicAdd(P.tile_counter, 1u);
    k = __shfl_sync(0xffffffffu, k, 0);
    if (k >= P.n_warp_tiles) break;
    int ix, iy;
    const int tile_slot = warp_tile_pixel(P, k, lane, ix, iy);
    const long long tile_t0 = clock64();
    unsigned int samples = 0, rays = 0;
    if (ix < P.width && iy < P.height) {
      const unsigned int pixel_index = (unsigned)ix + (unsigned)iy * (unsigned)P.width;
      Tea rng((unsigned)P.frame_index, pixel_index); // shaders_pathtracing.cu:644
      float out_a = 0.f;
      f3 out_c = mk3(0.f);
      for (int sidx = 0; sidx < P.spp; sidx++) {
        float sx = ((float)ix + .5f) * P.rw, sy = ((float)iy + .5f) * P.rh;
        if (P.spp > 1) {
          const float2 j = rng.draw();
          sx += (j.x - 0.5f) * P.rw;
          sy += (j.y - 0.5f) * P.rh;
        }
        const f3 cam_dir = normalize3(P.cam_dir + (sx - 0.5f) * P.cam_hor + (sy - 0.5f) * P.cam_ver);
        rays++;
        // pathtracing (shaders_pathtracing.cu:477-542), recursion unrolled
        f3 org = oo, dir = xfm_vector(P.wto.l, cam_dir);
        float t0 = 0.f, t1 = FLT_MAX;
        f3 le = mk3(0.f);
        if (intersect_box(t0, t1, org, dir, mk3(0.f), mk3(1.f))) {
          out_a += 1.f;
          f3 stack[PT_MAX_STACK];
          int depth = 0, scatter_index = 0;
          f3 terminal = mk3(0.f);
          for (;;) {
            float t;
            f3 albedo;
            if (!delta_tracking<SAMPLER>(P, tf, rng, org, dir, t0, t1, t, albedo, samples)) {
              if (scatter_index != 0) terminal = mk3(P.light_ambient);
              break;
            }
            scatter_index++;
            if (scatter_index > P.max_num_scatters || depth >= PT_MAX_STACK) break;
            const f3 dir_w = uniform_sphere(rng.draw());
            org = (depth == 0) ? xfm_point(P.wto, P.cam_pos + t * cam_dir) : org + t * dir;
            dir = xfm_vector(P.wto.l, dir_w);
            stack[depth++] = albedo;
            scatter_index += 1;
            t0 = 0.f;
            t1 = FLT_MAX;
            // q14: under OptiX the scattered ray passes through the instance transform again before the clip; the walk
            // itself runs along the ray as handed over, inside the clip's t range
            if (P.pt_child_instance_xfm) {
              if (!intersect_box(t0, t1, xfm_point(P.wto, org), xfm_vector(P.wto.l, dir), mk3(0.f), mk3(1.f))) break;
            }
            else if (!intersect_box(t0, t1, org, dir, mk3(0.f), mk3(1.f))) break;
          }
          le = terminal;
          for (int i = depth - 1; i >= 0; i--) le = stack[i] * le;
        }
        out_c = out_c + le;
      }
      const float rspp = 1.f / (float)P.spp;
      float4 px4 = make_float4(out_c.x * rspp, out_c.y * rspp, out_c.z * rspp, out_a * rspp);
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
      P.rgba[pixel_index] = px4;
      float* g = P.grad + 3 * (size_t)pixel_index; // output.gradient stays 0 in the path tracer
      g[0] = 0.f;
      g[1] = 0.f;
      g[2] = 0.f;
    }
    n_delta += __reduce_add_sync(0xffffffffu, samples);
    n_rays += __reduce_add_sync(0xffffffffu, rays);
    if (lane == 0) tile_cost_add(P, tile_slot, tile_t0);
  }
  if (lane == 0) {
    atomicAdd(&P.counters->delta_samples, n_delta);
    atomicAdd(&P.counters->rays, n_rays);
  }
}

} // namespace b200k
