// Fused ray-march kernel for sm_100a: ray generation + ray/box clip + fixed-step march + TF classification +
// gradient shading + shadow march + front-to-back compositing + frame write / progressive accumulation.
// Replaces the reference's OptiX pipeline K1-K5 (ovr/devices/optix7/shaders_raymarching.cu:44-171,178-253,260-413,
// shaders_common.h:156-215,379-451): one thread per pixel, one 8x4-pixel tile per warp, warps pull tiles from a global
// counter (persistent CTAs, grid = SMs x resident CTAs), the transfer function lives in shared memory.
//
// Empty-space skipping (SKIP): a sample whose macrocell majorant is 0 has alpha == 0 exactly and adds +0 to every
// accumulator, so it is never fetched; t still advances by the reference's running fp32 sum (`t.x = t.y;
// t.y = min(t.x + step, t_max)`, shaders_raymarching.cu:168-169) so the sample lattice is the reference's bit for bit.
#pragma once
#include "sampling.cuh"

namespace b200k {

enum { SHADING_NONE = 0, SHADING_REF = 1 };

struct RayStats { unsigned int primary, shadow, gradient; };

// raymarching_shadow + __closesthit__volume_shadow (shaders_raymarching.cu:44-85,205-229)
template<int SAMPLER, bool SKIP>
__device__ __forceinline__ float shadow_march(const LaunchParams& P, const TfShared& tf, f3 org, RayStats& st)
{
  const f3 dir = P.light_dir;
  const f3 oo = xfm_point(P.wto, org), od = xfm_vector(P.wto.l, dir);
  float t0 = 0.f, t1 = FLT_MAX;
  if (!intersect_box(t0, t1, oo, od, mk3(0.f), mk3(1.f))) return 0.f;
  const float sstep = (P.step * 10.f) * P.step; // q1: sampling_scale * self.step, sampling_scale = step * 10
  f3 od_rcp;
  if (SKIP) od_rcp = mk3(1.f / od.x, 1.f / od.y, 1.f / od.z);
  float alpha = 0.f;
  float tx = t0, ty = fminf(t1, t0 + sstep);
  while (ty > tx && alpha < 0.9999f) {
    const float tm = 0.5f * (tx + ty);
    const f3 p = xfm_point(P.wto, org + tm * dir);
    bool live = true;
    if (SKIP) {
      i3 c;
      const int ci = cell_of(P, p, c);
      if (__ldg(P.majorants + ci) == 0.f) {
        live = false;
        const float t_exit = cell_exit_t(P, c, oo, od_rcp, od);
        do {
          tx = ty;
          ty = fminf(tx + sstep, t1);
        } while (ty > tx && 0.5f * (tx + ty) < t_exit);
      }
    }
    if (live) {
      const float s = sample_volume<SAMPLER>(P, p);
      st.shadow++;
      const float a = correct_alpha(classify_alpha(P, tf, s), P.base, ty - tx);
      alpha += (1.f - alpha) * a;
      tx = ty;
      ty = fminf(tx + sstep, t1);
    }
  }
  return alpha;
}

// known-answer probe: the kernels' own sampling + classification at arbitrary object-space points
template<int SAMPLER>
__global__ void probe_kernel(const __grid_constant__ LaunchParams P, const float* __restrict__ pts, long long n,
                             float* __restrict__ values, float4* __restrict__ rgba)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const TfShared tf = stage_transfer_function(P, smem);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float s = sample_volume<SAMPLER>(P, mk3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
    values[i] = s;
    if (rgba) {
      const f3 c = classify_color(P, tf, s);
      rgba[i] = make_float4(c.x, c.y, c.z, classify_alpha(P, tf, s));
    }
  }
}

// 1-D known-answer probe: hardware tex1D vs the software filter used for the shared-memory transfer function
template<int CH>
__global__ void probe_tex1d_kernel(cudaTextureObject_t tex, const float* __restrict__ table, int n, const float* __restrict__ u,
                                   long long n_u, float* __restrict__ hw, float* __restrict__ sw)
{
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_u; i += (long long)gridDim.x * blockDim.x) {
    int i0, i1, a;
    hw_coord(u[i], n, i0, i1, a);
    if (CH == 1) {
      hw[i] = tex1D<float>(tex, u[i]);
      sw[i] = lerp_hw(table[i0], table[i1], a);
    }
    else {
      const float4 t = tex1D<float4>(tex, u[i]);
      hw[4 * i] = t.x, hw[4 * i + 1] = t.y, hw[4 * i + 2] = t.z, hw[4 * i + 3] = t.w;
      for (int c = 0; c < 4; c++) sw[4 * i + c] = lerp_hw(table[4 * i0 + c], table[4 * i1 + c], a);
    }
  }
}

// fetch-rate micro-benchmark (the P_fetch denominator of SURVEY 8d): every warp marches an 8x4 bundle of parallel rays,
// one voxel apart, `iters` unit steps through a cubic footprint of the resident volume; nothing but the fetch and an add.
template<int SAMPLER>
__global__ void __launch_bounds__(RENDER_THREADS)
fetch_bench_kernel(const __grid_constant__ LaunchParams P, int footprint, int iters, float dx, float dy, float dz, float* __restrict__ sink)
{
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  // bundle origin: hashed position inside the footprint, rays offset in the plane orthogonal-ish to the direction
  unsigned int hsh = (unsigned)warp * 2654435761u;
  const float fx = (float)(hsh % (unsigned)footprint), fy = (float)((hsh >> 8) % (unsigned)footprint), fz = (float)((hsh >> 16) % (unsigned)footprint);
  float x = fx + (float)(lane & 7), y = fy + (float)(lane >> 3), z = fz;
  const float fp = (float)footprint;
  float acc = 0.f;
  for (int i = 0; i < iters; i++) {
    acc += sample_volume<SAMPLER>(P, mk3((x + 0.37f) * P.rdim.x, (y + 0.41f) * P.rdim.y, (z + 0.29f) * P.rdim.z));
    x += dx; y += dy; z += dz;
    if (x >= fp) x -= fp;
    if (y >= fp) y -= fp;
    if (z >= fp) z -= fp;
  }
  if (acc == 123.456f) sink[0] = acc; // keeps the loop alive
}

struct MarchResult { float alpha; f3 color, gradient; };

// raymarching (shaders_raymarching.cu:87-171)
template<int SHADING, int SAMPLER, bool SKIP>
__device__ __forceinline__ void primary_march(const LaunchParams& P, const TfShared& tf, f3 org, f3 dir, f3 oo, f3 od,
                                              float t_min, float t_max, MarchResult& r, RayStats& st)
{
  if (t_min >= t_max) return;
  const float step = P.step;
  f3 od_rcp;
  if (SKIP) od_rcp = mk3(1.f / od.x, 1.f / od.y, 1.f / od.z);
  float tx = t_min, ty = fminf(t_max, t_min + step);
  while (ty > tx && r.alpha < 0.9999f) {
    const float tm = 0.5f * (tx + ty);
    const f3 pos = org + tm * dir;
    const f3 p = xfm_point(P.wto, pos);
    if (SKIP) {
      i3 c;
      const int ci = cell_of(P, p, c);
      if (__ldg(P.majorants + ci) == 0.f) {
        const float t_exit = cell_exit_t(P, c, oo, od_rcp, od);
        do {
          tx = ty;
          ty = fminf(tx + step, t_max);
        } while (ty > tx && 0.5f * (tx + ty) < t_exit);
        continue;
      }
    }
    const float s = sample_volume<SAMPLER>(P, p);
    st.primary++;
    const float a = correct_alpha(classify_alpha(P, tf, s), P.base, ty - tx);
    if (!SKIP || a != 0.f) {
      f3 rgb = classify_color(P, tf, s);
      f3 normal_c = mk3(0.f);
      if (SHADING == SHADING_REF) {
        // compute_volume_gradient_object_space (shaders_common.h:195-215): forward differences, step 1/dims
        f3 stp = P.rdim;
        if (p.x + stp.x > 1.f) stp.x *= -1.f;
        if (p.y + stp.y > 1.f) stp.y *= -1.f;
        if (p.z + stp.z > 1.f) stp.z *= -1.f;
        const float gx = sample_volume<SAMPLER>(P, mk3(p.x + stp.x, p.y, p.z)) - s;
        const float gy = sample_volume<SAMPLER>(P, mk3(p.x, p.y + stp.y, p.z)) - s;
        const float gz = sample_volume<SAMPLER>(P, mk3(p.x, p.y, p.z + stp.z)) - s;
        st.gradient += 3;
        const f3 g = mk3(gx, gy, gz) / stp;
        const f3 normal_o = -normalize3(g);
        // xfmNormal(otw, n) = transpose(inverse(otw.l)) * n = transpose(wto.l) * n
        const f3 normal_w = normalize3(xfm_vector(transposed(P.wto.l), normal_o));
        normal_c = normalize3(xfm_vector(P.wtc, normal_w)); // wtc holds transpose(inverse(world_to_camera.l))
        const float sa = shadow_march<SAMPLER, SKIP>(P, tf, pos, st);
        const float cosNL = fabsf(dot3(P.light_dir, normal_w));
        const float k = 0.5f + ((0.5f * cosNL) * 2.f) * (1.f - sa); // light_rgb = 2 (shaders_raymarching.cu:138,157)
        rgb = rgb * k;
      }
      const float tr = 1.f - r.alpha;
      r.color = r.color + (tr * clamp01(rgb)) * a;
      r.gradient = r.gradient + (tr * clamp01(normal_c)) * a;
      r.alpha += tr * a;
    }
    tx = ty;
    ty = fminf(tx + step, t_max);
  }
}

template<int SHADING, int SAMPLER, bool SKIP>
__global__ void __launch_bounds__(RENDER_THREADS)
raymarch_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const TfShared tf = stage_transfer_function(P, smem);
  const int lane = threadIdx.x & 31;
  const f3 oo = xfm_point(P.wto, P.cam_pos);
  unsigned long long n_primary = 0, n_shadow = 0, n_gradient = 0, n_rays = 0; // warp totals, kept by lane 0

  for (;;) {
    int k = 0;
    if (lane == 0) k = (int)atomicAdd(P.tile_counter, 1u);
    k = __shfl_sync(0xffffffffu, k, 0);
    if (k >= P.n_warp_tiles) break;
    int ix, iy;
    const int tile_slot = warp_tile_pixel(P, k, lane, ix, iy);
    const long long tile_t0 = clock64();
    RayStats st = { 0u, 0u, 0u };
    unsigned int rays = 0;
    if (ix < P.width && iy < P.height) {
    const unsigned int pixel_index = (unsigned)ix + (unsigned)iy * (unsigned)P.width;
    Tea rng((unsigned)P.frame_index, pixel_index);

    float out_a = 0.f;
    f3 out_c = mk3(0.f), out_g = mk3(0.f);
    for (int sidx = 0; sidx < P.spp; sidx++) {
      // compute_screen_position + pixel jitter (shaders_common.h:400-404, shaders_raymarching.cu:351-362)
      float sx = ((float)ix + .5f) * P.rw, sy = ((float)iy + .5f) * P.rh;
      if (P.spp > 1) {
        const float2 j = rng.draw();
        sx += (j.x - 0.5f) * P.rw;
        sy += (j.y - 0.5f) * P.rh;
      }
      const f3 dir = normalize3(P.cam_dir + (sx - 0.5f) * P.cam_hor + (sy - 0.5f) * P.cam_ver);
      rays++;
      const f3 od = xfm_vector(P.wto.l, dir);
      MarchResult r;
      r.alpha = 0.f;
      r.color = mk3(0.f);
      r.gradient = mk3(0.f);
      float t0 = 0.f, t1 = FLT_MAX;
      if (intersect_box(t0, t1, oo, od, P.clip_lo, P.clip_hi))
        primary_march<SHADING, SAMPLER, SKIP>(P, tf, P.cam_pos, dir, oo, od, t0, t1, r, st);
      // render_raymarching (shaders_raymarching.cu:260-321): background is always 0 -> alpha_blend = C / A
      out_a += r.alpha;
      if (P.premultiplied) {
        out_c = out_c + r.color;
        out_g = out_g + r.gradient;
      }
      else if (r.alpha > 0.f) {
        out_c = out_c + r.color / r.alpha;
        out_g = out_g + r.gradient / r.alpha;
      }
    }
    const float rspp = 1.f / (float)P.spp;
    float4 px4 = make_float4(out_c.x * rspp, out_c.y * rspp, out_c.z * rspp, out_a * rspp);
    out_g = out_g * rspp;
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
    P.rgba[pixel_index] = px4;
    float* g = P.grad + 3 * (size_t)pixel_index;
    g[0] = out_g.x;
    g[1] = out_g.y;
    g[2] = out_g.z;
    } // in-bounds pixel
    // executed-sample counters: reduced per tile (32-bit is ample for one tile), summed in 64 bits
    n_primary += __reduce_add_sync(0xffffffffu, st.primary);
    n_shadow += __reduce_add_sync(0xffffffffu, st.shadow);
    n_gradient += __reduce_add_sync(0xffffffffu, st.gradient);
    n_rays += __reduce_add_sync(0xffffffffu, rays);
    if (lane == 0) tile_cost_add(P, tile_slot, tile_t0);
  }

  if (lane == 0) { // one atomic set per warp
    atomicAdd(&P.counters->primary_samples, n_primary);
    atomicAdd(&P.counters->shadow_samples, n_shadow);
    atomicAdd(&P.counters->gradient_fetches, n_gradient);
    atomicAdd(&P.counters->rays, n_rays);
  }
}

} // namespace b200k
