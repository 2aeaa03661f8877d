// Ray-march kernel, second generation: per-lane STATE MACHINE (primary march <-> shadow march) instead of nested loops.
//
// Why (profiles/r1a_*: the nested-loop kernel ran C2 at 13 % of the texture roofline): it executed 145 thread-instructions
// per trilinear fetch at 55 % warp efficiency -- instruction-issue bound, not texture bound. A lane whose sample had
// alpha == 0 idled through its neighbours' whole shadow march, and every shadow sample paid two 3x4 transforms, six
// clamps, a cell lookup with divisions, a generic 64-bit texture-coordinate emulation and a pow.
//
// What changed, none of it moving a sample (the t lattices of primary and shadow rays are still the reference's running
// fp32 sums, shaders_raymarching.cu:64-66,168-169; parity bar: 1/255 max-abs, 45 dB):
//   * every lane is either in PRIMARY or in SHADOW state; one loop iteration advances whichever it is in, so lanes that
//     need no shadow ray keep marching while their neighbours shade (warp efficiency), and the loop ends when all 32 rays
//     of the 8x4 tile are done;
//   * rays live in object space: a sample position is 3 FMAs (o' + t d'), the shadow direction d'_L = wto.l * L is a
//     kernel-wide constant;
//   * the shadow march accumulates log2-transmittance: (1-A) = prod (1-a_k)^(adj_k) = exp2(sum adj_k * lg2(1-a_k)), the
//     same product the reference forms with one __powf per sample (shaders_raymarching.cu:73-79), minus the ex2;
//     early termination A >= 0.9999 becomes sum <= log2(1 - 0.9999f);
//   * tex3D is fetched without software clamps (positions lie in the box; the unit's clamp addressing does the rest);
//   * TF lookups use the power-of-two fast path of the measured texture-unit coordinate model and a shared-memory table
//     of (value, slope/256) pairs: one LDS.64 + one FMA per alpha lookup;
//   * empty-space test on a padded majorant grid (no clamps, no divisions); primary rays skip whole macrocells.
#pragma once
#include "sampling.cuh"

namespace b200k {

struct TfFast {
  const float2* apair;  // [n_alpha + 1]: (alpha[i], (alpha[i+1] - alpha[i]) / 256)
  const float4* color;  // [n_color + 1]: colour table, last entry replicated
};

__host__ __device__ __forceinline__ size_t tf_fast_shared_bytes(int n_color, int n_alpha)
{
  return sizeof(float4) * (size_t)(n_color + 1) + sizeof(float2) * (size_t)(n_alpha + 1);
}

__device__ __forceinline__ TfFast stage_tf_fast(const LaunchParams& P, unsigned char* smem)
{
  float4* c = reinterpret_cast<float4*>(smem);
  float2* a = reinterpret_cast<float2*>(smem + sizeof(float4) * (size_t)(P.n_color + 1));
  for (int i = threadIdx.x; i <= P.n_color; i += blockDim.x) c[i] = P.tf_color[min(i, P.n_color - 1)];
  for (int i = threadIdx.x; i <= P.n_alpha; i += blockDim.x) {
    const float v0 = P.tf_alpha[min(i, P.n_alpha - 1)], v1 = P.tf_alpha[min(i + 1, P.n_alpha - 1)];
    a[i] = make_float2(v0, (v1 - v0) * (1.f / 256.f));
  }
  __syncthreads();
  TfFast t;
  t.apair = a;
  t.color = c;
  return t;
}

// normalised TF coordinate v in [0, 1]: (clamp(s, lower, upper) - lower) * scale, then clamp01 (shaders_common.h:311-319,
// 356-367); the saturate modifier does both clamps
__device__ __forceinline__ float tf_unit(const LaunchParams& P, float s)
{
  return __saturatef((s - P.tf_lower) * P.tf_scale);
}

// texel index i and fraction a (of 256) of the texture unit for array1d_nodal(v) on a table of n entries
__device__ __forceinline__ void tf_coord(float v, int n, float mul, float addf, int addi, int shift, int& i, int& a)
{
  if (shift >= 0) { // n = 2^m: exact integer form of the unit's coordinate conversion
    const int q = (__float2int_rz(__fmaf_rn(v, mul, addf)) + addi) >> shift;
    i = q >> 8;
    a = q & 255;
  }
  else {
    int i1;
    hw_coord(nodal_coord(v, n), n, i, i1, a);
  }
}

__device__ __forceinline__ float alpha_fast(const LaunchParams& P, const TfFast& tf, float s)
{
  int i, a;
  tf_coord(tf_unit(P, s), P.n_alpha, P.tfa_mul, P.tfa_addf, P.tfa_addi, P.tfa_shift, i, a);
  const float2 e = tf.apair[i];
  return __fmaf_rn((float)a, e.y, e.x);
}

// the same lookup through a 32-bit shared-window address (no generic-to-shared conversion in the loop)
__device__ __forceinline__ float alpha_fast_s(const LaunchParams& P, unsigned int apair_addr, float s)
{
  int i, a;
  tf_coord(tf_unit(P, s), P.n_alpha, P.tfa_mul, P.tfa_addf, P.tfa_addi, P.tfa_shift, i, a);
  float ex, ey;
  asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(ex), "=f"(ey) : "r"(apair_addr + 8u * (unsigned int)i));
  return __fmaf_rn((float)a, ey, ex);
}

__device__ __forceinline__ f3 color_fast(const LaunchParams& P, const TfFast& tf, float s)
{
  int i, a;
  tf_coord(tf_unit(P, s), P.n_color, P.tfc_mul, P.tfc_addf, P.tfc_addi, P.tfc_shift, i, a);
  const float4 c0 = tf.color[i], c1 = tf.color[i + 1];
  const float w = (float)a * (1.f / 256.f);
  return mk3(__fmaf_rn(w, c1.x - c0.x, c0.x), __fmaf_rn(w, c1.y - c0.y, c0.y), __fmaf_rn(w, c1.z - c0.z, c0.z));
}

__device__ __forceinline__ float lg2_fast(float x)
{
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float ex2_fast(float x)
{
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// macrocell of an object-space position on the padded grid (cell c sits at c + 1 on every axis)
__device__ __forceinline__ int pad_cell(const LaunchParams& P, f3 p, i3& c)
{
  c.x = min((unsigned)__float2int_rz(p.x * P.cells_per_unit.x), (unsigned)P.mc.x);
  c.y = min((unsigned)__float2int_rz(p.y * P.cells_per_unit.y), (unsigned)P.mc.y);
  c.z = min((unsigned)__float2int_rz(p.z * P.cells_per_unit.z), (unsigned)P.mc.z);
  return (c.x + 1) + (c.y + 1) * P.mcp_sy + (c.z + 1) * P.mcp_sz;
}

// Same index for a position given in CELL units minus one half (x = p * cells_per_unit - 0.5), without any conversion
// instruction: x + 1.5 * 2^23 rounds x to the nearest integer n (= floor of the cell coordinate) and leaves 0x4B400000 + n
// in the float's bits; the three magic constants and the +1 pad offsets are folded into P.mcp_bias. n may be -1 .. mc
// (positions a rounding error outside the box): that is what the pad ring is for; one clamp guards everything else.
__device__ __forceinline__ unsigned int pad_index_magic(const LaunchParams& P, float xh, float yh, float zh)
{
  const unsigned int bx = __float_as_uint(xh + 12582912.f), by = __float_as_uint(yh + 12582912.f), bz = __float_as_uint(zh + 12582912.f);
  return min(bx + by * (unsigned)P.mcp_sy + bz * (unsigned)P.mcp_sz - P.mcp_bias, P.mcp_last);
}

constexpr int SM_MIN_BLOCKS = 4; // 128 threads x 4 CTAs/SM wanted at least (register cap 128/thread; ptxas reports the real use)

// kernel-wide constants of the shadow rays: direction L in object space and its slab-test reciprocals
struct ShadowDir {
  f3 d, rcp;
  bool sx, sy, sz; // |d| < FLT_MIN on that axis (intersect_box, shaders_common.h:163-171)
};

// intersect_box for a ray starting at `o` with the constant direction: same arithmetic as device_math.cuh::intersect_box
__device__ __forceinline__ bool clip_shadow(const ShadowDir& S, f3 o, float& t0, float& t1)
{
  const float lx = S.sx ? FLT_MAX : __fmul_rn(0.f - o.x, S.rcp.x), hx = S.sx ? -FLT_MAX : __fmul_rn(1.f - o.x, S.rcp.x);
  const float ly = S.sy ? FLT_MAX : __fmul_rn(0.f - o.y, S.rcp.y), hy = S.sy ? -FLT_MAX : __fmul_rn(1.f - o.y, S.rcp.y);
  const float lz = S.sz ? FLT_MAX : __fmul_rn(0.f - o.z, S.rcp.z), hz = S.sz ? -FLT_MAX : __fmul_rn(1.f - o.z, S.rcp.z);
  const float nr = fmaxf(fmaxf(fminf(lx, hx), fminf(ly, hy)), fminf(lz, hz));
  const float fr = fminf(fminf(fmaxf(lx, hx), fmaxf(ly, hy)), fmaxf(lz, hz));
  t0 = fmaxf(0.f, nr);
  t1 = fminf(FLT_MAX, fr);
  return t1 > t0;
}

// the same test with the host-derived constants of LaunchParams (constant-bank operands)
__device__ __forceinline__ bool clip_shadow_p(const LaunchParams& P, f3 o, float& t0, float& t1)
{
  const float lx = P.sh_deg_x ? FLT_MAX : __fmul_rn(0.f - o.x, P.sh_rcp.x), hx = P.sh_deg_x ? -FLT_MAX : __fmul_rn(1.f - o.x, P.sh_rcp.x);
  const float ly = P.sh_deg_y ? FLT_MAX : __fmul_rn(0.f - o.y, P.sh_rcp.y), hy = P.sh_deg_y ? -FLT_MAX : __fmul_rn(1.f - o.y, P.sh_rcp.y);
  const float lz = P.sh_deg_z ? FLT_MAX : __fmul_rn(0.f - o.z, P.sh_rcp.z), hz = P.sh_deg_z ? -FLT_MAX : __fmul_rn(1.f - o.z, P.sh_rcp.z);
  const float nr = fmaxf(fmaxf(fminf(lx, hx), fminf(ly, hy)), fminf(lz, hz));
  const float fr = fminf(fminf(fmaxf(lx, hx), fmaxf(ly, hy)), fmaxf(lz, hz));
  t0 = fmaxf(0.f, nr);
  t1 = fminf(FLT_MAX, fr);
  return t1 > t0;
}

template<int SHADING, bool SKIP>
__global__ void __launch_bounds__(RENDER_THREADS, SM_MIN_BLOCKS)
raymarch_sm_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const TfFast tf = stage_tf_fast(P, smem);
  const int lane = threadIdx.x & 31;
  const f3 oo = xfm_point(P.wto, P.cam_pos);
  const float step = P.step;
  const float sstep = (P.step * 10.f) * P.step; // q1: sampling_scale * self.step with sampling_scale = step * 10
  const float log_ert = lg2_fast(1.f - 0.9999f); // A >= 0.9999  <=>  log2(1 - A) <= log2(1 - 0.9999f)
  ShadowDir S;
  S.d = xfm_vector(P.wto.l, P.light_dir);
  S.rcp = mk3(__frcp_rn(S.d.x), __frcp_rn(S.d.y), __frcp_rn(S.d.z));
  S.sx = fabsf(S.d.x) < FLT_MIN, S.sy = fabsf(S.d.y) < FLT_MIN, S.sz = fabsf(S.d.z) < FLT_MIN;
  const lin3 wto_t = transposed(P.wto.l); // xfmNormal(otw, n) = transpose(inverse(otw.l)) * n
  unsigned long long n_primary = 0, n_shadow = 0, n_gradient = 0, n_rays = 0; // warp totals, kept by lane 0

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
    Tea rng((unsigned)P.frame_index, pixel_index);
    unsigned int c_primary = 0, c_shadow = 0, c_shaded = 0, c_rays = 0;
    float out_a = 0.f;
    f3 out_c = mk3(0.f), out_g = mk3(0.f);

    for (int sidx = 0; sidx < P.spp; sidx++) {
      // ---- ray generation + clip (shaders_common.h:400-404, shaders_raymarching.cu:351-362, shaders_common.h:156-184) ----
      float sx = ((float)ix + .5f) * P.rw, sy = ((float)iy + .5f) * P.rh;
      if (P.spp > 1) {
        const float2 j = rng.draw();
        sx += (j.x - 0.5f) * P.rw;
        sy += (j.y - 0.5f) * P.rh;
      }
      const f3 dir = normalize3(P.cam_dir + (sx - 0.5f) * P.cam_hor + (sy - 0.5f) * P.cam_ver);
      const f3 od = xfm_vector(P.wto.l, dir);
      float tx = 0.f, t1 = FLT_MAX;
      bool active;
      float own_exit = FLT_MAX; // sort-last brick: samples with midpoint in [b0, own_exit) are this handle's
      float b0 = 0.f;
      if (!P.brick) active = inside && intersect_box(tx, t1, oo, od, P.clip_lo, P.clip_hi);
      else {
        // Both intervals come from WORLD-space slab tests on the world ray: every rank then computes bit-identical
        // parameters for a shared cut plane, so the bricks partition the lattice samples exactly (no sample is rendered
        // twice or dropped), and all ranks walk the same lattice.
        active = inside && intersect_box(tx, t1, P.cam_pos, dir, P.lat_lo, P.lat_hi);
        active = intersect_box(b0, own_exit, P.cam_pos, dir, P.own_lo, P.own_hi) && active;
      }
      if (inside) c_rays++;
      float ty = fminf(t1, tx + step);
      active = active && ty > tx;
      if (P.brick && active && 0.5f * (tx + ty) < b0) { // walk the lattice up to the brick: the same running sum as the march itself
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
        active = ty > tx;
      }
      const f3 od_rcp = mk3(1.f / od.x, 1.f / od.y, 1.f / od.z);
      float A = 0.f;
      f3 C = mk3(0.f), G = mk3(0.f);
      // macrocell the primary ray is in: valid while the sample parameter stays below cell_exit
      float cell_exit = -1.f;
      bool cell_empty = false;
      // shadow state + the shading inputs waiting for it
      bool in_shadow = false;
      float stx = 0.f, sty = 0.f, st1 = 0.f, slog = 0.f;
      f3 so = mk3(0.f), pend_rgb = mk3(0.f);
      float pend_w = 0.f, pend_cos = 0.f;

      // shade + composite the pending primary sample once its shadow ray is done (shaders_raymarching.cu:156-169)
      auto finish_shading = [&]() {
        const float k_sh = 0.5f + ((0.5f * pend_cos) * 2.f) * ex2_fast(slog); // light_rgb = 2; 1 - shadow.alpha = 2^slog
        C = C + clamp01(pend_rgb * k_sh) * pend_w;
        in_shadow = false;
        tx = ty;
        ty = fminf(tx + step, t1);
        active = ty > tx && A < 0.9999f;
      };

      while (__any_sync(0xffffffffu, active)) {
        if (active) {
          do {
            if (SHADING == SHADING_REF && in_shadow) {
              // ---- one shadow sample (raymarching_shadow, shaders_raymarching.cu:44-85); invariant on entry:
              //      sty > stx and slog > log_ert, i.e. the reference's loop condition holds ----
              const float tm = 0.5f * (stx + sty);
              const f3 p = mk3(__fmaf_rn(tm, S.d.x, so.x), __fmaf_rn(tm, S.d.y, so.y), __fmaf_rn(tm, S.d.z, so.z));
              bool live = true;
              if (SKIP) {
                i3 c;
                live = __ldg(P.maj_pad + pad_cell(P, p, c)) != 0.f;
              }
              if (live) {
                const float s = tex3D<float>(P.tex, p.x, p.y, p.z);
                c_shadow++;
                const float a = alpha_fast(P, tf, s);
                // (1 - A) *= (1 - a)^(base * dt): the reference's opacity correction + compositing in the log domain
                slog = __fmaf_rn(P.base * (sty - stx), lg2_fast(fmaxf(1.f - a, 0.f)), slog);
              }
              stx = sty;
              sty = fminf(stx + sstep, st1);
              if (!(sty > stx && slog > log_ert)) finish_shading();
              break;
            }
            // ---- one primary sample (raymarching, shaders_raymarching.cu:87-171) ----
            const float tm = 0.5f * (tx + ty);
            if (!(tm < own_exit)) { // the rest of the lattice belongs to the bricks behind this one
              active = false;
              break;
            }
            const f3 p = mk3(__fmaf_rn(tm, od.x, oo.x), __fmaf_rn(tm, od.y, oo.y), __fmaf_rn(tm, od.z, oo.z));
            if (SKIP) {
              if (!(tm < cell_exit)) {
                i3 c;
                cell_empty = __ldg(P.maj_pad + pad_cell(P, p, c)) == 0.f;
                cell_exit = cell_exit_t(P, c, oo, od_rcp, od);
              }
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
                active = ty > tx;
                break;
              }
            }
            const float s = tex3D<float>(P.tex, p.x, p.y, p.z);
            c_primary++;
            float a = alpha_fast(P, tf, s);
            {
              const float adj = P.base * (ty - tx); // opacity correction (shaders_raymarching.cu:118-122)
              if (!(fabsf(adj - 1.f) < 1e-7f)) a = clamp01(1.f - __powf(1.f - a, adj));
            }
            if (!SKIP || a != 0.f) {
              const f3 rgb = color_fast(P, tf, s);
              const float w = (1.f - A) * a;
              if (SHADING == SHADING_REF) {
                // compute_volume_gradient_object_space (shaders_common.h:195-215): forward differences, step 1/dims
                f3 stp = P.rdim;
                if (p.x + stp.x > 1.f) stp.x = -stp.x;
                if (p.y + stp.y > 1.f) stp.y = -stp.y;
                if (p.z + stp.z > 1.f) stp.z = -stp.z;
                const float gx = tex3D<float>(P.tex, p.x + stp.x, p.y, p.z) - s;
                const float gy = tex3D<float>(P.tex, p.x, p.y + stp.y, p.z) - s;
                const float gz = tex3D<float>(P.tex, p.x, p.y, p.z + stp.z) - s;
                c_shaded++;
                const f3 normal_o = -normalize3(mk3(gx, gy, gz) / stp);
                const f3 normal_w = normalize3(xfm_vector(wto_t, normal_o));
                const f3 normal_c = normalize3(xfm_vector(P.wtc, normal_w)); // wtc = transpose(inverse(world_to_camera.l))
                G = G + clamp01(normal_c) * w;
                A += w;
                pend_rgb = rgb;
                pend_w = w;
                pend_cos = fabsf(dot3(P.light_dir, normal_w));
                // shadow ray from the sample toward the light (__closesthit__volume_shadow, shaders_raymarching.cu:205-229)
                so = p;
                slog = 0.f;
                in_shadow = true;
                bool any = clip_shadow(S, so, stx, st1);
                if (any) {
                  sty = fminf(st1, stx + sstep);
                  any = sty > stx;
                }
                if (!any) finish_shading(); // no interval: shadow alpha 0
                break;
              }
              C = C + clamp01(rgb) * w;
              A += w;
            }
            tx = ty;
            ty = fminf(tx + step, t1);
            active = ty > tx && A < 0.9999f;
          } while (0);
        }
      }
      // render_raymarching (shaders_raymarching.cu:260-321): background is always 0 -> alpha_blend = C / A
      out_a += A;
      if (P.premultiplied) {
        out_c = out_c + C;
        out_g = out_g + G;
      }
      else if (A > 0.f) {
        out_c = out_c + C / A;
        out_g = out_g + G / A;
      }
    }
    if (inside) {
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
    }
    n_primary += __reduce_add_sync(0xffffffffu, c_primary);
    n_shadow += __reduce_add_sync(0xffffffffu, c_shadow);
    n_gradient += __reduce_add_sync(0xffffffffu, 3u * c_shaded);
    n_rays += __reduce_add_sync(0xffffffffu, c_rays);
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
