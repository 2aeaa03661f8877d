// Ray-march kernel with gradient + shadow shading, fifth generation: the deferred-shading queue of march_q.cuh with
// DEPTH-BLOCKED shading passes -- a pass takes J x 32 tasks, J per lane, and walks the shadow lattice of all of them in
// lockstep: for every shadow step k the lane issues the J fetches of its tasks back to back, then classifies them.
//
// Why (profiles/r1c_ncu_full_raymarch_q_C3.txt: 1024^3, 55.5 GB of DRAM reads per frame for 4.3 GB of voxels, L1 hit 47 %, L2
// hit 27 %): with one task per lane a pass touches, per shadow step, one 8x4x1 sheet of samples -- at 1.6 voxels per pixel
// neighbouring footprints of 2x2x2 texels barely share a 32-byte sector, and the next pass of the same warp (one voxel
// further along the primary rays, i.e. in the SAME sectors half of the time) comes ~180 steps x 24 warps later, long after
// the sectors left L1. With J tasks per lane the J x 32 samples of a shadow step form an 8x4xJ block of neighbouring
// positions whose fetches are issued within a few instructions: sectors shared along the primary ray are hit in L1 (or
// merged while still in flight), and the lane has J independent fetches in flight instead of the two of march_q.cuh's
// software pipeline -- fewer resident warps (2 CTAs/SM) then cover the latency and leave L1 to the texels.
//
// The arithmetic of a task is shade_task()'s (march_q.cuh), sample for sample: same shadow lattice (running fp32 sum from
// the clip entry), same cell test, same log-domain transmittance, same early termination, same shading; contributions go
// back to the owning pixel in march order.
#pragma once
#include "march_q.cuh"

namespace b200k {

#ifndef B200VR_QBC_THREADS
#define B200VR_QBC_THREADS 256
#endif
constexpr int QB_THREADS = B200VR_QBC_THREADS; // threads per CTA of the depth-blocked kernels (march_qb.cuh, march_qc.cuh): A/B builds
#ifndef B200VR_QB_MIN_CTAS
#define B200VR_QB_MIN_CTAS 2
#endif
constexpr int QB_MIN_CTAS = B200VR_QB_MIN_CTAS; // 2 x 256 threads: 128 registers per thread for the J-fold pass state

template<int J> __host__ __device__ constexpr int qb_slots() { return (J + 1) * 32; } // a pass takes J x 32, one march step adds <= 32
template<int J> __host__ __device__ __forceinline__ size_t qb_queue_bytes()
{
  return sizeof(float) * Q_SLOT_FLOATS * (size_t)qb_slots<J>() * (QB_THREADS / 32);
}
template<int J> __host__ __device__ __forceinline__ size_t qb_shared_bytes(int n_alpha, int occ_words_staged)
{
  return q_alpha_bytes(n_alpha) + (occ_words_staged > 0 ? q_occ_bytes(occ_words_staged) : 0) + qb_queue_bytes<J>();
}

template<bool SKIP, bool POW2, bool OCCS, int J>
__global__ void __launch_bounds__(QB_THREADS, QB_MIN_CTAS)
raymarch_qb_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float2* const s_apair = reinterpret_cast<float2*>(smem);
  unsigned int* const s_occ = reinterpret_cast<unsigned int*>(smem + q_alpha_bytes(P.n_alpha));
  const unsigned int apair_base = pin_u32((unsigned int)__cvta_generic_to_shared(smem));
  const unsigned int occ_base = pin_u32(apair_base + (unsigned int)q_alpha_bytes(P.n_alpha));
  // this warp's task queue: slot q = 32 bytes at queue_base + 32 q
  const unsigned int queue_base = pin_u32(occ_base + (unsigned int)(OCCS ? q_occ_bytes(P.occ_words) : 0) +
                                          (unsigned int)warp * (unsigned int)(qb_slots<J>() * Q_SLOT_FLOATS * sizeof(float)));
  if (SKIP && OCCS)
    for (int i = threadIdx.x; i < P.occ_words; i += blockDim.x) s_occ[i] = P.occ_bits[i];
  for (int i = threadIdx.x; i <= P.n_alpha; i += blockDim.x) {
    const float v0 = P.tf_alpha[min(i, P.n_alpha - 1)], v1 = P.tf_alpha[min(i + 1, P.n_alpha - 1)];
    s_apair[i] = make_float2(v0, (v1 - v0) * (1.f / 256.f));
  }
  __syncthreads();
  const f3 oo = P.cam_oo;
  const float step = P.step;
  unsigned int c_primary = 0, c_shadow = 0, c_shaded = 0, c_rays = 0; // per lane, over all tiles of this warp
  const unsigned int lt_mask = (1u << lane) - 1u;

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
      f3 od = xfm_vector(P.wto.l, dir);
      float tx = 0.f, t1 = FLT_MAX;
      bool alive = inside && intersect_box(tx, t1, oo, od, P.clip_lo, P.clip_hi);
      if (inside) c_rays++;
      float ty = fminf(t1, tx + step);
      alive = alive && ty > tx;
      float A = 0.f;
      f3 C = mk3(0.f), G = mk3(0.f);
      float cell_exit = -1.f; // macrocell the ray is in: valid while the sample parameter stays below cell_exit
      bool cell_empty = false;
      float s_pre = 0.f;     // prefetched value of the sample at (tx, ty), valid while have_pre
      bool have_pre = false;
      unsigned int mine[J + 1]; // queue slots holding this pixel's tasks, 32 per word
#pragma unroll
      for (int g = 0; g <= J; g++) mine[g] = 0u;
      int count = 0;            // tasks waiting (warp-uniform)

      for (;;) {
        const bool any_alive = __any_sync(0xffffffffu, alive);
        if (count >= J * 32 || (!any_alive && count > 0)) {
          // =================== shading pass: lane takes tasks lane, lane + 32, ... ===================
          const int n = min(count, J * 32);
          __syncwarp();
          float px[J], py[J], pz[J], st1[J], stx[J], slog[J];
          unsigned int live_tasks = 0u; // bit j: task j of this lane still marches its shadow ray
#pragma unroll
          for (int j = 0; j < J; j++) {
            px[j] = py[j] = pz[j] = 0.f;
            st1[j] = stx[j] = 0.f;
            slog[j] = 0.f;
            if (j * 32 + lane < n) {
              const float4 q0 = lds_f32x4(queue_base + 32u * (unsigned)(j * 32 + lane));
              px[j] = q0.x, py[j] = q0.y, pz[j] = q0.z;
              // shadow ray from the sample toward the light (__closesthit__volume_shadow, shaders_raymarching.cu:205-229)
              if (clip_shadow_p(P, mk3(q0.x, q0.y, q0.z), stx[j], st1[j]) && fminf(st1[j], stx[j] + P.sstep) > stx[j]) live_tasks |= 1u << j;
            }
          }
          // ---- shadow march of all tasks in lockstep on their t lattices (raymarching_shadow, shaders_raymarching.cu:44-85) ----
          while (live_tasks != 0u) {
            float sv[J];
            unsigned int fetched = 0u;
#pragma unroll
            for (int j = 0; j < J; j++) {
              sv[j] = 0.f;
              if ((live_tasks >> j) & 1u) {
                const float sty = fminf(st1[j], stx[j] + P.sstep);
                const float tm = 0.5f * (stx[j] + sty);
                const f3 q = mk3(__fmaf_rn(tm, P.sh_d.x, px[j]), __fmaf_rn(tm, P.sh_d.y, py[j]), __fmaf_rn(tm, P.sh_d.z, pz[j]));
                // a sample in an empty macrocell has alpha == 0 exactly and adds 0 to the log sum: cell test only
                if (!SKIP || occ_bit<OCCS>(P, occ_base, pad_index_floor(P, q))) {
                  sv[j] = tex3D<float>(P.tex, q.x, q.y, q.z);
                  fetched |= 1u << j;
                }
              }
            }
#pragma unroll
            for (int j = 0; j < J; j++) {
              if ((live_tasks >> j) & 1u) {
                const float sty = fminf(st1[j], stx[j] + P.sstep);
                if ((fetched >> j) & 1u) {
                  c_shadow++;
                  const float a = alpha_lookup<POW2>(P, apair_base, sv[j]);
                  // (1 - A) *= (1 - a)^(base * dt): opacity correction + compositing of the reference, in the log domain
                  slog[j] = __fmaf_rn(P.base * (sty - stx[j]), lg2_fast(fmaxf(1.f - a, 0.f)), slog[j]);
                }
                stx[j] = sty;
                if (!(fminf(st1[j], sty + P.sstep) > sty && slog[j] > P.log_ert)) live_tasks &= ~(1u << j);
              }
            }
          }
          // ---- gradient, normals, shading (shaders_raymarching.cu:126-169): the 3 J gradient fetches first, then the arithmetic ----
          float gx[J], gy[J], gz[J];
#pragma unroll
          for (int j = 0; j < J; j++) {
            gx[j] = gy[j] = gz[j] = 0.f;
            if (j * 32 + lane < n) {
              // compute_volume_gradient_object_space (shaders_common.h:195-215): forward differences, step 1/dims
              const float sx_ = px[j] + P.rdim.x > 1.f ? -P.rdim.x : P.rdim.x;
              const float sy_ = py[j] + P.rdim.y > 1.f ? -P.rdim.y : P.rdim.y;
              const float sz_ = pz[j] + P.rdim.z > 1.f ? -P.rdim.z : P.rdim.z;
              gx[j] = tex3D<float>(P.tex, px[j] + sx_, py[j], pz[j]);
              gy[j] = tex3D<float>(P.tex, px[j], py[j] + sy_, pz[j]);
              gz[j] = tex3D<float>(P.tex, px[j], py[j], pz[j] + sz_);
            }
          }
          __syncwarp();
#pragma unroll
          for (int j = 0; j < J; j++) {
            if (j * 32 + lane < n) {
              const unsigned int slot = queue_base + 32u * (unsigned)(j * 32 + lane);
              const float s = lds_f32x4(slot).w, w = lds_f32x4(slot + 16u).x;
              c_shaded++;
              const f3 rgb = color_global<POW2>(P, s);
              const float sx_ = px[j] + P.rdim.x > 1.f ? -P.rdim.x : P.rdim.x;
              const float sy_ = py[j] + P.rdim.y > 1.f ? -P.rdim.y : P.rdim.y;
              const float sz_ = pz[j] + P.rdim.z > 1.f ? -P.rdim.z : P.rdim.z;
              // shading normals: v * rsqrt.approx(|v|^2) as in shade_task(); the zero-gradient case gives NaN -> clamp -> 0
              const f3 normal_o = -normalize3_fast(mk3((gx[j] - s) * rcp_fast(sx_), (gy[j] - s) * rcp_fast(sy_), (gz[j] - s) * rcp_fast(sz_)));
              const f3 normal_w = normalize3_fast(xfm_vector(P.wto_t, normal_o));
              const f3 normal_c = normalize3_fast(xfm_vector(P.wtc, normal_w));
              const float cosNL = fabsf(dot3(P.light_dir, normal_w));
              const float k_sh = 0.5f + ((0.5f * cosNL) * 2.f) * ex2_fast(slog[j]); // light_rgb = 2; 1 - shadow.alpha = 2^slog
              const f3 cC = clamp01(rgb * k_sh) * w, cG = clamp01(normal_c) * w;      // shaders_raymarching.cu:156-165
              sts_f32x4(slot, make_float4(cC.x, cC.y, cC.z, cG.x));
              sts_f32x4(slot + 16u, make_float4(cG.y, cG.z, 0.f, 0.f));
            }
          }
          __syncwarp();
          // every pixel adds its own samples' contributions in march order (= slot order)
#pragma unroll
          for (int g = 0; g < J; g++) {
            unsigned int own = mine[g];
            const int left = n - g * 32; // slots of this word that were shaded
            if (left < 32) own = left > 0 ? own & ((1u << left) - 1u) : 0u;
            while (own) {
              const int j = __ffs(own) - 1;
              own &= own - 1u;
              const unsigned int slot = queue_base + 32u * (unsigned)(g * 32 + j);
              const float4 r0 = lds_f32x4(slot), r1 = lds_f32x4(slot + 16u);
              C = C + mk3(r0.x, r0.y, r0.z);
              G = G + mk3(r0.w, r1.x, r1.y);
            }
          }
          __syncwarp();
          // move the tasks that did not fit (at most 31, all in word J) to the front
          const int rest = count - n;
          float4 m0, m1;
          if (lane < rest) m0 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane)), m1 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane) + 16u);
          __syncwarp();
          if (lane < rest) sts_f32x4(queue_base + 32u * (unsigned)(lane), m0), sts_f32x4(queue_base + 32u * (unsigned)(lane) + 16u, m1);
          const unsigned int carried = rest > 0 ? mine[J] : 0u;
#pragma unroll
          for (int g = 0; g <= J; g++) mine[g] = 0u;
          mine[0] = carried;
          count = rest;
          continue;
        }
        if (!any_alive) break;

        // =================== one primary sample per live ray (raymarching, shaders_raymarching.cu:87-171) ===================
        bool want = false;
        f3 tp = mk3(0.f);
        float ts = 0.f, tw = 0.f;
        if (alive) {
          do {
            const float tm = 0.5f * (tx + ty);
            const f3 p = mk3(__fmaf_rn(tm, od.x, oo.x), __fmaf_rn(tm, od.y, oo.y), __fmaf_rn(tm, od.z, oo.z));
            if (SKIP) {
              if (!(tm < cell_exit)) {
                i3 c;
                cell_empty = !occ_bit<OCCS>(P, occ_base, (unsigned)pad_cell(P, p, c));
                // approximate reciprocals: see march_q.cuh
                cell_exit = cell_exit_t(P, c, oo, mk3(rcp_fast(od.x), rcp_fast(od.y), rcp_fast(od.z)), od);
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
                alive = ty > tx;
                break;
              }
            }
            float s;
            if (have_pre) s = s_pre;
            else s = tex3D<float>(P.tex, p.x, p.y, p.z);
            {
              // the next lattice sample, when it stays in this (live) macrocell, is fetched before this one is classified
              const float nty = fminf(ty + step, t1), ntm = 0.5f * (ty + nty);
              have_pre = nty > ty && (!SKIP || ntm < cell_exit);
              if (have_pre) s_pre = tex3D<float>(P.tex, __fmaf_rn(ntm, od.x, oo.x), __fmaf_rn(ntm, od.y, oo.y), __fmaf_rn(ntm, od.z, oo.z));
            }
            c_primary++;
            float a = alpha_lookup<POW2>(P, apair_base, s);
            {
              const float adj = P.base * (ty - tx); // opacity correction (shaders_raymarching.cu:118-122)
              if (!(fabsf(adj - 1.f) < 1e-7f)) a = clamp01(1.f - __powf(1.f - a, adj));
            }
            if (!SKIP || a != 0.f) {
              tw = (1.f - A) * a;
              A += tw;
              tp = p;
              ts = s;
              want = true;
            }
            tx = ty;
            ty = fminf(tx + step, t1);
            alive = ty > tx && A < 0.9999f;
          } while (0);
        }
        const unsigned int m = __ballot_sync(0xffffffffu, want);
        if (want) {
          const int slot = count + __popc(m & lt_mask);
          sts_f32x4(queue_base + 32u * (unsigned)(slot), make_float4(tp.x, tp.y, tp.z, ts));
          sts_f32x4(queue_base + 32u * (unsigned)(slot) + 16u, make_float4(tw, 0.f, 0.f, 0.f));
          const unsigned int bit = 1u << (slot & 31);
#pragma unroll
          for (int g = 0; g <= J; g++) mine[g] |= (slot >> 5) == g ? bit : 0u; // selects, not a dynamic index: the words stay in registers
        }
        count += __popc(m);
      }
      // render_raymarching (shaders_raymarching.cu:260-321): background is always 0 -> alpha_blend = C / A
      if (!P.premultiplied) { // un-premultiply; A == 0 adds nothing, as alpha_blend does
        C = A > 0.f ? C / A : mk3(0.f);
        G = A > 0.f ? G / A : mk3(0.f);
      }
      if (inside) {
        float4 px4 = make_float4(C.x, C.y, C.z, A);
        float* g = P.grad + 3 * (size_t)pixel_index;
        if (sidx > 0) { // running sums over the samples of this pixel
          const float4 prev = P.rgba[pixel_index];
          px4 = make_float4(prev.x + px4.x, prev.y + px4.y, prev.z + px4.z, prev.w + px4.w);
          G = G + mk3(g[0], g[1], g[2]);
        }
        if (sidx == P.spp - 1) {
          const float rspp = 1.f / (float)P.spp;
          px4 = make_float4(px4.x * rspp, px4.y * rspp, px4.z * rspp, px4.w * rspp);
          G = G * rspp;
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
        g[0] = G.x;
        g[1] = G.y;
        g[2] = G.z;
      }
    }
    if (lane == 0) tile_cost_add(P, tile_slot, tile_t0);
  }

  // executed-sample counters: one atomic set per warp
  const unsigned int n_primary = __reduce_add_sync(0xffffffffu, c_primary), n_shadow = __reduce_add_sync(0xffffffffu, c_shadow);
  const unsigned int n_shaded = __reduce_add_sync(0xffffffffu, c_shaded), n_rays = __reduce_add_sync(0xffffffffu, c_rays);
  if (lane == 0) {
    atomicAdd(&P.counters->primary_samples, (unsigned long long)n_primary);
    atomicAdd(&P.counters->shadow_samples, (unsigned long long)n_shadow);
    atomicAdd(&P.counters->gradient_fetches, 3ull * n_shaded);
    atomicAdd(&P.counters->rays, (unsigned long long)n_rays);
  }
}

} // namespace b200k
