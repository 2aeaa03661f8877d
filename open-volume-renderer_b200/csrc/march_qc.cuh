// Ray-march kernel with gradient + shadow shading, sixth generation: the depth-blocked shading passes of march_qb.cuh
// (J tasks per lane, all shadow lattices walked in lockstep) on a register diet and software-pipelined.
//
// Why (profiles/r2d_ncu_metrics_C3_q_vs_qb.csv, 1024^3): depth blocking took the DRAM reads of a frame from 55.5 GB to 23.2 GB
// (J = 4) / 18.1 GB (J = 8) and the L1 hit rate from 47 % to 72-75 %, but the launch only went from 10.7 to 8.7 ms: at
// 2.7 TB/s it is no longer HBM-bound, it waits -- 16 resident warps per SM (128 registers per thread), each with J fetches in
// flight that mostly miss L2 (hit rate 34 %), issue slots 40 % busy. More fetches in flight per register is what is needed:
//   * a task's position p (3 registers per task in march_qb.cuh) is re-read from its queue slot in shared memory at every
//     shadow step (one LDS.128; the shared-memory pipe is idle otherwise);
//   * the lattice clock is shared: a shadow ray starts at t = 0 at its sample (the sample is inside the box), so the running
//     fp32 sums stx, sty = stx + sstep of raymarching_shadow (shaders_raymarching.cu:44-85) are the same numbers for every
//     task and are kept once per lane; only the last interval of a ray is cut by its own exit st1 (one register per task).
//     The rare task whose clip entry is not exactly 0 (a sample a rounding error outside the box) is marched on its own
//     after the lockstep loop with the plain per-task loop, so every task still sees exactly the reference's lattice;
//   * per task that leaves st1, the log-transmittance and the fetched value: 3 registers, so that the fetches of step k + 1
//     are issued BEFORE step k is classified (one more register per task) -- 2 J fetches in flight per lane, speculative
//     past early termination like the primary march's prefetch (an unused fetch is not counted as a sample).
// Everything a task computes is shade_task()'s (march_q.cuh) arithmetic, sample for sample; contributions return to the
// owning pixel in march order: frames are bit-identical to march_q.cuh / march_qb.cuh.
#pragma once
#include "march_qb.cuh"

namespace b200k {

#ifndef B200VR_QC_MIN_CTAS
#define B200VR_QC_MIN_CTAS 2
#endif
constexpr int QC_THREADS = QB_THREADS;
constexpr int QC_MIN_CTAS = B200VR_QC_MIN_CTAS;
// -DB200VR_PHASE_STATS=1 (diagnostic builds only): every warp adds up the SM clocks it spends marching primary samples, in the
// clip / lattice / gradient + shade + hand-back parts of its shading passes and in everything else; the sums come back through the
// debug fields of Counters and b200vr_wait prints them (option debug_tiles). DESIGN.md section 5.1b quotes one such run.
#ifndef B200VR_PHASE_STATS
#define B200VR_PHASE_STATS 0
#endif
#if B200VR_PHASE_STATS
#define QC_PHASE(var) do { const unsigned int now_ = (unsigned int)clock(); var += now_ - ph_t; ph_t = now_; } while (0)
#else
#define QC_PHASE(var) do { } while (0)
#endif

// slot layout (32 bytes): [0] p.x  [4] p.y  [8] p.z  [12] s  [16] w  [20] log2-transmittance of an off-lattice task  [24] [28] -
// after shading:          [0] C.x  [4] C.y  [8] C.z  [12] G.x [16] G.y [20] G.z

// SYNC != 0 -- an experiment that LOST, kept selectable for the record (profiles/r2k_*). The eight warps of
// a CTA take eight CONSECUTIVE warp tiles of the launch order -- in sheet order (build.cuh sheet_key_kernel) a strip of tiles along
// the screen projection of the light direction -- so that the shadow samples of step k of the eight tiles form one block of the
// volume that slides along the strip by one shadow step (10 voxels) per iteration: what a warp fetches at step k, a neighbour
// fetched at step k - 1. Without it every L1 miss of a shadow fetch goes to DRAM (profiles/r2h_ncu_full_raymarch_qc6_C3.txt: 753 M
// sector misses from L1, 620 M sectors read from DRAM).
//   SYNC = 1 (option qd6): passes start together and the shadow lattice is walked in CTA-wide lockstep, one barrier per step;
//   SYNC = 3 (option qg6): only the tile groups start together, no barrier inside a group.
// Measured on C3 against 6.16 ms / 19.8 GB read from DRAM for the unsynchronised kernel: 14.9 ms / 13.0 GB (SYNC 1), 12.3 ms /
// 12.4 GB (barrier per pass only), 7.7 ms / 9.7 GB (SYNC 3). The traffic does fall -- by half with nothing but a common start -- and
// the frame gets slower every time: at 3.2 of 6.5 TB/s the kernel is not waiting for DRAM bandwidth but for latency, and every
// barrier takes independent warps away from the four a scheduler has. Which warp shades which task and the order of every sum are
// unchanged: frames are bit-identical to the other queue kernels (test_every_shading_kernel_variant_matches_the_oracle).
template<bool SKIP, bool POW2, bool OCCS, int J, int SYNC = 0>
__global__ void __launch_bounds__(QC_THREADS, QC_MIN_CTAS)
raymarch_qc_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int s_tile_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float2* const s_apair = reinterpret_cast<float2*>(smem);
  unsigned int* const s_occ = reinterpret_cast<unsigned int*>(smem + q_alpha_bytes(P.n_alpha));
  const unsigned int apair_base = pin_u32((unsigned int)__cvta_generic_to_shared(smem));
  const unsigned int occ_base = pin_u32(apair_base + (unsigned int)q_alpha_bytes(P.n_alpha));
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
  const unsigned int my_slot0 = queue_base + 32u * (unsigned)lane; // task j of this lane: slot j * 32 + lane
#if B200VR_PHASE_STATS
  unsigned int ph_t = (unsigned int)clock(), ph_other = 0, ph_primary = 0, ph_clip = 0, ph_lattice = 0, ph_shade = 0;
  unsigned int ph_iters = 0, ph_lanes = 0, ph_passes = 0, ph_steps = 0; // primary iterations, live lanes in them, passes, lattice steps
#endif

  for (;;) {
    int k = 0;
    bool have_tile = true;
    if (SYNC) {
      __syncthreads(); // every warp has left the previous group (s_tile_base may be overwritten)
      if (threadIdx.x == 0) s_tile_base = (int)atomicAdd(P.tile_counter, (unsigned int)(QC_THREADS / 32));
      __syncthreads();
      const int base = s_tile_base;
      if (base >= P.n_warp_tiles) break; // CTA-uniform
      k = base + warp;
      have_tile = k < P.n_warp_tiles;    // the last group may be short: its spare warps only take part in the barriers
      k = min(k, P.n_warp_tiles - 1);
    }
    else {
      if (lane == 0) k = (int)atomicAdd(P.tile_counter, 1u);
      k = __shfl_sync(0xffffffffu, k, 0);
      if (k >= P.n_warp_tiles) break;
    }
    int ix, iy;
    const int tile_slot = warp_tile_pixel(P, k, lane, ix, iy);
    const long long tile_t0 = clock64();
    const bool inside = have_tile && ix < P.width && iy < P.height;
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
      unsigned int pre_widx = 0xffffffffu, pre_word = 0u; // occupancy word requested ahead of the ray (see the march)
      unsigned int mine[J + 1]; // queue slots holding this pixel's tasks, 32 per word
#pragma unroll
      for (int g = 0; g <= J; g++) mine[g] = 0u;
      int count = 0;            // tasks waiting (warp-uniform)
      QC_PHASE(ph_other);

      for (;;) {
        const bool any_alive = __any_sync(0xffffffffu, alive);
        if (count >= J * 32 || !any_alive) {
          // a warp arrives here with a full queue or with all its rays finished
          if (SYNC == 1 || SYNC == 2) {
            // all warps of the CTA start their passes together; a warp without tasks goes through the pass for its barriers
            // only, and the tile group ends when no warp has anything left
            if (!__syncthreads_or(count > 0)) break;
          }
          else if (count == 0) break;
          // =================== shading pass: lane takes tasks lane, lane + 32, ... ===================
          const int n = min(count, J * 32);
          __syncwarp();
          QC_PHASE(ph_primary);
#if B200VR_PHASE_STATS
          ph_passes++;
#endif
#include "march_qc_pass.inc"
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
          QC_PHASE(ph_shade);
          continue;
        }

        // =================== one primary sample per live ray (raymarching, shaders_raymarching.cu:87-171) ===================
#if B200VR_PHASE_STATS
        ph_iters++, ph_lanes += (unsigned int)__popc(__ballot_sync(0xffffffffu, alive));
#endif
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
                const unsigned int idx = (unsigned)pad_cell(P, p, c);
                if (OCCS) cell_empty = !occ_bit<OCCS>(P, occ_base, idx);
                else {
                  // the occupancy word of this cell was requested when the ray entered the previous one (below); only when the
                  // ray came out somewhere else is it loaded now (march_rq.cuh has the measurements)
                  unsigned int w;
                  if ((idx >> 5) == pre_widx) w = pre_word;
                  else w = __ldg(P.occ_bits + (idx >> 5));
                  cell_empty = ((w >> (idx & 31u)) & 1u) == 0u;
                }
                // approximate reciprocals: see march_q.cuh
                cell_exit = cell_exit_t(P, c, oo, mk3(rcp_fast(od.x), rcp_fast(od.y), rcp_fast(od.z)), od);
                if (!OCCS) {
                  const float tn = cell_exit + 0.5f * step; // about where the first sample behind this cell will be
                  pre_widx = pad_index_floor(P, mk3(__fmaf_rn(tn, od.x, oo.x), __fmaf_rn(tn, od.y, oo.y), __fmaf_rn(tn, od.z, oo.z))) >> 5;
                  pre_word = __ldg(P.occ_bits + pre_widx);
                }
              }
            }
            // the lanes in live cells issue their fetches first, then the lanes in empty cells run their skip arithmetic while
            // those fetches are in flight, then the live lanes classify
            const bool skip = SKIP && cell_empty;
            float s = 0.f;
            if (!skip) {
              if (have_pre) s = s_pre;
              else s = tex3D<float>(P.tex, p.x, p.y, p.z);
              // the next lattice sample, when it stays in this (live) macrocell, is fetched before this one is classified.
              // Measured and rejected (profiles/r2n_*, r2x_*): this sample and the next three fetched
              // together and consumed over the following steps (C3 6.41 vs 6.14 ms), and four samples fetched, classified and
              // corrected as independent chains with only the compositing recurrence and the queue stores in order (6.93 ms;
              // C2 3.54 vs 2.29) -- neither shortens a small partition's launch either (1/8 of C3: 2.26 ms all three ways).
              const float nty = fminf(ty + step, t1), ntm = 0.5f * (ty + nty);
              // (also when it is the first sample of the next macrocell, whose occupancy is not known yet: march_rq.cuh)
              have_pre = nty > ty;
              if (have_pre) s_pre = tex3D<float>(P.tex, __fmaf_rn(ntm, od.x, oo.x), __fmaf_rn(ntm, od.y, oo.y), __fmaf_rn(ntm, od.z, oo.z));
            }
            if (skip) {
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
              have_pre = false;
              break;
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
      QC_PHASE(ph_primary);
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
    if (lane == 0 && have_tile) tile_cost_add(P, tile_slot, tile_t0);
    QC_PHASE(ph_other);
  }
#if B200VR_PHASE_STATS
  if (lane == 0) {
    atomicAdd(&P.counters->tile_clocks_sum, (unsigned long long)ph_primary);
    atomicAdd(&P.counters->drain_clocks_sum, (unsigned long long)ph_clip);
    atomicAdd(&P.counters->blocks_published, (unsigned long long)ph_lattice);
    atomicAdd(&P.counters->blocks_stolen, (unsigned long long)ph_shade);
    atomicAdd(&P.counters->tile_clocks_max, (unsigned long long)ph_other);
    // four more sums packed into two words: iterations | live lanes, passes | lattice steps (each < 2^32 per launch at these sizes)
    atomicAdd(&P.counters->warp_clocks_max, ((unsigned long long)ph_iters << 40) + (unsigned long long)ph_lanes);
    atomicAdd(&P.counters->delta_samples, ((unsigned long long)ph_passes << 40) + (unsigned long long)ph_steps);
  }
#endif

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
