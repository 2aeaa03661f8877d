// Ray-march kernel with gradient + shadow shading, fourth generation: the deferred-shading tasks of march_q.cuh go through a
// GLOBAL ring of 32-task blocks instead of staying with the warp that produced them.
//
// Why (profiles/, tools/partition_probe.py): in the queue kernel a warp tile is one serial chain -- march 32 rays, shade 32
// samples, march ... -- and the heaviest 8x4-pixel tile of the C2 frame took 3.2-4.1 M clocks (1.6-2.1 ms) for a 2.3 ms
// frame: the launch could not finish before its longest tile, an eighth of the tiles (sort-first on 8 GPUs) still took
// 1.1-1.5 ms, and launching expensive tiles first did not help because the tile itself is the critical path.
// Here a marching warp with 32 staged tasks first looks at its CTA's ring: if fewer than `ring_high` blocks are waiting there
// it PUBLISHES the tasks as one 1 KB block and keeps marching, otherwise it shades them itself straight from shared memory,
// exactly like the queue kernel (best locality: its shadow rays run through texels the SM has just had in L1; shading
// everything through the ring measured 3.5-3.7 ms; keeping the locally shaded sums in registers and adding them to the
// accumulators once per pixel 2.93 against 2.65 ms -- register pressure). Warps that have run out of tiles CONSUME blocks from any ring. In the
// bulk of the launch nobody drains, the rings stay full and every warp works for itself; at the tail the idle warps empty
// the rings, the marching warps refill them instead of shading, and a heavy tile costs its primary march only.
//
// A consumer does not know the pixel's other samples, so contributions are accumulated per pixel with fire-and-forget
// 64-bit integer reductions in fixed point (two 32-bit fields per word, 30 fractional bits: every contribution is <= its
// compositing weight w and sum(w) <= 1, so a field never carries into its neighbour). Integer sums do not depend on the
// order of the additions: the frame is deterministic and identical for every sort-first partition. resolve_kernel
// (build.cuh) converts the sums, un-premultiplies with the ray's alpha and applies the spp / accumulation logic.
//
// Ring: bounded multi-producer multi-consumer queue with per-slot sequence numbers (tickets are never reset, all arithmetic
// is modulo 2^32): slot i is free for ticket t when seq[i] == t, holds its data when seq[i] == t + 1 and is released with
// seq[i] = t + cap. Every CTA has its own ring: (a) one ring for the launch meant 0.8 M fetch-adds per frame on each of two
// words, about 4 ns apiece at the L2 -- 6.7 ms for a 2.3 ms frame; (b) a producer above the high-water mark then shades
// blocks of its own CTA, i.e. of neighbouring tiles, whose shadow rays run through the texels the SM has just had in L1
// (64 rings shared by CTAs of different SMs: 3.67 ms). Only warps that have run out of tiles take blocks from other CTAs.
// All CTAs of the launch are resident (persistent grid), so the waits below always end; they are bounded anyway and set
// an error flag instead of hanging the device.
#pragma once
#include "march_q.cuh"

namespace b200k {

constexpr int RING_BLOCK_FLOAT4 = 64;          // 32 tasks x 2 float4: (p.xyz, s), (w, pixel index bits, -, -)
constexpr unsigned int RING_SPIN_LIMIT = 1u << 24;
constexpr unsigned int RING_NO_PIXEL = 0xffffffffu;
constexpr float ACC_SCALE = 1073741824.f;      // 2^30

// control words, 128 bytes apart (separate L2 lines). Per ring r at ctrl + 96 r: [0] head (publish tickets), [32] tail (claim
// tickets), [64] warps of this ring still marching. After the rings: [0] warps still marching in total, [32] error flag.
constexpr int RING_CTRL_STRIDE = 32, RING_CTRL_PER_RING = 3 * RING_CTRL_STRIDE;
__host__ __device__ __forceinline__ size_t ring_ctrl_words(int rings) { return (size_t)rings * RING_CTRL_PER_RING + 2 * RING_CTRL_STRIDE; }
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p)
{
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned int* p, unsigned int v)
{
  asm volatile("st.release.gpu.global.u32 [%0], %1;" : : "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_relaxed_u32(const unsigned int* p)
{
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// fixed-point pair: lo | hi << 32, both in [0, 1] scaled by 2^30
__device__ __forceinline__ unsigned long long acc_pack(float lo, float hi)
{
  return (unsigned long long)__float2uint_rn(lo * ACC_SCALE) | ((unsigned long long)__float2uint_rn(hi * ACC_SCALE) << 32);
}
__device__ __forceinline__ void red_add_u64(unsigned long long* p, unsigned long long v)
{
  asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" : : "l"(p), "l"(v) : "memory");
}

template<bool SKIP, bool POW2, bool OCCS>
__global__ void __launch_bounds__(Q_THREADS, Q_MIN_CTAS)
raymarch_ring_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float2* const s_apair = reinterpret_cast<float2*>(smem);
  unsigned int* const s_occ = reinterpret_cast<unsigned int*>(smem + q_alpha_bytes(P.n_alpha));
  const unsigned int apair_base = pin_u32((unsigned int)__cvta_generic_to_shared(smem));
  const unsigned int occ_base = pin_u32(apair_base + (unsigned int)q_alpha_bytes(P.n_alpha));
  // this warp's staging queue: slot j = 32 bytes at queue_base + 32 j; published 32 tasks at a time
  const unsigned int queue_base = pin_u32(occ_base + (unsigned int)(OCCS ? q_occ_bytes(P.occ_words) : 0) +
                                          (unsigned int)warp * (unsigned int)(Q_SLOTS * Q_SLOT_FLOATS * sizeof(float)));
  if (SKIP && OCCS)
    for (int i = threadIdx.x; i < P.occ_words; i += blockDim.x) s_occ[i] = P.occ_bits[i];
  for (int i = threadIdx.x; i <= P.n_alpha; i += blockDim.x) {
    const float v0 = P.tf_alpha[min(i, P.n_alpha - 1)], v1 = P.tf_alpha[min(i + 1, P.n_alpha - 1)];
    s_apair[i] = make_float2(v0, (v1 - v0) * (1.f / 256.f));
  }
  __syncthreads();
  const int my_ring = (int)blockIdx.x, n_rings = (int)gridDim.x;
  unsigned int* const r_head = P.ring_ctrl + my_ring * RING_CTRL_PER_RING, * const r_tail = r_head + RING_CTRL_STRIDE;
  unsigned int* const r_left_all = P.ring_ctrl + (size_t)n_rings * RING_CTRL_PER_RING, * const r_err = r_left_all + RING_CTRL_STRIDE;
  unsigned int* const r_avail = r_left_all + 1; // blocks published and not yet taken, over all rings (same sector as r_left_all)
  const unsigned int cap_mask = (unsigned int)P.ring_cap - 1u; // per ring
  unsigned int* const my_seq = P.ring_seq + (size_t)my_ring * P.ring_cap;
  float4* const my_data = P.ring_data + (size_t)my_ring * P.ring_cap * RING_BLOCK_FLOAT4;
  const f3 oo = P.cam_oo;
  const float step = P.step;
  unsigned int c_primary = 0, c_shadow = 0, c_shaded = 0, c_rays = 0; // per lane, over the whole launch
  const unsigned int lt_mask = (1u << lane) - 1u;

  // ---- warp state: role, tile, ray ----
  bool marching = true, have_tile = false;
  int tile_slot = 0;
  long long tile_t0 = 0;
  unsigned int pixel_index = 0;
  bool inside = false, alive = false, cell_empty = false, have_pre = false;
  f3 od = mk3(0.f);
  float tx = 0.f, ty = 0.f, t1 = 0.f, A = 0.f, cell_exit = -1.f, s_pre = 0.f;
  int count = 0; // tasks staged (warp-uniform)
  const long long warp_t0 = clock64();
  long long drain_t0 = 0;
  unsigned int n_pub = 0, n_stolen = 0, idle_spins = 0, idle_ns = 256u;

  for (;;) {
    if (marching && !have_tile) {
      // ---- next tile: one 8x4 warp tile per atomic ticket ----
      int k = 0;
      if (lane == 0) k = (int)atomicAdd(P.tile_counter, 1u);
      k = __shfl_sync(0xffffffffu, k, 0);
      if (k >= P.n_warp_tiles) {
        marching = false; // everything this warp published is ordered before the decrement by the fence
        drain_t0 = clock64();
        if (lane == 0) {
          __threadfence();
          atomicSub(r_head + 2 * RING_CTRL_STRIDE, 1u);
          atomicSub(r_left_all, 1u);
        }
        continue;
      }
      int ix, iy;
      tile_slot = warp_tile_pixel(P, k, lane, ix, iy);
      tile_t0 = clock64();
      inside = ix < P.width && iy < P.height;
      pixel_index = (unsigned)ix + (unsigned)iy * (unsigned)P.width;
      // ---- ray generation + clip (shaders_common.h:400-404, shaders_raymarching.cu:351-362, shaders_common.h:156-184) ----
      float sx = ((float)ix + .5f) * P.rw, sy = ((float)iy + .5f) * P.rh;
      if (P.spp > 1) {
        Tea rng((unsigned)P.frame_index, pixel_index); // RandomTEA(v0 = frame_index, v1 = pixel_index): draw number sidx + 1
        float2 j = rng.draw();
        for (int d = 0; d < P.sidx; d++) j = rng.draw();
        sx += (j.x - 0.5f) * P.rw;
        sy += (j.y - 0.5f) * P.rh;
      }
      const f3 dir = normalize3(P.cam_dir + (sx - 0.5f) * P.cam_hor + (sy - 0.5f) * P.cam_ver);
      od = xfm_vector(P.wto.l, dir);
      tx = 0.f, t1 = FLT_MAX;
      alive = inside && intersect_box(tx, t1, oo, od, P.clip_lo, P.clip_hi);
      if (inside) c_rays++;
      ty = fminf(t1, tx + step);
      alive = alive && ty > tx;
      A = 0.f;
      cell_exit = -1.f; // macrocell the ray is in: valid while the sample parameter stays below cell_exit
      cell_empty = false;
      have_pre = false;
      count = 0;
      have_tile = true;
      continue;
    }

    bool consume = !marching;
    if (marching) {
      // the march as a loop of its own (see march_rq.cuh: as one more branch of this state machine ptxas waited for the
      // prefetched sample at the back edge)
      bool any_alive = true;
      while ((any_alive = __any_sync(0xffffffffu, alive)) && count < 32) {
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
          sts_f32x4(queue_base + 32u * (unsigned)slot, make_float4(tp.x, tp.y, tp.z, ts));
          sts_f32x4(queue_base + 32u * (unsigned)slot + 16u, make_float4(tw, __uint_as_float(pixel_index), 0.f, 0.f));
        }
        __syncwarp();
        count += __popc(m);
      }
      if (count >= 32 || (!any_alive && count > 0)) {
        unsigned int fill = 0;
        if (lane == 0) fill = ld_relaxed_u32(r_head) - ld_relaxed_u32(r_tail);
        fill = __shfl_sync(0xffffffffu, fill, 0);
        if ((int)fill >= P.ring_high) consume = true; // nobody is waiting for work: shade the staged tasks here
        else {
          // =================== publish 32 staged tasks as one block ===================
          const int n = min(count, 32);
          unsigned int t = 0;
          if (lane == 0) {
            t = atomicAdd(r_head, 1u);
            unsigned int spins = 0;
            while (ld_acquire_u32(my_seq + (t & cap_mask)) != t) // the consumer of the previous lap still reads the slot
              if (++spins > RING_SPIN_LIMIT) {
                atomicExch(r_err, 1u);
                break;
              }
          }
          t = __shfl_sync(0xffffffffu, t, 0);
          n_pub++;
          float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = make_float4(0.f, __uint_as_float(RING_NO_PIXEL), 0.f, 0.f);
          if (lane < n) a = lds_f32x4(queue_base + 32u * (unsigned)lane), b = lds_f32x4(queue_base + 32u * (unsigned)lane + 16u);
          float4* const dst = my_data + (size_t)(t & cap_mask) * RING_BLOCK_FLOAT4 + 2 * lane;
          __stcg(dst, a);
          __stcg(dst + 1, b);
          __syncwarp();
          if (lane == 0) {
            __threadfence();
            st_release_u32(my_seq + (t & cap_mask), t + 1u);
            atomicAdd(r_avail, 1u);
          }
          // move the tasks that did not fit (at most 31) to the front
          const int rest = count - n;
          float4 m0, m1;
          if (lane < rest) m0 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane)), m1 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane) + 16u);
          __syncwarp();
          if (lane < rest) sts_f32x4(queue_base + 32u * (unsigned)lane, m0), sts_f32x4(queue_base + 32u * (unsigned)lane + 16u, m1);
          __syncwarp();
          count = rest;
          continue;
        }
      }
      else if (!any_alive) {
        // ---- tile done: the ray's opacity goes to resolve_kernel directly, colour and normals through the accumulators ----
        if (inside) P.ray_alpha[pixel_index] = A;
        if (lane == 0) tile_cost_add(P, tile_slot, tile_t0);
        have_tile = false;
        continue;
      }
    }

    if (consume) {
      // =================== shade one block from the ring: lane j takes task j ===================
      // Claims are fetch-adds, never compare-and-swap (thousands of warps retrying a CAS on one word took 300 ms a frame).
      // A draining warp looks at all
      // rings, 32 at a time starting at its own, and draws a ticket where it sees blocks; if it
      // loses that race the ticket may not be published yet: it then waits on that slot's own sequence word (no shared hot
      // spot) and gives the ticket up only when no warp of that ring marches any more and the ticket is beyond its last block.
      float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = make_float4(0.f, __uint_as_float(RING_NO_PIXEL), 0.f, 0.f);
      int got = 1;
      if (marching) {
        // ---- local: the 32 oldest staged tasks, straight from shared memory (the steady state of the launch) ----
        const int n = min(count, 32);
        if (lane < n) q0 = lds_f32x4(queue_base + 32u * (unsigned)lane), q1 = lds_f32x4(queue_base + 32u * (unsigned)lane + 16u);
        const int rest = count - n; // move the tasks that did not fit (at most 31) to the front
        float4 m0, m1;
        if (lane < rest) m0 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane)), m1 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane) + 16u);
        __syncwarp();
        if (lane < rest) sts_f32x4(queue_base + 32u * (unsigned)lane, m0), sts_f32x4(queue_base + 32u * (unsigned)lane + 16u, m1);
        __syncwarp();
        count = rest;
      }
      else {
        // ---- remote: a block some marching warp published because the rings were running dry ----
        // warps without work watch one word pair (blocks available anywhere, warps still marching) and look through the rings
        // only when there is something to find
        unsigned int left = 0u, avail = 0u;
        if (lane == 0) avail = ld_relaxed_u32(r_avail), left = ld_acquire_u32(r_left_all);
        left = __shfl_sync(0xffffffffu, left, 0);
        avail = __shfl_sync(0xffffffffu, avail, 0);
        int found = -1;
        if ((int)avail > 0 || left == 0u)
          for (int h = 0; h * 32 < n_rings && found < 0; h++) {
            const int i = lane + 32 * h; // ring my_ring + i, wrapping once around all rings
            const unsigned int* c = P.ring_ctrl + (size_t)((my_ring + i) % n_rings) * RING_CTRL_PER_RING;
            const unsigned int have = __ballot_sync(0xffffffffu, i < n_rings && (int)(ld_relaxed_u32(c) - ld_relaxed_u32(c + RING_CTRL_STRIDE)) > 0);
            if (have) found = (my_ring + (__ffs(have) - 1) + 32 * h) % n_rings;
          }
        if (found < 0) {
          // bounded like every other wait: r_left_all counts all warps of the grid, also those of CTAs that are not resident
          // yet -- the launch is cooperative (b200vr.cu), so they are, but a hang must never be the failure mode
          int done = left == 0u; // nobody marches any more (acquire: their last blocks are visible) and every block has a taker
          if (lane == 0 && !done && (++idle_spins & 63u) == 0u) {
            done = ld_relaxed_u32(r_err) != 0u;
            if (!done && idle_spins > RING_SPIN_LIMIT / 16) {
              atomicExch(r_err, 1u);
              done = 1;
            }
          }
          if (__shfl_sync(0xffffffffu, done, 0)) break;
          __nanosleep(idle_ns);
          if (idle_ns < 4096u) idle_ns <<= 1;
          continue;
        }
        idle_ns = 256u;
        idle_spins = 0u;
        unsigned int* const c_ctrl = P.ring_ctrl + (size_t)found * RING_CTRL_PER_RING;
        unsigned int* const c_seq = P.ring_seq + (size_t)found * P.ring_cap;
        unsigned int t = 0;
        if (lane == 0) {
          t = atomicAdd(c_ctrl + RING_CTRL_STRIDE, 1u);
          unsigned int spins = 0, late = 0, ns = 64;
          while (ld_acquire_u32(c_seq + (t & cap_mask)) != t + 1u) { // not published yet / its producer is still writing
            if ((++spins & 15u) == 0u) {
              if ((int)(ld_relaxed_u32(c_ctrl) - t) > 0) { // ticket handed to a producer: the block must appear shortly
                if (++late > RING_SPIN_LIMIT / 1024) {
                  atomicExch(r_err, 1u);
                  got = 0;
                  break;
                }
              }
              else if (ld_acquire_u32(c_ctrl + 2 * RING_CTRL_STRIDE) == 0u && (int)(ld_relaxed_u32(c_ctrl) - t) <= 0) {
                got = 0; // no warp of this ring marches any more and the ticket is beyond its last block
                break;
              }
              if (ld_relaxed_u32(r_err) != 0u) {
                got = 0;
                break;
              }
            }
            __nanosleep(ns);
            if (ns < 2048) ns <<= 1;
          }
        }
        got = __shfl_sync(0xffffffffu, got, 0);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (got) {
          n_stolen++;
          if (lane == 0) atomicSub(r_avail, 1u);
          const float4* const src = P.ring_data + ((size_t)found * P.ring_cap + (t & cap_mask)) * RING_BLOCK_FLOAT4 + 2 * lane;
          q0 = __ldcg(src), q1 = __ldcg(src + 1); // L2: the block was written by another SM
          __ballot_sync(0xffffffffu, __float_as_uint(q1.y) != RING_NO_PIXEL); // every lane holds its task: free the slot
          if (lane == 0) st_release_u32(c_seq + (t & cap_mask), t + (unsigned int)P.ring_cap);
        }
      }
      const unsigned int pix = __float_as_uint(q1.y);
      if (got && pix != RING_NO_PIXEL) {
        f3 cC, cG;
        shade_task<SKIP, POW2, OCCS>(P, apair_base, occ_base, mk3(q0.x, q0.y, q0.z), q0.w, q1.x, c_shaded, c_shadow, cC, cG);
        unsigned long long* const acc = P.acc + 3 * (size_t)pix;
        red_add_u64(acc, acc_pack(cC.x, cC.y));
        red_add_u64(acc + 1, acc_pack(cC.z, cG.x));
        red_add_u64(acc + 2, acc_pack(cG.y, cG.z));
      }
      continue;
    }

  }

  // executed-sample counters: one atomic set per warp
  const unsigned int n_primary = __reduce_add_sync(0xffffffffu, c_primary), n_shadow = __reduce_add_sync(0xffffffffu, c_shadow);
  const unsigned int n_shaded = __reduce_add_sync(0xffffffffu, c_shaded), n_rays = __reduce_add_sync(0xffffffffu, c_rays);
  if (lane == 0) {
    atomicAdd(&P.counters->primary_samples, (unsigned long long)n_primary);
    atomicAdd(&P.counters->shadow_samples, (unsigned long long)n_shadow);
    atomicAdd(&P.counters->gradient_fetches, 3ull * n_shaded);
    atomicAdd(&P.counters->rays, (unsigned long long)n_rays);
#if B200VR_TILE_STATS
    const long long now = clock64();
    atomicAdd(&P.counters->drain_clocks_sum, (unsigned long long)(now - drain_t0));
    atomicMax(&P.counters->warp_clocks_max, (unsigned long long)(now - warp_t0));
    atomicAdd(&P.counters->blocks_published, (unsigned long long)n_pub);
    atomicAdd(&P.counters->blocks_stolen, (unsigned long long)n_stolen);
#endif
  }
}

} // namespace b200k
