// Ray-march kernel with gradient + shadow shading, seventh generation: the work-sharing rings of march_ring.cuh carrying the
// depth-blocked shading passes of march_qc.cuh -- for sort-first partitions and sparse (foveated) frames.
//
// Why (profiles/r2s_partition_C3.jsonl, r2y_phase_stats.txt): a launch lasts about its mean load plus its longest tile, and a tile
// is one serial chain of ~2000-clock steps -- 25 march iterations to fill the queue, then a pass of 28 lattice steps, a dozen
// times and more for the heavy tiles of C3. An eighth of the C3 frame is worth 0.77 ms of fetches and took 2.23 ms with
// march_qc.cuh, and as long with the one-task-per-lane rings (they move the shading off the chain but shade at 10.96 instead of
// 6.14 ms per frame). Here a marching warp that has J x 32 tasks staged either shades them itself with the pass of march_qc.cuh
// (included textually: the same statements) or publishes them as ONE block of J KB to its CTA's ring and keeps marching. Warps
// that have run out of tiles take blocks from any ring, copy them into their own (idle) staging slots and run the same pass.
// Who shades: the warp itself while `ring_high` blocks are ready in its ring (nobody is short of work), but only the first
// `rq_self` blocks of a tile -- a tile that goes on filling blocks is one of the heavy ones, the launch ends when the heaviest has
// marched AND shaded, and its later blocks are better left in the ring (`rq_cap` slots) for the first warp that runs dry.
// Measured, an eighth of C3 / of C2 on one B200 (profiles/r3a_*, r3h_*): 1.12 / 0.57 ms, against 2.23 (march_qc.cuh) / 1.03 ms (rings
// of single tasks). What is left is the march of the longest tile itself: ~800 dependent iterations of ~1900 clocks.
//
// As in march_ring.cuh a consumer does not know the pixel's other samples: every task's contribution goes to per-pixel fixed-point
// sums with fire-and-forget 64-bit reductions (integers: the frame is deterministic, identical whoever shades what and for every
// partition) and resolve_kernel finishes the pixel. The ring protocol (tickets, per-slot sequence words, bounded waits, error
// word) is march_ring.cuh's with blocks of J KB; warps without work watch one word pair (blocks available anywhere, warps still
// marching) instead of scanning every ring. Task slot: [0] p.xyz [12] s [16] w [28] pixel index.
//
// Measured and not kept (profiles/r3a_rq_experiments.md): per-pixel sums of a warp's own tile in shared memory with one reduction
// set per tile (the 6 KB push the carve-out up a step, L1 shrinks: 7.55 against 7.18 ms per C3 frame) or in registers as in
// march_qc.cuh (102 B of spills: 8.80 ms) -- the reductions themselves cost 0.16 ms; only two helper warps per CTA; four lattice
// samples of a ray fetched together once the tiles have run out or for heavy tiles (2.1 useful steps per iteration on C3's cells,
// 25 % more time where it is used, no shorter launch); the march step as one more branch of the state machine instead of a loop
// of its own (ptxas then waits for the prefetched sample at the back edge: 3400 instead of 2100 clocks per iteration).
#pragma once
#include "march_qc.cuh"
#include "march_ring.cuh"

namespace b200k {

template<int J> __host__ __device__ constexpr int rq_block_float4() { return J * RING_BLOCK_FLOAT4; }
constexpr unsigned int RQ_PIXEL_MASK = 0x7fffffffu; // task word [28]: pixel index; RING_NO_PIXEL marks an empty slot of a block
// Warps per CTA that stay to shade published blocks once they have run out of tiles; the others leave (8 = all: measured best)
#ifndef B200VR_RQ_HELPERS
#define B200VR_RQ_HELPERS 8
#endif
constexpr int RQ_HELPERS = B200VR_RQ_HELPERS;

template<bool SKIP, bool POW2, bool OCCS, int J>
__global__ void __launch_bounds__(QC_THREADS, QC_MIN_CTAS)
raymarch_rq_kernel(const __grid_constant__ LaunchParams P)
{
  constexpr int SYNC = 0; // the pass body's CTA-barrier experiments do not apply here
  extern __shared__ __align__(16) unsigned char smem[];
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
  const int my_ring = (int)blockIdx.x, n_rings = (int)gridDim.x;
  unsigned int* const r_head = P.ring_ctrl + my_ring * RING_CTRL_PER_RING, * const r_tail = r_head + RING_CTRL_STRIDE;
  unsigned int* const r_left_all = P.ring_ctrl + (size_t)n_rings * RING_CTRL_PER_RING, * const r_err = r_left_all + RING_CTRL_STRIDE;
  unsigned int* const r_avail = r_left_all + 1; // blocks published and not yet taken, over all rings (same sector as r_left_all)
  const unsigned int cap_mask = (unsigned int)P.ring_cap - 1u; // per ring
  unsigned int* const my_seq = P.ring_seq + (size_t)my_ring * P.ring_cap;
  float4* const my_data = P.ring_data + (size_t)my_ring * P.ring_cap * rq_block_float4<J>();
  const f3 oo = P.cam_oo;
  const float step = P.step;
  unsigned int c_primary = 0, c_shadow = 0, c_shaded = 0, c_rays = 0; // per lane, over the whole launch
  const unsigned int lt_mask = (1u << lane) - 1u;
  const unsigned int my_slot0 = queue_base + 32u * (unsigned)lane; // task j of this lane: slot j * 32 + lane
#if B200VR_PHASE_STATS
  // as in march_qc.cuh; "other" = tile setup, ring fill checks and publishing; waiting for blocks is not counted
  unsigned int ph_t = (unsigned int)clock(), ph_other = 0, ph_primary = 0, ph_clip = 0, ph_lattice = 0, ph_shade = 0, ph_idle = 0;
  unsigned int ph_iters = 0, ph_lanes = 0, ph_passes = 0, ph_steps = 0;
#endif

  // ---- warp state: role, tile, ray ----
  bool marching = true, have_tile = false;
  int tile_slot = 0;
  long long tile_t0 = 0;
  unsigned int pixel_index = 0;
  bool inside = false, alive = false, cell_empty = false, have_pre = false;
  f3 od = mk3(0.f);
  float tx = 0.f, ty = 0.f, t1 = 0.f, A = 0.f, cell_exit = -1.f, s_pre = 0.f;
  int count = 0; // tasks staged (warp-uniform)
  unsigned int pre_widx = 0xffffffffu, pre_word = 0u; // occupancy word requested ahead of the ray (see the march)
  int self_blocks = 0; // blocks of the tile in flight this warp shaded itself
  unsigned int idle_spins = 0, idle_ns = 256u;
#if B200VR_TILE_STATS // diagnostic builds: where the tail of the launch goes (render_wait prints the sums)
  const long long warp_t0 = clock64();
  long long drain_t0 = warp_t0, idle_t0 = 0, idle_clk = 0;
  unsigned int n_pub = 0, n_stolen = 0;
  unsigned int t_iters = 0, t_pubs = 0, t_local = 0; // of the tile in flight: march iterations, blocks published / shaded here
  long long t_pub_clk = 0, t_local_clk = 0, t_mark = 0;
  unsigned long long t_sec[4] = { 0, 0, 0, 0 };
#endif

  for (;;) {
    if (marching && !have_tile) {
      // ---- next tile: one 8x4 warp tile per atomic ticket ----
      int k = 0;
      if (lane == 0) k = (int)atomicAdd(P.tile_counter, 1u);
      k = __shfl_sync(0xffffffffu, k, 0);
      if (k >= P.n_warp_tiles) {
        marching = false; // everything this warp published is ordered before the decrement by the fence
#if B200VR_TILE_STATS
        drain_t0 = idle_t0 = clock64();
#endif
        if (lane == 0) {
          __threadfence();
          atomicSub(r_head + 2 * RING_CTRL_STRIDE, 1u);
          atomicSub(r_left_all, 1u);
        }
        if (warp >= RQ_HELPERS) break; // see RQ_HELPERS
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
      cell_exit = -1.f;
      cell_empty = false;
      have_pre = false;
      pre_widx = 0xffffffffu;
      count = 0;
      self_blocks = 0;
      have_tile = true;
#if B200VR_TILE_STATS
      t_iters = t_pubs = t_local = 0, t_pub_clk = t_local_clk = 0;
      t_sec[0] = t_sec[1] = t_sec[2] = t_sec[3] = 0;
#endif
      QC_PHASE(ph_other);
      continue;
    }

    bool consume = !marching;
    if (marching) {
      // ---- primary samples until J x 32 tasks are staged or every ray of the tile has ended. A loop of its own: with the march
      // step as one more branch of the state machine below, ptxas waited for the prefetched sample at the loop's back edge and an
      // iteration took 3400 instead of 2000 clocks.
      bool any_alive = true;
      while ((any_alive = __any_sync(0xffffffffu, alive)) && count < J * 32) {
#if B200VR_TILE_STATS
        t_iters++;
#endif
#if B200VR_PHASE_STATS
        ph_iters++, ph_lanes += (unsigned int)__popc(__ballot_sync(0xffffffffu, alive));
#endif

        // =================== one primary sample per live ray (raymarching, shaders_raymarching.cu:87-171) ===================
        bool want = false;
        f3 tp = mk3(0.f);
        float ts = 0.f, tw = 0.f;
#if B200VR_TILE_STATS
        const unsigned int it_t0 = (unsigned int)clock();
        unsigned int it_t1 = it_t0, it_t2 = it_t0;
        bool it_entry = false;
#endif
        if (alive) {
          do {
            const float tm = 0.5f * (tx + ty);
            const f3 p = mk3(__fmaf_rn(tm, od.x, oo.x), __fmaf_rn(tm, od.y, oo.y), __fmaf_rn(tm, od.z, oo.z));
#if B200VR_TILE_STATS
            it_entry = SKIP && !(tm < cell_exit);
#endif
            if (SKIP) {
              if (!(tm < cell_exit)) {
                i3 c;
                const unsigned int idx = (unsigned)pad_cell(P, p, c);
                if (OCCS) cell_empty = !occ_bit<OCCS>(P, occ_base, idx);
                else {
                  // The occupancy word of this cell was requested when the ray entered the previous one (below); only when the
                  // ray came out somewhere else is it loaded now. With the mask in global memory this load, at L2 latency in
                  // front of the cell's first fetch, was half of a march iteration -- and with 32 rays in a warp some lane
                  // enters a new macrocell in nearly every iteration (5.6 lanes on the longest tile of C3).
                  unsigned int w;
                  if ((idx >> 5) == pre_widx) w = pre_word;
                  else w = __ldg(P.occ_bits + (idx >> 5));
                  cell_empty = ((w >> (idx & 31u)) & 1u) == 0u;
                }
                cell_exit = cell_exit_t(P, c, oo, mk3(rcp_fast(od.x), rcp_fast(od.y), rcp_fast(od.z)), od); // see march_q.cuh
                if (!OCCS) {
                  const float tn = cell_exit + 0.5f * step; // about where the first sample behind this cell will be
                  pre_widx = pad_index_floor(P, mk3(__fmaf_rn(tn, od.x, oo.x), __fmaf_rn(tn, od.y, oo.y), __fmaf_rn(tn, od.z, oo.z))) >> 5;
                  pre_word = __ldg(P.occ_bits + pre_widx);
                }
              }
            }
            // Order of the three parts below: the lanes in live cells issue their fetches first, then the lanes in empty cells run
            // their skip arithmetic (a chain of some twenty dependent additions) while those fetches are in flight, then the live
            // lanes classify. With the skip in front, every iteration of the warp paid for it before its fetches went out.
            const bool skip = SKIP && cell_empty;
#if B200VR_TILE_STATS
            it_t1 = it_t2 = (unsigned int)clock();
#endif
            float s = 0.f;
            if (!skip) {
              if (have_pre) s = s_pre;
              else s = tex3D<float>(P.tex, p.x, p.y, p.z);
              // the next lattice sample is fetched before this one is classified -- also when it is the first one of the next
              // macrocell, whose occupancy is not known yet: a wasted fetch if that cell turns out empty, instead of a fetch all
              // lanes of the warp wait for (longest tile of an eighth of C3: 945 -> 650 clocks per iteration until the sample
              // is there). The first sample behind an EMPTY cell fetched the same way at the end of the skip: slower (1.28 ms).
              const float nty = fminf(ty + step, t1), ntm = 0.5f * (ty + nty);
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
#if B200VR_TILE_STATS
            asm volatile("mov.u32 %0, %%clock;" : "=r"(it_t2) : "f"(a)); // after the sample has arrived
#endif
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
          sts_f32x4(queue_base + 32u * (unsigned)slot + 16u, make_float4(tw, 0.f, 0.f, __uint_as_float(pixel_index)));
        }
        __syncwarp();
        count += __popc(m);
#if B200VR_TILE_STATS
        {
          // the slowest lane of each section, and how many lanes entered a new macrocell in this iteration
          const unsigned int it_t3 = (unsigned int)clock();
          t_sec[0] += __reduce_max_sync(0xffffffffu, it_t1 - it_t0);
          t_sec[1] += __reduce_max_sync(0xffffffffu, it_t2 - it_t1);
          t_sec[2] += it_t3 - it_t0;
          t_sec[3] += (unsigned int)__popc(__ballot_sync(0xffffffffu, it_entry));
        }
#endif
      }
      QC_PHASE(ph_primary);
      if (count >= J * 32 || (!any_alive && count > 0)) {
        unsigned int fill = 0;
#if B200VR_TILE_STATS
        t_mark = clock64();
#endif
        if (lane == 0) fill = ld_relaxed_u32(r_head) - ld_relaxed_u32(r_tail);
        fill = __shfl_sync(0xffffffffu, fill, 0);
        // Shade the staged tasks here when nobody is waiting for work (ring_high blocks ready) -- but only the first rq_self blocks
        // of a tile: a tile that goes on producing blocks is one of the heavy ones, and the launch ends when the heaviest has
        // marched AND shaded (the longest tile of an eighth of C3: 1.2 Mclk of marching, 1.05 Mclk for the seven of its 21
        // blocks it shaded itself, of a launch of 2.6 Mclk). Its later blocks wait in the ring for the first warp without a tile.
        const int room = P.ring_cap - (QC_THREADS / 32) - 8; // every warp of the CTA may hold a ticket
        if ((int)fill >= room || ((int)fill >= P.ring_high && self_blocks < P.rq_self)) consume = true, self_blocks++;
        else {
          // =================== publish the J x 32 oldest staged tasks as one block ===================
          const int n = min(count, J * 32);
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
          float4* const dst = my_data + (size_t)(t & cap_mask) * rq_block_float4<J>();
#pragma unroll
          for (int j = 0; j < J; j++) {
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = make_float4(0.f, 0.f, 0.f, __uint_as_float(RING_NO_PIXEL));
            if (j * 32 + lane < n) a = lds_f32x4(my_slot0 + 1024u * (unsigned)j), b = lds_f32x4(my_slot0 + 1024u * (unsigned)j + 16u);
            __stcg(dst + 2 * (j * 32 + lane), a);
            __stcg(dst + 2 * (j * 32 + lane) + 1, b);
          }
          __syncwarp();
          if (lane == 0) {
            __threadfence();
            st_release_u32(my_seq + (t & cap_mask), t + 1u);
            atomicAdd(r_avail, 1u);
          }
#if B200VR_TILE_STATS
          n_pub++;
          t_pubs++;
#endif
          // move the tasks that did not fit (at most 32) to the front
          const int rest = count - n;
          float4 m0, m1;
          if (lane < rest) m0 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane)), m1 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane) + 16u);
          __syncwarp();
          if (lane < rest) sts_f32x4(queue_base + 32u * (unsigned)lane, m0), sts_f32x4(queue_base + 32u * (unsigned)lane + 16u, m1);
          __syncwarp();
          count = rest;
          QC_PHASE(ph_other);
#if B200VR_TILE_STATS
          t_pub_clk += clock64() - t_mark;
#endif
          continue;
        }
      }
      else if (!any_alive) {
        // ---- tile done: the ray's opacity goes to resolve_kernel directly, colour and normals through the accumulators ----
        if (inside) P.ray_alpha[pixel_index] = A;
#if B200VR_TILE_STATS
        if (lane == 0) {
          const unsigned long long dt = (unsigned long long)(clock64() - tile_t0);
          if (dt > atomicMax(&P.counters->dbg[0], dt)) { // the longest so far (last writer wins: good enough for a diagnostic)
            P.counters->dbg[1] = t_iters, P.counters->dbg[2] = t_pubs, P.counters->dbg[3] = t_local;
            P.counters->dbg[5] = (unsigned long long)t_pub_clk, P.counters->dbg[6] = (unsigned long long)t_local_clk;
            P.counters->dbg[7] = (unsigned long long)(tile_t0 - warp_t0);
            P.counters->dbg[4] = t_sec[0] | (t_sec[1] << 32);
            P.counters->dbg[8] = t_sec[2] | (t_sec[3] << 32);
          }
        }
#endif
        if (lane == 0) tile_cost_add(P, tile_slot, tile_t0);
        have_tile = false;
        QC_PHASE(ph_other);
        continue;
      }
    }

    if (consume) {
      // =================== one depth-blocked pass: own staged tasks, or a block taken from a ring ===================
      int n = 0;
      if (marching) n = min(count, J * 32); // local: the oldest staged tasks are already in slots [0, n)
      else {
        // ---- remote: claim a block (march_ring.cuh: fetch-add tickets, wait on the slot's own sequence word) ----
        // Warps without work watch ONE word pair (blocks available anywhere, warps still marching) and look through the rings
        // only when there is something to find: scanning 2 x 296 control words per poll from thousands of idle warps asked
        // more of the L2 than the marching warps' own fetches.
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
          int done = left == 0u; // nobody marches any more (acquire: their last blocks are visible) and every block has a taker
          if (lane == 0 && !done && (++idle_spins & 63u) == 0u) {
            done = ld_relaxed_u32(r_err) != 0u;
            if (!done && idle_spins > RING_SPIN_LIMIT / 16) { // bounded like every other wait
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
        int got = 1;
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
        if (!got) continue;
        if (lane == 0) atomicSub(r_avail, 1u);
        // the block goes into this warp's own staging slots (empty: it has no tile), then the ring slot is free again
        const float4* const src = P.ring_data + ((size_t)found * P.ring_cap + (t & cap_mask)) * rq_block_float4<J>();
#pragma unroll
        for (int j = 0; j < J; j++) {
          const float4 a = __ldcg(src + 2 * (j * 32 + lane)), b = __ldcg(src + 2 * (j * 32 + lane) + 1); // L2: written by another SM
          sts_f32x4(my_slot0 + 1024u * (unsigned)j, a);
          sts_f32x4(my_slot0 + 1024u * (unsigned)j + 16u, b);
          n += __popc(__ballot_sync(0xffffffffu, __float_as_uint(b.w) != RING_NO_PIXEL)); // valid tasks are contiguous from slot 0
        }
        __syncwarp();
        if (lane == 0) st_release_u32(c_seq + (t & cap_mask), t + (unsigned int)P.ring_cap);
        QC_PHASE(ph_idle);
#if B200VR_TILE_STATS
        n_stolen++;
        idle_clk += clock64() - idle_t0; // from the end of the previous block (or of the last tile) to having this one
#endif
      }
      __syncwarp();
      QC_PHASE(ph_other);
#if B200VR_PHASE_STATS
      ph_passes++;
#endif
#include "march_qc_pass.inc"
      __syncwarp();
      // every task's contribution to its pixel's fixed-point sums
      {
#pragma unroll
        for (int j = 0; j < J; j++) {
          if (j * 32 + lane < n) {
            const unsigned int slot = my_slot0 + 1024u * (unsigned)j;
            const float4 r0 = lds_f32x4(slot), r1 = lds_f32x4(slot + 16u);
            unsigned long long* const acc = P.acc + 3 * (size_t)__float_as_uint(r1.w);
            red_add_u64(acc, acc_pack(r0.x, r0.y));
            red_add_u64(acc + 1, acc_pack(r0.z, r0.w));
            red_add_u64(acc + 2, acc_pack(r1.x, r1.y));
          }
        }
      }
      __syncwarp();
      if (marching) {
        const int rest = count - n; // move the tasks that did not fit (at most 32) to the front
        float4 m0, m1;
        if (lane < rest) m0 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane)), m1 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane) + 16u);
        __syncwarp();
        if (lane < rest) sts_f32x4(queue_base + 32u * (unsigned)lane, m0), sts_f32x4(queue_base + 32u * (unsigned)lane + 16u, m1);
        __syncwarp();
        count = rest;
      }
      QC_PHASE(ph_shade);
#if B200VR_TILE_STATS
      idle_t0 = clock64();
      if (marching) t_local++, t_local_clk += idle_t0 - t_mark;
#endif
      continue;
    }
  }
#if B200VR_PHASE_STATS
  (void)ph_idle;
  if (lane == 0) { // the fields march_qc.cuh uses (render_wait prints them)
    atomicAdd(&P.counters->tile_clocks_sum, (unsigned long long)ph_primary);
    atomicAdd(&P.counters->drain_clocks_sum, (unsigned long long)ph_clip);
    atomicAdd(&P.counters->blocks_published, (unsigned long long)ph_lattice);
    atomicAdd(&P.counters->blocks_stolen, (unsigned long long)ph_shade);
    atomicAdd(&P.counters->tile_clocks_max, (unsigned long long)ph_other);
    atomicAdd(&P.counters->warp_clocks_max, ((unsigned long long)ph_iters << 40) + (unsigned long long)ph_lanes);
    atomicAdd(&P.counters->delta_samples, ((unsigned long long)ph_passes << 40) + (unsigned long long)ph_steps);
  }
#endif

#if B200VR_TILE_STATS
  if (lane == 0) {
    const long long now = clock64();
    idle_clk += now - idle_t0;
    atomicAdd(&P.counters->drain_clocks_sum, (unsigned long long)(now - drain_t0));
    atomicMax(&P.counters->warp_clocks_max, (unsigned long long)(now - warp_t0));
    atomicAdd(&P.counters->blocks_published, (unsigned long long)n_pub);
    atomicAdd(&P.counters->blocks_stolen, (unsigned long long)n_stolen);
    atomicAdd(&P.counters->delta_samples, (unsigned long long)idle_clk); // of the time after the last tile: waiting for blocks
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
