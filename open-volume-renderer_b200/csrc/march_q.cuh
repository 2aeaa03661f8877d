// Ray-march kernel with gradient + shadow shading, third generation: a warp marches the 32 primary rays of an 8x4 tile
// in lockstep and DEFERS shading -- every sample that needs a shadow ray becomes a task in a per-warp shared-memory
// queue; whenever 32 tasks are waiting, all 32 lanes take one each (gradient, normals, shadow march, shade) and hand the
// colour / normal contributions back to the pixel that owns the sample.
//
// Why this is legal: in the reference (shaders_raymarching.cu:87-171) the shadow ray only scales the sample's colour;
// the opacity A that drives compositing weights (1-A)*a and early termination never depends on it. So the primary march
// can run ahead, and a pixel's colour is the sum of its tasks' contributions, added here in the reference's order.
// Why it is faster (profiles/: nested loops ran at 55 % warp efficiency, a per-lane state machine at 26 %): the shadow
// march is 92 % of the fetches, and with one task per lane it always runs on full warps whose rays start next to each
// other (same step, neighbouring pixels) and so end together.
#pragma once
#include "march_sm.cuh"

namespace b200k {

#ifndef B200VR_Q_THREADS
#define B200VR_Q_THREADS 256
#endif
constexpr int Q_THREADS = B200VR_Q_THREADS;    // warps of a CTA share one staged alpha table
#ifndef B200VR_Q_MIN_CTAS
#define B200VR_Q_MIN_CTAS 3
#endif
constexpr int Q_MIN_CTAS = B200VR_Q_MIN_CTAS;  // resident CTAs/SM wanted: 3 -> 24 warps/SM at <= 80 registers (4 -> 64 registers spills and measured slower)
constexpr int Q_SLOTS = 64;                    // per-warp queue capacity: flushed at >= 32, one step adds <= 32
constexpr int Q_SLOT_FLOATS = 8;               // task: p.xyz, s, w, -, -, -   result: C.xyz, G.xyz, -, -

// dynamic shared memory: [alpha pairs (n_alpha + 1) x float2][macrocell occupancy bits][task queues, one per warp]
// (the alpha pairs sit at offset 0 so that the hot lookup is a plain LDS [i * 8 + base]; the colour table is read once
// per shaded sample and stays in global memory / L1)
#ifndef B200VR_Q_OCC_SMEM_WORDS
#define B200VR_Q_OCC_SMEM_WORDS 3072
#endif
constexpr int Q_OCC_SMEM_WORDS = B200VR_Q_OCC_SMEM_WORDS; // occupancy bit masks up to 12 KB (padded grids up to 46^3 cells) are staged per CTA
__host__ __device__ __forceinline__ size_t q_alpha_bytes(int n_alpha) { return (sizeof(float2) * (size_t)(n_alpha + 1) + 15) & ~(size_t)15; }
__host__ __device__ __forceinline__ size_t q_occ_bytes(int occ_words) { return (sizeof(unsigned int) * (size_t)occ_words + 15) & ~(size_t)15; }
__host__ __device__ __forceinline__ size_t q_queue_bytes() { return sizeof(float) * Q_SLOT_FLOATS * Q_SLOTS * (Q_THREADS / 32); }
__host__ __device__ __forceinline__ bool q_occ_in_smem(int occ_words) { return occ_words <= Q_OCC_SMEM_WORDS; }
__host__ __device__ __forceinline__ size_t q_shared_bytes(int n_alpha, int occ_words)
{
  return q_alpha_bytes(n_alpha) + (q_occ_in_smem(occ_words) ? q_occ_bytes(occ_words) : 0) + q_queue_bytes();
}

// One bit per padded macrocell: majorant != 0. A shadow ray tests the cell of EVERY sample (its step is about one cell),
// so the test has to be cheap and must not queue behind the texture fetches in L1TEX: the mask of a 512^3 volume is
// 4.9 KB and lives in shared memory; larger masks are read through the read-only path.
// Loads from the staged tables go through 32-bit shared-space addresses (ld.shared), not through generic pointers derived
// from the extern array: nvcc rebuilt the shared-window base (S2UR CgaCtaId + ULEA) at every such use inside the loops.
__device__ __forceinline__ unsigned int lds_u32(unsigned int addr)
{
  unsigned int v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ float2 lds_f32x2(unsigned int addr)
{
  float2 v;
  asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}

__device__ __forceinline__ float4 lds_f32x4(unsigned int addr)
{
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
// The same load as a volatile asm WITHOUT the memory clobber: it is neither removed, merged nor hoisted (the slots it reads
// change from pass to pass) and it keeps its order among the other volatile asm statements -- all the shared-memory loads and
// stores of these kernels -- but it is no barrier for the arithmetic and the texture fetches around it, so the loads of
// independent tasks can be interleaved with them. (A pure, non-volatile asm is wrong here: with loop-invariant operands the
// compiler reuses one pass's values in the next.)
__device__ __forceinline__ float4 lds_f32x4_nv(unsigned int addr)
{
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
// tex3D<float>(t, x, y, z) as ONE predicated instruction (the PTX nvcc emits for tex3D, under a guard predicate): 0 when `on`
// is false. nvcc puts a branch around a conditional tex3D(); in straight-line code that interleaves several independent
// tasks a branch per fetch is what keeps the tasks apart.
__device__ __forceinline__ float tex3d_if(cudaTextureObject_t t, float x, float y, float z, bool on)
{
  float v;
  asm("{\n\t.reg .pred p;\n\t.reg .f32 d1, d2, d3;\n\tsetp.ne.u32 p, %5, 0;\n\tmov.f32 %0, 0f00000000;\n\t"
      "@p tex.3d.v4.f32.f32 {%0, d1, d2, d3}, [%1, {%2, %3, %4, %4}];\n\t}"
      : "=f"(v) : "l"(t), "f"(x), "f"(y), "f"(z), "r"((unsigned int)on));
  return v;
}
__device__ __forceinline__ void sts_f32x4(unsigned int addr, float4 v)
{
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" : : "r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// keeps a loop-invariant address in a register: ptxas otherwise rebuilds it (5-7 instructions) at every use under the
// register cap of this kernel
__device__ __forceinline__ unsigned int pin_u32(unsigned int x)
{
  unsigned int y;
  asm volatile("mov.u32 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}

__device__ __forceinline__ float rcp_fast(float x)
{
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ f3 normalize3_fast(f3 v)
{
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(dot3(v, v)));
  return v * r;
}

template<bool OCCS>
__device__ __forceinline__ bool occ_bit(const LaunchParams& P, unsigned int occ_base, unsigned int idx)
{
  const unsigned int w = OCCS ? lds_u32(occ_base + ((idx >> 5) << 2)) : __ldg(P.occ_bits + (idx >> 5));
  return ((w >> (idx & 31u)) & 1u) != 0u;
}

// padded-grid index of the macrocell holding position q (object space, [0,1]^3 up to rounding): the FMA rounds toward
// -inf at 1.5 * 2^23, which leaves 0x4B400000 + floor(q * cells_per_unit) in the float's bits (floor = -1 .. mc are all
// addressable thanks to the pad ring; the one clamp guards against garbage positions). Cells carry a two-voxel apron
// (build.cuh MC_APRON), so a position that rounds into the neighbouring cell at a boundary is still classified safely.
__device__ __forceinline__ unsigned int pad_index_floor(const LaunchParams& P, f3 q)
{
  const unsigned int bx = __float_as_uint(__fmaf_rd(q.x, P.cells_per_unit.x, 12582912.f));
  const unsigned int by = __float_as_uint(__fmaf_rd(q.y, P.cells_per_unit.y, 12582912.f));
  const unsigned int bz = __float_as_uint(__fmaf_rd(q.z, P.cells_per_unit.z, 12582912.f));
  return min(bx + by * (unsigned)P.mcp_sy + bz * (unsigned)P.mcp_sz - P.mcp_bias, P.mcp_last);
}

// alpha lookup of one sample: texture-unit coordinate model (POW2: exact integer fast path for tables of 2^m entries),
// then value + fraction * slope from the pair table
template<bool POW2>
__device__ __forceinline__ float alpha_lookup(const LaunchParams& P, unsigned int apair_base, float s)
{
  const float v = tf_unit(P, s);
  int i, a;
  if (POW2) {
    const int q = (__float2int_rz(__fmaf_rn(v, P.tfa_mul, P.tfa_addf)) + P.tfa_addi) >> P.tfa_shift;
    i = q >> 8;
    a = q & 255;
  }
  else {
    int i1;
    hw_coord(nodal_coord(v, P.n_alpha), P.n_alpha, i, i1, a);
  }
  const float2 e = lds_f32x2(apair_base + ((unsigned)i << 3));
  return __fmaf_rn((float)a, e.y, e.x);
}

// colour lookup of a shaded sample straight from the global table (L1-resident, once per task)
template<bool POW2>
__device__ __forceinline__ f3 color_global(const LaunchParams& P, float s)
{
  int i, a;
  tf_coord(tf_unit(P, s), P.n_color, P.tfc_mul, P.tfc_addf, P.tfc_addi, P.tfc_shift, i, a);
  const float4 c0 = __ldg(P.tf_color + i), c1 = __ldg(P.tf_color + min(i + 1, P.n_color - 1));
  const float w = (float)a * (1.f / 256.f);
  return mk3(__fmaf_rn(w, c1.x - c0.x, c0.x), __fmaf_rn(w, c1.y - c0.y, c0.y), __fmaf_rn(w, c1.z - c0.z, c0.z));
}

// One trilinear fetch through the unit (the default everywhere) or through the software filter of sampling.cuh (eight global
// loads from the linear copy of the volume, integer weights as the unit applies them): the same queue kernel with either sampler
// is what the tex3D-versus-manual comparison of DESIGN.md section 5.6 runs (ncu captures under profiles/).
template<int SAMPLER>
__device__ __forceinline__ float fetch3(const LaunchParams& P, float x, float y, float z)
{
  if (SAMPLER == SAMPLER_TEX3D) return tex3D<float>(P.tex, x, y, z);
  return sample_volume<SAMPLER_MANUAL>(P, mk3(x, y, z));
}

// Shading of one deferred sample (position p in object space, scalar s, compositing weight w): colour lookup, forward-
// difference gradient, shadow march toward the light, Phong-less "0.5 + cos * transmittance" shading of the reference
// (shaders_raymarching.cu:126-169, 205-229, 44-85). Returns the colour and normal-layer contributions of the sample.
template<bool SKIP, bool POW2, bool OCCS, int SAMPLER = SAMPLER_TEX3D>
__device__ __forceinline__ void shade_task(const LaunchParams& P, unsigned int apair_base, unsigned int occ_base, f3 p, float s, float w,
                                           unsigned int& c_shaded, unsigned int& c_shadow, f3& cC, f3& cG)
{
  const f3 rgb = color_global<POW2>(P, s);
  // compute_volume_gradient_object_space (shaders_common.h:195-215): forward differences, step 1/dims
  f3 stp = P.rdim;
  if (p.x + stp.x > 1.f) stp.x = -stp.x;
  if (p.y + stp.y > 1.f) stp.y = -stp.y;
  if (p.z + stp.z > 1.f) stp.z = -stp.z;
  const float gx = fetch3<SAMPLER>(P, p.x + stp.x, p.y, p.z) - s;
  const float gy = fetch3<SAMPLER>(P, p.x, p.y + stp.y, p.z) - s;
  const float gz = fetch3<SAMPLER>(P, p.x, p.y, p.z + stp.z) - s;
  c_shaded++;
  // shadow ray from the sample toward the light (__closesthit__volume_shadow + raymarching_shadow,
  // shaders_raymarching.cu:205-229,44-85), transmittance accumulated as a sum of logs
  // Software pipeline, two samples deep: sample positions do not depend on fetched data (fixed t lattice), so
  // while sample k is classified the fetch of k+1 and the majorant of k+2 are already in flight.
  float slog = 0.f, stx, st1;
  if (clip_shadow_p(P, p, stx, st1)) {
    const float sstep = P.sstep;
    // All lanes walk the fixed t lattice in lockstep (neighbouring shadow rays then fetch neighbouring texels: letting
    // each lane run ahead to its next live sample measured 3.60 ms against 2.95 ms); a sample in an empty macrocell
    // (alpha == 0 exactly, adds 0 to the log sum) costs the cell test only.
    auto cell_live = [&](float tm) { // bit test in shared memory (reading the float majorants through L1TEX instead: 2.91 vs 2.61 ms)
      return occ_bit<OCCS>(P, occ_base, pad_index_floor(P, mk3(__fmaf_rn(tm, P.sh_d.x, p.x), __fmaf_rn(tm, P.sh_d.y, p.y),
                                                               __fmaf_rn(tm, P.sh_d.z, p.z))));
    };
    auto fetch = [&](float tm) {
      return fetch3<SAMPLER>(P, __fmaf_rn(tm, P.sh_d.x, p.x), __fmaf_rn(tm, P.sh_d.y, p.y), __fmaf_rn(tm, P.sh_d.z, p.z));
    };
    auto classify = [&](float sv, float dt) {
      c_shadow++;
      const float a = alpha_lookup<POW2>(P, apair_base, sv);
      // (1 - A) *= (1 - a)^(base * dt): opacity correction + compositing of the reference, in the log domain
      slog = __fmaf_rn(P.base * dt, lg2_fast(fmaxf(1.f - a, 0.f)), slog);
    };
    // Software pipeline: sample positions do not depend on fetched data (fixed t lattice), so while sample k is
    // classified the fetch of k+1 and the macrocell test of k+2 are already in flight.
    // intervals: current [stx, sty], next [sty, nty], the one after [nty, n2ty]
    // (Measured alternatives on C2: batches of 4/8 independent samples per ray 3.09/3.07 ms against 2.95 ms for this
    // pipeline; a second fetch in flight per ray 2.64 against 2.61 ms.)
    float sty = fminf(st1, stx + sstep);
    const float tm_c = 0.5f * (stx + sty);
    bool live_c = !SKIP || cell_live(tm_c);
    float s_c = 0.f;
    if (live_c && sty > stx) s_c = fetch(tm_c);
    float nty = fminf(sty + sstep, st1);
    float tm_n = 0.5f * (sty + nty);
    bool live_n = !SKIP || cell_live(tm_n);
#pragma unroll 2
    while (sty > stx && slog > P.log_ert) {
      float s_n = 0.f;
      if (live_n && nty > sty) s_n = fetch(tm_n); // prefetch; unused if the ray ends at this sample
      const float n2ty = fminf(nty + sstep, st1);
      const float tm_n2 = 0.5f * (nty + n2ty);
      const bool live_n2 = !SKIP || cell_live(tm_n2);
      if (live_c) classify(s_c, sty - stx);
      stx = sty, sty = nty, nty = n2ty;
      s_c = s_n, live_c = live_n, live_n = live_n2, tm_n = tm_n2;
    }
  }
  // shading normals: v * rsqrt.approx(|v|^2) (2 ulp) instead of the reference's v / sqrt(|v|^2) -- three IEEE
  // square roots and nine divisions per shaded sample otherwise; the zero-gradient case gives NaN -> clamp -> 0 in both
  const f3 normal_o = -normalize3_fast(mk3(gx * rcp_fast(stp.x), gy * rcp_fast(stp.y), gz * rcp_fast(stp.z)));
  const f3 normal_w = normalize3_fast(xfm_vector(P.wto_t, normal_o));
  const f3 normal_c = normalize3_fast(xfm_vector(P.wtc, normal_w)); // wtc = transpose(inverse(world_to_camera.l))
  const float cosNL = fabsf(dot3(P.light_dir, normal_w));
  const float k_sh = 0.5f + ((0.5f * cosNL) * 2.f) * ex2_fast(slog); // light_rgb = 2; 1 - shadow.alpha = 2^slog
  cC = clamp01(rgb * k_sh) * w; // shaders_raymarching.cu:156-165
  cG = clamp01(normal_c) * w;
}

template<bool SKIP, bool POW2, bool OCCS, int SAMPLER = SAMPLER_TEX3D>
__global__ void __launch_bounds__(Q_THREADS, Q_MIN_CTAS)
raymarch_q_kernel(const __grid_constant__ LaunchParams P)
{
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float2* const s_apair = reinterpret_cast<float2*>(smem);
  unsigned int* const s_occ = reinterpret_cast<unsigned int*>(smem + q_alpha_bytes(P.n_alpha));
  const unsigned int apair_base = pin_u32((unsigned int)__cvta_generic_to_shared(smem));
  const unsigned int occ_base = pin_u32(apair_base + (unsigned int)q_alpha_bytes(P.n_alpha));
  // this warp's task queue: slot j = 32 bytes at queue_base + 32 j
  const unsigned int queue_base = pin_u32(occ_base + (unsigned int)(OCCS ? q_occ_bytes(P.occ_words) : 0) +
                                          (unsigned int)warp * (unsigned int)(Q_SLOTS * Q_SLOT_FLOATS * sizeof(float)));
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
    // one 8x4 warp tile per atomic ticket (CTA-wide chunks of 16 tiles measured slower: tail imbalance)
    int k = 0;
    if (lane == 0) k = (int)atomicAdd(P.tile_counter, 1u);
    k = __shfl_sync(0xffffffffu, k, 0);
    if (k >= P.n_warp_tiles) break;
    int ix, iy;
    const int tile_slot = warp_tile_pixel(P, k, lane, ix, iy);
    const long long tile_t0 = clock64();
    const bool inside = ix < P.width && iy < P.height;
    const unsigned int pixel_index = (unsigned)ix + (unsigned)iy * (unsigned)P.width;
    // per-pixel sums over the spp loop are kept in the frame itself (same thread reads and writes its own pixel) and the
    // RNG state is re-derived per sample, so that neither occupies registers during the march

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
      unsigned long long mine = 0ull; // queue slots holding this pixel's tasks
      int count = 0;                  // tasks waiting (warp-uniform)

      for (;;) {
        const bool any_alive = __any_sync(0xffffffffu, alive);
        if (count >= 32 || (!any_alive && count > 0)) {
          // =================== shading pass: lane j takes task j ===================
          const int n = min(count, 32);
          __syncwarp();
          f3 cC = mk3(0.f), cG = mk3(0.f);
          if (lane < n) {
            const float4 q0 = lds_f32x4(queue_base + 32u * (unsigned)(lane)), q1 = lds_f32x4(queue_base + 32u * (unsigned)(lane) + 16u);
            const f3 p = mk3(q0.x, q0.y, q0.z);
            const float s = q0.w, w = q1.x;
            shade_task<SKIP, POW2, OCCS, SAMPLER>(P, apair_base, occ_base, p, s, w, c_shaded, c_shadow, cC, cG);
          }
          __syncwarp();
          if (lane < n) {
            sts_f32x4(queue_base + 32u * (unsigned)(lane), make_float4(cC.x, cC.y, cC.z, cG.x));
            sts_f32x4(queue_base + 32u * (unsigned)(lane) + 16u, make_float4(cG.y, cG.z, 0.f, 0.f));
          }
          __syncwarp();
          // every pixel adds its own samples' contributions in march order (= slot order)
          unsigned int own = (unsigned int)(mine & 0xffffffffull);
          if (n < 32) own &= (1u << n) - 1u;
          while (own) {
            const int j = __ffs(own) - 1;
            own &= own - 1u;
            const float4 r0 = lds_f32x4(queue_base + 32u * (unsigned)(j)), r1 = lds_f32x4(queue_base + 32u * (unsigned)(j) + 16u);
            C = C + mk3(r0.x, r0.y, r0.z);
            G = G + mk3(r0.w, r1.x, r1.y);
          }
          __syncwarp();
          // move the tasks that did not fit (at most 31) to the front
          const int rest = count - n;
          float4 m0, m1;
          if (lane < rest) m0 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane)), m1 = lds_f32x4(queue_base + 32u * (unsigned)(n + lane) + 16u);
          __syncwarp();
          if (lane < rest) sts_f32x4(queue_base + 32u * (unsigned)(lane), m0), sts_f32x4(queue_base + 32u * (unsigned)(lane) + 16u, m1);
          mine = n < 32 ? 0ull : (mine >> 32);
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
                // approximate reciprocals: cell_exit only decides which samples are skipped with their (empty) cell, and a
                // sample a rounding error past the boundary still has its footprint inside the cell's two-voxel apron
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
                have_pre = false;
                break;
              }
            }
#ifndef B200VR_Q_PRIMARY_PREFETCH
#define B200VR_Q_PRIMARY_PREFETCH 1
#endif
            float s;
            if (B200VR_Q_PRIMARY_PREFETCH && have_pre) s = s_pre;
            else s = fetch3<SAMPLER>(P, p.x, p.y, p.z);
            if (B200VR_Q_PRIMARY_PREFETCH) {
              // the next lattice sample, when it stays in this (live) macrocell, is fetched before this one is classified
              // (speculative past early termination): its latency overlaps the classification and the queue traffic
              const float nty = fminf(ty + step, t1), ntm = 0.5f * (ty + nty);
              // (also when it is the first sample of the next macrocell, whose occupancy is not known yet: march_rq.cuh; C2 2.30 -> 2.27 ms)
              have_pre = nty > ty;
              if (have_pre) s_pre = fetch3<SAMPLER>(P, __fmaf_rn(ntm, od.x, oo.x), __fmaf_rn(ntm, od.y, oo.y), __fmaf_rn(ntm, od.z, oo.z));
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
          mine |= 1ull << slot;
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
