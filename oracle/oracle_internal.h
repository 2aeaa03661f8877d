// TEST INFRASTRUCTURE ONLY -- internal declarations shared by the oracle translation units.
#pragma once
#include "oracle_math.h"
#include "ovr_oracle.h"

#include <algorithm>
#include <atomic>
#include <functional>
#include <thread>
#include <vector>

namespace ovr_oracle {

constexpr float float_large = FLT_MAX;   // ovr/common/math_def.h:56
constexpr float float_small = FLT_MIN;   // ovr/common/math_def.h:57
constexpr float float_epsilon = FLT_EPSILON;
constexpr int MC_SIZE = 16;              // accel/spatial_partition.h:24

// ---- RandomTEA (ovr/common/random/random.h:146-188) ----
struct TEA {
  uint32_t v0, v1;
  TEA(uint32_t idx, uint32_t seed) : v0(idx), v1(seed) {}
  inline void draw(float& a, float& b)
  {
    uint32_t s0 = v0, s1 = v1, sum = 0;
    for (int i = 0; i < 16; i++) {
      sum += 0x9e3779b9u;
      s0 += ((s1 << 4) + 0xa341316cu) ^ (s1 + sum) ^ ((s1 >> 5) + 0xc8013ea4u);
      s1 += ((s0 << 4) + 0xad90777du) ^ (s0 + sum) ^ ((s0 >> 5) + 0x7e95761eu);
    }
    v0 = s0;
    v1 = s1;
    const float tofloat = 2.3283064365386962890625e-10f; // 2^-32; uint->float is round-to-nearest
    a = float(v0) * tofloat;
    b = float(v1) * tofloat;
  }
};

// ---- texture-unit model ----
struct VolumeTex {
  const float* v;
  int nx, ny, nz;
  int mode;
};

// quantise an interpolation weight the way the texture unit stores it (1.8 fixed point)
static inline float quant_weight(float a, int mode)
{
  if (mode == OVR_ORACLE_FILTER_FIXED8_RN) return std::floor(a * 256.f + 0.5f) * (1.f / 256.f);
  if (mode == OVR_ORACLE_FILTER_FIXED8_TRUNC) return std::floor(a * 256.f) * (1.f / 256.f);
  return a;
}

float tex3d(const VolumeTex& t, float px, float py, float pz);          // normalised coords, clamp, linear
float tex1d_scalar(const float* tab, int n, float u, int mode);          // normalised coord, clamp, linear
void tex1d_float4(const float* tab, int n, float u, int mode, float out[4]);

// sample_volume_object_space (shaders_common.h:186-193)
static inline float sample_volume(const VolumeTex& t, vec3 p)
{
  return tex3d(t, cclamp(p.x, 0.f, 1.f), cclamp(p.y, 0.f, 1.f), cclamp(p.z, 0.f, 1.f));
}

struct RGBA { float r, g, b, a; };
// sample_transfer_function + array1d_nodal (shaders_common.h:311-319,356-367)
RGBA sample_tf(const ovr_oracle_scene& s, float sample);

// intersect_box (shaders_common.h:156-184)
bool intersect_box(float& t0, float& t1, vec3 org, vec3 dir, vec3 lower, vec3 upper);

// ---- 3-D DDA iterator (accel/dda.h:30-148): grid units, parametrised by the ray's own t ----
struct DDAIter {
  vec3 t_next;
  vec3i_t cell;
  float next_cell_begin = 0.f;

  void init(vec3 org, vec3 dir, float t_min, float /*t_max*/, vec3i_t grid)
  {
    const vec3 o = org + t_min * dir;
    const vec3 fc = v3(cmax(0.f, cmin(float(grid.x) - 1.f, std::floor(o.x))),
                       cmax(0.f, cmin(float(grid.y) - 1.f, std::floor(o.y))),
                       cmax(0.f, cmin(float(grid.z) - 1.f, std::floor(o.z))));
    const vec3 fe = v3(dir.x > 0.f ? fc.x + 1.f : fc.x, dir.y > 0.f ? fc.y + 1.f : fc.y,
                       dir.z > 0.f ? fc.z + 1.f : fc.z);
    const vec3 ts = v3(std::fabs(1.f / dir.x), std::fabs(1.f / dir.y), std::fabs(1.f / dir.z));
    t_next = v3(dir.x == 0.f ? float_large : std::fabs(fe.x - o.x) * ts.x,
                dir.y == 0.f ? float_large : std::fabs(fe.y - o.y) * ts.y,
                dir.z == 0.f ? float_large : std::fabs(fe.z - o.z) * ts.z);
    cell = vec3i_t{ int(fc.x), int(fc.y), int(fc.z) };
    next_cell_begin = 0.f;
  }

  // lambda(cell, t0, t1) -> bool "move on to the next cell"
  template<typename F>
  bool next(vec3 /*org*/, vec3 dir, float t_min, float t_max, vec3i_t grid, const F& lambda)
  {
    const vec3i_t stop = { dir.x > 0.f ? grid.x : -1, dir.y > 0.f ? grid.y : -1, dir.z > 0.f ? grid.z : -1 };
    if (cell.x == stop.x) return false;
    if (cell.y == stop.y) return false;
    if (cell.z == stop.z) return false;
    const vec3 ts = v3(std::fabs(1.f / dir.x), std::fabs(1.f / dir.y), std::fabs(1.f / dir.z));
    const vec3i_t delta = { dir.x > 0.f ? 1 : -1, dir.y > 0.f ? 1 : -1, dir.z > 0.f ? 1 : -1 };
    const float t_closest = cmin(cmin(t_next.x, t_next.y), t_next.z);
    const float cell_t0 = cmax(t_min + next_cell_begin, t_min);
    const float cell_t1 = cmin(t_min + t_closest, t_max);
    if (cell_t0 >= cell_t1) return false;
    const bool go_on = lambda(cell, cell_t0, cell_t1);
    if (go_on || cmax(t_min + next_cell_begin, t_min) >= cell_t1) {
      if (t_next.x == t_closest) { t_next.x += ts.x; cell.x += delta.x; if (cell.x == stop.x) return false; }
      if (t_next.y == t_closest) { t_next.y += ts.y; cell.y += delta.y; if (cell.y == stop.y) return false; }
      if (t_next.z == t_closest) { t_next.z += ts.z; cell.z += delta.z; if (cell.z == stop.z) return false; }
      next_cell_begin = t_closest;
    }
    return go_on;
  }
};

// row-tile parallel-for over pixel rows
void parallel_rows(int y0, int y1, int stride_y, int num_threads, const std::function<void(int)>& row_fn);

} // namespace ovr_oracle
