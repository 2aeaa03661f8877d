// TEST INFRASTRUCTURE ONLY -- CPU oracle: delta-tracking path tracing over the macrocell majorant grid.
// Restates ovr/devices/optix7/shaders_pathtracing.cu:269-289,385-404 (delta_tracking, use_dda==1 branch),
// :477-542 (pathtracing, recursion unrolled into a loop + albedo stack so the products keep the reference's
// inside-out order), :603-721 (ray generation) and accel/spatial_partition.h:56-96 (DeltaTrackingIter<true>).
#include "oracle_internal.h"

namespace ovr_oracle {

struct PathCtx {
  const ovr_oracle_scene& s;
  VolumeTex tex;
  affine3 wto;
  vec3i_t mc;
  vec3 spac_rcp;
};

// DeltaTrackingIter<true>::hashit (spatial_partition.h:70-95). Note the reference does not advance `t` across
// skipped (majorant <= eps) cells, so the next cell's majorant is applied from the stale `t`; reproduced.
static bool hashit(const PathCtx& c, DDAIter& it, vec3 org, vec3 dir, float tnear, float tfar, float& rayt,
                   float& majorant, float& tau)
{
  const vec3 m_org = org * c.spac_rcp, m_dir = dir * c.spac_rcp;
  bool found = false;
  float t = it.next_cell_begin + tnear;
  while (it.next(m_org, m_dir, tnear, tfar, c.mc, [&](const vec3i_t& cell, float /*t0*/, float t1) {
    majorant = c.s.mc_majorants[size_t(cell.x) + size_t(c.mc.x) * (size_t(cell.y) + size_t(c.mc.y) * size_t(cell.z))];
    if (std::fabs(majorant) <= float_epsilon) return true;
    tau -= (t1 - t) * majorant;
    t = t1;
    if (tau > 0.f) return true;
    t = t + tau / majorant;
    found = true;
    it.next_cell_begin = t - tnear;
    rayt = t;
    return false;
  })) {
  }
  return found;
}

// delta_tracking, use_dda == 1 (:385-404); org/dir already in object space
static bool delta_tracking(const PathCtx& c, TEA& rng, vec3 org, vec3 dir, float t_min, float t_max, float& t_hit,
                           vec3& albedo, ovr_oracle_counters& n)
{
  DDAIter it;
  it.init(org * c.spac_rcp, dir * c.spac_rcp, t_min, t_max, c.mc);
  float xa, xb;
  rng.draw(xa, xb);
  float tau = -std::log(1.f - xa);
  float t = t_min, majorant = 0.f;
  while (hashit(c, it, org, dir, t_min, t_max, t, majorant, tau)) {
    const vec3 p = org + t * dir;
    const float sample = sample_volume(c.tex, p);
    n.delta_samples++;
    const RGBA rgba = sample_tf(c.s, sample); // no opacity correction: sigma_t = alpha per unit length
    rng.draw(xa, xb);
    if (xa * majorant < rgba.a) {
      albedo = v3(rgba.r, rgba.g, rgba.b);
      t_hit = t;
      return true;
    }
    rng.draw(xa, xb);
    tau = -std::log(1.f - xa);
  }
  return false;
}

} // namespace ovr_oracle

// Known-answer entry point for the free-path iteration alone: repeated hashit calls along one ray, taus[k] handed to call k
// (the caller of the reference re-draws tau after every tentative collision). out[4k..] = (found, rayt, majorant, tau after).
// Pinned against the reference's own DeltaTrackingIter<true> (oracle/ref_shim.cpp: ref_delta_hashit).
extern "C" int ovr_oracle_delta_hashit(const float* majorants, const int32_t grid[3], const float spac_rcp[3], const float org[3],
                                       const float dir[3], float tnear, float tfar, const float* taus, int n_taus, float* out)
{
  using namespace ovr_oracle;
  ovr_oracle_scene s{};
  s.mc_majorants = majorants;
  PathCtx c{ s, VolumeTex{ nullptr, 0, 0, 0, 0 }, affine3{}, vec3i_t{ grid[0], grid[1], grid[2] },
             v3(spac_rcp[0], spac_rcp[1], spac_rcp[2]) };
  const vec3 o = v3(org[0], org[1], org[2]), d = v3(dir[0], dir[1], dir[2]);
  DDAIter it;
  it.init(o * c.spac_rcp, d * c.spac_rcp, tnear, tfar, c.mc);
  float rayt = tnear, majorant = 0.f;
  int k = 0;
  for (; k < n_taus; k++) {
    float tau = taus[k];
    const bool found = hashit(c, it, o, d, tnear, tfar, rayt, majorant, tau);
    out[4 * k] = found ? 1.f : 0.f, out[4 * k + 1] = rayt, out[4 * k + 2] = majorant, out[4 * k + 3] = tau;
    if (!found) return k + 1;
  }
  return k;
}

namespace ovr_oracle {

// uniform_sample_sphere (shaders_common.h:339-354); 2*M_PI*s.x is evaluated in double by the reference
static inline vec3 uniform_sphere(float sx, float sy)
{
  const float phi = float(2 * M_PI * double(sx));
  const float cos_t = 1.f * (1.f - 2.f * sy);
  const float sin_t = 2.f * 1.f * std::sqrt(sy * (1.f - sy));
  return v3(std::cos(phi) * sin_t, std::sin(phi) * sin_t, cos_t);
}

} // namespace ovr_oracle

extern "C" void ovr_oracle_uniform_sphere(float sx, float sy, float out[3]) /* known-answer entry point */
{
  const ovr_oracle::vec3 v = ovr_oracle::uniform_sphere(sx, sy);
  out[0] = v.x, out[1] = v.y, out[2] = v.z;
}

namespace ovr_oracle {

// pathtracing (:477-542) for one camera ray; returns (Le, alpha)
static void trace_path(const PathCtx& c, TEA& rng, vec3 cam_org, vec3 cam_dir, vec3& color, float& alpha,
                       ovr_oracle_counters& n)
{
  color = v3(0.f);
  alpha = 0.f;
  // primary: world ray carried into object space by the instance transform, t shared (q8: xfmVector)
  vec3 org = xfmPoint(c.wto, cam_org);
  vec3 dir = xfmVector(c.wto.l, cam_dir);
  float t0 = 0.f, t1 = float_large;
  if (!intersect_box(t0, t1, org, dir, v3(0.f), v3(1.f))) return;
  alpha = 1.f;
  vec3 stack[40];
  int depth = 0;
  int scatter_index = 0;
  vec3 terminal = v3(0.f);
  for (;;) {
    float t;
    vec3 albedo;
    if (!delta_tracking(c, rng, org, dir, t0, t1, t, albedo, n)) {
      if (scatter_index != 0) terminal = v3(c.s.light_ambient);
      break;
    }
    scatter_index++;
    if (scatter_index > c.s.max_num_scatters) break; // Le stays 0
    float xa, xb;
    rng.draw(xa, xb);
    const vec3 dir_w = uniform_sphere(xa, xb);
    // collision point: the primary ray is still a world-space ray in the reference (:510-513)
    org = (depth == 0) ? xfmPoint(c.wto, cam_org + t * cam_dir) : org + t * dir;
    dir = xfmVector(c.wto.l, dir_w);
    stack[depth++] = albedo;
    scatter_index += 1; // child payload starts at scatter_index + 1
    t0 = 0.f;
    t1 = float_large;
    if (c.s.pt_child_instance_xfm) {
      // q14: OptiX carries the (already object-space) scattered ray through the instance transform once more before the clip;
      // the march itself runs along the ray as passed (optixGetWorldRayOrigin / Direction), with the clip's t range
      if (!intersect_box(t0, t1, xfmPoint(c.wto, org), xfmVector(c.wto.l, dir), v3(0.f), v3(1.f))) break;
    }
    else if (!intersect_box(t0, t1, org, dir, v3(0.f), v3(1.f))) break; // child misses: its colour stays 0
  }
  vec3 le = terminal;
  for (int i = depth - 1; i >= 0; i--) le = stack[i] * le; // Le = albedo * child.color, inside out
  color = le;
}

} // namespace ovr_oracle

using namespace ovr_oracle;

extern "C" void ovr_oracle_render_path(const ovr_oracle_scene* sp, int x0, int y0, int x1, int y1, int stride_x,
                                       int stride_y, float* rgba, ovr_oracle_counters* counters, int num_threads)
{
  const ovr_oracle_scene& s = *sp;
  PathCtx c{ s };
  c.tex = VolumeTex{ s.voxels, s.dims[0], s.dims[1], s.dims[2], s.filter_mode };
  c.wto = inverse(from_3x4(s.obj_to_world));
  c.mc = vec3i_t{ s.mc_dims[0], s.mc_dims[1], s.mc_dims[2] };
  {
    // sp_singlemc.cu:99-111: spac = 16 / dims, spac_rcp = 1 / spac
    const vec3 spac = v3(float(MC_SIZE)) / v3(float(s.dims[0]), float(s.dims[1]), float(s.dims[2]));
    c.spac_rcp = v3(1.f) / spac;
  }
  const vec3 ch = v3(s.cam_horizontal[0], s.cam_horizontal[1], s.cam_horizontal[2]);
  const vec3 cv = v3(s.cam_vertical[0], s.cam_vertical[1], s.cam_vertical[2]);
  const vec3 cd = v3(s.cam_direction[0], s.cam_direction[1], s.cam_direction[2]);
  const vec3 cp = v3(s.cam_position[0], s.cam_position[1], s.cam_position[2]);
  const int W = s.width, H = s.height;
  const float rw = 1.f / float(W), rh = 1.f / float(H);
  if (stride_x < 1) stride_x = 1;
  std::vector<ovr_oracle_counters> per_row(size_t(std::max(H, 1)));
  for (auto& r : per_row) r = ovr_oracle_counters{};

  parallel_rows(y0, y1, stride_y, num_threads, [&](int iy) {
    ovr_oracle_counters n{};
    for (int ix = x0; ix < x1; ix += stride_x) {
      const uint32_t pixel_index = uint32_t(ix) + uint32_t(iy) * uint32_t(W);
      TEA rng(uint32_t(s.frame_index), pixel_index); // :644
      float out_a = 0.f;
      vec3 out_c = v3(0.f);
      for (int k = 0; k < s.spp; k++) {
        float sx = (float(ix) + .5f) * rw, sy = (float(iy) + .5f) * rh;
        if (s.spp > 1) {
          float j0, j1;
          rng.draw(j0, j1);
          sx += (j0 - 0.5f) * rw;
          sy += (j1 - 0.5f) * rh;
        }
        const vec3 dir = normalize(cd + (sx - 0.5f) * ch + (sy - 0.5f) * cv);
        n.rays++;
        vec3 col;
        float a;
        trace_path(c, rng, cp, dir, col, a, n);
        out_a += a;
        out_c = out_c + col;
      }
      const float rspp = 1.f / float(s.spp);
      float* px = rgba + 4 * size_t(pixel_index);
      px[0] = out_c.x * rspp, px[1] = out_c.y * rspp, px[2] = out_c.z * rspp, px[3] = out_a * rspp;
    }
    per_row[size_t(iy)] = n;
  });
  if (counters) {
    for (const auto& r : per_row) {
      counters->delta_samples += r.delta_samples;
      counters->rays += r.rays;
    }
  }
}
