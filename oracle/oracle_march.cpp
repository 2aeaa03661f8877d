// TEST INFRASTRUCTURE ONLY -- CPU oracle: fixed-step ray marching with gradient shading and shadow rays.
// Restates ovr/devices/optix7/shaders_raymarching.cu:44-171 (raymarching_shadow, raymarching), :178-229 (closest-hit
// wrappers), :260-321 (render_raymarching), :323-413 (ray generation / frame write) -- SURVEY Appendix A.2-A.5.
#include "oracle_internal.h"

namespace ovr_oracle {

struct MarchCtx {
  const ovr_oracle_scene& s;
  VolumeTex tex;
  affine3 otw, wto;
  linear3 wtc;
  vec3 rdim;
  vec3 light_dir;
  vec3 clip_lo, clip_hi;
  vec3i_t mc;
  bool skip;
};

// macrocell majorant at an object-space position (skip_empty mode only)
static inline bool cell_is_empty(const MarchCtx& c, vec3 p)
{
  const int cx = std::min(int(cclamp(p.x, 0.f, 1.f) * float(c.tex.nx)) / MC_SIZE, c.mc.x - 1);
  const int cy = std::min(int(cclamp(p.y, 0.f, 1.f) * float(c.tex.ny)) / MC_SIZE, c.mc.y - 1);
  const int cz = std::min(int(cclamp(p.z, 0.f, 1.f) * float(c.tex.nz)) / MC_SIZE, c.mc.z - 1);
  return c.s.mc_majorants[size_t(cx) + size_t(c.mc.x) * (size_t(cy) + size_t(c.mc.y) * size_t(cz))] == 0.f;
}

// opacity correction (shaders_raymarching.cu:73-77,118-122); __powf(x,y) = exp2f(y * log2f(x))
static inline float correct_alpha(float a, float base, float dt)
{
  const float adj = base * dt;
  if (!(std::fabs(adj - 1.f) < 1e-7f)) a = clamp01(1.f - std::exp2(adj * std::log2(1.f - a)));
  return a;
}

// raymarching_shadow + __closesthit__volume_shadow (:44-85, :205-229)
static float shadow_march(const MarchCtx& c, vec3 org, vec3 dir, ovr_oracle_counters& n)
{
  // shadow optixTrace (:139-155): tmin = 0, tmax = inf, clipped in object space against the unit cube
  float t0 = 0.f, t1 = float_large;
  if (!intersect_box(t0, t1, xfmPoint(c.wto, org), xfmVector(c.wto.l, dir), v3(0.f), v3(1.f))) return 0.f;
  if (t0 >= t1) return 0.f;
  const float sstep = (c.s.step * 10.f) * c.s.step; // sampling_scale * self.step with sampling_scale = step*10
  float alpha = 0.f;
  float tx = t0, ty = cmin(t1, t0 + sstep);
  while (ty > tx && alpha < 0.9999f) {
    const vec3 pos = org + (0.5f * (tx + ty)) * dir;
    const vec3 p = xfmPoint(c.wto, pos);
    if (!(c.skip && cell_is_empty(c, p))) {
      const float sample = sample_volume(c.tex, p);
      n.shadow_samples++;
      float a = sample_tf(c.s, sample).a;
      a = correct_alpha(a, c.s.base, ty - tx);
      alpha += (1.f - alpha) * a;
    }
    tx = ty;
    ty = cmin(tx + sstep, t1);
  }
  return alpha;
}

struct Payload {
  float alpha = 0.f;
  vec3 color = v3(0.f);
  vec3 gradient = v3(0.f);
};

// raymarching (:87-171)
static void primary_march(const MarchCtx& c, vec3 org, vec3 dir, float t_min, float t_max, Payload& pl,
                          ovr_oracle_counters& n)
{
  if (t_min >= t_max) return;
  const float step = c.s.step;
  float tx = t_min, ty = cmin(t_max, t_min + step);
  while (ty > tx && pl.alpha < 0.9999f) {
    const vec3 pos = org + (0.5f * (tx + ty)) * dir;
    const vec3 p = xfmPoint(c.wto, pos);
    bool live = !(c.skip && cell_is_empty(c, p));
    if (live) {
      const float sample = sample_volume(c.tex, p);
      n.primary_samples++;
      RGBA rgba = sample_tf(c.s, sample);
      rgba.a = correct_alpha(rgba.a, c.s.base, ty - tx);
      if (c.skip && rgba.a == 0.f) live = false; // contributes exactly +0 to every accumulator
      if (live) {
        vec3 rgb = v3(rgba.r, rgba.g, rgba.b);
        vec3 normal_c = v3(0.f);
        if (c.s.shading == OVR_ORACLE_SHADING_REF) {
          // compute_volume_gradient_object_space (shaders_common.h:195-215)
          vec3 stp = c.rdim;
          if (p.x + stp.x > 1.f) stp.x *= -1.f;
          if (p.y + stp.y > 1.f) stp.y *= -1.f;
          if (p.z + stp.z > 1.f) stp.z *= -1.f;
          const vec3 g = v3(sample_volume(c.tex, p + v3(stp.x, 0.f, 0.f)) - sample,
                            sample_volume(c.tex, p + v3(0.f, stp.y, 0.f)) - sample,
                            sample_volume(c.tex, p + v3(0.f, 0.f, stp.z)) - sample) / stp;
          n.gradient_fetches += 3;
          const vec3 normal_o = -normalize(g);
          const vec3 normal_w = normalize(xfmNormal(c.otw.l, normal_o));
          normal_c = normalize(xfmNormal(c.wtc, normal_w));
          const float sa = shadow_march(c, pos, c.light_dir, n);
          const float cosNL = std::fabs(dot(c.light_dir, normal_w));
          const vec3 k = v3(0.5f) + ((0.5f * cosNL) * v3(2.f)) * (1.f - sa);
          rgb = rgb * k;
        }
        const float tr = 1.f - pl.alpha;
        pl.color = pl.color + (tr * clamp01(rgb)) * rgba.a;
        pl.gradient = pl.gradient + (tr * clamp01(normal_c)) * rgba.a;
        pl.alpha += tr * rgba.a;
      }
    }
    tx = ty;
    ty = cmin(tx + step, t_max);
  }
}

} // namespace ovr_oracle

using namespace ovr_oracle;

extern "C" void ovr_oracle_render_march(const ovr_oracle_scene* sp, int x0, int y0, int x1, int y1, int stride_x,
                                        int stride_y, float* rgba, float* grad, int premultiplied,
                                        ovr_oracle_counters* counters, int num_threads)
{
  const ovr_oracle_scene& s = *sp;
  MarchCtx c{ s };
  c.tex = VolumeTex{ s.voxels, s.dims[0], s.dims[1], s.dims[2], s.filter_mode };
  c.otw = from_3x4(s.obj_to_world);
  c.wto = inverse(c.otw);
  const vec3 ch = v3(s.cam_horizontal[0], s.cam_horizontal[1], s.cam_horizontal[2]);
  const vec3 cv = v3(s.cam_vertical[0], s.cam_vertical[1], s.cam_vertical[2]);
  const vec3 cd = v3(s.cam_direction[0], s.cam_direction[1], s.cam_direction[2]);
  const vec3 cp = v3(s.cam_position[0], s.cam_position[1], s.cam_position[2]);
  {
    // get_xfm_world_to_camera (shaders_common.h:276-289)
    const vec3 x = normalize(ch), y = normalize(cv), z = -normalize(cd);
    c.wtc = linear3{ v3(x.x, y.x, z.x), v3(x.y, y.y, z.y), v3(x.z, y.z, z.z) };
  }
  c.rdim = v3(1.f) / v3(float(s.dims[0]), float(s.dims[1]), float(s.dims[2]));
  c.light_dir = normalize(v3(s.light_pos[0], s.light_pos[1], s.light_pos[2]));
  c.clip_lo = v3(s.clip_lo[0], s.clip_lo[1], s.clip_lo[2]);
  c.clip_hi = v3(s.clip_hi[0], s.clip_hi[1], s.clip_hi[2]);
  c.mc = vec3i_t{ s.mc_dims[0], s.mc_dims[1], s.mc_dims[2] };
  c.skip = s.skip_empty != 0 && s.mc_majorants != nullptr;
  const int W = s.width, H = s.height;
  const float rw = 1.f / float(W), rh = 1.f / float(H);
  if (stride_x < 1) stride_x = 1;
  std::vector<ovr_oracle_counters> per_row(size_t(std::max(H, 1)));
  for (auto& r : per_row) r = ovr_oracle_counters{};

  parallel_rows(y0, y1, stride_y, num_threads, [&](int iy) {
    ovr_oracle_counters n{};
    for (int ix = x0; ix < x1; ix += stride_x) {
      const uint32_t pixel_index = uint32_t(ix) + uint32_t(iy) * uint32_t(W);
      TEA rng(uint32_t(s.frame_index), pixel_index); // :336
      float out_a = 0.f;
      vec3 out_c = v3(0.f), out_g = v3(0.f);
      for (int k = 0; k < s.spp; k++) {
        float sx = (float(ix) + .5f) * rw, sy = (float(iy) + .5f) * rh; // compute_screen_position
        if (s.spp > 1) {
          float j0, j1;
          rng.draw(j0, j1);
          sx += (j0 - 0.5f) * rw;
          sy += (j1 - 0.5f) * rh;
        }
        const vec3 dir = normalize(cd + (sx - 0.5f) * ch + (sy - 0.5f) * cv);
        n.rays++;
        Payload pl;
        float t0 = 0.f, t1 = float_large;
        if (intersect_box(t0, t1, xfmPoint(c.wto, cp), xfmVector(c.wto.l, dir), c.clip_lo, c.clip_hi))
          primary_march(c, cp, dir, t0, t1, pl, n);
        // render_raymarching (:260-321): the background trace never hits, so bg = 0
        const float alpha = pl.alpha + (1.f - pl.alpha) * 0.f;
        out_a += alpha;
        if (premultiplied) {
          out_c = out_c + pl.color;
          out_g = out_g + pl.gradient;
        }
        else if (alpha > 0.f) {
          out_c = out_c + pl.color / alpha;
          out_g = out_g + pl.gradient / alpha;
        }
      }
      const float rspp = 1.f / float(s.spp);
      out_a *= rspp;
      out_c = out_c * rspp;
      out_g = out_g * rspp;
      float* px = rgba + 4 * size_t(pixel_index);
      px[0] = out_c.x, px[1] = out_c.y, px[2] = out_c.z, px[3] = out_a;
      if (grad) {
        float* pg = grad + 3 * size_t(pixel_index);
        pg[0] = out_g.x, pg[1] = out_g.y, pg[2] = out_g.z;
      }
    }
    per_row[size_t(iy)] = n;
  });
  if (counters) {
    for (const auto& r : per_row) {
      counters->primary_samples += r.primary_samples;
      counters->shadow_samples += r.shadow_samples;
      counters->gradient_fetches += r.gradient_fetches;
      counters->delta_samples += r.delta_samples;
      counters->rays += r.rays;
    }
  }
}
