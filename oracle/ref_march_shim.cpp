// TEST INFRASTRUCTURE ONLY -- the REFERENCE's own ray-marching shader functions executed on the host.
// oracle/extract_march.py cuts raymarching(), raymarching_shadow() (ovr/devices/optix7/shaders_raymarching.cu:43-171) and the
// helpers they call (shaders_common.h: payload structs, corrected_value, sample_volume_object_space,
// compute_volume_gradient_object_space, get_xfm_*, project_to_screen, compute_optical_flow, array1d_nodal, nearly_equal,
// alpha_blend, sample_transfer_function) out of the reference tree into oracle/_ref/m_*.h at build time; this file supplies
// what the OptiX / CUDA environment supplies on the device and nothing else:
//   * the texture unit: tex3D / tex1D are the oracle's model of the B200 unit (oracle_sample.cpp, pinned against hardware
//     fetches in tests/golden/tex_b200_hw.npz) -- the one thing that cannot be run from the reference's source;
//   * the OptiX plumbing around the shadow ray: optixTrace -> __intersection__volume (shaders_common.h:379-392: object-space
//     clip against the unit cube) -> __closesthit__volume_shadow (shaders_raymarching.cu:205-229: step * 10, then
//     raymarching_shadow), restated in optixTrace() below with the reference's own intersect_box;
//   * launch parameters, transform getters' matrix source, __frcp_rn (IEEE), __powf (its CUDA definition exp2f(y * log2f(x)) with the host's exp2 / log2: only reached when
//     step * base is not within 1e-7 of 1, i.e. at the last partial step of a ray or when the running fp32 sum drifts by an ulp).
// tests/golden/make_golden.py drives ref_march_rays() to produce tests/golden/ref_march.npz and ref_march_spp.npz, and
// ref_path_rays() (the reference's path tracer, namespace pt below) to produce ref_path.npz; they pin the oracle.
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <atomic>
#include <thread>
#include <vector>

#include <cuda_runtime.h> // float2 / make_float2 (host-side vector types only)
#include <gdt/math/mat.h>
#include <gdt/math/vec.h>

#undef __device__
#define __device__
#undef __forceinline__
#define __forceinline__ inline
#include <dda.h> // ovr/devices/optix7/accel/dda.h (ovr::dda::DDAIter), used by the path tracer's macrocell iterator

#include "ovr_oracle.h" // the texture-unit model: ovr_oracle_sample_volume, ovr_oracle_tex1d

namespace ref_march {
using namespace gdt;
using std::fabs;
using std::fmaf;

constexpr float float_large = std::numeric_limits<float>::max(); // ovr/common/math_def.h:56-58
constexpr float float_small = std::numeric_limits<float>::min();
constexpr float inf = float_large;                                // shaders_common.h uses `inf` for the ray's t_max
static inline float __frcp_rn(float x) { return 1.f / x; }
static inline float __frcp_rn(int x) { return 1.f / (float)x; }
static inline float __powf(float x, float y) { return std::exp2(y * std::log2(x)); } // CUDA: __powf(x, y) = exp2f(y * __log2f(x))

// ---- textures ----
struct Tex3 { const float* voxels; int32_t dims[3]; };
struct Tex1 { const float* table; int n; int channels; };
struct float4 { // shadows CUDA's float4 inside this namespace: the device code converts it to vec4f implicitly
  float x, y, z, w;
  operator vec4f() const { return vec4f(x, y, z, w); }
};
static thread_local int g_filter_mode = OVR_ORACLE_FILTER_HW; // per thread: ref_march_rays_mt runs one copy of the launch state per worker
template<typename T> static inline T tex3D(const Tex3& t, float x, float y, float z)
{
  const float p[3] = { x, y, z };
  float out;
  ovr_oracle_sample_volume(t.voxels, t.dims, g_filter_mode, p, 1, &out);
  return out;
}
template<typename T> static inline T tex1D(const Tex1& t, float u);
template<> inline float tex1D<float>(const Tex1& t, float u)
{
  float out;
  ovr_oracle_tex1d(t.table, t.n, 1, g_filter_mode, &u, 1, &out);
  return out;
}
template<> inline float4 tex1D<float4>(const Tex1& t, float u)
{
  float out[4];
  ovr_oracle_tex1d(t.table, t.n, 4, g_filter_mode, &u, 1, out);
  return float4{ out[0], out[1], out[2], out[3] };
}

// ---- path tracing only: RandomTEA (ovr/common/random/random.h:146-188), the macrocell grid and its delta-tracking iterator
// (accel/spatial_partition.h:23-121), cut out at build time by oracle/Makefile and oracle/extract_march.py ----
namespace tea {
struct float2 { // CUDA's float2 with the conversion gdt::vec2f has under nvcc
  float x, y;
  operator vec2f() const { return vec2f(x, y); }
};
static inline float2 make_float2(float x, float y) { return float2{ x, y }; }
#include <random_tea_extract.h>
}
typedef tea::RandomTEA RandomTEA;
namespace dda = ovr::dda;
struct range1f { float lower, upper; };
constexpr float float_epsilon = std::numeric_limits<float>::epsilon(); // ovr/common/math_def.h:58
#define __NVCC__ 1
#include <singlemc_extract.h>
template<>
#include <delta_iter_extract.h>
inline DeviceSpacePartiton_SingleMC::DeltaTrackingIter<true>
#include <p_sp_iter.h>
#undef __NVCC__

// ---- device-side structs, members named as the shaders use them (array.h, volume.h:40-62, params.h:55-99) ----
struct Array3DScalarOptix7 {
  Tex3 data;
  vec3i dims;
  struct { float v; } lower, upper, scale;
};
template<int D, int N> struct ArrayOptix7 {
  Tex1 data;
  struct { int v; } dims;
};
typedef ArrayOptix7<1, 4> Array1DFloat4Optix7;
typedef ArrayOptix7<1, 1> Array1DScalarOptix7;
struct DeviceStructuredRegularVolume {
  Array3DScalarOptix7 volume;
  struct { Array1DFloat4Optix7 color; Array1DScalarOptix7 opacity; } tfn;
  float step, base;
  DeviceSpacePartiton_SingleMC sp; // volume.h:46
  int use_dda = 1;                 // volume.h:54
};
struct LaunchParams {
  struct DeviceCamera { vec3f position, direction, horizontal, vertical; };
  DeviceCamera camera, last_camera;
  vec3f light_directional_pos;
  float light_ambient_intensity = 1.f; // params.h:80
  int max_num_scatters = 24;           // params.h:86
  int traversable = 0;
};
static thread_local LaunchParams optix_launch_params;
static thread_local float g_otw[12], g_wto[12];
static thread_local const DeviceStructuredRegularVolume* g_self = nullptr;
static inline void optixGetWorldToObjectTransformMatrix(float m[12]) { std::memcpy(m, g_wto, sizeof(g_wto)); }
static inline void optixGetObjectToWorldTransformMatrix(float m[12]) { std::memcpy(m, g_otw, sizeof(g_otw)); }

// ---- OptiX plumbing of the shadow ray ----
enum RayType { RAYMARCHING_RAY_TYPE = 0, SHADOW_RAY_TYPE, RAY_TYPE_COUNT }; // shaders_raymarching.cu:41
struct OptixVisibilityMask {
  unsigned m;
  explicit OptixVisibilityMask(unsigned long long v) : m((unsigned)(v & 255u)) {} // 8-bit masks in OptiX
};
enum { OPTIX_RAY_FLAG_DISABLE_ANYHIT = 1 };
#define VISIBILITY_VOLUME (1ULL << 0) // params.h:52-53
#define VISIBILITY_GEOMETRY (1ULL << 1)
static inline void pack_pointer(void* ptr, uint32_t& i0, uint32_t& i1)
{
  const uint64_t u = reinterpret_cast<uint64_t>(ptr);
  i0 = (uint32_t)(u >> 32), i1 = (uint32_t)(u & 0xffffffffu);
}
static inline void* unpack_pointer(uint32_t i0, uint32_t i1) { return reinterpret_cast<void*>(((uint64_t)i0 << 32) | i1); }
static void optixTrace(int, vec3f org, vec3f dir, float tmin, float tmax, float, OptixVisibilityMask, int, int, int, int, uint32_t& u0,
                       uint32_t& u1);

#include <m_shadow_payload.h>
#include <m_march_payload.h>
static inline float
#include <m_corrected_f.h>
static inline vec3f
#include <m_corrected_v.h>
static bool
#include <intersect_box_extract.h>
static inline float
#include <m_sample_volume.h>
static inline vec3f
#include <m_gradient.h>
static inline affine3f
#include <m_xfm_wto.h>
static inline affine3f
#include <m_xfm_otw.h>
static inline affine3f
#include <m_xfm_wtc.h>
static inline vec2f
#include <m_project.h>
static inline vec2f
#include <m_flow.h>
template<typename T, int N> static inline T
#include <m_nodal.h>
static inline bool
#include <m_nearly_equal.h>
template<typename T> static inline T
#include <m_alpha_blend.h>
static inline vec4f
#include <m_sample_tf.h>
static inline void
#include <m_raymarching_shadow.h>
static inline void
#include <m_raymarching.h>
static inline void
#include <m_render_raymarching.h>

// what the pipeline does with a traced ray. SHADOW_RAY_TYPE: __intersection__volume, then __closesthit__volume_shadow
// (shaders_raymarching.cu:205-229) or __miss__shadow (:249-253). RAYMARCHING_RAY_TYPE (render_raymarching's two traces):
// the scene holds one volume instance (visibility VISIBILITY_VOLUME) and no geometry, so the "non-volume" trace misses;
// the volume trace runs __intersection__volume on the instance-transformed ray, then __closesthit__volume_raymarching
// (:178-203, ray jitter compiled out) or __miss__raymarching (:237-246).
static void optixTrace(int, vec3f org, vec3f dir, float tmin, float tmax, float, OptixVisibilityMask mask, int, int ray_type, int, int,
                       uint32_t& u0, uint32_t& u1)
{
  const affine3f wto = get_xfm_wto();
  float t0 = tmin, t1 = tmax;
  if (ray_type == SHADOW_RAY_TYPE) {
    ShadowPayload& data = *(ShadowPayload*)unpack_pointer(u0, u1);
    if (intersect_box(t0, t1, xfmPoint(wto, org), xfmVector(wto, dir), vec3f(0.f), vec3f(1.f))) {
      const float step = g_self->step * 10.f; // "we use a lower sampling rate for shadow"
      raymarching_shadow(*g_self, org, dir, t0, t1, step, data);
      data.t_max = t1 + float_small;
    }
    else
      data.t_max = tmax; // __miss__shadow
    return;
  }
  RayMarchingPayload& data = *(RayMarchingPayload*)unpack_pointer(u0, u1);
  if ((mask.m & VISIBILITY_VOLUME) && intersect_box(t0, t1, xfmPoint(wto, org), xfmVector(wto, dir), vec3f(0.f), vec3f(1.f))) {
    raymarching(*g_self, org, dir, t0, t1, data);
    data.t_max = t1 + float_small;
  }
  else { // __miss__raymarching
    data.t_max = tmax;
    if (mask.m & ~VISIBILITY_VOLUME) {
      data.color = vec3f(0.0, 0.0, 0.0);
      data.alpha = 0.f;
    }
  }
}

// ---- the delta-tracking path tracer: delta_tracking(), pathtracing(), render_pathtracing() (shaders_pathtracing.cu:269-542,
// 603-631) compiled from the reference's text; this namespace supplies that file's ray-type enum and what the pipeline does
// with a PATHTRACING_RAY_TYPE ray ----
namespace pt {
enum RayType { PATHTRACING_RAY_TYPE = 0, RAY_TYPE_COUNT }; // shaders_pathtracing.cu:45
#define USE_DELTA_TRACKING_ITER 1                          // shaders_pathtracing.cu:38-39
#define VARYING_MAJORANT 1
// 1: OptiX semantics -- the instance's world-to-object transform is applied to EVERY traced ray before __intersection__volume
//    clips it, also to the scattered rays, which pathtracing() hands to optixTrace already in object space (:510-513);
// 0: scattered rays are clipped as the object-space rays they are.
// Same meaning as ovr_oracle_scene::pt_child_instance_xfm, which ref_path_rays reads.
static thread_local int g_child_instance_xfm = 1;
static inline int optixGetGASTraversableHandle() { return 0; }
struct OptixVisibilityMask { explicit OptixVisibilityMask(int) {} }; // own types: argument-dependent lookup must not reach the
enum { OPTIX_RAY_FLAG_DISABLE_ANYHIT = 1 };                          // shadow-ray optixTrace of the enclosing namespace
static void optixTrace(int, vec3f org, vec3f dir, float tmin, float tmax, float, OptixVisibilityMask, int, int, int, int, uint32_t& u0,
                       uint32_t& u1);
#include <p_payload.h>
static inline vec3f
#include <spherical_extract.h>
static inline vec3f
#include <uniform_sphere_extract.h>
static inline bool
#include <p_delta_tracking.h>
static inline void
#include <p_pathtracing.h>
static inline void
#include <p_render_pathtracing.h>

// __intersection__volume (shaders_common.h:379-392) on the object-space ray, then __closesthit__volume_pathtracing
// (shaders_pathtracing.cu:548-565: t0 = the reported hit distance, t1 = attribute 1, world ray as passed to optixTrace);
// __miss__pathtracing leaves the payload untouched.
static void optixTrace(int, vec3f org, vec3f dir, float tmin, float tmax, float, OptixVisibilityMask, int, int, int, int, uint32_t& u0,
                       uint32_t& u1)
{
  PathTracingPayload& data = *(PathTracingPayload*)unpack_pointer(u0, u1);
  const affine3f wto = get_xfm_wto();
  const bool as_is = g_child_instance_xfm == 0 && data.wto != nullptr;
  const vec3f o = as_is ? org : xfmPoint(wto, org), d = as_is ? dir : xfmVector(wto, dir);
  float t0 = tmin, t1 = tmax;
  if (intersect_box(t0, t1, o, d, vec3f(0.f), vec3f(1.f))) pathtracing(*g_self, org, dir, t0, t1, data);
}
} // namespace pt

} // namespace ref_march

// One frame's worth of primary rays through the reference's raymarching(): for every listed pixel the ray of
// __raygen__render_frame with one sample per pixel (shaders_raymarching.cu:351-362), the volume hit of
// __intersection__volume / __closesthit__volume_raymarching (:176-203), render_raymarching's final blend over the empty
// background (:260-321). s: the oracle's scene description (same struct the oracle renders from). out: 7 floats per pixel
// (r, g, b, a, gradient xyz).
extern "C" void ref_march_rays(const ovr_oracle_scene* s, const int32_t* pixels_xy, int64_t n_pixels, float* out)
{
  using namespace ref_march;
  g_filter_mode = s->filter_mode;
  DeviceStructuredRegularVolume self;
  self.volume.data = Tex3{ s->voxels, { s->dims[0], s->dims[1], s->dims[2] } };
  self.volume.dims = vec3i(s->dims[0], s->dims[1], s->dims[2]);
  self.volume.lower.v = s->tf_lower, self.volume.upper.v = s->tf_upper;
  self.volume.scale.v = 1.f / (s->tf_upper - s->tf_lower); // volume.cpp:145
  self.tfn.color.data = Tex1{ s->tf_color, s->n_color, 4 }, self.tfn.color.dims.v = s->n_color;
  self.tfn.opacity.data = Tex1{ s->tf_alpha, s->n_alpha, 1 }, self.tfn.opacity.dims.v = s->n_alpha;
  self.step = s->step, self.base = s->base;
  g_self = &self;
  std::memcpy(g_otw, s->obj_to_world, sizeof(g_otw));
  {
    affine3f a;
    a.l.vx = vec3f(g_otw[0], g_otw[4], g_otw[8]), a.l.vy = vec3f(g_otw[1], g_otw[5], g_otw[9]), a.l.vz = vec3f(g_otw[2], g_otw[6], g_otw[10]);
    a.p = vec3f(g_otw[3], g_otw[7], g_otw[11]);
    const affine3f inv = rcp(a);
    const float m[12] = { inv.l.vx.x, inv.l.vy.x, inv.l.vz.x, inv.p.x, inv.l.vx.y, inv.l.vy.y, inv.l.vz.y, inv.p.y,
                          inv.l.vx.z, inv.l.vy.z, inv.l.vz.z, inv.p.z };
    std::memcpy(g_wto, m, sizeof(m));
  }
  LaunchParams& lp = optix_launch_params;
  lp.camera.position = vec3f(s->cam_position[0], s->cam_position[1], s->cam_position[2]);
  lp.camera.direction = vec3f(s->cam_direction[0], s->cam_direction[1], s->cam_direction[2]);
  lp.camera.horizontal = vec3f(s->cam_horizontal[0], s->cam_horizontal[1], s->cam_horizontal[2]);
  lp.camera.vertical = vec3f(s->cam_vertical[0], s->cam_vertical[1], s->cam_vertical[2]);
  lp.last_camera = lp.camera;
  lp.light_directional_pos = vec3f(s->light_pos[0], s->light_pos[1], s->light_pos[2]);
  const vec2f rsize = vec2f(1.f) / vec2f((float)s->width, (float)s->height);
  const int spp = s->spp;
  for (int64_t i = 0; i < n_pixels; i++) {
    // __raygen__render_frame (:323-413): RandomTEA(frame_index, pixel_index), pixel jitter when spp > 1, mean over spp
    const int ix = pixels_xy[2 * i], iy = pixels_xy[2 * i + 1];
    const vec2f screen_coord = vec2f((float)ix + .5f, (float)iy + .5f) * rsize; // compute_screen_position, method 1
    const uint32_t pixel_index = (uint32_t)ix + (uint32_t)iy * (uint32_t)s->width;
    RandomTEA rng_state((uint32_t)s->frame_index, pixel_index);
    float alpha = 0.f;
    vec3f c = 0.f, g = 0.f;
    vec2f flow = 0.f;
    for (int k = 0; k < spp; k++) {
      vec2f screen = screen_coord;
      if (spp > 1) {
        const vec2f pixel_jitter = vec2f(rng_state.get_floats()) - 0.5f;
        screen += pixel_jitter * rsize;
      }
      const vec3f dir = normalize(lp.camera.direction + (screen.x - 0.5f) * lp.camera.horizontal + (screen.y - 0.5f) * lp.camera.vertical);
      render_raymarching(lp.camera.position, dir, &rng_state, alpha, c, g, flow);
    }
    const float rspp = 1.f / spp;
    alpha *= rspp;
    c *= rspp;
    g *= rspp;
    float* o = out + 7 * i;
    o[0] = c.x, o[1] = c.y, o[2] = c.z, o[3] = alpha, o[4] = g.x, o[5] = g.y, o[6] = g.z;
  }
}

// The same over `threads` host threads (0: all hardware threads), pixels handed out in chunks of 64 through an atomic counter:
// the timed CPU arm of bench.py (`--impl reference`, cpu_baseline.kind "reference"). Every worker sets up its own copy of
// the launch state (thread_local above), so the result is ref_march_rays' pixel for pixel.
template<typename Fn> static void run_chunks_mt(int64_t n_pixels, int threads, const Fn& fn)
{
  if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
  const int64_t chunk = 64;
  std::atomic<int64_t> next{ 0 };
  auto work = [&]() {
    for (;;) {
      const int64_t b = next.fetch_add(chunk);
      if (b >= n_pixels) break;
      fn(b, std::min(chunk, n_pixels - b));
    }
  };
  std::vector<std::thread> pool;
  for (int i = 1; i < threads; i++) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
}

extern "C" void ref_march_rays_mt(const ovr_oracle_scene* s, const int32_t* pixels_xy, int64_t n_pixels, float* out, int threads)
{
  run_chunks_mt(n_pixels, threads, [&](int64_t b, int64_t n) { ref_march_rays(s, pixels_xy + 2 * b, n, out + 7 * b); });
}

// One path-traced frame through the reference's render_pathtracing(): for every listed pixel the loop of
// __raygen__render_frame (shaders_pathtracing.cu:633-676: RandomTEA(frame_index, pixel_index), spp samples with pixel jitter when
// spp > 1, mean). s->mc_majorants / mc_dims: the macrocell majorants (ovr_oracle_build_macrocells); spac as
// SpacePartiton_SingleMC::allocate derives it (sp_singlemc.cu:99-104). out: 4 floats per pixel (r, g, b, a), one frame, no
// accumulation. s->pt_child_instance_xfm: see pt::g_child_instance_xfm.
extern "C" void ref_path_rays(const ovr_oracle_scene* s, const int32_t* pixels_xy, int64_t n_pixels, float* out)
{
  using namespace ref_march;
  g_filter_mode = s->filter_mode;
  pt::g_child_instance_xfm = s->pt_child_instance_xfm;
  DeviceStructuredRegularVolume self;
  self.volume.data = Tex3{ s->voxels, { s->dims[0], s->dims[1], s->dims[2] } };
  self.volume.dims = vec3i(s->dims[0], s->dims[1], s->dims[2]);
  self.volume.lower.v = s->tf_lower, self.volume.upper.v = s->tf_upper;
  self.volume.scale.v = 1.f / (s->tf_upper - s->tf_lower); // volume.cpp:145
  self.tfn.color.data = Tex1{ s->tf_color, s->n_color, 4 }, self.tfn.color.dims.v = s->n_color;
  self.tfn.opacity.data = Tex1{ s->tf_alpha, s->n_alpha, 1 }, self.tfn.opacity.dims.v = s->n_alpha;
  self.step = s->step, self.base = s->base;
  self.sp.majorants = const_cast<float*>(s->mc_majorants);
  self.sp.dims = vec3i(s->mc_dims[0], s->mc_dims[1], s->mc_dims[2]);
  self.sp.spac = vec3f((float)DeviceSpacePartiton_SingleMC::MACROCELL_SIZE) / vec3f(self.volume.dims);
  self.sp.spac_rcp = 1.f / self.sp.spac;
  g_self = &self;
  std::memcpy(g_otw, s->obj_to_world, sizeof(g_otw));
  {
    affine3f a;
    a.l.vx = vec3f(g_otw[0], g_otw[4], g_otw[8]), a.l.vy = vec3f(g_otw[1], g_otw[5], g_otw[9]), a.l.vz = vec3f(g_otw[2], g_otw[6], g_otw[10]);
    a.p = vec3f(g_otw[3], g_otw[7], g_otw[11]);
    const affine3f inv = rcp(a);
    const float m[12] = { inv.l.vx.x, inv.l.vy.x, inv.l.vz.x, inv.p.x, inv.l.vx.y, inv.l.vy.y, inv.l.vz.y, inv.p.y,
                          inv.l.vx.z, inv.l.vy.z, inv.l.vz.z, inv.p.z };
    std::memcpy(g_wto, m, sizeof(m));
  }
  LaunchParams& lp = optix_launch_params;
  lp.camera.position = vec3f(s->cam_position[0], s->cam_position[1], s->cam_position[2]);
  lp.camera.direction = vec3f(s->cam_direction[0], s->cam_direction[1], s->cam_direction[2]);
  lp.camera.horizontal = vec3f(s->cam_horizontal[0], s->cam_horizontal[1], s->cam_horizontal[2]);
  lp.camera.vertical = vec3f(s->cam_vertical[0], s->cam_vertical[1], s->cam_vertical[2]);
  lp.last_camera = lp.camera;
  lp.light_ambient_intensity = s->light_ambient;
  lp.max_num_scatters = s->max_num_scatters;
  const vec2f rsize = vec2f(1.f) / vec2f((float)s->width, (float)s->height);
  const int spp = s->spp;
  for (int64_t i = 0; i < n_pixels; i++) {
    const int ix = pixels_xy[2 * i], iy = pixels_xy[2 * i + 1];
    const vec2f screen_coord = vec2f((float)ix + .5f, (float)iy + .5f) * rsize; // compute_screen_position, method 1
    const uint32_t pixel_index = (uint32_t)ix + (uint32_t)iy * (uint32_t)s->width;
    RandomTEA rng_state((uint32_t)s->frame_index, pixel_index);
    float alpha = 0.f;
    vec3f color = 0.f, gradient = 0.f;
    vec2f flow = 0.f;
    for (int k = 0; k < spp; k++) {
      vec2f screen = screen_coord;
      if (spp > 1) {
        const vec2f pixel_jitter = vec2f(rng_state.get_floats()) - 0.5f;
        screen += pixel_jitter * rsize;
      }
      const vec3f ray_dir = normalize(lp.camera.direction + (screen.x - 0.5f) * lp.camera.horizontal + (screen.y - 0.5f) * lp.camera.vertical);
      pt::render_pathtracing(lp.camera.position, ray_dir, &rng_state, alpha, color, gradient, flow);
    }
    const float rspp = 1.f / spp;
    alpha *= rspp;
    color *= rspp;
    float* o = out + 4 * i;
    o[0] = color.x, o[1] = color.y, o[2] = color.z, o[3] = alpha;
  }
}

// ref_path_rays over `threads` host threads (0: all), same chunking as ref_march_rays_mt; pixel results do not depend on it
// (one RandomTEA stream per pixel).
extern "C" void ref_path_rays_mt(const ovr_oracle_scene* s, const int32_t* pixels_xy, int64_t n_pixels, float* out, int threads)
{
  run_chunks_mt(n_pixels, threads, [&](int64_t b, int64_t n) { ref_path_rays(s, pixels_xy + 2 * b, n, out + 4 * b); });
}
