// TEST INFRASTRUCTURE ONLY -- known-answer shim around the REFERENCE's own headers, compiled as host code.
// Nothing here is a copy: the reference headers are #included where they lie under /root/reference
// (see oracle/Makefile, target `ref`); the output goes to oracle/_ref/ (git-ignored).
//   ovr/devices/optix7/accel/dda.h          dda::DDAIter, dda::dda3
//   ovr/common/random/random.h              random::RandomTEA
//   extern/gdt/gdt/math/{vec,mat}.h         normalize, xfmPoint/Vector/Normal, affine inverse
//   ovr/common/generate_mask.cu:55-86       generate_sparse_samples  } kernel bodies cut out at build time and run as
//   ovr/common/random/blue_noise.h:81-101   generate_blue_noise_kernel } host loops (threadIdx/blockIdx are plain structs here)
// tests/golden/make_golden.py drives it to produce the fixtures that pin the oracle.
#include <cuda_runtime.h> // float2 / make_float2 for random.h (host-side CUDA headers only)
#include <dda.h>
namespace ovr { namespace random {
#include <random_tea_extract.h> // struct RandomTEA, cut from ovr/common/random/random.h:146-188 at build time
} }

#include <cstdint>

using namespace ovr;
using namespace gdt;

extern "C" void ref_tea(uint32_t v0, uint32_t v1, int n_draws, float* out_pairs)
{
  ovr::random::RandomTEA rng(v0, v1);
  for (int i = 0; i < n_draws; i++) {
    const float2 f = rng.get_floats();
    out_pairs[2 * i] = f.x;
    out_pairs[2 * i + 1] = f.y;
  }
}

extern "C" int ref_dda3(const float org[3], const float dir[3], float tmin, float tmax, const int32_t grid[3],
                        int max_cells, int32_t* cells, float* t0t1)
{
  int n = 0;
  dda::dda3(vec3f(org[0], org[1], org[2]), vec3f(dir[0], dir[1], dir[2]), tmin, tmax,
            vec3i(grid[0], grid[1], grid[2]), false, [&](const vec3i& c, float t0, float t1) {
              if (n < max_cells) {
                cells[3 * n] = c.x, cells[3 * n + 1] = c.y, cells[3 * n + 2] = c.z;
                t0t1[2 * n] = t0, t0t1[2 * n + 1] = t1;
              }
              n++;
              return true;
            });
  return n;
}

extern "C" int ref_dda_iter(const float org[3], const float dir[3], float tmin, float tmax, const int32_t grid[3],
                            int max_cells, int32_t* cells, float* t0t1)
{
  dda::DDAIter it;
  const vec3f o(org[0], org[1], org[2]), d(dir[0], dir[1], dir[2]);
  const vec3i g(grid[0], grid[1], grid[2]);
  it.init(o, d, tmin, tmax, g);
  int n = 0;
  while (it.next(o, d, tmin, tmax, g, false, [&](const vec3i& c, float t0, float t1) {
    if (n < max_cells) {
      cells[3 * n] = c.x, cells[3 * n + 1] = c.y, cells[3 * n + 2] = c.z;
      t0t1[2 * n] = t0, t0t1[2 * n + 1] = t1;
    }
    n++;
    return true;
  })) {
  }
  return n;
}

// gdt transforms: m = 3x4 row-major; out[0..2] = xfmPoint, [3..5] = xfmVector, [6..8] = xfmNormal,
// [9..11] = normalize(v), [12..23] = inverse(m) as 3x4 row-major
extern "C" void ref_xfm(const float m[12], const float v[3], float* out)
{
  affine3f a;
  a.l.vx = vec3f(m[0], m[4], m[8]);
  a.l.vy = vec3f(m[1], m[5], m[9]);
  a.l.vz = vec3f(m[2], m[6], m[10]);
  a.p = vec3f(m[3], m[7], m[11]);
  const vec3f x(v[0], v[1], v[2]);
  const vec3f p = xfmPoint(a, x), w = xfmVector(a, x), n = xfmNormal(a, x), u = normalize(x);
  out[0] = p.x, out[1] = p.y, out[2] = p.z;
  out[3] = w.x, out[4] = w.y, out[5] = w.z;
  out[6] = n.x, out[7] = n.y, out[8] = n.z;
  out[9] = u.x, out[10] = u.y, out[11] = u.z;
  const affine3f inv = rcp(a);
  out[12] = inv.l.vx.x, out[13] = inv.l.vy.x, out[14] = inv.l.vz.x, out[15] = inv.p.x;
  out[16] = inv.l.vx.y, out[17] = inv.l.vy.y, out[18] = inv.l.vz.y, out[19] = inv.p.y;
  out[20] = inv.l.vx.z, out[21] = inv.l.vy.z, out[22] = inv.l.vz.z, out[23] = inv.p.z;
}

// ---- sparse-sampling mask: the reference's two kernels, executed one "thread" at a time on the host ----
#include <cassert>
#include <cmath>
#include <vector>
namespace ref_mask {
struct Idx { size_t x; };
static Idx threadIdx, blockIdx, blockDim;
static int tile_xy = 128;
static inline float __expf(float x) { return std::exp(x); } // the device intrinsic is approximate; see the tolerance in the tests
#define __restrict__
#define OVR_NOISE_TILE_SIZE_XY ref_mask::tile_xy
#define OVR_NOISE_TILE_SIZE_T 64
static void
#include <sparse_samples_extract.h>
template<typename T> static void
#include <blue_noise_extract.h>
} // namespace ref_mask

// generate_sparse_sampling_mask_d (generate_mask.cu:100-124) without the compaction: xy[2i], xy[2i+1] = pixel or -1
extern "C" void ref_sparse_mask(int width, int height, int frame_index, float* noise, int tile_xy, const float mean[2],
                                float sigma, float base_noise, int32_t* xy)
{
  using namespace ref_mask;
  ref_mask::tile_xy = tile_xy;
  const size_t n = (size_t)width * height;
  std::vector<float> dist(n);
  blockIdx.x = 0, blockDim.x = 0;
  for (size_t i = 0; i < n; i++) {
    threadIdx.x = i;
    generate_blue_noise_kernel<float>((uint32_t)n, noise, dist.data(), (uint32_t)height, (uint32_t)width, frame_index);
  }
  for (size_t i = 0; i < n; i++) {
    threadIdx.x = i;
    generate_sparse_samples(n, xy, dist.data(), vec2i(width, height), 1, vec2f(mean[0], mean[1]), 1.f / (sigma * sigma), base_noise);
  }
}

// ---- delta tracking over the macrocell grid: the reference's DeltaTrackingIter<true> (accel/spatial_partition.h:23-96), the
// struct and the iterator cut out at build time; RandomTEA comes from the extract above ----
#include <cfloat>
namespace ovr { namespace optix7 {
struct range1f { float lower, upper; };
constexpr float float_epsilon = std::numeric_limits<float>::epsilon(); // ovr/common/math_def.h:58
typedef ovr::random::RandomTEA RandomTEA;
#define __NVCC__ 1
#include <singlemc_extract.h>
template<>
#include <delta_iter_extract.h>
#undef __NVCC__
} }

// Repeated hashit calls along one ray, tau re-drawn by the caller after every tentative collision (delta_tracking,
// shaders_pathtracing.cu:385-404): taus[k] is the free path handed to call k. out[4k..] = (found, rayt, majorant, tau after).
extern "C" int ref_delta_hashit(const float* majorants, const int32_t grid[3], const float spac_rcp[3], const float org[3],
                                const float dir[3], float tnear, float tfar, const float* taus, int n_taus, float* out)
{
  using namespace ovr::optix7;
  DeviceSpacePartiton_SingleMC sp;
  sp.majorants = const_cast<float*>(majorants);
  sp.dims = vec3i(grid[0], grid[1], grid[2]);
  sp.spac_rcp = vec3f(spac_rcp[0], spac_rcp[1], spac_rcp[2]);
  const vec3f o(org[0], org[1], org[2]), d(dir[0], dir[1], dir[2]);
  DeviceSpacePartiton_SingleMC::DeltaTrackingIter<true> it(sp, o, d, tnear, tfar);
  RandomTEA rng(0, 0);
  float rayt = tnear, majorant = 0.f;
  int k = 0;
  for (; k < n_taus; k++) {
    float tau = taus[k];
    const bool found = it.hashit(sp, o, d, tnear, tfar, rng, rayt, majorant, tau);
    out[4 * k] = found ? 1.f : 0.f, out[4 * k + 1] = rayt, out[4 * k + 2] = majorant, out[4 * k + 3] = tau;
    if (!found) return k + 1;
  }
  return k;
}

// ---- small shader helpers cut out of ovr/devices/optix7/shaders_common.h at build time: intersect_box (:156-184),
// spherical_to_cartesian + uniform_sample_sphere (:339-354). __frcp_rn is the IEEE reciprocal. ----
namespace ref_shader {
using namespace gdt;
constexpr float float_large = std::numeric_limits<float>::max(); // ovr/common/math_def.h:56-57
constexpr float float_small = std::numeric_limits<float>::min();
static inline float __frcp_rn(float x) { return 1.f / x; }
using std::fabs;
using std::sqrt;
static bool
#include <intersect_box_extract.h>
static vec3f
#include <spherical_extract.h>
static vec3f
#include <uniform_sphere_extract.h>
} // namespace ref_shader

extern "C" int ref_intersect_box(const float org[3], const float dir[3], const float lo[3], const float hi[3], float* t0, float* t1)
{
  return ref_shader::intersect_box(*t0, *t1, vec3f(org[0], org[1], org[2]), vec3f(dir[0], dir[1], dir[2]), vec3f(lo[0], lo[1], lo[2]),
                                   vec3f(hi[0], hi[1], hi[2])) ? 1 : 0;
}

extern "C" void ref_uniform_sphere(float sx, float sy, float out[3])
{
  const vec3f v = ref_shader::uniform_sample_sphere(1.f, vec2f(sx, sy));
  out[0] = v.x, out[1] = v.y, out[2] = v.z;
}

// ---- host-side derivations of DeviceOptix7::Impl, statement ranges cut out of ovr/devices/optix7/device_impl.cpp at build time ----
// commit(): the device camera from the look-at camera (:128-136). out = position, direction, horizontal, vertical.
extern "C" void ref_camera(const float from[3], const float at[3], const float up[3], float fovy_deg, int width, int height, float out[12])
{
  struct { vec3f from, at, up; struct { float fovy; } perspective; } camera;
  camera.from = vec3f(from[0], from[1], from[2]), camera.at = vec3f(at[0], at[1], at[2]), camera.up = vec3f(up[0], up[1], up[2]);
  camera.perspective.fovy = fovy_deg;
  struct DeviceCamera { vec3f position, direction, horizontal, vertical; };
  struct { struct { vec2i size; } frame; DeviceCamera camera, last_camera; } params;
  params.frame.size = vec2i(width, height);
#include <camera_extract.h>
  const vec3f v[4] = { params.camera.position, params.camera.direction, params.camera.horizontal, params.camera.vertical };
  for (int i = 0; i < 4; i++) out[3 * i] = v[i].x, out[3 * i + 1] = v[i].y, out[3 * i + 2] = v[i].z;
}

// buildScene(): object-to-world matrix of the volume instance (:288-296), then its inverse as get_xfm_wto() rebuilds it from
// the instance's 3x4 (gdt rcp). out = otw 3x4 row major, wto 3x4 row major.
extern "C" void ref_volume_matrix(const float grid_origin[3], const float grid_spacing[3], const int32_t dims[3], float out[24])
{
  struct { vec3f grid_origin, grid_spacing; struct D { vec3i dims; } d, *data = &d; } scene_volume;
  scene_volume.grid_origin = vec3f(grid_origin[0], grid_origin[1], grid_origin[2]);
  scene_volume.grid_spacing = vec3f(grid_spacing[0], grid_spacing[1], grid_spacing[2]);
  scene_volume.d.dims = vec3i(dims[0], dims[1], dims[2]);
  struct { affine3f matrix; } v;
#include <volume_matrix_extract.h>
  const affine3f a = v.matrix, inv = rcp(a);
  const affine3f* m[2] = { &a, &inv };
  for (int k = 0; k < 2; k++) {
    const affine3f& x = *m[k];
    const float r[12] = { x.l.vx.x, x.l.vy.x, x.l.vz.x, x.p.x, x.l.vx.y, x.l.vy.y, x.l.vz.y, x.p.y, x.l.vx.z, x.l.vy.z, x.l.vz.z, x.p.z };
    for (int i = 0; i < 12; i++) out[12 * k + i] = r[i];
  }
}

// ---- macrocell value ranges: the reference's value_range_kernel (accel/sp_singlemc.cu:10-54), body cut out at build time and run
// one "thread" per macrocell on the host. The texture is point-read at voxel centres there ((i + 0.5) / dims with a linear
// filter: weight 0), so tex3D is the voxel itself. (majorant_kernel is not run here: its float -> uint32 conversion of -1.f is
// undefined behaviour and gives 0xffffffff on x86 where the GPU saturates to 0 -- SURVEY A.7 q7.) ----
#include <gdt/math/box.h>
namespace ref_mc {
using gdt::range1f;
struct Idx3 { int x, y, z; };
static Idx3 threadIdx, blockIdx, blockDim;
struct Tex { const float* v; int nx, ny, nz; };
typedef const Tex* cudaTextureObject_t;
static inline void tex3D(float* f, cudaTextureObject_t t, float x, float y, float z)
{
  const int ix = std::min(t->nx - 1, std::max(0, (int)std::floor(x * (float)t->nx)));
  const int iy = std::min(t->ny - 1, std::max(0, (int)std::floor(y * (float)t->ny)));
  const int iz = std::min(t->nz - 1, std::max(0, (int)std::floor(z * (float)t->nz)));
  *f = t->v[(size_t)ix + (size_t)t->nx * ((size_t)iy + (size_t)t->ny * (size_t)iz)];
}
static void
#include <value_range_extract.h>
} // namespace ref_mc

extern "C" void ref_value_ranges(const float* voxels, const int32_t dims[3], float* ranges)
{
  using namespace ref_mc;
  const Tex tex{ voxels, dims[0], dims[1], dims[2] };
  const vec3i mc((dims[0] + 15) / 16, (dims[1] + 15) / 16, (dims[2] + 15) / 16); // div_round_up, sp_singlemc.cu:102
  std::vector<range1f> out((size_t)mc.x * mc.y * mc.z);
  ref_mc::blockDim = { 1, 1, 1 }, ref_mc::threadIdx = { 0, 0, 0 };
  for (int z = 0; z < mc.z; z++)
    for (int y = 0; y < mc.y; y++)
      for (int x = 0; x < mc.x; x++) {
        ref_mc::blockIdx = { x, y, z };
        value_range_kernel(vec3i(dims[0], dims[1], dims[2]), &tex, mc, 16u, out.data());
      }
  for (size_t i = 0; i < out.size(); i++) ranges[2 * i] = out[i].lower, ranges[2 * i + 1] = out[i].upper;
}

// ---- integer_normalize (ovr/devices/optix7/array.h:64-106): how a transfer-function range bound in file units is carried into
// texture units for each element type; templates and the switch cut out at build time, ValueType from ovr/scene.h ----
#include <scene.h>
#include <stdexcept>
#include <type_traits>
namespace ref_norm {
using namespace ovr;
#include <integer_normalize_extract.h>
}
// type: 0..7 = UINT8, INT8, UINT16, INT16, UINT32, INT32, FLOAT, DOUBLE (the order of B200VR_TYPE_*)
extern "C" float ref_integer_normalize(float value, int type)
{
  static const ovr::ValueType map[8] = { ovr::VALUE_TYPE_UINT8, ovr::VALUE_TYPE_INT8, ovr::VALUE_TYPE_UINT16, ovr::VALUE_TYPE_INT16,
                                         ovr::VALUE_TYPE_UINT32, ovr::VALUE_TYPE_INT32, ovr::VALUE_TYPE_FLOAT, ovr::VALUE_TYPE_DOUBLE };
  return ref_norm::integer_normalize(value, map[type]);
}
