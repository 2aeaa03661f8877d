/* TEST INFRASTRUCTURE ONLY -- CPU oracle for the B200 volume device. The reference tree has no tests, golden images or
 * known-answer vectors for this path; the oracle is pinned by outputs of the reference's own code
 * compiled and run on the host (oracle/_ref: whole ray-marched frames from its raymarching() / raymarching_shadow(), whole
 * path-traced frames from its delta_tracking() / pathtracing() / render_pathtracing(), its dda.h, RandomTEA, DeltaTrackingIter, intersect_box, uniform_sample_sphere, gdt transforms, mask kernels, tfn loader --
 * fixtures under tests/golden/, all bit-exact) and, for the texture unit, by fetches captured on the B200.
 *
 * A CPU restatement (fp32, C++17, no CUDA) of the reference's optix7 shader arithmetic:
 *   ovr/devices/optix7/shaders_raymarching.cu:44-171,260-413      primary + shadow march, pixel finalisation
 *   ovr/devices/optix7/shaders_common.h:96-120,156-236,311-367    clip, sampling, gradient, TF lookup, blend
 *   ovr/devices/optix7/shaders_pathtracing.cu:269-289,385-404,477-542,603-721   delta tracking / path tracing
 *   ovr/devices/optix7/accel/{dda.h:30-148, spatial_partition.h:23-96, sp_singlemc.cu:10-137}
 *   ovr/common/random/random.h:146-188                            RandomTEA
 *   ovr/devices/optix7/device_impl.cpp:125-136                    camera derivation
 * Recipes: oracle/Makefile (target `ref`), oracle/ref_shim.cpp, oracle/ref_march_shim.cpp, oracle/ref_tfn_shim.cpp,
 * oracle/extract_march.py, tests/golden/make_golden.py; also analytic closed forms (tests/test_oracle_analytic.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may load this library.
 */
#ifndef OVR_ORACLE_H
#define OVR_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { OVR_ORACLE_SHADING_NONE = 0, OVR_ORACLE_SHADING_REF = 1 };
/* texture-filter model. HW = the B200 texture unit as measured on the GPU box (integer 8-tap weights split z -> x -> y,
 * oracle_sample.cpp); the other three follow the CUDA Programming Guide text ("Linear Filtering": separable, weight in
 * 9-bit fixed point with 8 fractional bits) and exist for the calibration tests. */
enum { OVR_ORACLE_FILTER_FIXED8_RN = 0, OVR_ORACLE_FILTER_FIXED8_TRUNC = 1, OVR_ORACLE_FILTER_FP32 = 2,
       OVR_ORACLE_FILTER_HW = 3 };
/* macrocell value-range footprint */
enum { OVR_ORACLE_MC_REFERENCE = 0 /* sp_singlemc.cu:36-42 as written (upper apron missing) */,
       OVR_ORACLE_MC_CONSERVATIVE = 1 /* voxels 16c-2 .. 16c+17, what the B200 device builds */ };

typedef struct ovr_oracle_scene {
  /* volume (a2, a7): fp32 voxels, x fastest; object space is the unit cube */
  int32_t dims[3];
  const float* voxels;
  float obj_to_world[12]; /* 3x4 row major = translate(origin) * scale(spacing*dims), device_impl.cpp:288-296 */
  /* transfer function (a8): colour table float4[n_color], opacity table float[n_alpha], lookup window */
  const float* tf_color;
  int32_t n_color;
  const float* tf_alpha;
  int32_t n_alpha;
  float tf_lower, tf_upper;
  /* sampling (volume.cpp:156-179) */
  float step; /* = 1 / volume_sampling_rate */
  float base; /* = 1 */
  /* camera in device form (device_impl.cpp:125-136) */
  float cam_position[3], cam_direction[3], cam_horizontal[3], cam_vertical[3];
  int32_t width, height;
  int32_t spp;
  int32_t frame_index;      /* value seen by the launch (already incremented, device_impl.cpp:241) */
  int32_t shading;          /* OVR_ORACLE_SHADING_* */
  int32_t skip_empty;       /* 0 = the reference algorithm verbatim; 1 = bit-exact macrocell skipping */
  int32_t filter_mode;      /* OVR_ORACLE_FILTER_* */
  int32_t max_num_scatters; /* 24, params.h:86 */
  float light_pos[3];       /* params.h:79 */
  float light_ambient;      /* params.h:80 */
  /* optional sub-box of the unit cube to clip against (sort-last bricks); {0,0,0,1,1,1} for the reference */
  float clip_lo[3], clip_hi[3];
  /* macrocells, built with ovr_oracle_build_macrocells (needed by skip_empty and by path tracing) */
  const float* mc_majorants;
  int32_t mc_dims[3];
  /* path tracing, SURVEY A.7 q14: 0 = scattered rays are clipped against the unit cube as the object-space rays they are;
   * 1 (default, ovr_oracle_default_scene) = as the reference behaves under OptiX: pathtracing() hands the scattered ray to optixTrace already
   * in object space (shaders_pathtracing.cu:510-513) and the instance transform is applied to it a second time before
   * __intersection__volume clips it, so its far end is bounded by the macrocell grid only (differs when dims % 16 != 0). */
  int32_t pt_child_instance_xfm;
} ovr_oracle_scene;

typedef struct ovr_oracle_counters {
  uint64_t primary_samples; /* TF-classified primary-march fetches executed */
  uint64_t shadow_samples;  /* shadow-march fetches executed */
  uint64_t gradient_fetches;/* extra trilinear fetches for gradients */
  uint64_t delta_samples;   /* delta-tracking tentative collisions */
  uint64_t rays;
} ovr_oracle_counters;

void ovr_oracle_default_scene(ovr_oracle_scene* s);

/* a4: camera derivation, fovy in degrees */
void ovr_oracle_camera(const float from[3], const float at[3], const float up[3], float fovy, int width, int height,
                       float position[3], float direction[3], float horizontal[3], float vertical[3]);

/* K1-K5: ray-marched frame. Pixels (x0+i*stride_x, y0+j*stride_y) inside [x0,x1)x[y0,y1) are rendered; the
 * outputs are full-frame arrays (width*height*4 and *3 floats, row 0 = bottom); untouched pixels are left as is.
 * premultiplied != 0 stores (C, A) before the un-premultiply of alpha_blend (used by the sort-last tests). */
void ovr_oracle_render_march(const ovr_oracle_scene* s, int x0, int y0, int x1, int y1, int stride_x, int stride_y,
                             float* rgba, float* grad, int premultiplied, ovr_oracle_counters* counters,
                             int num_threads);

/* K6-K8: one delta-tracking frame (spp samples per pixel); returns per-frame pixel values (no accumulation) */
void ovr_oracle_render_path(const ovr_oracle_scene* s, int x0, int y0, int x1, int y1, int stride_x, int stride_y,
                            float* rgba, ovr_oracle_counters* counters, int num_threads);

/* K9/K10: macrocell value ranges (2 floats per cell) and majorants */
void ovr_oracle_macrocell_dims(const int32_t dims[3], int32_t mc_dims[3]);
void ovr_oracle_build_ranges(const float* voxels, const int32_t dims[3], int mode, float* ranges);
void ovr_oracle_build_majorants(const float* ranges, int64_t n_cells, const float* tf_alpha, int n_alpha,
                                float win_lo, float win_hi, float* majorants);

/* known-answer entry points */
void ovr_oracle_tea(uint32_t v0, uint32_t v1, int n_draws, float* out_pairs, uint32_t* out_state);
int ovr_oracle_dda(const float org[3], const float dir[3], float tmin, float tmax, const int32_t grid[3],
                   int max_cells, int32_t* cells, float* t0t1);
int ovr_oracle_clip(const float org[3], const float dir[3], const float lo[3], const float hi[3], float* t0,
                    float* t1);
void ovr_oracle_sample_volume(const float* voxels, const int32_t dims[3], int filter_mode, const float* pts,
                              int64_t n, float* out);
void ovr_oracle_sample_tf(const ovr_oracle_scene* s, const float* values, int64_t n, float* rgba);
/* 1-D texture fetch model (normalised coordinate, clamp, linear): table of n entries x channels (1 or 4) floats */
void ovr_oracle_tex1d(const float* table, int n, int channels, int filter_mode, const float* u, int64_t n_u, float* out);
/* synthetic procedural volume of SURVEY 8(d): brick [org, org+bdims) of noise(N, seed) */
void ovr_oracle_noise(int32_t n, uint32_t seed, const int32_t org[3], const int32_t bdims[3], float* out,
                      int num_threads);
void ovr_oracle_xfm(const float m[12], const float v[3], float* out24); /* gdt transforms, layout of ref_xfm */
/* sparse-sampling mask (generate_mask.cu:55-86 + blue_noise.h:81-101): keep[x + y*width] in {0,1}, p_out optional */
void ovr_oracle_sparse_mask(int width, int height, int frame_index, const float* noise, int tile_xy, int tile_t,
                            const float mean[2], float sigma, float base_noise, uint8_t* keep, float* p_out);
/* free-path iteration of delta tracking alone (spatial_partition.h:56-96): see oracle_path.cpp */
int ovr_oracle_delta_hashit(const float* majorants, const int32_t grid[3], const float spac_rcp[3], const float org[3],
                            const float dir[3], float tnear, float tfar, const float* taus, int n_taus, float* out);
void ovr_oracle_uniform_sphere(float sx, float sy, float out[3]); /* uniform_sample_sphere(1, s), shaders_common.h:339-354 */
int ovr_oracle_hardware_threads(void);

#ifdef __cplusplus
}
#endif
#endif
