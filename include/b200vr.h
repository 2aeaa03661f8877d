/* b200vr -- thin C ABI of the B200-native volume-rendering device for Open Volume Renderer (OVR).
 *
 * This is the drop-in boundary beneath OVR's C++ plugin interface: `DeviceB200 : ovr::MainRenderer`
 * (open-volume-renderer_b200/device/device_b200.cpp, loaded by create_renderer("b200") through
 * ovr/renderer.cpp:42-61 + ovr/common/dylink/ObjectFactory.h:35-85) translates each MainRenderer call into the
 * entry points below; the Python host mirror (open-volume-renderer_b200/renderer.py) binds the same ABI with
 * ctypes. Plain pointers and sizes only; every call returns 0 on success or a negative B200VR_E* code, with a
 * message available from b200vr_last_error(). Calls on one handle must be serialised by the caller
 * (same rule as MainRenderer's worker-thread methods, ovr/renderer.h:112-116). There is no CPU fallback: every
 * entry point that computes runs hand-written sm_100a CUDA kernels and fails if no CUDA device is present.
 *
 * Citations (file:line) are relative to the reference tree.
 */
#ifndef B200VR_H
#define B200VR_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B200VR_ABI_VERSION 4

enum {
  B200VR_OK = 0,
  B200VR_EINVAL = -1,  /* bad argument / call order (maps to std::runtime_error in DeviceB200) */
  B200VR_ECUDA = -2,   /* CUDA runtime error, text in b200vr_last_error */
  B200VR_ENODEV = -3,  /* no usable CUDA device: the product path never falls back to the CPU */
  B200VR_ESTATE = -4   /* render before volume / transfer function / framebuffer size are set */
};

/* scalar types a volume can arrive in: ovr::ValueType, ovr/scene.h:32-53 */
enum {
  B200VR_TYPE_UINT8 = 0, B200VR_TYPE_INT8, B200VR_TYPE_UINT16, B200VR_TYPE_INT16,
  B200VR_TYPE_UINT32, B200VR_TYPE_INT32, B200VR_TYPE_FLOAT, B200VR_TYPE_DOUBLE
};

enum { B200VR_MODE_RAYMARCH = 0, B200VR_MODE_PATHTRACE = 1 };        /* set_path_tracing, ovr/renderer.h:187 */
enum { B200VR_SHADING_NONE = 0, B200VR_SHADING_REF = 1 };            /* REF = gradient + shadow rays of the reference */
enum { B200VR_SAMPLER_TEX3D = 0, B200VR_SAMPLER_MANUAL = 1 };        /* hardware filtered 3-D texture | 8-tap ld.global */

typedef struct b200vr b200vr_t;

/* replaces LaunchParams' per-frame switches (ovr/devices/optix7/params.h:55-99) and
 * StructuredRegularVolume::set_sampling_rate (volume.cpp:156-179) */
typedef struct b200vr_params {
  float sampling_rate;      /* step = 1 / sampling_rate (world units) */
  float base;               /* opacity-correction base; <= 0 keeps the default 1 */
  int32_t spp;              /* set_sample_per_pixel */
  int32_t mode;             /* B200VR_MODE_* */
  int32_t shading;          /* B200VR_SHADING_* */
  int32_t sampler;          /* B200VR_SAMPLER_* */
  int32_t skip_empty;       /* 1 = macrocell empty-space skipping (bit-exact), 0 = visit every sample */
  int32_t accumulate;       /* set_frame_accumulation */
  int32_t max_num_scatters; /* 24, params.h:86 */
  int32_t premultiplied;    /* 1 = store (C, A) before the final divide (sort-last partial images) */
  float light_pos[3];       /* fixed directional light, params.h:79 */
  float light_ambient;      /* params.h:80 */
  /* Path tracing, scattered rays. pathtracing() hands them to optixTrace already in object space
   * (shaders_pathtracing.cu:510-513) and OptiX applies the instance's world-to-object transform to them once more before
   * __intersection__volume (shaders_common.h:379-392) clips them, so in the reference their far end is bounded by the
   * macrocell grid, not by the volume: where dims % 16 != 0 they go on colliding with clamp-to-edge voxels in the
   * overhang of the last macrocells. 1 (default) = the reference's behaviour (pinned by the reference's
   * own code run on the host, oracle/ref_march_shim.cpp); 0 = clip them against the volume as the object-space rays they
   * are. Frames are identical under both where every dim is a multiple of 16. */
  int32_t pt_child_instance_xfm;
} b200vr_params;

typedef struct b200vr_stats {
  float ms;                  /* CUDA-event time of the render kernel(s) of the last b200vr_render */
  uint64_t primary_samples;  /* TF-classified primary-march fetches executed */
  uint64_t shadow_samples;   /* shadow-march fetches executed */
  uint64_t gradient_fetches; /* extra trilinear fetches for gradients */
  uint64_t delta_samples;    /* delta-tracking tentative collisions */
  uint64_t rays;
  int32_t frame_index;       /* value the launch saw (device_impl.cpp:241) */
  int32_t kernel_launches;   /* kernels launched by this render call */
} b200vr_stats;

/* lifetime: replaces DeviceOptix7::Impl::init's context creation, device_impl.cpp:70-100,344-387 */
int b200vr_abi_version(void);
int b200vr_create(b200vr_t** out, int cuda_device);
void b200vr_destroy(b200vr_t* h);
const char* b200vr_last_error(const b200vr_t* h); /* h may be NULL: error of the last failed create */
void b200vr_default_params(b200vr_params* p);
/* launch on a caller-owned cudaStream_t (framebuffer_stream, device_impl.cpp:102-111); NULL restores the private stream */
/* NULL selects the handle's own non-blocking stream again. The legacy default stream has the handle 0 = NULL: pass
 * cudaStreamLegacy ((cudaStream_t)0x1) to run on it (this is what torch.cuda.current_stream() is unless a stream context is active). */
int b200vr_set_stream(b200vr_t* h, void* cuda_stream);

/* volume: replaces StructuredRegularVolume::load_from_array3d_scalar (volume.cpp:181-257) =
 * CreateArray3DScalarOptix7 (array.cpp:287-351: min/max, u16/i16/f64 -> f32 conversion, 3-D array + texture)
 * + SpacePartiton_SingleMC::allocate/compute_value_range (accel/sp_singlemc.cu:99-120).
 * voxels: x fastest, nx*ny*nz elements of `type`; host pointer, or device pointer when on_device != 0.
 * obj_to_world: 3x4 row-major = translate(grid_origin) * scale(grid_spacing * dims) (device_impl.cpp:288-296). */
int b200vr_set_volume(b200vr_t* h, const void* voxels, int type, int nx, int ny, int nz,
                      const float obj_to_world[12], int on_device);
/* synthetic volume noise(seed) generated on the device; the brick [org, org+dims) of the infinite
 * lattice field is produced, so sort-last ranks can each create their own brick of a larger volume */
int b200vr_set_volume_noise(b200vr_t* h, uint32_t seed, const int32_t org[3], int nx, int ny, int nz,
                            const float obj_to_world[12]);
int b200vr_get_value_range(b200vr_t* h, float* lo, float* hi); /* original_value_range, volume.cpp:187-189 */
int b200vr_get_macrocells(b200vr_t* h, int32_t mc_dims[3], float* ranges_or_null, float* majorants_or_null);
int b200vr_download_volume(b200vr_t* h, float* host_voxels); /* fp32 copy of the resident volume (tests) */
/* known-answer probe: sample the resident volume at n object-space points (host arrays, 3 floats each) with the given
 * B200VR_SAMPLER_* and classify through the transfer function exactly as the render kernels do
 * (sample_volume_object_space + sample_transfer_function, shaders_common.h:186-193,356-367).
 * values_out: n floats; rgba_out: 4n floats or NULL. */
int b200vr_probe(b200vr_t* h, const float* points, int64_t n, int sampler, float* values_out, float* rgba_out);

/* known-answer probe of the 1-D filter: builds the reference's hardware TF texture (1-D CUDA array, linear filter, clamp,
 * normalised coordinates: CreateArray1DScalarOptix7 / CreateArray1DFloat4Optix7, array.cpp:116-141,217-245) from `table`
 * (n entries of `channels` = 1 or 4 floats) and fetches it at n_u coordinates with tex1D (hw_out), and evaluates the
 * kernels' shared-memory software filter at the same coordinates (sw_out). Both outputs: n_u * channels floats. */
int b200vr_probe_tex1d(b200vr_t* h, const float* table, int n, int channels, const float* u, int64_t n_u,
                       float* hw_out, float* sw_out);

/* fetch-rate micro-benchmark on the resident volume (roofline denominator P_fetch): n_warps bundles of 8x4
 * parallel rays march `iters` unit steps along dir (voxels/step) inside a footprint^3-voxel cube; reports the CUDA-event
 * time and the number of trilinear fetches issued. */
int b200vr_bench_fetch(b200vr_t* h, int sampler, int footprint, int iters, const float dir[3], int ctas_per_sm,
                       float* ms_out, uint64_t* fetches_out);

/* replaces StructuredRegularVolume::set_transfer_function(vector<float> c, vector<float> o, vec2f r)
 * (volume.cpp:104-129, fed by MainRenderer::set_scene, ovr/renderer.h:299-341): colors = 3*n_colors floats RGB,
 * alphas = 2*n_alphas floats (position, alpha) pairs of which only alpha is read, [range_lo, range_hi] = TF value
 * range (ignored when range_hi < range_lo, like set_value_range, volume.cpp:131-145). Rebuilds the majorants
 * (accel/sp_singlemc.cu:122-137). */
int b200vr_set_transfer_function(b200vr_t* h, const float* colors, int n_colors, const float* alphas, int n_alphas,
                                 float range_lo, float range_hi);

/* camera: replaces the camera block of Impl::commit (device_impl.cpp:125-144). set_camera_lookat performs that
 * derivation itself from ovr::Camera {from, at, up, perspective.fovy} and the current framebuffer size. */
int b200vr_set_camera(b200vr_t* h, const float position[3], const float direction[3], const float horizontal[3],
                      const float vertical[3]);
int b200vr_set_camera_lookat(b200vr_t* h, const float from[3], const float at[3], const float up[3], float fovy);
int b200vr_get_camera(b200vr_t* h, float position[3], float direction[3], float horizontal[3], float vertical[3]);

int b200vr_set_params(b200vr_t* h, const b200vr_params* p);
int b200vr_get_params(b200vr_t* h, b200vr_params* p);
int b200vr_resize(b200vr_t* h, int width, int height);   /* framebuffer.resize, device_impl.cpp:116-122 */
int b200vr_reset_accumulation(b200vr_t* h);               /* framebuffer_reset, device_impl.cpp:225-233 */

/* sort-last: clip rays against a sub-box of the object-space unit cube (one brick), default [0,1]^3 */
int b200vr_set_clip_box(b200vr_t* h, const float lo[3], const float hi[3]);
/* sort-last with pixel-matched results: this handle holds ONE BRICK of a larger volume (its own voxels plus ghost layers,
 * placed by obj_to_world). vol_lo..vol_hi is the world-space box of the whole volume, own_lo..own_hi the world-space box of
 * the voxels this brick is responsible for. Every ray walks the sample lattice t_k of the WHOLE volume (it starts where the
 * ray enters vol_lo..vol_hi) and the handle renders exactly the samples whose midpoint lies inside own_lo..own_hi. Both
 * intervals are computed in world space, so that all bricks agree bit for bit on a shared cut plane: the union over the
 * bricks is the sample set of the single-device march (shaders_raymarching.cu:87-171), only early termination is per
 * brick. All four NULL restores the default. Ray marching with SHADING_NONE only (shadow rays leave the brick). */
int b200vr_set_brick(b200vr_t* h, const float own_lo_or_null[3], const float own_hi_or_null[3], const float vol_lo_or_null[3],
                     const float vol_hi_or_null[3]);
/* sort-first: render only the image tiles owned by `rank` of `nranks` (64x32-pixel tiles, round robin) */
int b200vr_set_tile_partition(b200vr_t* h, int rank, int nranks);

/* sparse (foveated) sampling: replaces MainRenderer::set_sparse_sampling / set_focus as the optix7 device consumes them
 * (device_impl.cpp:176-183) and generate_sparse_sampling_mask_d + stream compaction (ovr/common/generate_mask.cu:55-124,
 * device_impl.cpp:327-341). While enabled, every b200vr_render keeps pixel (x, y) with probability
 * p = (1 - base_noise) * exp(-0.5 * (fx^2 + fy^2) / focus_scale^2) + base_noise, fx = x/W - cx, fy = (y/H - cy) / (W/H),
 * decided against noise[(y % tile_xy), (x % tile_xy), frame_index % tile_t] (blue_noise.h:81-101), renders only the kept
 * pixels (compute_screen_position method 3, shaders_common.h:437-449) and, when not accumulating, clears both frame
 * buffers first (device_impl.cpp:234-239). Single-device mode: refused with a tile partition, brick or frame target. */
int b200vr_set_sparse_sampling(b200vr_t* h, int enable, const float focus_center[2], float focus_scale, float base_noise);
/* noise table, tile_xy * tile_xy * tile_t floats laid out [y][x][t] like the reference's embedded stbn_128x128x64 /
 * blue_64x64x64 tables (blue_noise.h:60-101; data/noise/blue_64x64x64.bin). NULL: built-in 128x128x64 white-noise table. */
int b200vr_set_sparse_noise(b200vr_t* h, const float* table_or_null, int tile_xy, int tile_t);
/* pixel list of the last sparse frame: returns its length (or a negative error); xy_out (may be NULL) receives (x, y) int32
 * pairs in launch order -- image-tile order, not the reference's row-major order; the set is the reference's */
int64_t b200vr_get_sparse_samples(b200vr_t* h, int32_t* xy_out_or_null, int64_t capacity_pairs);

/* Allocation-free variant of the sparse-sample list above. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). Same mask, same compaction kernels and same tile-ordered list as a sparse b200vr_render produces
 * (b200vr_get_sparse_samples), but nothing is allocated, freed, copied back or waited for: the three kernels are enqueued on
 * the handle's stream and the call returns. The per-block scratch of the compaction lives in a static device array of
 * B200VR_NOALLOC_MAX_ENTRIES / 1024 words (64 KiB, held by every GPU that loads the library until the process exits), so the
 * frame, padded to whole 64x32-pixel tiles, may hold at most B200VR_NOALLOC_MAX_ENTRIES pixels (a 3840x2160 frame has 8.4 M);
 * larger frames are refused with B200VR_EINVAL. All pointers except focus_center are DEVICE pointers the caller owns:
 * noise_dev = the [y][x][t] table (tile_xy * tile_xy * tile_t floats), xy_out_dev = room for width * height (x, y) int32 pairs
 * (capacity_pairs says how many; the count is only known on the device), count_out_dev = one uint32 that receives the number of
 * pairs written. The handle supplies the device and the stream only; its own sparse-sampling state is neither read nor
 * changed. All calls in a process share the one scratch array of a GPU: calls that overlap on that GPU (different streams or
 * handles) race; calls on one stream are ordered and need no clearing in between. */
#define B200VR_NOALLOC_MAX_ENTRIES (1 << 24)
int b200vr_sparse_samples_noalloc(b200vr_t* h, const float* noise_dev, int tile_xy, int tile_t, int width, int height,
                                  int frame_index, const float focus_center[2], float focus_scale, float base_noise,
                                  int32_t* xy_out_dev, int64_t capacity_pairs, uint32_t* count_out_dev);

/* replaces Impl::render (device_impl.cpp:199-269): one synchronous frame; stats may be NULL */
int b200vr_render(b200vr_t* h, b200vr_stats* stats);
/* The same frame without the wait: b200vr_render_async enqueues the frame on the handle's stream and returns;
 * b200vr_wait blocks until the stream is idle and reports the counters summed over the frames enqueued since the last wait
 * (normally one) and the kernel time and frame index of the last of them (stats may be NULL).
 * b200vr_render == b200vr_render_async + b200vr_wait. For callers that order their own work behind the frame on the same
 * stream (multi-GPU exchange, frame copies) instead of stalling the host once per frame as DeviceOptix7::render does
 * (ovr/devices/optix7/device.cpp:34-43). */
int b200vr_render_async(b200vr_t* h);
int b200vr_wait(b200vr_t* h, b200vr_stats* stats);
/* name of the render kernel the last frame was enqueued with ("raymarch_q_kernel", "raymarch_qc_kernel<J=6>", "raymarch_ring_kernel",
 * "raymarch_ns_kernel", "pathtrace_kernel", ...; "" before the first frame): what the measurement tools put beside a time, the
 * counterpart of reading the OptiX program names off the reference's pipeline (device_impl.cpp:380-460). Owned by the handle. */
const char* b200vr_last_kernel(const b200vr_t* h);
/* Tuning switches of a handle; none changes what is computed per sample, they pick between kernels / launch orders that
 * render the same frame (A/B runs, tests of every kernel variant). Defaults come from the environment variable
 * B200VR_<NAME IN CAPITALS> at b200vr_create. Names: "kernel" (auto | q | qb2 | qb4 | qb8 | qc4 | qc6 | qc8 | qd6 | qg6 | rq2 | rq4 | rq6 | ring | legacy | sm),
 * "tile_order" (-1 auto, 0 rows, 1 centre-out, 2 previous cost, 3 Z-order, 4 | 5 columns, 6 middle rows first, 7 sheets of the
 * light direction), "sheet_px", "max_ctas_per_sm" (0 = all that fit), "ring_high", "rq_self", "rq_cap" (work-sharing rings, csrc/march_rq.cuh), "occ_smem" (0 = macrocell occupancy mask
 * read from global memory even when it would fit in shared memory), "carveout" (shared-memory carve-out in percent, -1 =
 * driver's choice), "tex_native8" (0 = 8-bit volumes held as fp32 arrays), "debug_tiles". Values are decimal integers
 * except for "kernel". */
int b200vr_set_option(b200vr_t* h, const char* name, const char* value);
/* replaces Impl::mapframe / swap (device_impl.cpp:102-111,271-281): device pointers of the current buffer
 * (rgba: W*H float4, grad: W*H float3, row 0 = bottom), valid until the next render/resize on this handle */
int b200vr_map(b200vr_t* h, void** rgba_dev, void** grad_dev);
int b200vr_swap(b200vr_t* h);
/* the same pointers for buffer 0 or 1 of the double buffer, whichever is current (b200vr_current_buffer): the direct-send
 * compositor writes the final frame into the buffer nobody rendered into this frame */
int b200vr_map_index(b200vr_t* h, int buffer, void** rgba_dev, void** grad_dev);
int b200vr_download(b200vr_t* h, float* rgba_host, float* grad_host_or_null);
/* same copy enqueued on a caller-owned cudaStream_t without waiting (pinned host memory): lets the application read frame
 * k while frame k+1 renders into the other buffer after b200vr_swap, the loop of apps/main_app.cpp:244-263
 * (commit -> mapframe -> to_cpu -> swap -> render). The caller synchronises the stream before touching the host arrays. */
int b200vr_download_async(b200vr_t* h, float* rgba_host, float* grad_host_or_null, void* cuda_stream);

/* sort-first exchange helpers: pack the tiles owned by (rank, nranks) of the current frame into a dense slice /
 * scatter a slice received from `rank` back into this handle's frame. Slices are device pointers. */
int64_t b200vr_tile_slice_floats(b200vr_t* h, int rank, int nranks);
int b200vr_pack_tiles(b200vr_t* h, float* slice_dev);
int b200vr_unpack_tiles(b200vr_t* h, const float* slice_dev, int rank, int nranks);
/* sort-last exchange: out = front OVER back on premultiplied float4 pixels (device pointers; `back` may be a peer
 * GPU's buffer mapped over NVLink); finalize divides colour by alpha (alpha_blend, shaders_common.h:329-337) */
int b200vr_composite_over(b200vr_t* h, const float* front_dev, const float* back_dev, float* out_dev,
                          int64_t n_pixels);
int b200vr_finalize_premultiplied(b200vr_t* h, float* rgba_dev, int64_t n_pixels);
/* sort-last, direct-send schedule: out = layers[0] OVER layers[1] OVER ... OVER layers[n-1] (front to back, the same blend
 * applied layer by layer) over one pixel range of 1..16 premultiplied float4 layers in ONE pass -- `layers_dev` is a HOST array of
 * device pointers, normally n-1 of them peer mappings of the other GPUs' frames (read once over NVLink) -- written to out_dev,
 * which may itself be a peer mapping (the display GPU's frame); finalize != 0 un-premultiplies on the way out
 * (alpha_blend, shaders_common.h:329-337). No counterpart in the single-device reference. */
int b200vr_composite_layers(b200vr_t* h, const float* const* layers_dev, int n_layers, float* out_dev, int64_t n_pixels, int finalize);
/* fused sort-last exchange over NVLink (one process per GPU): a rank exports its two rgba frame buffers as CUDA IPC handles
 * (64 bytes each, sent to the peers by any means), opens the peers' handles and passes the resulting device pointers as
 * `back_dev` / `front_dev` of b200vr_composite_over or as `src_dev` of b200vr_copy_device: the blend then reads the
 * partner's half image straight through the peer mapping, no staging copy, no send/recv. Handles stay valid until the next
 * b200vr_resize of the exporting handle. On a caller-provided stream (b200vr_set_stream) composite / finalize / copy are
 * stream-ordered and do not block the host. */
int b200vr_ipc_export_frame(b200vr_t* h, int buffer /* 0 | 1: rgba, 2 | 3: grad */, unsigned char handle_out[64]);
int b200vr_ipc_open(b200vr_t* h, const unsigned char handle[64], void** dev_ptr_out);
int b200vr_ipc_close(b200vr_t* h, void* dev_ptr);
int b200vr_current_buffer(b200vr_t* h); /* index of the buffer b200vr_map returns */
/* Pure host arithmetic of the device set-up, callable without a handle or a GPU (known-answer tests against the reference's own
 * statements, tests/golden/host_derivations.npz): the device camera b200vr_set_camera_lookat derives
 * (DeviceOptix7::Impl::commit, device_impl.cpp:128-136; out = position, direction, horizontal, vertical) and the mapping of a
 * transfer-function range bound given in file units into texture units (integer_normalize(float, ValueType), array.h:92-106,
 * for the element type the texture has after CreateArray3DScalarOptix7, array.cpp:322-351). */
int b200vr_host_camera(const float from[3], const float at[3], const float up[3], float fovy, int width, int height, float out[12]);
float b200vr_host_integer_normalize(float value, int value_type /* B200VR_TYPE_* */);
/* Device group: several GPUs of ONE process behind one renderer, split sort-first (the reference is single-device,
 * device_impl.cpp:370-372 -- this is what a MainRenderer plugin needs to use more than one GPU; the route for device-private
 * options is MainRenderer::init(argc, argv), ovr/renderer.h:252, as the OSPRay device shows, ospray/device_impl.cpp:419).
 * b200vr_group_create makes one handle per entry of cuda_devices (the same device may be named twice), gives member i the image
 * tiles of rank i of n and enables peer access to member 0's GPU. The caller applies every setter (volume, transfer function,
 * camera, params, resize) to EVERY member (b200vr_group_member); b200vr_group_render enqueues the frame on all members -- each
 * renders its tiles straight into member 0's current frame buffer over NVLink peer memory -- and waits for all of them; stats
 * are sums over the members (ms: the slowest member). b200vr_map / b200vr_swap / b200vr_download on member 0 then see the
 * whole frame. Errors of group calls: b200vr_group_last_error (NULL group: the error of a failed b200vr_group_create). */
typedef struct b200vr_group b200vr_group_t;
int b200vr_group_create(b200vr_group_t** out, const int* cuda_devices, int n_devices);
void b200vr_group_destroy(b200vr_group_t* g);
int b200vr_group_size(const b200vr_group_t* g);
b200vr_t* b200vr_group_member(b200vr_group_t* g, int i);
int b200vr_group_render(b200vr_group_t* g, b200vr_stats* stats_or_null);
const char* b200vr_group_last_error(const b200vr_group_t* g_or_null);

/* fused sort-first exchange: render the owned tiles straight into these buffers (W*H float4 and W*H*3 float, normally the
 * display rank's frame opened with b200vr_ipc_open) -- the kernel's own pixel stores are the transfer, there is no pack /
 * gather / unpack step. Both NULL: back to the handle's own double buffer. */
int b200vr_set_frame_target(b200vr_t* h, float* rgba_dev_or_null, float* grad_dev_or_null);
int b200vr_copy_device(b200vr_t* h, void* dst_dev, const void* src_dev, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200VR_H */
