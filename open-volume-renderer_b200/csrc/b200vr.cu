// b200vr -- host side of the thin C ABI declared in include/b200vr.h. Owns all device state of one renderer
// (volume array + texture, linear copy, macrocells, TF tables, double-buffered frame) and launches the sm_100a
// kernels in build.cuh / raymarch.cuh / pathtrace.cuh. No CPU compute path exists: without a CUDA device
// b200vr_create fails with B200VR_ENODEV.
#include "../../include/b200vr.h"

#include "build.cuh"
#include "pathtrace.cuh"
#include "raymarch.cuh"
#include "march_sm.cuh"
#include "march_q.cuh"
#include "march_qb.cuh"
#include "march_qc.cuh"
#include "march_ring.cuh"
#include "march_rq.cuh"
#include "march_ns.cuh"
#include "sparse.cuh"

#include <cub/device/device_radix_sort.cuh>

#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

using namespace b200k;

namespace {
thread_local std::string g_create_error;

// Tuning switches of one handle (b200vr_set_option; defaults from the environment at b200vr_create). None of them changes
// what is computed per sample; they choose between kernels / launch orders that render the same frame.
enum { KERNEL_AUTO = 0, KERNEL_Q, KERNEL_QB2, KERNEL_QB4, KERNEL_QB8, KERNEL_QC4, KERNEL_QC6, KERNEL_QC8, KERNEL_QG6, KERNEL_QD6, KERNEL_RQ6, KERNEL_RQ2, KERNEL_RQ4, KERNEL_RING, KERNEL_LEGACY,
       KERNEL_SM };
enum { TILE_ORDER_AUTO = -1, TILE_ORDER_ROWS = 0, TILE_ORDER_CENTRE = 1, TILE_ORDER_COST = 2, TILE_ORDER_Z = 3, TILE_ORDER_COLS_UP = 4,
       TILE_ORDER_COLS_DOWN = 5, TILE_ORDER_MIDDLE_ROWS = 6, TILE_ORDER_SHEET = 7 };
struct Tuning {
  int kernel = KERNEL_AUTO;
  int tile_order = TILE_ORDER_AUTO;
  int sheet_px = 8;         // width of a sheet strip in pixels (TILE_ORDER_SHEET)
  int max_ctas_per_sm = 0;  // 0: as many as fit
  int ring_high = 4;        // blocks kept ready per CTA for warps that have run out of tiles (march_ring.cuh)
  int rq_self = 1;          // march_rq.cuh: blocks of one tile its warp shades itself before it only publishes
  int rq_cap = 128;          // march_rq.cuh: slots per ring (power of two)
  int occ_smem = -1;        // macrocell occupancy mask in shared memory: -1 when it fits (march_q.cuh Q_OCC_SMEM_WORDS), 0 never
  int carveout = -1;        // cudaFuncAttributePreferredSharedMemoryCarveout in percent; -1: just what the launch needs
  int tex_native8 = 1;      // 8-bit volumes stay 8-bit 3-D arrays
  int debug_tiles = 0;
};

struct LaunchConfig {
  int per_sm = 0; // resident CTAs per SM (occupancy API), cached per (kernel, threads, shared bytes, carve-out)
};
}

struct b200vr {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string error;

  // volume
  int nx = 0, ny = 0, nz = 0;
  float* lin = nullptr;
  cudaArray_t array = nullptr;
  cudaTextureObject_t tex = 0;
  aff3 otw, wto;
  float data_lo = 0.f, data_hi = 0.f; // original_value_range (volume.cpp:187-189)
  bool has_volume = false;
  int vol_type = B200VR_TYPE_FLOAT; // texture element type as the reference would hold it (array.cpp:322-351)
  bool tex_native8 = false;         // the 3-D array holds the original 8-bit voxels (cudaReadModeNormalizedFloat)
  // macrocells
  i3 mc = { 0, 0, 0 };
  float2* ranges = nullptr;
  float* majorants = nullptr;
  float* maj_pad = nullptr;
  unsigned int* occ_bits = nullptr;
  int occ_words = 0;
  int* minmax = nullptr;
  // transfer function
  float4* tf_color = nullptr;
  float* tf_alpha = nullptr;
  int n_color = 0, n_alpha = 0;
  float tf_lower = 0.f, tf_upper = 1.f; // lookup window volume.lower/upper (volume.cpp:131-145)
  bool has_tf = false;
  // camera
  f3 cam_pos, cam_dir, cam_hor, cam_ver;
  bool has_camera = false;
  bool lookat_valid = false;
  float la_from[3], la_at[3], la_up[3], la_fovy = 45.f;
  // params
  b200vr_params prm;
  f3 clip_lo = { 0.f, 0.f, 0.f }, clip_hi = { 1.f, 1.f, 1.f };
  bool has_lattice = false; // sort-last brick: the t lattice starts where the ray enters lat_lo..lat_hi, not the clip box
  f3 lat_lo = { 0.f, 0.f, 0.f }, lat_hi = { 1.f, 1.f, 1.f }; // world space: the whole volume
  f3 own_lo = { 0.f, 0.f, 0.f }, own_hi = { 1.f, 1.f, 1.f }; // world space: the part this handle renders
  int rank = 0, nranks = 1;
  // frame: double buffer {rgba, grad} + accumulation (optix7_common.h:256-414, device_impl.cpp:204-241)
  int width = 0, height = 0;
  float4* rgba[2] = { nullptr, nullptr };
  float* grad[2] = { nullptr, nullptr };
  float4* accum = nullptr;
  int current = 0;
  int frame_index = 0;
  bool reset = true;
  unsigned int* tile_counter = nullptr;
  // launch order of the super tiles (build.cuh tile_order_kernel): cost of the previous frame, cost being collected, order
  unsigned int* tile_cost[2] = { nullptr, nullptr };
  int* tile_order = nullptr;
  int tile_slots = 0;      // allocated slots (super tiles of the whole image)
  int cost_cur = 0;        // tile_cost[cost_cur] is written by the next launch, tile_cost[cost_cur ^ 1] orders it
  float4* target_rgba = nullptr; // sort-first over peer memory: render into these buffers (the display rank's) instead of own
  float* target_grad = nullptr;
  bool cost_valid = false; // false: no history for this image size / partition -> centre-out order
  // deferred-shading ring of the ring kernel (march_ring.cuh), allocated on first use
  unsigned int* ring_ctrl = nullptr;
  unsigned int* ring_seq = nullptr;
  float4* ring_data = nullptr;
  int ring_cap = 0, ring_count = 0, ring_block = 0; // slots per ring, rings, float4 per slot
  unsigned long long* acc = nullptr; // per pixel fixed-point sums (allocated with the frame)
  float* ray_alpha = nullptr;
  Counters* counters = nullptr;
  // sparse (foveated) sampling, sparse.cuh (device_impl.cpp:180-183,236-238,327-341)
  bool sparse = false;
  float focus_x = 0.5f, focus_y = 0.5f, focus_scale = 0.06f, base_noise = 0.07f; // apps/main_app.cpp:122-124
  float* sp_noise = nullptr;
  int sp_tile_xy = 0, sp_tile_t = 0;
  int2* sp_xy = nullptr;           // compacted pixel list of the last sparse frame
  unsigned int* sp_blocks = nullptr; // per 1024-entry block: count, then exclusive offset
  unsigned int* sp_count = nullptr;
  int sp_alloc_px = 0;             // pixels sp_xy was allocated for
  int sp_last_count = -1;          // entries of sp_xy after the last sparse render (-1: none yet)
  // tuning + cached launch configurations
  Tuning tune;
  std::map<std::tuple<const void*, int, size_t, int>, LaunchConfig> launch_cache;
  // sheet order of the warp tiles (build.cuh sheet_key_kernel): keys / values before and after the sort, sort scratch
  unsigned int* wt_keys[2] = { nullptr, nullptr };
  unsigned int* wt_vals[2] = { nullptr, nullptr };
  void* wt_tmp = nullptr;
  size_t wt_tmp_bytes = 0;
  int wt_alloc = 0;                // entries allocated
  int wt_count = 0;                // entries of the valid table (0: none)
  f3 wt_dir, wt_hor, wt_ver, wt_light; // camera / light the valid table was built for
  int wt_rank = -1, wt_nranks = -1, wt_w = 0, wt_h = 0, wt_px = 0;
  // last frame (b200vr_render_async -> b200vr_wait)
  int last_launches = 0, last_grid = 0, last_warps_per_cta = 1;
  int order_mode_done = -1;        // super-tile order mode h->tile_order currently holds,
  int order_n_done = 0;            // and for how many owned super tiles
  bool last_ring = false;
  std::string last_kernel;         // name of the render kernel of the last enqueued frame
  bool pending = false;            // a frame was enqueued and its counters have not been fetched
};

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      h->error = std::string(#call) + ": " + cudaGetErrorString(e_);                               \
      return B200VR_ECUDA;                                                                         \
    }                                                                                              \
  } while (0)

static int fail(b200vr_t* h, int code, const char* msg)
{
  h->error = msg;
  return code;
}

static int grid_for(size_t n, int threads, int sm_count)
{
  const size_t want = (n + threads - 1) / threads;
  const size_t cap = (size_t)sm_count * 16;
  return (int)std::max<size_t>(1, std::min(want, cap));
}

// ---- tuning options -------------------------------------------------------------------------------------------------
static const char* const kOptionNames[] = { "kernel", "tile_order", "sheet_px", "max_ctas_per_sm", "ring_high", "rq_self", "rq_cap", "occ_smem", "carveout",
                                            "tex_native8", "debug_tiles" };

static int apply_option(b200vr_t* h, const std::string& name, const char* value)
{
  Tuning& t = h->tune;
  if (name == "kernel") {
    static const std::pair<const char*, int> names[] = { { "auto", KERNEL_AUTO }, { "q", KERNEL_Q }, { "qb2", KERNEL_QB2 }, { "qb4", KERNEL_QB4 },
                                                         { "qb8", KERNEL_QB8 }, { "qc4", KERNEL_QC4 }, { "qc6", KERNEL_QC6 }, { "qc8", KERNEL_QC8 }, { "qg6", KERNEL_QG6 }, { "qd6", KERNEL_QD6 }, { "rq6", KERNEL_RQ6 }, { "rq2", KERNEL_RQ2 }, { "rq4", KERNEL_RQ4 }, { "ring", KERNEL_RING }, { "legacy", KERNEL_LEGACY }, { "sm", KERNEL_SM } };
    for (const auto& n : names)
      if (std::strcmp(value, n.first) == 0) {
        t.kernel = n.second;
        return B200VR_OK;
      }
    return fail(h, B200VR_EINVAL, "set_option: kernel is one of auto, q, qb2, qb4, qb8, qc4, qc6, qc8, qd6, qg6, rq2, rq4, rq6, ring, legacy, sm");
  }
  char* end = nullptr;
  const long v = std::strtol(value, &end, 10);
  if (end == value || *end != 0) return fail(h, B200VR_EINVAL, "set_option: value is not an integer");
  if (name == "tile_order") {
    if (v < TILE_ORDER_AUTO || v > TILE_ORDER_SHEET) return fail(h, B200VR_EINVAL, "set_option: tile_order is -1 (auto) or 0..7");
    t.tile_order = (int)v;
    h->wt_count = 0;
  }
  else if (name == "sheet_px") {
    if (v < 1 || v > 4096) return fail(h, B200VR_EINVAL, "set_option: sheet_px out of range");
    t.sheet_px = (int)v;
    h->wt_count = 0;
  }
  else if (name == "max_ctas_per_sm") t.max_ctas_per_sm = (int)std::max(0l, v);
  else if (name == "ring_high") t.ring_high = (int)std::max(1l, v);
  else if (name == "rq_self") t.rq_self = (int)std::max(0l, v);
  else if (name == "rq_cap") {
    if (v < 32 || v > 1024 || (v & (v - 1))) return fail(h, B200VR_EINVAL, "set_option: rq_cap is a power of two in 32..1024");
    t.rq_cap = (int)v;
  }
  else if (name == "occ_smem") t.occ_smem = v == 0 ? 0 : -1;
  else if (name == "carveout") {
    if (v < -1 || v > 100) return fail(h, B200VR_EINVAL, "set_option: carveout is -1 or a percentage");
    t.carveout = (int)v;
  }
  else if (name == "tex_native8") t.tex_native8 = v != 0;
  else if (name == "debug_tiles") t.debug_tiles = v != 0;
  else return fail(h, B200VR_EINVAL, "set_option: unknown option");
  return B200VR_OK;
}

extern "C" int b200vr_set_option(b200vr_t* h, const char* name, const char* value)
{
  if (!h) return B200VR_EINVAL;
  if (!name || !value) return fail(h, B200VR_EINVAL, "set_option: NULL argument");
  return apply_option(h, name, value);
}

// defaults of a new handle: B200VR_<OPTION> in the environment (A/B runs without touching the caller)
static void options_from_env(b200vr_t* h)
{
  for (const char* name : kOptionNames) {
    std::string env = "B200VR_";
    for (const char* c = name; *c; c++) env += (char)std::toupper((unsigned char)*c);
    if (const char* v = std::getenv(env.c_str()))
      if (apply_option(h, name, v) != B200VR_OK) std::fprintf(stderr, "[b200vr] ignoring %s=%s: %s\n", env.c_str(), v, h->error.c_str());
  }
  h->error.clear();
}

extern "C" int b200vr_abi_version(void) { return B200VR_ABI_VERSION; }

extern "C" void b200vr_default_params(b200vr_params* p)
{
  std::memset(p, 0, sizeof(*p));
  p->sampling_rate = 1.f;
  p->base = 1.f;
  p->spp = 1;
  p->mode = B200VR_MODE_RAYMARCH;
  p->shading = B200VR_SHADING_REF;
  p->sampler = B200VR_SAMPLER_TEX3D;
  p->skip_empty = 1;
  p->accumulate = 0;
  p->max_num_scatters = 24;
  p->premultiplied = 0;
  p->light_pos[0] = -907.108f; // params.h:79
  p->light_pos[1] = 2205.875f;
  p->light_pos[2] = -400.0267f;
  p->light_ambient = 1.f;
  p->pt_child_instance_xfm = 1; // the reference's behaviour
}

extern "C" const char* b200vr_last_error(const b200vr_t* h) { return h ? h->error.c_str() : g_create_error.c_str(); }

extern "C" int b200vr_create(b200vr_t** out, int cuda_device)
{
  if (!out) return B200VR_EINVAL;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    g_create_error = std::string("no CUDA device available (") + cudaGetErrorString(e) + "); b200vr has no CPU fallback";
    return B200VR_ENODEV;
  }
  if (cuda_device < 0 || cuda_device >= n) {
    g_create_error = "cuda_device out of range";
    return B200VR_EINVAL;
  }
  b200vr_t* h = new b200vr();
  h->device = cuda_device;
  b200vr_default_params(&h->prm);
  options_from_env(h);
  auto bail = [&](cudaError_t err, const char* what) {
    g_create_error = std::string(what) + ": " + cudaGetErrorString(err);
    delete h;
    return B200VR_ECUDA;
  };
  if ((e = cudaSetDevice(cuda_device)) != cudaSuccess) return bail(e, "cudaSetDevice");
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, cuda_device)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
  h->sm_count = prop.multiProcessorCount;
  if ((e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
  h->stream = h->own_stream;
  if ((e = cudaEventCreate(&h->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if ((e = cudaEventCreate(&h->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if ((e = cudaMalloc(&h->tile_counter, sizeof(unsigned int))) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = cudaMalloc(&h->counters, sizeof(Counters))) != cudaSuccess) return bail(e, "cudaMalloc");
  if ((e = cudaMalloc(&h->minmax, 2 * sizeof(int))) != cudaSuccess) return bail(e, "cudaMalloc");
  *out = h;
  return B200VR_OK;
}

static void free_volume(b200vr_t* h)
{
  if (h->tex) cudaDestroyTextureObject(h->tex);
  if (h->array) cudaFreeArray(h->array);
  cudaFree(h->lin);
  cudaFree(h->ranges);
  cudaFree(h->majorants);
  cudaFree(h->maj_pad);
  h->maj_pad = nullptr;
  cudaFree(h->occ_bits);
  h->occ_bits = nullptr;
  h->tex = 0;
  h->array = nullptr;
  h->lin = nullptr;
  h->ranges = nullptr;
  h->majorants = nullptr;
  h->has_volume = false;
}
static void free_frame(b200vr_t* h)
{
  for (int i = 0; i < 2; i++) {
    cudaFree(h->rgba[i]);
    cudaFree(h->grad[i]);
    h->rgba[i] = nullptr;
    h->grad[i] = nullptr;
  }
  cudaFree(h->accum);
  h->accum = nullptr;
  cudaFree(h->sp_xy);
  cudaFree(h->sp_blocks);
  h->sp_xy = nullptr, h->sp_blocks = nullptr, h->sp_alloc_px = 0, h->sp_last_count = -1;
}

extern "C" void b200vr_destroy(b200vr_t* h)
{
  if (!h) return;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  free_volume(h);
  free_frame(h);
  cudaFree(h->tf_color);
  cudaFree(h->tf_alpha);
  cudaFree(h->tile_counter);
  cudaFree(h->tile_cost[0]);
  cudaFree(h->tile_cost[1]);
  cudaFree(h->tile_order);
  cudaFree(h->ring_ctrl);
  cudaFree(h->ring_seq);
  cudaFree(h->ring_data);
  cudaFree(h->acc);
  cudaFree(h->ray_alpha);
  cudaFree(h->counters);
  cudaFree(h->minmax);
  cudaFree(h->sp_noise);
  cudaFree(h->sp_count);
  for (int i = 0; i < 2; i++) cudaFree(h->wt_keys[i]), cudaFree(h->wt_vals[i]);
  cudaFree(h->wt_tmp);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  delete h;
}

// ---------------------------------------------------------------------------------------------------------------------
// volume
// ---------------------------------------------------------------------------------------------------------------------

static int rebuild_majorants(b200vr_t* h);

// h->lin holds nx*ny*nz fp32 voxels: build the 3-D array + texture (array.cpp:287-309, cuda_buffer.h:248-331) and the
// macrocell ranges / global range (sp_singlemc.cu:99-120, array.cpp:27-66)
// native8: device copy of the original 8-bit voxels (or nullptr). The reference keeps 8-bit volumes as 8-bit 3-D arrays read
// through cudaReadModeNormalizedFloat (array.cpp:300-306); so does this device: a 32-byte sector of the texture then holds four
// times as many texels, which is what the L1TEX tag stage (4 sectors/clk/SM, the limiter of the march kernels, DESIGN.md
// section 5) and HBM see. Option tex_native8 = 0 forces the fp32 array (A/B runs). The fp32 copy `lin` is kept either way.
static int finish_volume(b200vr_t* h, const float obj_to_world[12], const void* native8 = nullptr, bool native8_signed = false)
{
  if (!h->tune.tex_native8) native8 = nullptr;
  const cudaChannelFormatDesc desc = !native8 ? cudaCreateChannelDesc<float>()
                                     : native8_signed ? cudaCreateChannelDesc<signed char>() : cudaCreateChannelDesc<unsigned char>();
  const cudaExtent ext = make_cudaExtent((size_t)h->nx, (size_t)h->ny, (size_t)h->nz);
  CK(cudaMalloc3DArray(&h->array, &desc, ext));
  cudaMemcpy3DParms cp;
  std::memset(&cp, 0, sizeof(cp));
  if (native8)
    cp.srcPtr = make_cudaPitchedPtr(const_cast<void*>(native8), (size_t)h->nx, (size_t)h->nx, (size_t)h->ny);
  else
    cp.srcPtr = make_cudaPitchedPtr(h->lin, (size_t)h->nx * sizeof(float), (size_t)h->nx, (size_t)h->ny);
  cp.dstArray = h->array;
  cp.extent = ext;
  cp.kind = cudaMemcpyDeviceToDevice;
  CK(cudaMemcpy3DAsync(&cp, h->stream));
  cudaResourceDesc rd;
  std::memset(&rd, 0, sizeof(rd));
  rd.resType = cudaResourceTypeArray;
  rd.res.array.array = h->array;
  cudaTextureDesc td;
  std::memset(&td, 0, sizeof(td));
  td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModeLinear;
  td.readMode = native8 ? cudaReadModeNormalizedFloat : cudaReadModeElementType;
  td.normalizedCoords = 1;
  CK(cudaCreateTextureObject(&h->tex, &rd, &td, nullptr));
  h->tex_native8 = native8 != nullptr;

  h->otw = from_3x4(obj_to_world);
  h->wto = inverse(h->otw);

  h->mc.x = (h->nx + MC_SIZE - 1) / MC_SIZE;
  h->mc.y = (h->ny + MC_SIZE - 1) / MC_SIZE;
  h->mc.z = (h->nz + MC_SIZE - 1) / MC_SIZE;
  const size_t n_cells = (size_t)h->mc.x * h->mc.y * h->mc.z;
  CK(cudaMalloc(&h->ranges, n_cells * sizeof(float2)));
  CK(cudaMalloc(&h->majorants, n_cells * sizeof(float)));
  CK(cudaMemsetAsync(h->majorants, 0, n_cells * sizeof(float), h->stream));
  const size_t n_pad = (size_t)(h->mc.x + 2) * (h->mc.y + 2) * (h->mc.z + 2);
  CK(cudaMalloc(&h->maj_pad, n_pad * sizeof(float)));
  CK(cudaMemsetAsync(h->maj_pad, 0, n_pad * sizeof(float), h->stream));
  h->occ_words = (int)((n_pad + 31) / 32);
  CK(cudaMalloc(&h->occ_bits, (size_t)h->occ_words * sizeof(unsigned int)));
  CK(cudaMemsetAsync(h->occ_bits, 0, (size_t)h->occ_words * sizeof(unsigned int), h->stream));
  const int init[2] = { 0x7fffffff, (int)0x80000000 };
  CK(cudaMemcpyAsync(h->minmax, init, sizeof(init), cudaMemcpyHostToDevice, h->stream));
  macrocell_range_kernel<<<(unsigned)(h->mc.y * h->mc.z), RANGE_THREADS, 2 * (size_t)h->mc.x * sizeof(int), h->stream>>>(
    h->lin, h->nx, h->ny, h->nz, h->mc.x, h->mc.y, h->mc.z, h->ranges, h->minmax);
  CK(cudaGetLastError());
  int mm[2];
  CK(cudaMemcpyAsync(mm, h->minmax, sizeof(mm), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  h->data_lo = ordered_to_float(mm[0]);
  h->data_hi = ordered_to_float(mm[1]);
  // load_from_array3d_scalar calls set_value_range(1,-1): lookup window = data range until a TF arrives
  if (!h->has_tf) {
    h->tf_lower = h->data_lo;
    h->tf_upper = h->data_hi;
  }
  h->has_volume = true;
  h->reset = true;
  if (h->has_tf) return rebuild_majorants(h);
  return B200VR_OK;
}

// keep != nullptr: the device copy of the typed voxels is handed back (*keep, *owned says whether the caller frees it)
template<typename T>
static int upload_typed(b200vr_t* h, const void* voxels, size_t n, int on_device, const void** keep = nullptr, bool* owned = nullptr)
{
  const T* src = (const T*)voxels;
  T* staged = nullptr;
  cudaError_t e = cudaSuccess;
  if (!on_device) {
    e = cudaMalloc(&staged, n * sizeof(T));
    if (e == cudaSuccess) e = cudaMemcpyAsync(staged, voxels, n * sizeof(T), cudaMemcpyHostToDevice, h->stream);
    src = staged;
  }
  if (e == cudaSuccess) {
    convert_kernel<T><<<grid_for(n, 256, h->sm_count), 256, 0, h->stream>>>(h->lin, src, n);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) {
    cudaFree(staged);
    h->error = std::string("upload_typed: ") + cudaGetErrorString(e);
    return B200VR_ECUDA;
  }
  if (keep) {
    *keep = src;
    *owned = staged != nullptr;
  }
  else if (staged) CK(cudaFree(staged));
  return B200VR_OK;
}

extern "C" int b200vr_set_volume(b200vr_t* h, const void* voxels, int type, int nx, int ny, int nz,
                                 const float obj_to_world[12], int on_device)
{
  if (!h) return B200VR_EINVAL;
  if (!voxels || !obj_to_world || nx <= 0 || ny <= 0 || nz <= 0) return fail(h, B200VR_EINVAL, "set_volume: bad argument");
  CK(cudaSetDevice(h->device));
  free_volume(h);
  h->nx = nx, h->ny = ny, h->nz = nz;
  const size_t n = (size_t)nx * ny * nz;
  CK(cudaMalloc(&h->lin, n * sizeof(float)));
  int rc = B200VR_OK;
  const void* native8 = nullptr;
  bool native8_owned = false;
  // 16-bit integers and doubles are converted to plain floats by the reference; 8/32-bit integers stay normalised
  h->vol_type = (type == B200VR_TYPE_UINT16 || type == B200VR_TYPE_INT16 || type == B200VR_TYPE_DOUBLE) ? B200VR_TYPE_FLOAT : type;
  switch (type) {
  case B200VR_TYPE_FLOAT:
    CK(cudaMemcpyAsync(h->lin, voxels, n * sizeof(float), on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                       h->stream));
    break;
  case B200VR_TYPE_UINT8: rc = upload_typed<uint8_t>(h, voxels, n, on_device, &native8, &native8_owned); break;
  case B200VR_TYPE_INT8: rc = upload_typed<int8_t>(h, voxels, n, on_device, &native8, &native8_owned); break;
  case B200VR_TYPE_UINT16: rc = upload_typed<uint16_t>(h, voxels, n, on_device); break;
  case B200VR_TYPE_INT16: rc = upload_typed<int16_t>(h, voxels, n, on_device); break;
  case B200VR_TYPE_UINT32: rc = upload_typed<uint32_t>(h, voxels, n, on_device); break;
  case B200VR_TYPE_INT32: rc = upload_typed<int32_t>(h, voxels, n, on_device); break;
  case B200VR_TYPE_DOUBLE: rc = upload_typed<double>(h, voxels, n, on_device); break;
  default: return fail(h, B200VR_EINVAL, "set_volume: unknown value type"); // array.cpp:347
  }
  if (rc != B200VR_OK) return rc;
  rc = finish_volume(h, obj_to_world, native8, type == B200VR_TYPE_INT8); // synchronises the stream before it returns
  if (native8_owned) cudaFree(const_cast<void*>(native8));
  return rc;
}

extern "C" int b200vr_set_volume_noise(b200vr_t* h, uint32_t seed, const int32_t org[3], int nx, int ny, int nz,
                                       const float obj_to_world[12])
{
  if (!h) return B200VR_EINVAL;
  if (!org || !obj_to_world || nx <= 0 || ny <= 0 || nz <= 0) return fail(h, B200VR_EINVAL, "set_volume_noise: bad argument");
  CK(cudaSetDevice(h->device));
  free_volume(h);
  h->nx = nx, h->ny = ny, h->nz = nz;
  const size_t n = (size_t)nx * ny * nz;
  CK(cudaMalloc(&h->lin, n * sizeof(float)));
  h->vol_type = B200VR_TYPE_FLOAT;
  noise_kernel<<<grid_for(n, 256, h->sm_count), 256, 0, h->stream>>>(h->lin, seed, org[0], org[1], org[2], nx, ny, nz);
  CK(cudaGetLastError());
  return finish_volume(h, obj_to_world);
}

extern "C" int b200vr_get_value_range(b200vr_t* h, float* lo, float* hi)
{
  if (!h || !lo || !hi) return B200VR_EINVAL;
  if (!h->has_volume) return fail(h, B200VR_ESTATE, "get_value_range: no volume");
  *lo = h->data_lo;
  *hi = h->data_hi;
  return B200VR_OK;
}

extern "C" int b200vr_get_macrocells(b200vr_t* h, int32_t mc_dims[3], float* ranges, float* majorants)
{
  if (!h || !mc_dims) return B200VR_EINVAL;
  if (!h->has_volume) return fail(h, B200VR_ESTATE, "get_macrocells: no volume");
  CK(cudaSetDevice(h->device));
  mc_dims[0] = h->mc.x, mc_dims[1] = h->mc.y, mc_dims[2] = h->mc.z;
  const size_t n_cells = (size_t)h->mc.x * h->mc.y * h->mc.z;
  if (ranges) CK(cudaMemcpy(ranges, h->ranges, n_cells * sizeof(float2), cudaMemcpyDeviceToHost));
  if (majorants) CK(cudaMemcpy(majorants, h->majorants, n_cells * sizeof(float), cudaMemcpyDeviceToHost));
  return B200VR_OK;
}

extern "C" int b200vr_download_volume(b200vr_t* h, float* host_voxels)
{
  if (!h || !host_voxels) return B200VR_EINVAL;
  if (!h->has_volume) return fail(h, B200VR_ESTATE, "download_volume: no volume");
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpy(host_voxels, h->lin, (size_t)h->nx * h->ny * h->nz * sizeof(float), cudaMemcpyDeviceToHost));
  return B200VR_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// transfer function, camera, parameters
// ---------------------------------------------------------------------------------------------------------------------

// integer_normalize(float, ValueType) (array.h:64-106): the TF range arrives in file units
static float integer_normalize(float v, int type)
{
  switch (type) {
  case B200VR_TYPE_UINT8: return (float)(uint8_t)v / 255.f;
  case B200VR_TYPE_INT8: return std::max((float)(int8_t)v / 127.f, -1.f);
  case B200VR_TYPE_UINT32: return (float)(uint32_t)v / 4294967295.f;
  case B200VR_TYPE_INT32: return std::max((float)(int32_t)v / 2147483647.f, -1.f);
  default: return v;
  }
}

static int rebuild_majorants(b200vr_t* h)
{
  const int n_cells = h->mc.x * h->mc.y * h->mc.z;
  const int threads = 256;
  const int grid = std::max(1, std::min((n_cells + threads - 1) / threads, h->sm_count * 8));
  const size_t maj_smem = (size_t)h->n_alpha * sizeof(float);
  if (maj_smem > 48 * 1024) CK(cudaFuncSetAttribute((const void*)majorant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)maj_smem));
  majorant_kernel<<<grid, threads, maj_smem, h->stream>>>(h->ranges, n_cells, h->tf_alpha, h->n_alpha, h->tf_lower, h->tf_upper, h->majorants);
  CK(cudaGetLastError());
  const int n_pad = (h->mc.x + 2) * (h->mc.y + 2) * (h->mc.z + 2);
  majorant_pad_kernel<<<std::max(1, std::min((n_pad + threads - 1) / threads, h->sm_count * 8)), threads, 0, h->stream>>>(
    h->majorants, h->mc.x, h->mc.y, h->mc.z, h->maj_pad);
  CK(cudaGetLastError());
  occupancy_bits_kernel<<<std::max(1, (h->occ_words + threads - 1) / threads), threads, 0, h->stream>>>(h->maj_pad, n_pad, h->occ_bits,
                                                                                                            h->occ_words);
  CK(cudaGetLastError());
  return B200VR_OK; // stream-ordered: the next render on this stream sees the new majorants
}

extern "C" int b200vr_set_transfer_function(b200vr_t* h, const float* colors, int n_colors, const float* alphas, int n_alphas,
                                            float range_lo, float range_hi)
{
  if (!h) return B200VR_EINVAL;
  if (!colors || !alphas || n_colors <= 0 || n_alphas <= 0) return fail(h, B200VR_EINVAL, "set_transfer_function: empty table");
  // every march kernel stages the opacity table in shared memory (8 B per entry as value / slope pairs next to 16-72 KB of task
  // queues, or 4 B + 16 B per colour entry in the older kernels): refuse here what no kernel could launch with
  if (q_alpha_bytes(n_alphas) + qb_queue_bytes<8>() > 200 * 1024 || tf_fast_shared_bytes(n_colors, n_alphas) > 200 * 1024)
    return fail(h, B200VR_EINVAL, "set_transfer_function: table too large for shared memory (at most ~16000 opacity / ~8000 colour entries)");
  CK(cudaSetDevice(h->device));
  // volume.cpp:104-121: colours -> float4(r,g,b,1); alphas = every second float of the (position, alpha) pairs
  std::vector<float4> c((size_t)n_colors);
  for (int i = 0; i < n_colors; i++) c[i] = make_float4(colors[3 * i], colors[3 * i + 1], colors[3 * i + 2], 1.f);
  std::vector<float> a((size_t)n_alphas);
  for (int i = 0; i < n_alphas; i++) a[i] = alphas[2 * i + 1];
  if (n_colors != h->n_color || n_alphas != h->n_alpha || !h->tf_color) { // same-size edits (the common case) reuse the tables
    cudaFree(h->tf_color);
    cudaFree(h->tf_alpha);
    h->tf_color = nullptr;
    h->tf_alpha = nullptr;
    CK(cudaMalloc(&h->tf_color, c.size() * sizeof(float4)));
    CK(cudaMalloc(&h->tf_alpha, a.size() * sizeof(float)));
  }
  // pageable sources: the runtime stages them before returning, the copies themselves run in stream order
  CK(cudaMemcpyAsync(h->tf_color, c.data(), c.size() * sizeof(float4), cudaMemcpyHostToDevice, h->stream));
  CK(cudaMemcpyAsync(h->tf_alpha, a.data(), a.size() * sizeof(float), cudaMemcpyHostToDevice, h->stream));
  h->n_color = n_colors;
  h->n_alpha = n_alphas;
  // set_value_range (volume.cpp:131-145): only a non-empty range replaces the lookup window
  if (range_hi >= range_lo) {
    h->tf_upper = integer_normalize(range_hi, h->vol_type);
    h->tf_lower = integer_normalize(range_lo, h->vol_type);
  }
  h->has_tf = true;
  h->reset = true;
  if (h->has_volume) return rebuild_majorants(h);
  return B200VR_OK;
}

static void derive_camera_values(const float la_from[3], const float la_at[3], const float la_up[3], float fovy, int width, int height,
                                 f3& pos, f3& dir, f3& hor, f3& ver)
{
  // device_impl.cpp:125-136; tan(float) is the float overload (tanf) in the reference's host code (tests/golden/host_derivations.npz)
  const float t = 2.f * std::tan(fovy * 0.5f * (float)M_PI / 180.f);
  const float aspect = (float)width / (float)height;
  const f3 from = mk3(la_from[0], la_from[1], la_from[2]), at = mk3(la_at[0], la_at[1], la_at[2]);
  const f3 up = mk3(la_up[0], la_up[1], la_up[2]);
  pos = from;
  dir = normalize3(at - from);
  hor = (t * aspect) * normalize3(cross3(dir, up));
  ver = cross3(hor, dir) / aspect;
}

static void derive_camera(b200vr_t* h)
{
  derive_camera_values(h->la_from, h->la_at, h->la_up, h->la_fovy, h->width, h->height, h->cam_pos, h->cam_dir, h->cam_hor, h->cam_ver);
  h->has_camera = true;
}

extern "C" int b200vr_host_camera(const float from[3], const float at[3], const float up[3], float fovy, int width, int height, float out[12])
{
  if (!from || !at || !up || !out || width <= 0 || height <= 0) return B200VR_EINVAL;
  f3 v[4];
  derive_camera_values(from, at, up, fovy, width, height, v[0], v[1], v[2], v[3]);
  for (int i = 0; i < 4; i++) out[3 * i] = v[i].x, out[3 * i + 1] = v[i].y, out[3 * i + 2] = v[i].z;
  return B200VR_OK;
}

extern "C" float b200vr_host_integer_normalize(float value, int value_type) { return integer_normalize(value, value_type); }

extern "C" int b200vr_set_camera(b200vr_t* h, const float position[3], const float direction[3], const float horizontal[3],
                                 const float vertical[3])
{
  if (!h || !position || !direction || !horizontal || !vertical) return B200VR_EINVAL;
  h->cam_pos = mk3(position[0], position[1], position[2]);
  h->cam_dir = mk3(direction[0], direction[1], direction[2]);
  h->cam_hor = mk3(horizontal[0], horizontal[1], horizontal[2]);
  h->cam_ver = mk3(vertical[0], vertical[1], vertical[2]);
  h->has_camera = true;
  h->lookat_valid = false;
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_set_camera_lookat(b200vr_t* h, const float from[3], const float at[3], const float up[3], float fovy)
{
  if (!h || !from || !at || !up) return B200VR_EINVAL;
  for (int i = 0; i < 3; i++) h->la_from[i] = from[i], h->la_at[i] = at[i], h->la_up[i] = up[i];
  h->la_fovy = fovy;
  h->lookat_valid = true;
  if (h->width > 0 && h->height > 0) derive_camera(h);
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_get_camera(b200vr_t* h, float position[3], float direction[3], float horizontal[3], float vertical[3])
{
  if (!h) return B200VR_EINVAL;
  if (!h->has_camera) return fail(h, B200VR_ESTATE, "get_camera: no camera");
  const f3 v[4] = { h->cam_pos, h->cam_dir, h->cam_hor, h->cam_ver };
  float* o[4] = { position, direction, horizontal, vertical };
  for (int i = 0; i < 4; i++)
    if (o[i]) o[i][0] = v[i].x, o[i][1] = v[i].y, o[i][2] = v[i].z;
  return B200VR_OK;
}

extern "C" int b200vr_set_params(b200vr_t* h, const b200vr_params* p)
{
  if (!h || !p) return B200VR_EINVAL;
  if (!(p->sampling_rate > 0.f)) return fail(h, B200VR_EINVAL, "set_params: sampling_rate must be positive");
  if (p->spp < 1) return fail(h, B200VR_EINVAL, "set_params: spp must be >= 1");
  if (p->mode != B200VR_MODE_RAYMARCH && p->mode != B200VR_MODE_PATHTRACE) return fail(h, B200VR_EINVAL, "set_params: bad mode");
  if (p->shading != B200VR_SHADING_NONE && p->shading != B200VR_SHADING_REF) return fail(h, B200VR_EINVAL, "set_params: bad shading");
  if (p->sampler != B200VR_SAMPLER_TEX3D && p->sampler != B200VR_SAMPLER_MANUAL) return fail(h, B200VR_EINVAL, "set_params: bad sampler");
  // the path tracer keeps one albedo per bounce (scatter_index advances by 2 per bounce) in a stack of PT_MAX_STACK
  if (p->max_num_scatters > 2 * PT_MAX_STACK) return fail(h, B200VR_EINVAL, "set_params: max_num_scatters above 32 is not supported");
  const float base = p->base > 0.f ? p->base : h->prm.base; // set_sampling_rate(r, b): b <= 0 keeps base (volume.cpp:156-162)
  h->prm = *p;
  h->prm.base = base;
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_get_params(b200vr_t* h, b200vr_params* p)
{
  if (!h || !p) return B200VR_EINVAL;
  *p = h->prm;
  return B200VR_OK;
}

extern "C" int b200vr_resize(b200vr_t* h, int width, int height)
{
  if (!h) return B200VR_EINVAL;
  if (width <= 0 || height <= 0) return fail(h, B200VR_EINVAL, "resize: size must be positive");
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  if (width != h->width || height != h->height) {
    free_frame(h);
    const size_t n = (size_t)width * height;
    for (int i = 0; i < 2; i++) {
      CK(cudaMalloc(&h->rgba[i], n * sizeof(float4)));
      CK(cudaMalloc(&h->grad[i], n * 3 * sizeof(float)));
      CK(cudaMemset(h->rgba[i], 0, n * sizeof(float4)));
      CK(cudaMemset(h->grad[i], 0, n * 3 * sizeof(float)));
    }
    CK(cudaMalloc(&h->accum, n * sizeof(float4)));
    CK(cudaMemset(h->accum, 0, n * sizeof(float4)));
    h->width = width;
    h->height = height;
    cudaFree(h->tile_cost[0]);
    cudaFree(h->tile_cost[1]);
    cudaFree(h->tile_order);
    h->tile_slots = ((width + SUPER_TILE_W - 1) / SUPER_TILE_W) * ((height + SUPER_TILE_H - 1) / SUPER_TILE_H);
    for (int i = 0; i < 2; i++) CK(cudaMalloc(&h->tile_cost[i], (size_t)h->tile_slots * sizeof(unsigned int)));
    CK(cudaMalloc(&h->tile_order, (size_t)h->tile_slots * sizeof(int)));
    h->cost_valid = false;
    h->order_mode_done = -1;
    h->wt_count = 0;
    cudaFree(h->acc);
    cudaFree(h->ray_alpha);
    CK(cudaMalloc(&h->acc, n * 3 * sizeof(unsigned long long)));
    CK(cudaMemset(h->acc, 0, n * 3 * sizeof(unsigned long long)));
    CK(cudaMalloc(&h->ray_alpha, n * sizeof(float)));
    CK(cudaMemset(h->ray_alpha, 0, n * sizeof(float)));
  }
  if (h->lookat_valid) derive_camera(h); // "camera.update() || framebuffer_size_updated", device_impl.cpp:125
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_reset_accumulation(b200vr_t* h)
{
  if (!h) return B200VR_EINVAL;
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_set_clip_box(b200vr_t* h, const float lo[3], const float hi[3])
{
  if (!h || !lo || !hi) return B200VR_EINVAL;
  for (int i = 0; i < 3; i++)
    if (!(hi[i] > lo[i])) return fail(h, B200VR_EINVAL, "set_clip_box: empty box");
  h->clip_lo = mk3(lo[0], lo[1], lo[2]);
  h->clip_hi = mk3(hi[0], hi[1], hi[2]);
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_set_brick(b200vr_t* h, const float own_lo[3], const float own_hi[3], const float vol_lo[3], const float vol_hi[3])
{
  if (!h) return B200VR_EINVAL;
  if (!own_lo || !own_hi || !vol_lo || !vol_hi) { // back to the default: one volume, lattice starts at the clip box
    h->has_lattice = false;
    h->reset = true;
    return B200VR_OK;
  }
  for (int i = 0; i < 3; i++)
    if (!(own_hi[i] > own_lo[i]) || !(vol_hi[i] > vol_lo[i])) return fail(h, B200VR_EINVAL, "set_brick: empty box");
  h->own_lo = mk3(own_lo[0], own_lo[1], own_lo[2]);
  h->own_hi = mk3(own_hi[0], own_hi[1], own_hi[2]);
  h->lat_lo = mk3(vol_lo[0], vol_lo[1], vol_lo[2]);
  h->lat_hi = mk3(vol_hi[0], vol_hi[1], vol_hi[2]);
  h->has_lattice = true;
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_set_tile_partition(b200vr_t* h, int rank, int nranks)
{
  if (!h) return B200VR_EINVAL;
  if (nranks < 1 || rank < 0 || rank >= nranks) return fail(h, B200VR_EINVAL, "set_tile_partition: bad rank");
  if (rank != h->rank || nranks != h->nranks) h->cost_valid = false, h->order_mode_done = -1, h->wt_count = 0; // slots mean other super tiles now
  h->rank = rank;
  h->nranks = nranks;
  h->reset = true;
  return B200VR_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// render
// ---------------------------------------------------------------------------------------------------------------------

typedef void (*render_kernel_t)(const LaunchParams);

// Which kernel renders a frame. Texture sampler: deferred-shading queue kernels (march_q.cuh one task per lane and pass,
// march_qb.cuh J tasks per lane and pass, march_ring.cuh with work sharing between warps) for the reference's gradient + shadow
// shading, the lean kernel (march_ns.cuh) without shading. Software sampler: the first-generation nested-loop kernel
// (raymarch.cuh). Option kernel = legacy | sm forces the older ones for A/B runs.
struct KernelChoice {
  render_kernel_t fn;
  int threads;
  size_t smem;
  bool ring = false; // march_ring.cuh / march_rq.cuh: needs the task rings, one launch per sample + resolve_kernel
  int ring_block = RING_BLOCK_FLOAT4; // float4 per ring slot
  bool cta_sync = false; // march_qc.cuh SYNC: the warps of a CTA take consecutive tiles of the (sheet) launch order and shade in lockstep
  const char* name = "";
};

// the six instantiations of a march kernel template: no skipping | skipping with the occupancy mask in shared / global memory,
// each with the power-of-two transfer-function fast path or the generic one
#define B200VR_PICK(KERNEL, skip, pow2, occs)                                                                                     \
  (!(skip) ? ((pow2) ? KERNEL(false, true, false) : KERNEL(false, false, false))                                                  \
   : (occs) ? ((pow2) ? KERNEL(true, true, true) : KERNEL(true, false, true))                                                     \
            : ((pow2) ? KERNEL(true, true, false) : KERNEL(true, false, false)))
#define B200VR_K_Q(a, b, c) raymarch_q_kernel<a, b, c>
#define B200VR_K_QM(a, b, c) raymarch_q_kernel<a, b, c, SAMPLER_MANUAL>
#define B200VR_K_QB2(a, b, c) raymarch_qb_kernel<a, b, c, 2>
#define B200VR_K_QB4(a, b, c) raymarch_qb_kernel<a, b, c, 4>
#define B200VR_K_QB8(a, b, c) raymarch_qb_kernel<a, b, c, 8>
#define B200VR_K_QC4(a, b, c) raymarch_qc_kernel<a, b, c, 4>
#define B200VR_K_QC6(a, b, c) raymarch_qc_kernel<a, b, c, 6>
#define B200VR_K_QC8(a, b, c) raymarch_qc_kernel<a, b, c, 8>
#define B200VR_K_QG6(a, b, c) raymarch_qc_kernel<a, b, c, 6, 3>
#define B200VR_K_QD6(a, b, c) raymarch_qc_kernel<a, b, c, 6, 1>
#define B200VR_K_RING(a, b, c) raymarch_ring_kernel<a, b, c>
#define B200VR_K_RQ6(a, b, c) raymarch_rq_kernel<a, b, c, 6>
#define B200VR_K_RQ4(a, b, c) raymarch_rq_kernel<a, b, c, 4>
#define B200VR_K_RQ2(a, b, c) raymarch_rq_kernel<a, b, c, 2>
#define B200VR_K_NS(a, b, c) raymarch_ns_kernel<a, b, c>

static KernelChoice pick_kernel(const b200vr_t* h)
{
  const b200vr_params& p = h->prm;
  const int force = h->tune.kernel;
  const bool legacy = force == KERNEL_LEGACY, sm = force == KERNEL_SM;
  // Whole frames: per-warp queue kernels. march_q.cuh (one task per lane and pass) while the macrocell occupancy mask fits into
  // shared memory; beyond that (edge > 704 voxels: far beyond the 126 MB L2) the frame is bound by the shadow rays' L1 misses
  // and the depth-blocked passes of march_qc.cuh win (C3: 6.2 ms against 10.7 ms, profiles/r2g_ab_C3_qc.json).
  // Sort-first partitions: a launch lasts about its mean load plus its heaviest warp tile, and the heaviest tile is the same
  // however many devices share the frame. The work-sharing rings take the shading off that chain; with depth-blocked passes
  // (march_rq.cuh) they beat both the queue kernels and the one-task-per-lane rings (march_ring.cuh) from two devices on, for
  // both volume classes (profiles/r3a_partition_rq.jsonl: an eighth of C3 1.38 against 2.23 ms, of C2 0.68 against 1.03 ms).
  // Sparse (foveated) frames keep the heavy tiles around the focus and drop most of the light ones, so they take the rings too
  // (C2-sparse: 1.25 ms a frame against 1.46 ms with march_ring.cuh and 1.59 ms with the queue kernel) -- not while
  // accumulating: resolve_kernel rewrites every pixel, the reference leaves unsampled pixels untouched then.
  const bool sparse_ring = h->sparse && !p.accumulate;
  const bool large = p.skip_empty && !q_occ_in_smem(h->occ_words);
  const bool ring = force == KERNEL_RING;
  // Blocks of 6 x 32 tasks for the large volumes, 4 x 32 for the small ones (half / an eighth of C2: 1.35 / 0.60 against
  // 1.56 / 0.61 ms; of C3: 3.37 / 1.27 against 3.36 / 1.23 ms; profiles/r3a_partition_c2_kernels.jsonl).
  const bool auto_rq = force == KERNEL_AUTO && (h->nranks != 1 || sparse_ring) && (size_t)h->width * h->height <= (size_t)RQ_PIXEL_MASK;
  const bool rq6 = force == KERNEL_RQ6 || (auto_rq && large), rq4 = force == KERNEL_RQ4 || (auto_rq && !large);
  const bool auto_qc6 = force == KERNEL_AUTO && large && !ring && !rq6 && !rq4;
  KernelChoice k;
  k.threads = RENDER_THREADS;
  k.smem = tf_shared_bytes(h->n_color, h->n_alpha);
  if (p.mode == B200VR_MODE_PATHTRACE) {
    // one 8x4 tile per warp. Measured and not kept (profiles/r3d_pathtrace_regen.md): every lane taking the next pixel when its
    // path ends, as a flat loop of one walk-to-the-next-event per iteration -- bit-identical frames, 1.6 to 2.7 times the time.
    k.fn = p.sampler == B200VR_SAMPLER_MANUAL ? pathtrace_kernel<SAMPLER_MANUAL> : pathtrace_kernel<SAMPLER_TEX3D>;
    k.name = "pathtrace_kernel";
    return k;
  }
  if (p.sampler == B200VR_SAMPLER_MANUAL && !legacy && !sm && p.shading == B200VR_SHADING_REF) {
    // the software sampler (eight global loads + the unit's integer weights, sampling.cuh) inside the current queue kernel: the
    // like-for-like side of the tex3D-versus-manual comparison; sort-first partitions included
    const bool pow2 = (h->n_alpha & (h->n_alpha - 1)) == 0 && h->n_alpha >= 2 && h->n_alpha <= 4096;
    const bool occs = p.skip_empty && h->tune.occ_smem != 0 && q_occ_in_smem(h->occ_words);
    k.fn = B200VR_PICK(B200VR_K_QM, p.skip_empty != 0, pow2, occs);
    k.threads = Q_THREADS, k.smem = q_alpha_bytes(h->n_alpha) + (occs ? q_occ_bytes(h->occ_words) : 0) + q_queue_bytes();
    k.name = "raymarch_q_kernel<SAMPLER_MANUAL>";
    return k;
  }
  if (p.sampler == B200VR_SAMPLER_TEX3D && !legacy) {
    const bool pow2 = (h->n_alpha & (h->n_alpha - 1)) == 0 && h->n_alpha >= 2 && h->n_alpha <= 4096;
    const bool occs = p.skip_empty && h->tune.occ_smem != 0 && q_occ_in_smem(h->occ_words);
    const bool skip = p.skip_empty != 0;
    if (p.shading == B200VR_SHADING_REF && !sm) {
      const size_t tables = q_alpha_bytes(h->n_alpha) + (occs ? q_occ_bytes(h->occ_words) : 0);
      if (ring) {
        k.fn = B200VR_PICK(B200VR_K_RING, skip, pow2, occs);
        k.ring = true;
        k.threads = Q_THREADS, k.smem = tables + q_queue_bytes(), k.name = "raymarch_ring_kernel";
      }
      else if (force == KERNEL_RQ2) {
        k.fn = B200VR_PICK(B200VR_K_RQ2, skip, pow2, occs);
        k.ring = true, k.ring_block = rq_block_float4<2>();
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<2>(), k.name = "raymarch_rq_kernel<J=2>";
      }
      else if (rq4) {
        k.fn = B200VR_PICK(B200VR_K_RQ4, skip, pow2, occs);
        k.ring = true, k.ring_block = rq_block_float4<4>();
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<4>(), k.name = "raymarch_rq_kernel<J=4>";
      }
      else if (rq6) {
        k.fn = B200VR_PICK(B200VR_K_RQ6, skip, pow2, occs);
        k.ring = true, k.ring_block = rq_block_float4<6>();
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<6>(), k.name = "raymarch_rq_kernel<J=6>";
      }
      else if (force == KERNEL_QB2) {
        k.fn = B200VR_PICK(B200VR_K_QB2, skip, pow2, occs);
        k.threads = QB_THREADS, k.smem = tables + qb_queue_bytes<2>(), k.name = "raymarch_qb_kernel";
      }
      else if (force == KERNEL_QB4) {
        k.fn = B200VR_PICK(B200VR_K_QB4, skip, pow2, occs);
        k.threads = QB_THREADS, k.smem = tables + qb_queue_bytes<4>(), k.name = "raymarch_qb_kernel";
      }
      else if (force == KERNEL_QB8) {
        k.fn = B200VR_PICK(B200VR_K_QB8, skip, pow2, occs);
        k.threads = QB_THREADS, k.smem = tables + qb_queue_bytes<8>(), k.name = "raymarch_qb_kernel";
      }
      else if (force == KERNEL_QC4) {
        k.fn = B200VR_PICK(B200VR_K_QC4, skip, pow2, occs);
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<4>(), k.name = "raymarch_qc_kernel<J=4>";
      }
      else if (force == KERNEL_QC6 || auto_qc6) {
        k.fn = B200VR_PICK(B200VR_K_QC6, skip, pow2, occs);
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<6>(), k.name = "raymarch_qc_kernel<J=6>";
      }
      else if (force == KERNEL_QC8) {
        k.fn = B200VR_PICK(B200VR_K_QC8, skip, pow2, occs);
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<8>(), k.name = "raymarch_qc_kernel<J=8>";
      }
      else if (force == KERNEL_QG6) {
        k.fn = B200VR_PICK(B200VR_K_QG6, skip, pow2, occs);
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<6>(), k.name = "raymarch_qc_kernel<J=6,SYNC=3>", k.cta_sync = true;
      }
      else if (force == KERNEL_QD6) {
        k.fn = B200VR_PICK(B200VR_K_QD6, skip, pow2, occs);
        k.threads = QC_THREADS, k.smem = tables + qb_queue_bytes<6>(), k.name = "raymarch_qc_kernel<J=6,SYNC>", k.cta_sync = true;
      }
      else {
        k.fn = B200VR_PICK(B200VR_K_Q, skip, pow2, occs);
        k.threads = Q_THREADS, k.smem = tables + q_queue_bytes(), k.name = "raymarch_q_kernel";
      }
      return k;
    }
    if (p.shading == B200VR_SHADING_NONE && !sm) { // lean classification + compositing kernel (march_ns.cuh)
      k.fn = B200VR_PICK(B200VR_K_NS, skip, pow2, occs);
      k.threads = NS_THREADS;
      k.smem = ns_shared_bytes(h->n_color, h->n_alpha, occs ? h->occ_words : 0);
      k.name = "raymarch_ns_kernel";
      return k;
    }
    k.smem = tf_fast_shared_bytes(h->n_color, h->n_alpha);
    k.name = "raymarch_sm_kernel";
    if (p.shading == B200VR_SHADING_REF)
      k.fn = p.skip_empty ? raymarch_sm_kernel<SHADING_REF, true> : raymarch_sm_kernel<SHADING_REF, false>;
    else
      k.fn = p.skip_empty ? raymarch_sm_kernel<SHADING_NONE, true> : raymarch_sm_kernel<SHADING_NONE, false>;
    return k;
  }
  k.name = "raymarch_kernel";
  const int key = (p.shading == B200VR_SHADING_REF ? 4 : 0) | (p.sampler == B200VR_SAMPLER_MANUAL ? 2 : 0) | (p.skip_empty ? 1 : 0);
  switch (key) {
  case 0: k.fn = raymarch_kernel<SHADING_NONE, SAMPLER_TEX3D, false>; break;
  case 1: k.fn = raymarch_kernel<SHADING_NONE, SAMPLER_TEX3D, true>; break;
  case 2: k.fn = raymarch_kernel<SHADING_NONE, SAMPLER_MANUAL, false>; break;
  case 3: k.fn = raymarch_kernel<SHADING_NONE, SAMPLER_MANUAL, true>; break;
  case 4: k.fn = raymarch_kernel<SHADING_REF, SAMPLER_TEX3D, false>; break;
  case 5: k.fn = raymarch_kernel<SHADING_REF, SAMPLER_TEX3D, true>; break;
  case 6: k.fn = raymarch_kernel<SHADING_REF, SAMPLER_MANUAL, false>; break;
  default: k.fn = raymarch_kernel<SHADING_REF, SAMPLER_MANUAL, true>; break;
  }
  return k;
}

static void fill_launch_params(b200vr_t* h, LaunchParams& P)
{
  std::memset(&P, 0, sizeof(P));
  P.tex = h->tex;
  P.lin = h->lin;
  P.nx = h->nx, P.ny = h->ny, P.nz = h->nz;
  P.rdim = mk3(1.f) / mk3((float)h->nx, (float)h->ny, (float)h->nz);
  P.otw = h->otw;
  P.wto = h->wto;
  P.clip_lo = h->clip_lo;
  P.clip_hi = h->clip_hi;
  P.brick = h->has_lattice ? 1 : 0;
  P.lat_lo = h->lat_lo;
  P.lat_hi = h->lat_hi;
  P.own_lo = h->own_lo;
  P.own_hi = h->own_hi;
  P.tf_color = h->tf_color;
  P.tf_alpha = h->tf_alpha;
  P.n_color = h->n_color, P.n_alpha = h->n_alpha;
  P.tf_lower = h->tf_lower, P.tf_upper = h->tf_upper;
  P.tf_scale = 1.f / (h->tf_upper - h->tf_lower); // volume.scale (volume.cpp:145)
  P.majorants = h->majorants;
  P.maj_pad = h->maj_pad;
  P.occ_bits = h->occ_bits;
  P.occ_words = h->occ_words;
  P.mcp_sy = h->mc.x + 2;
  P.mcp_sz = (h->mc.x + 2) * (h->mc.y + 2);
  P.mcp_bias = (0x4B400000u - 1u) * (1u + (unsigned)P.mcp_sy + (unsigned)P.mcp_sz); // see pad_index_magic()
  P.mcp_last = (unsigned)((h->mc.x + 2) * (h->mc.y + 2) * (h->mc.z + 2)) - 1u;
  {
    // texture-unit coordinate of array1d_nodal for a table of n = 2^m entries:
    // u*2^21 = fma(v, n-1, 0.5) * 2^(21-m); t = ((xi << (m+8)) + 2^20 >> 21) - 128 = (xi + 2^(12-m) - 128*2^(13-m)) >> (13-m)
    auto setup = [](int n, float& mul, float& addf, int& addi, int& shift) {
      int m = 0;
      while ((1 << m) < n) m++;
      if ((1 << m) != n || m > 12 || m < 1) {
        mul = addf = 0.f;
        addi = 0;
        shift = -1;
        return;
      }
      const float sc = (float)(1 << (21 - m));
      mul = (float)(n - 1) * sc;
      addf = 0.5f * sc;
      shift = 13 - m;
      addi = (1 << (12 - m)) - (128 << shift);
    };
    setup(h->n_alpha, P.tfa_mul, P.tfa_addf, P.tfa_addi, P.tfa_shift);
    setup(h->n_color, P.tfc_mul, P.tfc_addf, P.tfc_addi, P.tfc_shift);
  }
  P.mc = h->mc;
  P.cells_per_unit = mk3((float)h->nx, (float)h->ny, (float)h->nz) / (float)MC_SIZE;
  {
    const f3 spac = mk3((float)MC_SIZE) / mk3((float)h->nx, (float)h->ny, (float)h->nz); // sp_singlemc.cu:102-103
    P.spac_rcp = mk3(1.f) / spac;
  }
  P.step = 1.f / h->prm.sampling_rate; // volume.cpp:172-179
  P.base = h->prm.base;
  P.cam_pos = h->cam_pos, P.cam_dir = h->cam_dir, P.cam_hor = h->cam_hor, P.cam_ver = h->cam_ver;
  {
    // get_xfm_world_to_camera (shaders_common.h:276-289); kernels want transpose(inverse(l)) for xfmNormal
    const f3 x = normalize3(h->cam_hor), y = normalize3(h->cam_ver), z = -normalize3(h->cam_dir);
    lin3 l;
    l.vx = mk3(x.x, y.x, z.x);
    l.vy = mk3(x.y, y.y, z.y);
    l.vz = mk3(x.z, y.z, z.z);
    P.wtc = transposed(inverse(l));
  }
  P.light_dir = normalize3(mk3(h->prm.light_pos[0], h->prm.light_pos[1], h->prm.light_pos[2]));
  P.cam_oo = xfm_point(P.wto, P.cam_pos);
  P.sh_d = xfm_vector(P.wto.l, P.light_dir);
  P.sh_rcp = mk3(1.f / P.sh_d.x, 1.f / P.sh_d.y, 1.f / P.sh_d.z); // IEEE division = __frcp_rn
  P.sh_deg_x = std::fabs(P.sh_d.x) < FLT_MIN, P.sh_deg_y = std::fabs(P.sh_d.y) < FLT_MIN, P.sh_deg_z = std::fabs(P.sh_d.z) < FLT_MIN;
  P.sh_dcell = P.sh_d * P.cells_per_unit;
  P.sstep = (P.step * 10.f) * P.step;
  P.log_ert = std::log2(1.f - 0.9999f);
  P.wto_t = transposed(P.wto.l);
  P.light_ambient = h->prm.light_ambient;
  P.max_num_scatters = h->prm.max_num_scatters;
  P.pt_child_instance_xfm = h->prm.pt_child_instance_xfm;
  P.rgba = h->target_rgba ? h->target_rgba : h->rgba[h->current];
  P.grad = h->target_grad ? h->target_grad : h->grad[h->current];
  P.accum = h->accum;
  P.width = h->width, P.height = h->height;
  P.rw = 1.f / (float)h->width, P.rh = 1.f / (float)h->height;
  P.frame_index = h->frame_index;
  P.spp = h->prm.spp;
  P.accumulate = h->prm.accumulate;
  P.premultiplied = h->prm.premultiplied;
  P.super_w = (h->width + SUPER_TILE_W - 1) / SUPER_TILE_W;
  P.super_h = (h->height + SUPER_TILE_H - 1) / SUPER_TILE_H;
  P.rank = h->rank, P.nranks = h->nranks;
  P.n_warp_tiles = owned_super_tiles(h->rank, h->nranks, P.super_w * P.super_h) * WARP_TILES_PER_SUPER;
  P.tile_counter = h->tile_counter;
  P.tile_order = h->tile_order;
  P.tile_cost = h->tile_cost[h->cost_cur];
  P.counters = h->counters;
  P.ring_ctrl = h->ring_ctrl;
  P.ring_seq = h->ring_seq;
  P.ring_data = h->ring_data;
  P.ring_cap = h->ring_cap;
  P.ring_high = 0;
  P.acc = h->acc;
  P.ray_alpha = h->ray_alpha;
  P.sidx = 0;
}

// createSparseSamples (device_impl.cpp:327-341) + generate_sparse_sampling_mask_d (generate_mask.cu:100-124): mask, compaction,
// and the frame clear that goes with a sparse frame (device_impl.cpp:234-239). Needs the pixel count on the host to size the
// launch, like the reference (thrust::remove returns the new end): one 4-byte read-back.
static int build_sparse_list(b200vr_t* h, LaunchParams& P, int& launches)
{
  const int npx = h->width * h->height;
  if (!h->sp_noise) { // no table supplied: built-in white noise, 128 x 128 x 64 like the reference's STBN table
    const int n = 128 * 128 * 64;
    CK(cudaMalloc(&h->sp_noise, (size_t)n * sizeof(float)));
    sparse_default_noise_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(h->sp_noise, n, 1337u); // generate_mask.cu:35
    CK(cudaGetLastError());
    h->sp_tile_xy = 128, h->sp_tile_t = 64;
    launches++;
  }
  SparseParams S;
  S.noise = h->sp_noise;
  S.tile_xy = h->sp_tile_xy, S.tile_t = h->sp_tile_t;
  S.width = h->width, S.height = h->height;
  S.super_w = P.super_w;
  S.n_entries = P.super_w * P.super_h * WARP_TILES_PER_SUPER * 32;
  S.frame_index = h->frame_index; // already incremented (device_impl.cpp:241,252)
  S.mean_x = h->focus_x, S.mean_y = h->focus_y;
  S.sigma_rcp2 = 1.f / (h->focus_scale * h->focus_scale); // generate_mask.cu:92
  S.base_noise = h->base_noise;
  const int n_blocks = (S.n_entries + SPARSE_BLOCK - 1) / SPARSE_BLOCK;
  if (h->sp_alloc_px != npx) {
    cudaFree(h->sp_xy);
    cudaFree(h->sp_blocks);
    h->sp_xy = nullptr, h->sp_blocks = nullptr, h->sp_alloc_px = 0;
    CK(cudaMalloc(&h->sp_xy, (size_t)npx * sizeof(int2)));
    CK(cudaMalloc(&h->sp_blocks, (size_t)n_blocks * sizeof(unsigned int)));
    h->sp_alloc_px = npx;
  }
  if (!h->sp_count) CK(cudaMalloc(&h->sp_count, sizeof(unsigned int)));
  sparse_count_kernel<<<n_blocks, SPARSE_BLOCK, 0, h->stream>>>(S, h->sp_blocks);
  sparse_scan_kernel<<<1, 1024, 0, h->stream>>>(h->sp_blocks, n_blocks, h->sp_count);
  sparse_scatter_kernel<<<n_blocks, SPARSE_BLOCK, 0, h->stream>>>(S, h->sp_blocks, h->sp_xy);
  CK(cudaGetLastError());
  launches += 3;
  unsigned int count = 0;
  CK(cudaMemcpyAsync(&count, h->sp_count, sizeof(count), cudaMemcpyDeviceToHost, h->stream));
  if (!h->prm.accumulate) { // "in non-sparse sampling mode, the entire frame will be re-written anyway": clear both buffers
    for (int i = 0; i < 2; i++) {
      CK(cudaMemsetAsync(h->rgba[i], 0, (size_t)npx * sizeof(float4), h->stream));
      CK(cudaMemsetAsync(h->grad[i], 0, (size_t)npx * 3 * sizeof(float), h->stream));
    }
  }
  CK(cudaStreamSynchronize(h->stream));
  h->sp_last_count = (int)count;
  P.sparse_xy = h->sp_xy;
  P.sparse_count = (int)count;
  P.n_warp_tiles = ((int)count + 31) / 32;
  return B200VR_OK;
}

extern "C" int b200vr_set_sparse_sampling(b200vr_t* h, int enable, const float focus_center[2], float focus_scale, float base_noise)
{
  if (!h) return B200VR_EINVAL;
  if (enable) {
    if (!focus_center) return fail(h, B200VR_EINVAL, "set_sparse_sampling: focus_center is NULL");
    if (!(focus_scale > 0.f)) return fail(h, B200VR_EINVAL, "set_sparse_sampling: focus_scale must be positive");
    h->focus_x = focus_center[0], h->focus_y = focus_center[1];
    h->focus_scale = focus_scale;
    h->base_noise = base_noise;
  }
  h->sparse = enable != 0;
  h->reset = true; // framebuffer_reset on every focus / sparse change (device_impl.cpp:154-183)
  return B200VR_OK;
}

extern "C" int b200vr_set_sparse_noise(b200vr_t* h, const float* table, int tile_xy, int tile_t)
{
  if (!h) return B200VR_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  cudaFree(h->sp_noise);
  h->sp_noise = nullptr, h->sp_tile_xy = h->sp_tile_t = 0;
  if (!table) return B200VR_OK; // back to the built-in table
  if (tile_xy <= 0 || tile_t <= 0 || (size_t)tile_xy * tile_xy * tile_t > (size_t)1 << 28)
    return fail(h, B200VR_EINVAL, "set_sparse_noise: bad table size");
  const size_t n = (size_t)tile_xy * tile_xy * tile_t;
  CK(cudaMalloc(&h->sp_noise, n * sizeof(float)));
  CK(cudaMemcpy(h->sp_noise, table, n * sizeof(float), cudaMemcpyHostToDevice));
  h->sp_tile_xy = tile_xy, h->sp_tile_t = tile_t;
  return B200VR_OK;
}

extern "C" int64_t b200vr_get_sparse_samples(b200vr_t* h, int32_t* xy_out, int64_t capacity)
{
  if (!h) return B200VR_EINVAL;
  if (h->sp_last_count < 0) return fail(h, B200VR_ESTATE, "get_sparse_samples: no sparse frame rendered yet");
  if (xy_out) {
    if (capacity < h->sp_last_count) return fail(h, B200VR_EINVAL, "get_sparse_samples: capacity too small");
    if (cudaSetDevice(h->device) != cudaSuccess ||
        cudaMemcpy(xy_out, h->sp_xy, (size_t)h->sp_last_count * sizeof(int2), cudaMemcpyDeviceToHost) != cudaSuccess)
      return fail(h, B200VR_ECUDA, "get_sparse_samples: copy failed");
  }
  return h->sp_last_count;
}

// Sheet order of the owned warp tiles (build.cuh sheet_key_kernel): key per tile on the device, one radix sort, the sorted tile
// origins become P.wt_order. The table is kept while camera basis, light, image size and partition stay (nearly) the same:
// the order is a locality heuristic, any permutation renders the same frame.
static int build_sheet_order(b200vr_t* h, LaunchParams& P, int& launches)
{
  const int n = P.n_warp_tiles;
  if (n <= 0) return B200VR_OK;
  const f3 L = P.light_dir;
  auto close3 = [](f3 a, f3 b, float tol) { return std::fabs(a.x - b.x) <= tol && std::fabs(a.y - b.y) <= tol && std::fabs(a.z - b.z) <= tol; };
  const float tol = 1e-3f * std::sqrt(dot3(h->cam_hor, h->cam_hor));
  if (h->wt_count == n && h->wt_rank == h->rank && h->wt_nranks == h->nranks && h->wt_w == h->width && h->wt_h == h->height &&
      h->wt_px == h->tune.sheet_px && close3(h->wt_dir, h->cam_dir, 1e-3f) && close3(h->wt_hor, h->cam_hor, tol) &&
      close3(h->wt_ver, h->cam_ver, tol) && close3(h->wt_light, L, 1e-4f)) {
    P.wt_order = h->wt_vals[1];
    return B200VR_OK;
  }
  if (h->wt_alloc < n) {
    CK(cudaStreamSynchronize(h->stream));
    for (int i = 0; i < 2; i++) {
      cudaFree(h->wt_keys[i]);
      cudaFree(h->wt_vals[i]);
      h->wt_keys[i] = h->wt_vals[i] = nullptr;
    }
    cudaFree(h->wt_tmp);
    h->wt_tmp = nullptr, h->wt_alloc = 0, h->wt_count = 0;
    for (int i = 0; i < 2; i++) {
      CK(cudaMalloc(&h->wt_keys[i], (size_t)n * sizeof(unsigned int)));
      CK(cudaMalloc(&h->wt_vals[i], (size_t)n * sizeof(unsigned int)));
    }
    size_t bytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, h->wt_keys[0], h->wt_keys[1], h->wt_vals[0], h->wt_vals[1], n, 0, 32, h->stream));
    CK(cudaMalloc(&h->wt_tmp, std::max<size_t>(bytes, 16)));
    h->wt_tmp_bytes = bytes;
    h->wt_alloc = n;
  }
  SheetParams S;
  S.cam_dir = h->cam_dir, S.cam_hor = h->cam_hor, S.cam_ver = h->cam_ver;
  S.light = L;
  // basis of the plane perpendicular to the light: e1 = the view direction's component in it (phi = 0 at the image centre)
  f3 e1 = h->cam_dir - dot3(h->cam_dir, L) * L;
  if (dot3(e1, e1) < 1e-6f * dot3(h->cam_dir, h->cam_dir)) e1 = h->cam_hor - dot3(h->cam_hor, L) * L; // looking along the light
  e1 = normalize3(e1);
  S.e1 = e1;
  S.e2 = cross3(L, e1);
  S.rw = P.rw, S.rh = P.rh;
  // strip width: sheet_px pixels where the planes are spread widest on the screen (smallest |grad phi| over a coarse pixel grid)
  auto phi_at = [&](double x, double y) {
    const double sx = x / h->width - 0.5, sy = y / h->height - 0.5;
    const double d[3] = { h->cam_dir.x + sx * h->cam_hor.x + sy * h->cam_ver.x, h->cam_dir.y + sx * h->cam_hor.y + sy * h->cam_ver.y,
                          h->cam_dir.z + sx * h->cam_hor.z + sy * h->cam_ver.z };
    return std::atan2(d[0] * S.e2.x + d[1] * S.e2.y + d[2] * S.e2.z, d[0] * S.e1.x + d[1] * S.e1.y + d[2] * S.e1.z);
  };
  auto wrap = [](double a) {
    while (a > M_PI) a -= 2 * M_PI;
    while (a < -M_PI) a += 2 * M_PI;
    return a;
  };
  double gmin = 1e30;
  for (int gy = 0; gy <= 8; gy++)
    for (int gx = 0; gx <= 16; gx++) {
      const double x = (gx / 16.0) * (h->width - 1) + 0.5, y = (gy / 8.0) * (h->height - 1) + 0.5;
      const double p0 = phi_at(x, y), dx = wrap(phi_at(x + 1.0, y) - p0), dy = wrap(phi_at(x, y + 1.0) - p0);
      const double g = std::sqrt(dx * dx + dy * dy);
      if (g > 0.0) gmin = std::min(gmin, g);
    }
  if (!(gmin < 1e29)) gmin = 1.0 / std::max(h->width, h->height);
  const double dphi = std::max(gmin * (double)h->tune.sheet_px, 2.0 * M_PI / (double)(1 << 18));
  S.inv_dphi = (float)(1.0 / dphi);
  S.super_w = P.super_w, S.rank = h->rank, S.nranks = h->nranks, S.n_warp_tiles = n;
  sheet_key_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(S, h->wt_keys[0], h->wt_vals[0]);
  CK(cudaGetLastError());
  size_t bytes = h->wt_tmp_bytes;
  CK(cub::DeviceRadixSort::SortPairs(h->wt_tmp, bytes, h->wt_keys[0], h->wt_keys[1], h->wt_vals[0], h->wt_vals[1], n, 0, 32, h->stream));
  launches += 4; // key kernel + the passes of the radix sort (histogram, scan, onesweep: version dependent; counted as 3)
  h->wt_count = n;
  h->wt_dir = h->cam_dir, h->wt_hor = h->cam_hor, h->wt_ver = h->cam_ver, h->wt_light = L;
  h->wt_rank = h->rank, h->wt_nranks = h->nranks, h->wt_w = h->width, h->wt_h = h->height, h->wt_px = h->tune.sheet_px;
  P.wt_order = h->wt_vals[1];
  return B200VR_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// allocation-free variant of the sparse-sample compaction (no counterpart in the reference):
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it)
// ---------------------------------------------------------------------------------------------------------------------
// The compaction kernels of sparse.cuh need one word of scratch per 1024 enumeration entries (per-block counts, turned into
// exclusive offsets in place). b200vr_render keeps that scratch in cudaMalloc'ed memory of the handle (build_sparse_list); here
// it lives in a file-scope device array of fixed size, which every GPU that loads the library holds until the process exits.
constexpr int NOALLOC_MAX_BLOCKS = B200VR_NOALLOC_MAX_ENTRIES / SPARSE_BLOCK; // 16384 words = 64 KiB
static __device__ __align__(256) unsigned int g_noalloc_block_counts[NOALLOC_MAX_BLOCKS];

namespace {
// In host code the array's name is only a stand-in for the symbol, not a device address: the address is asked of the runtime
// once per device, with that device current, and handed to the kernels as an ordinary pointer.
constexpr int NOALLOC_MAX_DEVICES = 64;
unsigned int* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};
std::mutex g_noalloc_mutex;
}

static int noalloc_scratch(b200vr_t* h, unsigned int** out)
{
  if (h->device < 0 || h->device >= NOALLOC_MAX_DEVICES) return fail(h, B200VR_EINVAL, "sparse_samples_noalloc: device index above 63");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[h->device]) {
    unsigned int* ptr = nullptr;
    CK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_block_counts)); // the caller made h->device current
    g_noalloc_ptr[h->device] = ptr;
  }
  *out = g_noalloc_ptr[h->device];
  return B200VR_OK;
}

extern "C" int b200vr_sparse_samples_noalloc(b200vr_t* h, const float* noise_dev, int tile_xy, int tile_t, int width, int height,
                                             int frame_index, const float focus_center[2], float focus_scale, float base_noise,
                                             int32_t* xy_out_dev, int64_t capacity_pairs, uint32_t* count_out_dev)
{
  if (!h) return B200VR_EINVAL;
  if (!noise_dev || !focus_center || !xy_out_dev || !count_out_dev) return fail(h, B200VR_EINVAL, "sparse_samples_noalloc: NULL argument");
  if (tile_xy <= 0 || tile_t <= 0 || width <= 0 || height <= 0) return fail(h, B200VR_EINVAL, "sparse_samples_noalloc: sizes must be positive");
  if (!(focus_scale > 0.f)) return fail(h, B200VR_EINVAL, "sparse_samples_noalloc: focus_scale must be positive");
  const long long super_w = (width + SUPER_TILE_W - 1) / SUPER_TILE_W, super_h = (height + SUPER_TILE_H - 1) / SUPER_TILE_H;
  const long long n_entries = super_w * super_h * WARP_TILES_PER_SUPER * 32;
  if (n_entries > (long long)B200VR_NOALLOC_MAX_ENTRIES)
    return fail(h, B200VR_EINVAL, "sparse_samples_noalloc: frame too large for the static scratch (the image padded to 64x32-pixel tiles "
                                  "may hold at most B200VR_NOALLOC_MAX_ENTRIES = 16777216 pixels); use b200vr_render with sparse sampling instead");
  if (capacity_pairs < (int64_t)width * height)
    return fail(h, B200VR_EINVAL, "sparse_samples_noalloc: xy_out_dev must have room for width * height pairs (the count is not known on the host)");
  CK(cudaSetDevice(h->device));
  unsigned int* blocks = nullptr;
  {
    const int rc = noalloc_scratch(h, &blocks);
    if (rc != B200VR_OK) return rc;
  }
  SparseParams S;
  S.noise = noise_dev;
  S.tile_xy = tile_xy, S.tile_t = tile_t;
  S.width = width, S.height = height;
  S.super_w = (int)super_w;
  S.n_entries = (int)n_entries;
  S.frame_index = frame_index;
  S.mean_x = focus_center[0], S.mean_y = focus_center[1];
  S.sigma_rcp2 = 1.f / (focus_scale * focus_scale); // generate_mask.cu:92
  S.base_noise = base_noise;
  const int n_blocks = (S.n_entries + SPARSE_BLOCK - 1) / SPARSE_BLOCK;
  // Nothing has to start at zero: sparse_count_kernel overwrites the first n_blocks words, sparse_scan_kernel rewrites them in
  // place and writes *count_out_dev; words beyond n_blocks keep an earlier call's contents and are never read.
  sparse_count_kernel<<<n_blocks, SPARSE_BLOCK, 0, h->stream>>>(S, blocks);
  sparse_scan_kernel<<<1, 1024, 0, h->stream>>>(blocks, n_blocks, count_out_dev);
  sparse_scatter_kernel<<<n_blocks, SPARSE_BLOCK, 0, h->stream>>>(S, blocks, (int2*)xy_out_dev);
  CK(cudaGetLastError());
  return B200VR_OK; // stream-ordered on the handle's stream: no wait, no read-back, no allocation
}

// resident CTAs per SM of a kernel at (threads, shared bytes), with the shared-memory attributes set once per configuration
static int launch_config(b200vr_t* h, const KernelChoice& kc, int& per_sm)
{
  const auto key = std::make_tuple((const void*)kc.fn, kc.threads, kc.smem, h->tune.carveout);
  auto it = h->launch_cache.find(key);
  if (it == h->launch_cache.end()) {
    if (kc.smem > 48 * 1024) CK(cudaFuncSetAttribute((const void*)kc.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kc.smem));
    if (h->tune.carveout >= 0)
      CK(cudaFuncSetAttribute((const void*)kc.fn, cudaFuncAttributePreferredSharedMemoryCarveout, h->tune.carveout));
    LaunchConfig lc;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&lc.per_sm, (const void*)kc.fn, kc.threads, kc.smem));
    it = h->launch_cache.emplace(key, lc).first;
  }
  per_sm = it->second.per_sm;
  return B200VR_OK;
}

// Enqueues one frame on the handle's stream (Impl::render, device_impl.cpp:199-269, minus the wait): accumulation bookkeeping,
// launch order, the render kernel(s), events around them. Nothing here blocks the host except a sparse frame's 4-byte
// pixel-count read (the reference waits for thrust::remove the same way) and a reallocation of the task rings.
static int render_enqueue(b200vr_t* h)
{
  if (!h->has_volume) return fail(h, B200VR_ESTATE, "render: no volume");
  if (!h->has_tf) return fail(h, B200VR_ESTATE, "render: no transfer function");
  if (!h->has_camera) return fail(h, B200VR_ESTATE, "render: no camera");
  if (h->width <= 0 || h->height <= 0) return fail(h, B200VR_ESTATE, "render: framebuffer size not set");
  CK(cudaSetDevice(h->device));
  // accumulation bookkeeping (device_impl.cpp:225-241)
  const bool accum_reset = h->prm.accumulate && h->reset;
  if (accum_reset) h->frame_index = 0;
  h->reset = false;
  h->frame_index++;

  LaunchParams P;
  fill_launch_params(h, P);
  if (h->has_lattice && !(h->prm.mode == B200VR_MODE_RAYMARCH && h->prm.shading == B200VR_SHADING_NONE &&
                           h->prm.sampler == B200VR_SAMPLER_TEX3D && (h->tune.kernel == KERNEL_AUTO || h->tune.kernel == KERNEL_SM)))
    return fail(h, B200VR_ESTATE, "render: a brick (sort-last) needs ray marching with SHADING_NONE and the tex3D sampler: "
                                  "shadow rays and random walks leave the brick");
  const KernelChoice kc = pick_kernel(h);
  int extra_launches = 0;
  if (h->sparse) {
    if ((kc.ring && h->prm.accumulate) || h->nranks != 1 || h->has_lattice || h->target_rgba)
      return fail(h, B200VR_ESTATE, "render: sparse sampling is a single-device mode (no tile partition, brick or frame target)");
    if (accum_reset) {
      // framebuffer.reset() on an accumulation reset zeroes both buffers (device_impl.cpp:225-233, optix7_common.h:250-253,
      // 409-413): pixels the new view does not sample must not keep the previous view's content
      const size_t npx = (size_t)h->width * h->height;
      for (int i = 0; i < 2; i++) {
        CK(cudaMemsetAsync(h->rgba[i], 0, npx * sizeof(float4), h->stream));
        CK(cudaMemsetAsync(h->grad[i], 0, npx * 3 * sizeof(float), h->stream));
      }
      CK(cudaMemsetAsync(h->accum, 0, npx * sizeof(float4), h->stream));
    }
    const int rc = build_sparse_list(h, P, extra_launches);
    if (rc != B200VR_OK) return rc;
  }
  if (kc.ring_block != RING_BLOCK_FLOAT4 && (size_t)h->width * h->height > (size_t)RQ_PIXEL_MASK)
    return fail(h, B200VR_EINVAL, "render: kernel rq6 takes frames of up to 2^31 - 1 pixels");
  const render_kernel_t kernel = kc.fn;
  const size_t smem = kc.smem;
  int per_sm = 0;
  {
    const int rc = launch_config(h, kc, per_sm);
    if (rc != B200VR_OK) return rc;
  }
  if (per_sm < 1) return fail(h, B200VR_ECUDA, "render: kernel does not fit on an SM (transfer-function table too large for shared memory?)");
  if (h->tune.max_ctas_per_sm > 0) per_sm = std::min(per_sm, h->tune.max_ctas_per_sm); // fewer warps in flight = shorter reuse distance in L2
  const int warps_per_cta = kc.threads / 32;
  const int want = (P.n_warp_tiles + warps_per_cta - 1) / warps_per_cta;
  int grid = std::max(1, std::min(want, h->sm_count * per_sm)); // persistent: SMs x resident CTAs
  if (kc.ring && P.sparse_xy) grid = h->sm_count * per_sm;      // few tiles, all heavy: warps without a tile shade published blocks
  CK(cudaMemsetAsync(h->tile_counter, 0, sizeof(unsigned int), h->stream));
  // the executed-sample counters add up over the frames enqueued since the last wait
  if (!h->pending) CK(cudaMemsetAsync(h->counters, 0, sizeof(Counters), h->stream));
  {
    // launch order of the owned tiles. Super-tile orders (tile_order_kernel): 0 row-major (default), 1 centre-out, 2 by the
    // previous frame's cost, 3 Z-order, 4/5 columns, 6 middle rows first. Measured on C2 with the queue kernel: 2.29 / 2.52 /
    // 2.42 ms for 0 / 1 / 2 -- launching neighbouring tiles together keeps their texels in L1/L2, and a heavy tile is itself
    // the critical path whenever it starts (tools/partition_probe.py), so "expensive first" buys nothing. 7: warp tiles in
    // sheet order (build_sheet_order), for shaded frames of volumes much larger than L2.
    const int n_owned = P.n_warp_tiles / WARP_TILES_PER_SUPER;
    int order_mode = h->tune.tile_order == TILE_ORDER_AUTO ? (kc.cta_sync ? TILE_ORDER_SHEET : TILE_ORDER_ROWS) : h->tune.tile_order;
    const bool shaded_march = h->prm.mode == B200VR_MODE_RAYMARCH && h->prm.shading == B200VR_SHADING_REF;
    if (order_mode == TILE_ORDER_SHEET && (!shaded_march || P.sparse_xy)) order_mode = TILE_ORDER_ROWS;
    if (!h->cost_valid) CK(cudaMemsetAsync(h->tile_cost[h->cost_cur ^ 1], 0, (size_t)h->tile_slots * sizeof(unsigned int), h->stream));
    if (order_mode == TILE_ORDER_SHEET && !P.sparse_xy) {
      const int rc = build_sheet_order(h, P, extra_launches);
      if (rc != B200VR_OK) return rc;
    }
    const int super_mode = order_mode == TILE_ORDER_SHEET ? TILE_ORDER_ROWS : order_mode;
    // the table is kept from frame to frame; sparse frames walk the pixel list and neither use nor touch it
    if (n_owned > 0 && !P.sparse_xy &&
        (super_mode != h->order_mode_done || n_owned != h->order_n_done || super_mode == TILE_ORDER_COST || !h->cost_valid)) {
      tile_order_kernel<<<(n_owned + 127) / 128, 128, 0, h->stream>>>(h->tile_cost[h->cost_cur ^ 1], n_owned, h->rank, h->nranks,
                                                                       P.super_w, P.super_h, h->tile_order, super_mode);
      CK(cudaGetLastError());
      h->order_mode_done = super_mode, h->order_n_done = n_owned;
      extra_launches++;
    }
    CK(cudaMemsetAsync(h->tile_cost[h->cost_cur], 0, (size_t)h->tile_slots * sizeof(unsigned int), h->stream));
    h->cost_valid = true;
    h->cost_cur ^= 1; // P.tile_cost was taken before the flip: this launch writes it, the next one reads it
  }
  int launches = 0;
  if (kc.ring && P.n_warp_tiles > 0) {
    // one ring per CTA: high-water mark + one block in flight per warp of the CTA + slack, as a power of two
    P.ring_high = h->tune.ring_high;
    P.rq_self = h->tune.rq_self;
    int cap = kc.ring_block != RING_BLOCK_FLOAT4 ? h->tune.rq_cap : 32;
    while (cap < P.ring_high + warps_per_cta + 8) cap <<= 1;
    if (cap != h->ring_cap || grid != h->ring_count || kc.ring_block != h->ring_block) {
      CK(cudaStreamSynchronize(h->stream));
      cudaFree(h->ring_ctrl);
      cudaFree(h->ring_seq);
      cudaFree(h->ring_data);
      h->ring_ctrl = nullptr, h->ring_seq = nullptr, h->ring_data = nullptr, h->ring_cap = 0, h->ring_count = 0, h->ring_block = 0;
      CK(cudaMalloc(&h->ring_ctrl, ring_ctrl_words(grid) * sizeof(unsigned int)));
      CK(cudaMalloc(&h->ring_seq, (size_t)cap * grid * sizeof(unsigned int)));
      CK(cudaMalloc(&h->ring_data, (size_t)cap * grid * kc.ring_block * sizeof(float4)));
      h->ring_cap = cap, h->ring_count = grid, h->ring_block = kc.ring_block;
      ring_init_kernel<<<64, 256, 0, h->stream>>>(h->ring_seq, cap, grid, h->ring_ctrl, (int)ring_ctrl_words(grid));
      CK(cudaGetLastError());
      extra_launches++;
    }
    P.ring_ctrl = h->ring_ctrl, P.ring_seq = h->ring_seq, P.ring_data = h->ring_data, P.ring_cap = h->ring_cap;
  }
  CK(cudaEventRecord(h->ev0, h->stream));
  if (P.n_warp_tiles > 0 && !kc.ring) {
    void* args[] = { (void*)&P };
    CK(cudaLaunchKernel((const void*)kernel, dim3((unsigned)grid), dim3((unsigned)kc.threads), args, smem, h->stream));
    launches = 1;
  }
  else if (P.n_warp_tiles > 0) {
    // sparse frame: the list covers the image, resolve_kernel rewrites every pixel (unsampled ones: sums 0, opacity 0 -> 0)
    const int n_owned = P.sparse_xy ? P.super_w * P.super_h : P.n_warp_tiles / WARP_TILES_PER_SUPER;
    for (int sidx = 0; sidx < P.spp; sidx++) { // one march + one resolve per sample of the pixel
      if (P.sparse_xy) CK(cudaMemsetAsync(h->ray_alpha, 0, (size_t)h->width * h->height * sizeof(float), h->stream));
      P.sidx = sidx;
      if (sidx > 0) CK(cudaMemsetAsync(h->tile_counter, 0, sizeof(unsigned int), h->stream));
      ring_start_kernel<<<(grid + 127) / 128, 128, 0, h->stream>>>(h->ring_ctrl, RING_CTRL_STRIDE, grid, warps_per_cta);
      CK(cudaGetLastError());
      // The warps of the ring kernel wait on one another (draining warps for marching ones): every CTA has to be resident.
      // A cooperative launch fails with an error instead of hanging when the grid cannot be (another handle or stream on
      // the same GPU, an MPS limit).
      void* args[] = { (void*)&P };
      CK(cudaLaunchCooperativeKernel((const void*)kernel, dim3((unsigned)grid), dim3((unsigned)kc.threads), args, smem, h->stream));
      const size_t npx = (size_t)n_owned * SUPER_TILE_W * SUPER_TILE_H;
      resolve_kernel<<<(unsigned)std::min<size_t>((npx + 255) / 256, (size_t)h->sm_count * 16), 256, 0, h->stream>>>(P, n_owned);
      CK(cudaGetLastError());
      launches += 3;
    }
  }
  CK(cudaEventRecord(h->ev1, h->stream));
  h->last_launches = launches + extra_launches;
  h->last_ring = kc.ring;
  h->last_kernel = kc.name;
  h->last_grid = grid, h->last_warps_per_cta = warps_per_cta;
  h->pending = true;
  return B200VR_OK;
}

// Waits for the frame enqueued last and fetches its counters (the cudaDeviceSynchronize of DeviceOptix7::render, device.cpp:34-43)
static int render_wait(b200vr_t* h, b200vr_stats* stats)
{
  CK(cudaSetDevice(h->device));
  Counters c;
  std::memset(&c, 0, sizeof(c));
  unsigned int ring_err = 0;
  if (h->pending) {
    CK(cudaMemcpyAsync(&c, h->counters, sizeof(c), cudaMemcpyDeviceToHost, h->stream));
    if (h->last_ring && h->ring_ctrl)
      CK(cudaMemcpyAsync(&ring_err, h->ring_ctrl + (size_t)h->ring_count * RING_CTRL_PER_RING + RING_CTRL_STRIDE, sizeof(ring_err),
                         cudaMemcpyDeviceToHost, h->stream));
  }
  CK(cudaStreamSynchronize(h->stream));
  const bool had_frame = h->pending;
  h->pending = false;
  if (ring_err) {
    h->ring_cap = 0; // rebuilt on the next render
    return fail(h, B200VR_ECUDA, "render: the deferred-shading ring timed out (internal error)");
  }
#if B200VR_PHASE_STATS
  if (h->tune.debug_tiles && had_frame) { // diagnostic build -DB200VR_PHASE_STATS=1: where the warps of march_qc.cuh spend their clocks
    const double tot = (double)(c.tile_clocks_sum + c.drain_clocks_sum + c.blocks_published + c.blocks_stolen + c.tile_clocks_max);
    const int warps = h->last_grid * h->last_warps_per_cta;
    std::fprintf(stderr, "[b200] phases (%s, %d warps, %.0f kclk per warp): primary march %.1f %%, pass: clip %.1f %%, lattice loop %.1f %%, "
                         "gradient + shade + hand-back %.1f %%, other %.1f %%\n", h->last_kernel.c_str(), warps, tot * 1e-3 / warps,
                 100.0 * c.tile_clocks_sum / tot, 100.0 * c.drain_clocks_sum / tot, 100.0 * c.blocks_published / tot, 100.0 * c.blocks_stolen / tot,
                 100.0 * c.tile_clocks_max / tot);
    const double iters = (double)(c.warp_clocks_max >> 40), lanes = (double)(c.warp_clocks_max & ((1ull << 40) - 1));
    const double passes = (double)(c.delta_samples >> 40), steps = (double)(c.delta_samples & ((1ull << 40) - 1));
    std::fprintf(stderr, "[b200] primary: %.0f k iterations per warp, %.0f clocks each, %.1f live lanes, %.1f samples executed per iteration; passes: %.0f per "
                         "warp, %.1f lattice steps each, %.0f clocks per step\n", iters * 1e-3 / warps, c.tile_clocks_sum / std::max(iters, 1.0), lanes / std::max(iters, 1.0),
                 (double)c.primary_samples / std::max(iters, 1.0), passes / warps, steps / std::max(passes, 1.0), c.blocks_published / std::max(steps, 1.0));
    c.delta_samples = 0;
  }
#endif
  if (B200VR_TILE_STATS && h->tune.debug_tiles && had_frame) { // build with -DB200VR_TILE_STATS=1
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev0, h->ev1);
    const int warps = h->last_grid * h->last_warps_per_cta;
    std::fprintf(stderr, "[b200] tiles: kernel %.3f ms, longest warp tile %.0f kclk, all tiles %.0f Mclk over %d warps = %.0f kclk each\n", ms,
                 c.tile_clocks_max * 1e-3, c.tile_clocks_sum * 1e-6, warps, c.tile_clocks_sum * 1e-3 / warps);
    std::fprintf(stderr, "[b200] ring: longest warp %.0f kclk, after the last tile %.0f kclk per warp, blocks published %llu, taken by draining warps %llu\n",
                 c.warp_clocks_max * 1e-3, c.drain_clocks_sum * 1e-3 / warps, c.blocks_published, c.blocks_stolen);
    if (h->last_kernel.rfind("raymarch_rq_kernel", 0) == 0) {
      std::fprintf(stderr, "[b200] rq: longest tile %.0f kclk, started %.0f kclk into the launch: %llu march iterations, %llu blocks published (%.0f kclk), "
                           "%llu shaded by itself (%.0f kclk)\n", c.dbg[0] * 1e-3, c.dbg[7] * 1e-3, c.dbg[1], c.dbg[2], c.dbg[5] * 1e-3, c.dbg[3], c.dbg[6] * 1e-3);
      std::fprintf(stderr, "[b200] rq: its march iterations: macrocell section %.0f kclk, until the sample has arrived %.0f kclk, whole iterations %.0f kclk; "
                           "%llu lane-entries into new macrocells\n", (c.dbg[4] & 0xffffffffull) * 1e-3, (c.dbg[4] >> 32) * 1e-3, (c.dbg[8] & 0xffffffffull) * 1e-3, c.dbg[8] >> 32);
      std::fprintf(stderr, "[b200] rq: of the time after the last tile, waiting for blocks %.0f kclk per warp\n", c.delta_samples * 1e-3 / warps);
      c.delta_samples = 0;
    }
  }
  if (stats) {
    float ms = 0.f;
    if (had_frame) CK(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    stats->ms = ms;
    stats->primary_samples = c.primary_samples;
    stats->shadow_samples = c.shadow_samples;
    stats->gradient_fetches = c.gradient_fetches;
    stats->delta_samples = c.delta_samples;
    stats->rays = c.rays;
    stats->frame_index = h->frame_index;
    stats->kernel_launches = h->last_launches;
  }
  return B200VR_OK;
}

extern "C" int b200vr_render(b200vr_t* h, b200vr_stats* stats)
{
  if (!h) return B200VR_EINVAL;
  const int rc = render_enqueue(h);
  if (rc != B200VR_OK) return rc;
  return render_wait(h, stats);
}

extern "C" int b200vr_render_async(b200vr_t* h)
{
  if (!h) return B200VR_EINVAL;
  return render_enqueue(h);
}

extern "C" int b200vr_wait(b200vr_t* h, b200vr_stats* stats)
{
  if (!h) return B200VR_EINVAL;
  return render_wait(h, stats);
}

// ---------------------------------------------------------------------------------------------------------------------
// device group: several GPUs of one process behind one renderer (sort-first). The host layer that the plugin
// (device/device_b200.cpp) drives when OVR_B200_GPUS / --b200-gpus names more than one device: every member holds the whole
// scene (the caller applies each setter to every member), owns the image tiles rank i of n (b200vr_set_tile_partition) and
// renders them STRAIGHT INTO member 0's current frame buffer through a peer mapping -- the kernels' own pixel stores are the
// transfer over NVLink; map / swap / download then address member 0 alone. One host thread enqueues on all members' streams
// and waits for all of them, as DeviceOptix7::render waits for its one device (device.cpp:34-43).
// ---------------------------------------------------------------------------------------------------------------------
struct b200vr_group {
  std::vector<b200vr_t*> member;
  std::string error;
};

extern "C" int b200vr_group_create(b200vr_group_t** out, const int* cuda_devices, int n_devices)
{
  if (!out) return B200VR_EINVAL;
  *out = nullptr;
  if (!cuda_devices || n_devices < 1 || n_devices > 64) {
    g_create_error = "group_create: 1 to 64 devices";
    return B200VR_EINVAL;
  }
  b200vr_group_t* g = new b200vr_group();
  for (int i = 0; i < n_devices; i++) {
    b200vr_t* h = nullptr;
    const int rc = b200vr_create(&h, cuda_devices[i]);
    if (rc != B200VR_OK) {
      for (b200vr_t* m : g->member) b200vr_destroy(m);
      delete g;
      return rc; // g_create_error was set by b200vr_create
    }
    g->member.push_back(h);
    b200vr_set_tile_partition(h, i, n_devices);
  }
  // members 1.. store into member 0's frame: peer access from their device to its device (the same device twice is allowed:
  // several members on one GPU, which is how a one-GPU box exercises this path)
  const int dev0 = g->member[0]->device;
  for (int i = 1; i < n_devices; i++) {
    const int dev = g->member[i]->device;
    if (dev == dev0) continue;
    int can = 0;
    cudaError_t e = cudaDeviceCanAccessPeer(&can, dev, dev0);
    if (e == cudaSuccess && can) {
      e = cudaSetDevice(dev);
      if (e == cudaSuccess) e = cudaDeviceEnablePeerAccess(dev0, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) e = cudaSuccess, (void)cudaGetLastError();
    }
    if (e != cudaSuccess || !can) {
      g_create_error = "group_create: device " + std::to_string(dev) + " cannot access device " + std::to_string(dev0) +
                       " as a peer" + (e != cudaSuccess ? std::string(": ") + cudaGetErrorString(e) : std::string());
      for (b200vr_t* m : g->member) b200vr_destroy(m);
      delete g;
      return B200VR_ECUDA;
    }
  }
  *out = g;
  return B200VR_OK;
}

extern "C" void b200vr_group_destroy(b200vr_group_t* g)
{
  if (!g) return;
  for (size_t i = g->member.size(); i-- > 0;) b200vr_destroy(g->member[i]); // member 0 (whose frame the others store into) last
  delete g;
}

extern "C" int b200vr_group_size(const b200vr_group_t* g) { return g ? (int)g->member.size() : 0; }

extern "C" b200vr_t* b200vr_group_member(b200vr_group_t* g, int i)
{
  return g && i >= 0 && i < (int)g->member.size() ? g->member[i] : nullptr;
}

extern "C" const char* b200vr_group_last_error(const b200vr_group_t* g) { return g ? g->error.c_str() : g_create_error.c_str(); }

extern "C" int b200vr_group_render(b200vr_group_t* g, b200vr_stats* stats)
{
  if (!g) return B200VR_EINVAL;
  const int n = (int)g->member.size();
  b200vr_t* h0 = g->member[0];
  if (h0->width <= 0 || h0->height <= 0) {
    g->error = "group_render: framebuffer size not set";
    return B200VR_ESTATE;
  }
  for (int i = 1; i < n; i++) {
    b200vr_t* h = g->member[i];
    if (h->width != h0->width || h->height != h0->height) {
      g->error = "group_render: members have different framebuffer sizes (apply every setter to every member)";
      return B200VR_ESTATE;
    }
    // not through b200vr_set_frame_target: the target flips with every swap of member 0, which is no reason to restart a
    // progressive accumulation (the sums live in the member's own accum buffer)
    h->target_rgba = h0->rgba[h0->current];
    h->target_grad = h0->grad[h0->current];
  }
  int rc = B200VR_OK;
  int enq = 0;
  for (; enq < n && rc == B200VR_OK; enq++) {
    rc = render_enqueue(g->member[enq]);
    if (rc != B200VR_OK) g->error = "member " + std::to_string(enq) + ": " + g->member[enq]->error;
  }
  b200vr_stats sum;
  std::memset(&sum, 0, sizeof(sum));
  for (int i = 0; i < n; i++) { // every member that has a frame in flight is waited for, also after an error
    b200vr_stats st;
    std::memset(&st, 0, sizeof(st));
    const int rw = render_wait(g->member[i], &st);
    if (rw != B200VR_OK && rc == B200VR_OK) rc = rw, g->error = "member " + std::to_string(i) + ": " + g->member[i]->error;
    sum.ms = std::max(sum.ms, st.ms);
    sum.primary_samples += st.primary_samples, sum.shadow_samples += st.shadow_samples, sum.gradient_fetches += st.gradient_fetches;
    sum.delta_samples += st.delta_samples, sum.rays += st.rays, sum.kernel_launches += st.kernel_launches;
    sum.frame_index = st.frame_index;
  }
  if (stats) *stats = sum;
  return rc;
}

extern "C" const char* b200vr_last_kernel(const b200vr_t* h) { return h ? h->last_kernel.c_str() : ""; }

extern "C" int b200vr_map(b200vr_t* h, void** rgba_dev, void** grad_dev)
{
  if (!h) return B200VR_EINVAL;
  if (h->width <= 0) return fail(h, B200VR_ESTATE, "map: framebuffer size not set");
  if (rgba_dev) *rgba_dev = h->rgba[h->current];
  if (grad_dev) *grad_dev = h->grad[h->current];
  return B200VR_OK;
}

extern "C" int b200vr_map_index(b200vr_t* h, int buffer, void** rgba_dev, void** grad_dev)
{
  if (!h || buffer < 0 || buffer > 1) return B200VR_EINVAL;
  if (h->width <= 0) return fail(h, B200VR_ESTATE, "map_index: framebuffer size not set");
  if (rgba_dev) *rgba_dev = h->rgba[buffer];
  if (grad_dev) *grad_dev = h->grad[buffer];
  return B200VR_OK;
}

extern "C" int b200vr_swap(b200vr_t* h)
{
  if (!h) return B200VR_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  h->current ^= 1;
  return B200VR_OK;
}

extern "C" int b200vr_download_async(b200vr_t* h, float* rgba_host, float* grad_host, void* cuda_stream)
{
  if (!h || !rgba_host) return B200VR_EINVAL;
  if (h->width <= 0) return fail(h, B200VR_ESTATE, "download_async: framebuffer size not set");
  CK(cudaSetDevice(h->device));
  const size_t n = (size_t)h->width * h->height;
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
  CK(cudaMemcpyAsync(rgba_host, h->rgba[h->current], n * sizeof(float4), cudaMemcpyDeviceToHost, st));
  if (grad_host) CK(cudaMemcpyAsync(grad_host, h->grad[h->current], n * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
  return B200VR_OK;
}

extern "C" int b200vr_download(b200vr_t* h, float* rgba_host, float* grad_host)
{
  if (!h || !rgba_host) return B200VR_EINVAL;
  if (h->width <= 0) return fail(h, B200VR_ESTATE, "download: framebuffer size not set");
  CK(cudaSetDevice(h->device));
  const size_t n = (size_t)h->width * h->height;
  CK(cudaMemcpyAsync(rgba_host, h->rgba[h->current], n * sizeof(float4), cudaMemcpyDeviceToHost, h->stream));
  if (grad_host) CK(cudaMemcpyAsync(grad_host, h->grad[h->current], n * 3 * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return B200VR_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// multi-GPU exchange helpers
// ---------------------------------------------------------------------------------------------------------------------

extern "C" int64_t b200vr_tile_slice_floats(b200vr_t* h, int rank, int nranks)
{
  if (!h || h->width <= 0 || nranks < 1 || rank < 0 || rank >= nranks) return -1;
  const int sw = (h->width + SUPER_TILE_W - 1) / SUPER_TILE_W, sh = (h->height + SUPER_TILE_H - 1) / SUPER_TILE_H;
  return (int64_t)owned_super_tiles(rank, nranks, sw * sh) * SUPER_TILE_W * SUPER_TILE_H * 4;
}

static int tile_copy(b200vr_t* h, float* slice, int rank, int nranks, bool pack)
{
  if (h->width <= 0) return fail(h, B200VR_ESTATE, "tile copy: framebuffer size not set");
  CK(cudaSetDevice(h->device));
  const int sw = (h->width + SUPER_TILE_W - 1) / SUPER_TILE_W, sh = (h->height + SUPER_TILE_H - 1) / SUPER_TILE_H;
  const int n_owned = owned_super_tiles(rank, nranks, sw * sh);
  if (n_owned == 0) return B200VR_OK;
  const size_t n = (size_t)n_owned * SUPER_TILE_W * SUPER_TILE_H;
  const int grid = grid_for(n, 256, h->sm_count);
  if (pack)
    tile_copy_kernel<true><<<grid, 256, 0, h->stream>>>(h->rgba[h->current], (float4*)slice, h->width, h->height, sw, rank, nranks, n_owned);
  else
    tile_copy_kernel<false><<<grid, 256, 0, h->stream>>>(h->rgba[h->current], (float4*)slice, h->width, h->height, sw, rank, nranks, n_owned);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  return B200VR_OK;
}

extern "C" int b200vr_pack_tiles(b200vr_t* h, float* slice_dev)
{
  if (!h || !slice_dev) return B200VR_EINVAL;
  return tile_copy(h, slice_dev, h->rank, h->nranks, true);
}

extern "C" int b200vr_unpack_tiles(b200vr_t* h, const float* slice_dev, int rank, int nranks)
{
  if (!h || !slice_dev || nranks < 1 || rank < 0 || rank >= nranks) return B200VR_EINVAL;
  return tile_copy(h, const_cast<float*>(slice_dev), rank, nranks, false);
}

extern "C" int b200vr_composite_over(b200vr_t* h, const float* front_dev, const float* back_dev, float* out_dev, int64_t n_pixels)
{
  if (!h || !front_dev || !back_dev || !out_dev || n_pixels < 0) return B200VR_EINVAL;
  if (n_pixels == 0) return B200VR_OK;
  CK(cudaSetDevice(h->device));
  composite_over_kernel<<<grid_for((size_t)n_pixels, 256, h->sm_count), 256, 0, h->stream>>>(
    (const float4*)front_dev, (const float4*)back_dev, (float4*)out_dev, (size_t)n_pixels);
  CK(cudaGetLastError());
  if (h->stream == h->own_stream) CK(cudaStreamSynchronize(h->stream)); // on a caller's stream the call is stream-ordered
  return B200VR_OK;
}

extern "C" int b200vr_composite_layers(b200vr_t* h, const float* const* layers_dev, int n_layers, float* out_dev, int64_t n_pixels,
                                       int finalize)
{
  if (!h || !layers_dev || !out_dev || n_pixels < 0) return B200VR_EINVAL;
  if (n_layers < 1 || n_layers > COMPOSITE_MAX_LAYERS) return fail(h, B200VR_EINVAL, "composite_layers: 1 to 16 layers");
  CompositeLayers L;
  std::memset(&L, 0, sizeof(L));
  for (int k = 0; k < n_layers; k++) {
    if (!layers_dev[k]) return fail(h, B200VR_EINVAL, "composite_layers: NULL layer");
    L.layer[k] = (const float4*)layers_dev[k];
  }
  L.n = n_layers;
  if (n_pixels == 0) return B200VR_OK;
  CK(cudaSetDevice(h->device));
  composite_layers_kernel<<<grid_for((size_t)n_pixels, 256, h->sm_count), 256, 0, h->stream>>>(L, (float4*)out_dev, (size_t)n_pixels, finalize);
  CK(cudaGetLastError());
  if (h->stream == h->own_stream) CK(cudaStreamSynchronize(h->stream));
  return B200VR_OK;
}

// ---- peer access for the fused sort-last exchange: one process per GPU, frames shared through CUDA IPC ----
extern "C" int b200vr_ipc_export_frame(b200vr_t* h, int buffer, unsigned char handle_out[64])
{
  if (!h || !handle_out || buffer < 0 || buffer > 3) return B200VR_EINVAL;
  if (h->width <= 0) return fail(h, B200VR_ESTATE, "ipc_export_frame: framebuffer size not set");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  CK(cudaSetDevice(h->device));
  cudaIpcMemHandle_t mh;
  CK(cudaIpcGetMemHandle(&mh, buffer < 2 ? (void*)h->rgba[buffer] : (void*)h->grad[buffer - 2]));
  std::memcpy(handle_out, &mh, 64);
  return B200VR_OK;
}

extern "C" int b200vr_ipc_open(b200vr_t* h, const unsigned char handle[64], void** dev_ptr_out)
{
  if (!h || !handle || !dev_ptr_out) return B200VR_EINVAL;
  CK(cudaSetDevice(h->device));
  cudaIpcMemHandle_t mh;
  std::memcpy(&mh, handle, 64);
  CK(cudaIpcOpenMemHandle(dev_ptr_out, mh, cudaIpcMemLazyEnablePeerAccess));
  return B200VR_OK;
}

extern "C" int b200vr_ipc_close(b200vr_t* h, void* dev_ptr)
{
  if (!h || !dev_ptr) return B200VR_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaIpcCloseMemHandle(dev_ptr));
  return B200VR_OK;
}

extern "C" int b200vr_set_frame_target(b200vr_t* h, float* rgba_dev, float* grad_dev)
{
  if (!h || ((rgba_dev == nullptr) != (grad_dev == nullptr))) return B200VR_EINVAL;
  h->target_rgba = (float4*)rgba_dev;
  h->target_grad = grad_dev;
  h->reset = true;
  return B200VR_OK;
}

extern "C" int b200vr_current_buffer(b200vr_t* h)
{
  return h ? h->current : -1;
}

extern "C" int b200vr_copy_device(b200vr_t* h, void* dst_dev, const void* src_dev, int64_t bytes)
{
  if (!h || !dst_dev || !src_dev || bytes < 0) return B200VR_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaMemcpyAsync(dst_dev, src_dev, (size_t)bytes, cudaMemcpyDeviceToDevice, h->stream)); // UVA: src may be a peer's memory
  if (h->stream == h->own_stream) CK(cudaStreamSynchronize(h->stream));
  return B200VR_OK;
}

extern "C" int b200vr_finalize_premultiplied(b200vr_t* h, float* rgba_dev, int64_t n_pixels)
{
  if (!h || !rgba_dev || n_pixels < 0) return B200VR_EINVAL;
  if (n_pixels == 0) return B200VR_OK;
  CK(cudaSetDevice(h->device));
  finalize_premultiplied_kernel<<<grid_for((size_t)n_pixels, 256, h->sm_count), 256, 0, h->stream>>>((float4*)rgba_dev, (size_t)n_pixels);
  CK(cudaGetLastError());
  if (h->stream == h->own_stream) CK(cudaStreamSynchronize(h->stream));
  return B200VR_OK;
}

extern "C" int b200vr_probe(b200vr_t* h, const float* points, int64_t n, int sampler, float* values_out, float* rgba_out)
{
  if (!h || !points || !values_out || n < 0) return B200VR_EINVAL;
  if (!h->has_volume) return fail(h, B200VR_ESTATE, "probe: no volume");
  if (rgba_out && !h->has_tf) return fail(h, B200VR_ESTATE, "probe: no transfer function");
  if (n == 0) return B200VR_OK;
  CK(cudaSetDevice(h->device));
  float *d_pts = nullptr, *d_val = nullptr;
  float4* d_rgba = nullptr;
  CK(cudaMalloc(&d_pts, (size_t)n * 3 * sizeof(float)));
  CK(cudaMalloc(&d_val, (size_t)n * sizeof(float)));
  if (rgba_out) CK(cudaMalloc(&d_rgba, (size_t)n * sizeof(float4)));
  CK(cudaMemcpyAsync(d_pts, points, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, h->stream));
  LaunchParams P;
  const bool cam = h->has_camera;
  if (!cam) h->cam_pos = mk3(0.f), h->cam_dir = mk3(0.f, 0.f, 1.f), h->cam_hor = mk3(1.f, 0.f, 0.f), h->cam_ver = mk3(0.f, 1.f, 0.f);
  fill_launch_params(h, P);
  const size_t smem = rgba_out ? tf_shared_bytes(h->n_color, h->n_alpha) : 0;
  if (!rgba_out) P.n_color = P.n_alpha = 0;
  const void* k = sampler == B200VR_SAMPLER_MANUAL ? (const void*)probe_kernel<SAMPLER_MANUAL> : (const void*)probe_kernel<SAMPLER_TEX3D>;
  if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long nn = n;
  void* args[] = { (void*)&P, (void*)&d_pts, (void*)&nn, (void*)&d_val, (void*)&d_rgba };
  CK(cudaLaunchKernel(k, dim3((unsigned)grid_for((size_t)n, 128, h->sm_count)), dim3(128), args, smem, h->stream));
  CK(cudaMemcpyAsync(values_out, d_val, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  if (rgba_out) CK(cudaMemcpyAsync(rgba_out, d_rgba, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  cudaFree(d_pts);
  cudaFree(d_val);
  cudaFree(d_rgba);
  return B200VR_OK;
}

extern "C" int b200vr_probe_tex1d(b200vr_t* h, const float* table, int n, int channels, const float* u, int64_t n_u, float* hw_out,
                                  float* sw_out)
{
  if (!h || !table || !u || !hw_out || !sw_out || n <= 0 || n_u < 0 || (channels != 1 && channels != 4)) return B200VR_EINVAL;
  if (n_u == 0) return B200VR_OK;
  CK(cudaSetDevice(h->device));
  const cudaChannelFormatDesc desc = channels == 1 ? cudaCreateChannelDesc<float>() : cudaCreateChannelDesc<float4>();
  cudaArray_t arr = nullptr;
  CK(cudaMallocArray(&arr, &desc, (size_t)n, 0));
  const size_t row = (size_t)n * channels * sizeof(float);
  CK(cudaMemcpy2DToArray(arr, 0, 0, table, row, row, 1, cudaMemcpyHostToDevice));
  cudaResourceDesc rd;
  std::memset(&rd, 0, sizeof(rd));
  rd.resType = cudaResourceTypeArray;
  rd.res.array.array = arr;
  cudaTextureDesc td;
  std::memset(&td, 0, sizeof(td));
  td.addressMode[0] = cudaAddressModeClamp;
  td.filterMode = cudaFilterModeLinear;
  td.readMode = cudaReadModeElementType;
  td.normalizedCoords = 1;
  cudaTextureObject_t tex = 0;
  CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
  float *d_tab = nullptr, *d_u = nullptr, *d_hw = nullptr, *d_sw = nullptr;
  CK(cudaMalloc(&d_tab, row));
  CK(cudaMalloc(&d_u, (size_t)n_u * sizeof(float)));
  CK(cudaMalloc(&d_hw, (size_t)n_u * channels * sizeof(float)));
  CK(cudaMalloc(&d_sw, (size_t)n_u * channels * sizeof(float)));
  CK(cudaMemcpyAsync(d_tab, table, row, cudaMemcpyHostToDevice, h->stream));
  CK(cudaMemcpyAsync(d_u, u, (size_t)n_u * sizeof(float), cudaMemcpyHostToDevice, h->stream));
  const int grid = grid_for((size_t)n_u, 128, h->sm_count);
  if (channels == 1)
    probe_tex1d_kernel<1><<<grid, 128, 0, h->stream>>>(tex, d_tab, n, d_u, (long long)n_u, d_hw, d_sw);
  else
    probe_tex1d_kernel<4><<<grid, 128, 0, h->stream>>>(tex, d_tab, n, d_u, (long long)n_u, d_hw, d_sw);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(hw_out, d_hw, (size_t)n_u * channels * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaMemcpyAsync(sw_out, d_sw, (size_t)n_u * channels * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  cudaFree(d_tab);
  cudaFree(d_u);
  cudaFree(d_hw);
  cudaFree(d_sw);
  cudaDestroyTextureObject(tex);
  cudaFreeArray(arr);
  return B200VR_OK;
}

extern "C" int b200vr_bench_fetch(b200vr_t* h, int sampler, int footprint, int iters, const float dir[3], int ctas_per_sm,
                                  float* ms_out, uint64_t* fetches_out)
{
  if (!h || !dir || !ms_out || !fetches_out || iters <= 0 || ctas_per_sm <= 0) return B200VR_EINVAL;
  if (!h->has_volume) return fail(h, B200VR_ESTATE, "bench_fetch: no volume");
  const int min_dim = std::min(h->nx, std::min(h->ny, h->nz));
  if (footprint < 8 || footprint > min_dim) return fail(h, B200VR_EINVAL, "bench_fetch: footprint must be in [8, min(dims)]");
  CK(cudaSetDevice(h->device));
  LaunchParams P;
  if (!h->has_camera) h->cam_pos = mk3(0.f), h->cam_dir = mk3(0.f, 0.f, 1.f), h->cam_hor = mk3(1.f, 0.f, 0.f), h->cam_ver = mk3(0.f, 1.f, 0.f);
  fill_launch_params(h, P);
  const void* k = sampler == B200VR_SAMPLER_MANUAL ? (const void*)fetch_bench_kernel<SAMPLER_MANUAL> : (const void*)fetch_bench_kernel<SAMPLER_TEX3D>;
  float* sink = (float*)h->counters;
  float dx = dir[0], dy = dir[1], dz = dir[2];
  const int grid = h->sm_count * ctas_per_sm;
  void* args[] = { (void*)&P, (void*)&footprint, (void*)&iters, (void*)&dx, (void*)&dy, (void*)&dz, (void*)&sink };
  CK(cudaLaunchKernel(k, dim3((unsigned)grid), dim3(RENDER_THREADS), args, 0, h->stream)); // warm-up
  CK(cudaEventRecord(h->ev0, h->stream));
  CK(cudaLaunchKernel(k, dim3((unsigned)grid), dim3(RENDER_THREADS), args, 0, h->stream));
  CK(cudaEventRecord(h->ev1, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaEventElapsedTime(ms_out, h->ev0, h->ev1));
  *fetches_out = (uint64_t)grid * RENDER_THREADS * (uint64_t)iters;
  return B200VR_OK;
}

extern "C" int b200vr_set_stream(b200vr_t* h, void* cuda_stream)
{
  if (!h) return B200VR_EINVAL;
  CK(cudaSetDevice(h->device));
  CK(cudaStreamSynchronize(h->stream));
  h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->own_stream;
  return B200VR_OK;
}
