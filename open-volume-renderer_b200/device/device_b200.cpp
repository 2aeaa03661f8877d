// DeviceB200: ovr::MainRenderer back-end on top of the b200vr C ABI (see device_b200.h).
// Host logic mirrors what DeviceOptix7::Impl does around its launch (ovr/devices/optix7/device_impl.cpp:70-302): which
// parameter changes are applied in commit(), what resets the accumulation, what mapframe hands out. The arithmetic of
// the frame itself lives in the CUDA kernels behind b200vr_render.
#include "device_b200.h"

#include "b200vr.h"

#include <ovr/common/dylink/ObjectFactory.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

namespace ovr::b200 {

namespace {

int
to_b200vr_type(ValueType t)
{
  switch (t) {
  case VALUE_TYPE_UINT8: return B200VR_TYPE_UINT8;
  case VALUE_TYPE_INT8: return B200VR_TYPE_INT8;
  case VALUE_TYPE_UINT16: return B200VR_TYPE_UINT16;
  case VALUE_TYPE_INT16: return B200VR_TYPE_INT16;
  case VALUE_TYPE_UINT32: return B200VR_TYPE_UINT32;
  case VALUE_TYPE_INT32: return B200VR_TYPE_INT32;
  case VALUE_TYPE_FLOAT: return B200VR_TYPE_FLOAT;
  case VALUE_TYPE_DOUBLE: return B200VR_TYPE_DOUBLE;
  default: throw std::runtime_error("b200: unsupported volume value type"); // array.cpp:347
  }
}

int
env_int(const char* name, int fallback)
{
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : fallback;
}

// "0,1,2,3" or "4" (= devices 0..3) -> CUDA device indices; empty / unparsable -> {fallback}
std::vector<int>
parse_devices(const std::string& text, int fallback)
{
  std::vector<int> out;
  if (text.find(',') == std::string::npos) {
    const int n = text.empty() ? 0 : std::atoi(text.c_str());
    if (n > 1) {
      for (int i = 0; i < n; i++) out.push_back(i);
      return out;
    }
    return { fallback };
  }
  size_t pos = 0;
  while (pos <= text.size()) {
    const size_t end = std::min(text.find(',', pos), text.size());
    if (end > pos) out.push_back(std::atoi(text.substr(pos, end - pos).c_str()));
    pos = end + 1;
  }
  if (out.empty()) out.push_back(fallback);
  return out;
}

} // namespace

struct DeviceB200::Impl {
  DeviceB200* parent = nullptr;
  // One handle of the C ABI per GPU. With several GPUs (OVR_B200_GPUS / --b200-gpus) they form a b200vr group: every setter
  // goes to every member, the frame is split sort-first and assembled in member 0's buffers over NVLink (include/b200vr.h).
  b200vr_group_t* group = nullptr;
  std::vector<b200vr_t*> members;
  b200vr_t* handle = nullptr; // members[0]: the display GPU (map / swap / download)
  std::vector<int> devices;
  b200vr_params prm;
  bool initialized = false;
  bool prm_dirty = false;
  vec2i fbsize = vec2i(0, 0);
  // sparse sampling state (defaults of apps/main_app.cpp:122-126)
  vec2f focus_center = vec2f(0.5f, 0.5f);
  float focus_scale = 0.06f, base_noise = 0.07f;
  bool sparse_sampling = false, noise_loaded = false;
#ifndef OVR_BUILD_CUDA_DEVICES
  std::vector<float> host_rgba, host_grad; // host executable built without CUDA buffers: frames are handed out on the CPU
#endif

  ~Impl()
  {
    if (group) b200vr_group_destroy(group);
  }

  // applies one C-ABI setter to every member of the group
  template<typename F>
  void each(const char* what, F&& call) const
  {
    for (b200vr_t* m : members) {
      const int rc = call(m);
      if (rc != B200VR_OK) throw std::runtime_error(std::string("b200: ") + what + ": " + b200vr_last_error(m));
    }
  }

  void check(int rc, const char* what) const
  {
    if (rc != B200VR_OK)
      throw std::runtime_error(std::string("b200: ") + what + ": " + b200vr_last_error(handle));
  }

  // Device-private options arrive the way the reference passes them to its devices: the raw command line handed to
  // MainRenderer::init (ovr/renderer.h:252; the OSPRay device forwards it to ospInit, ospray/device_impl.cpp:419).
  // --b200-gpus <n | i,j,k> (or OVR_B200_GPUS, which wins when the host program rejects unknown flags): the GPUs to render on.
  void init(DeviceB200* p, int argc, const char** argv)
  {
    if (initialized) throw std::runtime_error("b200: device already initialized!"); // device_impl.cpp:73-75
    parent = p;
    std::string gpus;
    for (int i = 1; argv && i < argc; i++) {
      const std::string a = argv[i] ? argv[i] : "";
      if (a == "--b200-gpus" && i + 1 < argc && argv[i + 1]) gpus = argv[i + 1];
      else if (a.rfind("--b200-gpus=", 0) == 0) gpus = a.substr(12);
    }
    if (const char* e = std::getenv("OVR_B200_GPUS")) gpus = e;
    devices = parse_devices(gpus, env_int("OVR_B200_CUDA_DEVICE", 0));
    const int rc = b200vr_group_create(&group, devices.data(), (int)devices.size());
    if (rc != B200VR_OK) throw std::runtime_error(std::string("b200: cannot create device: ") + b200vr_group_last_error(nullptr));
    for (int i = 0; i < b200vr_group_size(group); i++) members.push_back(b200vr_group_member(group, i));
    handle = members[0];
    b200vr_default_params(&prm);
    // device-private switches that MainRenderer has no setter for (they default to the reference's behaviour)
    prm.shading = env_int("OVR_B200_SHADING", B200VR_SHADING_REF);
    prm.sampler = env_int("OVR_B200_SAMPLER", B200VR_SAMPLER_TEX3D);
    prm.skip_empty = env_int("OVR_B200_SKIP_EMPTY", 1);
    prm.pt_child_instance_xfm = env_int("OVR_B200_PT_CHILD_INSTANCE_XFM", 1);
    build_scene(parent->current_scene);
    initialized = true;
  }

  // buildScene (device_impl.cpp:283-302): exactly the scenes parse_single_volume_scene accepts
  void build_scene(const Scene& scene)
  {
    const auto& v = parse_single_volume_scene(scene, scene::Volume::STRUCTURED_REGULAR_VOLUME).structured_regular;
    const vec3f scale = v.grid_spacing * vec3f(v.data->dims);
    const vec3f translate = v.grid_origin;
    const float otw[12] = { scale.x, 0.f, 0.f, translate.x, 0.f, scale.y, 0.f, translate.y, 0.f, 0.f, scale.z, translate.z };
    each("set_volume", [&](b200vr_t* m) {
      return b200vr_set_volume(m, v.data->data(), to_b200vr_type(v.data->type), v.data->dims.x, v.data->dims.y, v.data->dims.z, otw,
                               /*on_device=*/0);
    });
    prm.sampling_rate = scene.volume_sampling_rate;
    prm_dirty = true;
  }

  // The reference embeds its noise table at build time (noise_files.h generated from stbn_128x128x64.bin, which the
  // repository does not ship; ovr/CMakeLists.txt:60-75). Here OVR_B200_NOISE_FILE names a raw little-endian f32 table laid
  // out [y][x][t] with t = 64 (blue_noise.h:39,95-99), e.g. data/noise/blue_64x64x64.bin; unset = the built-in table.
  void load_noise_table()
  {
    noise_loaded = true;
    const char* path = std::getenv("OVR_B200_NOISE_FILE");
    if (!path || !*path) return;
    std::FILE* f = std::fopen(path, "rb");
    if (!f) throw std::runtime_error(std::string("b200: cannot open noise table ") + path);
    std::vector<float> table;
    float buf[4096];
    size_t n;
    while ((n = std::fread(buf, sizeof(float), 4096, f)) > 0) table.insert(table.end(), buf, buf + n);
    std::fclose(f);
    int tile = 0;
    while ((size_t)(tile + 1) * (tile + 1) * 64 <= table.size()) tile++;
    if (tile == 0 || (size_t)tile * tile * 64 != table.size())
      throw std::runtime_error(std::string("b200: noise table is not tile x tile x 64 floats: ") + path);
    each("set_sparse_noise", [&](b200vr_t* m) { return b200vr_set_sparse_noise(m, table.data(), tile, 64); });
  }

  void swap() { check(b200vr_swap(handle), "swap"); }

  // Impl::commit (device_impl.cpp:113-197): only changed parameters are applied; every change resets the accumulation
  // (the C ABI setters do that themselves)
  void commit()
  {
    auto& pp = parent->params;
    bool size_changed = false;
    if (pp.fbsize.update()) {
      fbsize = pp.fbsize.ref();
      if (fbsize.x > 0 && fbsize.y > 0) each("resize", [&](b200vr_t* m) { return b200vr_resize(m, fbsize.x, fbsize.y); });
      size_changed = true;
    }
    if (pp.camera.update() || size_changed) {
      const Camera& c = pp.camera.ref();
      const float from[3] = { c.from.x, c.from.y, c.from.z }, at[3] = { c.at.x, c.at.y, c.at.z }, up[3] = { c.up.x, c.up.y, c.up.z };
      each("set_camera", [&](b200vr_t* m) { return b200vr_set_camera_lookat(m, from, at, up, c.perspective.fovy); });
    }
    if (pp.tfn.update()) {
      const auto& t = pp.tfn.ref();
      if (!t.tfn_colors.empty() && !t.tfn_alphas.empty())
        each("set_transfer_function", [&](b200vr_t* m) {
          return b200vr_set_transfer_function(m, t.tfn_colors.data(), (int)(t.tfn_colors.size() / 3), t.tfn_alphas.data(),
                                              (int)(t.tfn_alphas.size() / 2), t.tfn_value_range.x, t.tfn_value_range.y);
        });
    }
    if (pp.sample_per_pixel.update()) {
      prm.spp = pp.sample_per_pixel.ref();
      prm_dirty = true;
    }
    if (pp.path_tracing.update()) {
      prm.mode = pp.path_tracing.ref() ? B200VR_MODE_PATHTRACE : B200VR_MODE_RAYMARCH;
      prm_dirty = true;
    }
    if (pp.frame_accumulation.update()) {
      prm.accumulate = pp.frame_accumulation.ref() ? 1 : 0;
      prm_dirty = true;
    }
    if (pp.volume_sampling_rate.update()) {
      prm.sampling_rate = pp.volume_sampling_rate.get();
      prm_dirty = true;
    }
    // sparse (foveated) sampling, device_impl.cpp:154-183; lights / materials are accepted and ignored like the optix7
    // device ignores them
    bool sparse_dirty = false;
    if (pp.focus_center.update()) focus_center = pp.focus_center.ref(), sparse_dirty = true;
    if (pp.focus_scale.update()) focus_scale = pp.focus_scale.ref(), sparse_dirty = true;
    if (pp.base_noise.update()) base_noise = pp.base_noise.ref(), sparse_dirty = true;
    if (pp.sparse_sampling.update()) sparse_sampling = pp.sparse_sampling.ref(), sparse_dirty = true;
    if (sparse_dirty) {
      if (sparse_sampling && !noise_loaded) load_noise_table();
      const float fc[2] = { focus_center.x, focus_center.y };
      if (sparse_sampling && members.size() > 1)
        throw std::runtime_error("b200: sparse sampling is a single-device mode (the reference's too): run with one GPU");
      each("set_sparse_sampling", [&](b200vr_t* m) { return b200vr_set_sparse_sampling(m, sparse_sampling ? 1 : 0, fc, focus_scale, base_noise); });
    }
    if (prm_dirty) {
      each("set_params", [&](b200vr_t* m) { return b200vr_set_params(m, &prm); });
      prm_dirty = false;
    }
  }

  void render()
  {
    if (fbsize.x <= 0 || fbsize.y <= 0) return; // device_impl.cpp:215-216
    // synchronous, like DeviceOptix7::render (device.cpp:34-43); with several GPUs every member renders its image tiles
    // straight into member 0's frame (b200vr_group_render)
    if (members.size() > 1) {
      if (b200vr_group_render(group, nullptr) != B200VR_OK)
        throw std::runtime_error(std::string("b200: render: ") + b200vr_group_last_error(group));
    }
    else
      check(b200vr_render(handle, nullptr), "render");
    parent->variance = 0.f;                          // device_impl.cpp:266
  }

  void mapframe(FrameBufferData* fb)
  {
    const size_t n = (size_t)fbsize.x * (size_t)fbsize.y;
#ifdef OVR_BUILD_CUDA_DEVICES
    void *rgba = nullptr, *grad = nullptr;
    check(b200vr_map(handle, &rgba, &grad), "map");
    fb->rgba->set_data(rgba, n * sizeof(vec4f), CrossDeviceBuffer::DEVICE_CUDA); // non-owning (device_impl.cpp:271-281)
    fb->grad->set_data(grad, n * sizeof(vec3f), CrossDeviceBuffer::DEVICE_CUDA);
#else
    host_rgba.resize(n * 4);
    host_grad.resize(n * 3);
    check(b200vr_download(handle, host_rgba.data(), host_grad.data()), "download");
    fb->rgba->set_data(host_rgba.data(), n * sizeof(vec4f), CrossDeviceBuffer::DEVICE_CPU);
    fb->grad->set_data(host_grad.data(), n * sizeof(vec3f), CrossDeviceBuffer::DEVICE_CPU);
#endif
  }
};

DeviceB200::~DeviceB200()
{
  pimpl.reset();
}

DeviceB200::DeviceB200() : MainRenderer(), pimpl(new Impl()) {}

void
DeviceB200::init(int argc, const char** argv)
{
  pimpl->init(this, argc, argv);
  pimpl->commit();
}

void
DeviceB200::swap()
{
  pimpl->swap();
}

void
DeviceB200::commit()
{
  pimpl->commit();
}

void
DeviceB200::render()
{
  const auto start = std::chrono::high_resolution_clock::now();
  pimpl->render();
  const auto end = std::chrono::high_resolution_clock::now();
  render_time += std::chrono::duration<double, std::milli>(end - start).count();
}

void
DeviceB200::mapframe(FrameBufferData* fb)
{
  pimpl->mapframe(fb);
}

} // namespace ovr::b200

OVR_REGISTER_OBJECT(ovr::MainRenderer, renderer, ovr::b200::DeviceB200, b200)
