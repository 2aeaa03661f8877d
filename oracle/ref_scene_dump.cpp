// TEST INFRASTRUCTURE ONLY -- prints what the REFERENCE's own scene loader (ovr::scene::create_scene: ovr/scene.cpp,
// ovr/serializer/serializer_vidi3d.cpp: the reference's host objects in build/ovr/libovr_host.a; linked into
// build/ovr/scenedump by oracle/pyoracle.py build_scenedump) hands to a device for each scene file on the command line, one JSON object per line:
// dims, value type, grid origin / spacing, FNV-1a of the voxel bytes, voxel min / max, transfer-function colours, opacities
// and value range, camera, sampling rate. tests/golden/make_golden.py feeds it synthetic VIDI3D configs; the output
// (tests/golden/scene_dump.json) pins open-volume-renderer_b200/scene_io.py, the Python mirror of that loader.
#include <cstdint>
#include <cstdio>
#include <string>

#include "ovr/scene.h"
#include "ovr/serializer/serializer.h"

template<typename T> static void minmax(const char* p, size_t n, double& lo, double& hi)
{
  const T* v = (const T*)p;
  lo = hi = (double)v[0];
  for (size_t i = 1; i < n; i++) {
    if ((double)v[i] < lo) lo = (double)v[i];
    if ((double)v[i] > hi) hi = (double)v[i];
  }
}

int main(int argc, char** argv)
{
  using namespace ovr;
  for (int a = 1; a < argc; a++) {
    try {
      const Scene s = scene::create_scene(std::string(argv[a]));
      const auto& v = parse_single_volume_scene(s).structured_regular;
      const auto& tf = s.instances[0].models[0].volume_model.transfer_function;
      const size_t n = (size_t)v.data->dims.long_product(), bytes = n * value_type_size(v.data->type);
      uint64_t h = 1469598103934665603ull;
      const unsigned char* p = (const unsigned char*)v.data->data();
      for (size_t i = 0; i < bytes; i++) h = (h ^ p[i]) * 1099511628211ull;
      double lo = 0, hi = 0;
      switch (v.data->type) {
      case VALUE_TYPE_UINT8: minmax<uint8_t>(v.data->data(), n, lo, hi); break;
      case VALUE_TYPE_INT8: minmax<int8_t>(v.data->data(), n, lo, hi); break;
      case VALUE_TYPE_UINT16: minmax<uint16_t>(v.data->data(), n, lo, hi); break;
      case VALUE_TYPE_INT16: minmax<int16_t>(v.data->data(), n, lo, hi); break;
      case VALUE_TYPE_UINT32: minmax<uint32_t>(v.data->data(), n, lo, hi); break;
      case VALUE_TYPE_INT32: minmax<int32_t>(v.data->data(), n, lo, hi); break;
      case VALUE_TYPE_FLOAT: minmax<float>(v.data->data(), n, lo, hi); break;
      case VALUE_TYPE_DOUBLE: minmax<double>(v.data->data(), n, lo, hi); break;
      default: break;
      }
      std::printf("{\"file\": \"%s\", \"dims\": [%d, %d, %d], \"type_bytes\": %d, \"origin\": [%.9g, %.9g, %.9g], \"spacing\": [%.9g, %.9g, %.9g], "
                  "\"fnv\": \"%016llx\", \"min\": %.17g, \"max\": %.17g, \"tf_range\": [%.9g, %.9g], \"rate\": %.9g, "
                  "\"eye\": [%.9g, %.9g, %.9g], \"center\": [%.9g, %.9g, %.9g], \"up\": [%.9g, %.9g, %.9g], \"fovy\": %.9g, ",
                  argv[a], v.data->dims.x, v.data->dims.y, v.data->dims.z, (int)value_type_size(v.data->type), v.grid_origin.x,
                  v.grid_origin.y, v.grid_origin.z, v.grid_spacing.x, v.grid_spacing.y, v.grid_spacing.z, (unsigned long long)h, lo, hi,
                  tf.value_range.x, tf.value_range.y, s.volume_sampling_rate, s.camera.from.x, s.camera.from.y, s.camera.from.z,
                  s.camera.at.x, s.camera.at.y, s.camera.at.z, s.camera.up.x, s.camera.up.y, s.camera.up.z, s.camera.perspective.fovy);
      const float* c = (const float*)tf.color->data();
      std::printf("\"color\": [");
      for (size_t i = 0; i < (size_t)tf.color->dims.long_product() * 4; i++) std::printf("%s%.9g", i ? ", " : "", c[i]);
      const float* o = (const float*)tf.opacity->data();
      std::printf("], \"opacity\": [");
      for (size_t i = 0; i < (size_t)tf.opacity->dims.long_product(); i++) std::printf("%s%.9g", i ? ", " : "", o[i]);
      std::printf("]}\n");
    }
    catch (const std::exception& e) {
      std::printf("{\"file\": \"%s\", \"error\": \"%s\"}\n", argv[a], e.what());
    }
  }
  return 0;
}
