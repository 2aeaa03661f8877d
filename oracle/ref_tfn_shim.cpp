// TEST INFRASTRUCTURE ONLY -- the reference's transfer-function loader (extern/tfn/core.h: loadTransferFunction,
// updateColorMap, GaussianObject, updateFromAlphaControls) compiled as host code from where it lies under /root/reference
// (oracle/Makefile, target `ref`), with the vector types the reference's serializer gives it
// (ovr/serializer/serializer_vidi3d.cpp:30-40). tests/golden/make_golden.py drives it to pin scene_io.read_vidi3d_scene.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include <gdt/math/vec.h>
namespace tfn {
typedef gdt::vec2f vec2f;
typedef gdt::vec2i vec2i;
typedef gdt::vec3f vec3f;
typedef gdt::vec3i vec3i;
typedef gdt::vec4f vec4f;
typedef gdt::vec4i vec4i;
} // namespace tfn
#define TFN_MODULE_EXTERNAL_VECTOR_TYPES
#include <tfn/core.h>

// json_text: the "transferFunction" object of a VIDI3D scene file; rgba_out: capacity x 4 floats. Returns the table's
// resolution (negative: capacity too small). The table is what create_scene_tfn reads (serializer_vidi3d.cpp:210-221).
extern "C" int ref_tfn_table(const char* json_text, float* rgba_out, int capacity)
{
  const nlohmann::json j = nlohmann::json::parse(json_text, nullptr, true, true);
  tfn::TransferFunctionCore tf;
  tfn::loadTransferFunction(j, tf);
  const int n = tf.resolution();
  if (n > capacity) return -n;
  std::memcpy(rgba_out, tf.data(), sizeof(float) * 4 * (size_t)n);
  return n;
}
