"""TEST INFRASTRUCTURE ONLY -- cuts the reference's ray-marching and path-tracing functions out of its shader sources into the git-ignored
oracle/_ref/ directory, one file per function (name line + body, without the `inline __device__ <type>` line above it), so
that oracle/ref_march_shim.cpp can compile them as host code. Nothing is copied into the repository.

Usage: python extract_march.py <reference root> <output dir>"""
import os
import re
import sys

REF, OUT = sys.argv[1], sys.argv[2]
COMMON = "ovr/devices/optix7/shaders_common.h"
MARCH = "ovr/devices/optix7/shaders_raymarching.cu"
PATH = "ovr/devices/optix7/shaders_pathtracing.cu"
SPART = "ovr/devices/optix7/accel/spatial_partition.h"

# (source, first line regex, last line regex, output file)
ITEMS = [
    (COMMON, r"^struct ShadowPayload \{", r"^\};", "m_shadow_payload.h"),
    (COMMON, r"^struct RayMarchingPayload \{", r"^\};", "m_march_payload.h"),
    (COMMON, r"^corrected_value\(const float in\)", r"^\}", "m_corrected_f.h"),
    (COMMON, r"^corrected_value\(const vec3f in\)", r"^\}", "m_corrected_v.h"),
    (COMMON, r"^sample_volume_object_space\(", r"^\}", "m_sample_volume.h"),
    (COMMON, r"^compute_volume_gradient_object_space\(", r"^\}", "m_gradient.h"),
    (COMMON, r"^get_xfm_wto\(\)", r"^\}", "m_xfm_wto.h"),
    (COMMON, r"^get_xfm_otw\(\)", r"^\}", "m_xfm_otw.h"),
    (COMMON, r"^get_xfm_world_to_camera\(\)", r"^\}", "m_xfm_wtc.h"),
    (COMMON, r"^project_to_screen\(", r"^\}", "m_project.h"),
    (COMMON, r"^compute_optical_flow\(", r"^\}", "m_flow.h"),
    (COMMON, r"^array1d_nodal\(", r"^\}", "m_nodal.h"),
    (COMMON, r"^nearly_equal\(", r"^\}", "m_nearly_equal.h"),
    (COMMON, r"^alpha_blend\(", r"^\}", "m_alpha_blend.h"),
    (COMMON, r"^sample_transfer_function\(", r"^\}", "m_sample_tf.h"),
    (MARCH, r"^raymarching_shadow\(", r"^\}", "m_raymarching_shadow.h"),
    (MARCH, r"^raymarching\(", r"^\}", "m_raymarching.h"),
    (MARCH, r"^render_raymarching\(vec3f org,", r"^\}", "m_render_raymarching.h"),
    # delta-tracking path tracer (ref_march_shim.cpp, namespace pt)
    (COMMON, r"^struct PathTracingPayload \{", r"^\};", "p_payload.h"),
    (SPART, r"^  DeviceSpacePartiton_SingleMC::iter\(", r"^  \}", "p_sp_iter.h"),
    (PATH, r"^delta_tracking\(const DeviceStructuredRegularVolume& self,", r"^\}", "p_delta_tracking.h"),
    (PATH, r"^pathtracing\(const DeviceStructuredRegularVolume& self,", r"^\}", "p_pathtracing.h"),
    (PATH, r"^render_pathtracing\(vec3f org,", r"^\}", "p_render_pathtracing.h"),
]

os.makedirs(OUT, exist_ok=True)
for src, first, last, out in ITEMS:
    lines = open(os.path.join(REF, src)).read().split("\n")
    start = next(i for i, l in enumerate(lines) if re.match(first, l))
    end = next(i for i in range(start + 1, len(lines)) if re.match(last, lines[i]))
    with open(os.path.join(OUT, out), "w") as f:
        f.write("\n".join(lines[start:end + 1]) + "\n")
print(f"extracted {len(ITEMS)} pieces of the reference's march shaders into {OUT}")
