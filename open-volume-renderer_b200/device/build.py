"""Builds the OVR plugin and the drop-in driver (needs the reference tree; the products travel to the GPU box prebuilt):

  build/ovr/libdevice_b200.so   DeviceB200 : ovr::MainRenderer (device_b200.cpp) compiled against the reference's UNCHANGED
                                headers, linked to open-volume-renderer_b200/libb200vr.so
  build/ovr/libovr_host.a       the same reference objects without main_batch.o, for other host tools built from the reference
  build/ovr/renderbatch         the reference's own apps/main_batch.cpp + host sources (renderer.cpp, scene.cpp,
                                serializer_{vidi3d,diva}.cpp, imageio.cpp, dylink/Library.cpp) compiled where they lie under
                                /root/reference with OVR_BUILD_OPTIX7 / OVR_BUILD_OSPRAY undefined -- nothing is copied
                                into this repo. `renderbatch --device b200 ...` then loads the plugin through
                                create_renderer (ovr/renderer.cpp:42-61).

  build/ovr/renderloop          device/renderloop.cpp: the render loop of the reference's GUI (apps/main_app.cpp:233-263: commit ->
                                mapframe -> to_cpu -> swap -> render) without a window, linked against the same host objects

The reference's CMake build is not used (it needs OptiX / OSPRay / TBB / GLFW); this is the short recipe the task allows.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("OVR_REFERENCE", "/root/reference")
OUT = os.path.join(ROOT, "build", "ovr")
PKG = os.path.join(ROOT, "open-volume-renderer_b200")

INC = [REF, f"{REF}/ovr", f"{REF}/ovr/common", f"{REF}/ovr/common/cuda", f"{REF}/extern", f"{REF}/extern/gdt",
       f"{REF}/extern/tfn/colormaps", f"{REF}/extern/tfn", f"{REF}/apps", "/usr/local/cuda/include", f"{ROOT}/include"]
# OVR_BUILD_CUDA_DEVICES must have the same value in the host executable and the plugin: it changes the layout of
# CrossDeviceBuffer (ovr/common/cross_device_buffer.h:22-40)
DEFS = ["-DOVR_BUILD_CUDA_DEVICES"]
CXX = ["g++", "-std=c++17", "-fPIC", "-w"] + DEFS + [f"-I{i}" for i in INC]

HOST_SOURCES = [
    ("apps/main_batch.cpp", ["-O1"]),
    ("ovr/renderer.cpp", ["-O1"]),
    ("ovr/scene.cpp", ["-O1"]),
    ("ovr/serializer/serializer_vidi3d.cpp", ["-O1"]),
    # serializer_usda.cpp is optional in the reference (OVR_BUILD_SCENE_USD, ovr/CMakeLists.txt:90-94) and needs Pixar USD
    # (pxr/...), which is not installed here: USDA scenes stay disabled exactly as in a reference build without USD
    # create_json_scene is `inline` in its .cpp (serializer_diva.cpp:13) yet called from main_batch.cpp:245
    ("ovr/serializer/serializer_diva.cpp", ["-O0", "-fkeep-inline-functions"]),
    ("ovr/common/imageio.cpp", ["-O1"]),
    ("ovr/common/dylink/Library.cpp", ["-O1"]),
]


def colormap_sources():
    """extern/tfn/colormaps/**/*.cpp: the embedded colour tables scene.cpp's CreateColorMap links against
    (extern/tfn/colormaps/CMakeLists.txt lists the same files)."""
    out = []
    base = os.path.join(REF, "extern", "tfn", "colormaps")
    for d, _, files in os.walk(base):
        for f in sorted(files):
            if f.endswith(".cpp"):
                out.append((os.path.relpath(os.path.join(d, f), REF), ["-O0"]))
    return out


def run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: " + " ".join(cmd) + "\n" + r.stdout[-4000:] + r.stderr[-4000:])


def newer(target, deps):
    return os.path.exists(target) and all(os.path.getmtime(target) >= os.path.getmtime(d) for d in deps if os.path.exists(d))


def build(force=False):
    if not os.path.isdir(os.path.join(REF, "ovr")):
        print(f"[device/build.py] {REF} not present: keeping prebuilt build/ovr products")
        return None
    os.makedirs(OUT, exist_ok=True)
    plugin = os.path.join(OUT, "libdevice_b200.so")
    srcs = [os.path.join(HERE, "device_b200.cpp"), os.path.join(HERE, "device_b200.h"), os.path.join(ROOT, "include", "b200vr.h")]
    if force or not newer(plugin, srcs):
        run(CXX + ["-O2", "-shared", "-o", plugin, srcs[0], f"-L{PKG}", "-lb200vr",
                   "-Wl,-rpath,$ORIGIN:$ORIGIN/../../open-volume-renderer_b200", "-L/usr/local/cuda/lib64", "-lcudart"])
    exe = os.path.join(OUT, "renderbatch")
    hostlib = os.path.join(OUT, "libovr_host.a")
    if force or not os.path.exists(exe) or not os.path.exists(hostlib):
        def compile_one(item):
            rel, flags = item
            obj = os.path.join(OUT, rel.replace("/", "_").replace(".cpp", ".o"))
            run(CXX + flags + ["-c", os.path.join(REF, rel), "-o", obj])
            return obj
        with ThreadPoolExecutor(max_workers=8) as ex:
            objs = list(ex.map(compile_one, HOST_SOURCES + colormap_sources()))
        # -rdynamic: objectFactory also searches the executable itself (Library.cpp:169-191)
        run(["g++", "-rdynamic", "-o", exe] + objs + ["-L/usr/local/cuda/lib64", "-lcudart", "-ldl", "-lpthread",
                                                       "-Wl,-rpath,$ORIGIN:/usr/local/cuda/lib64"])
        # the reference's host objects without main(): other host tools built from the reference link against them
        if os.path.exists(hostlib):
            os.remove(hostlib)
        run(["ar", "rcs", hostlib] + [o for o in objs if "main_batch" not in o])
        for o in objs:
            os.remove(o)
    # renderloop: the GUI's render loop (apps/main_app.cpp:233-263) without a window, against the same host objects
    loop, loop_src = os.path.join(OUT, "renderloop"), os.path.join(HERE, "renderloop.cpp")
    if force or not newer(loop, [loop_src, hostlib]):
        run(CXX + ["-O1", "-rdynamic", "-o", loop, loop_src, "-Wl,--whole-archive", hostlib, "-Wl,--no-whole-archive",
                   "-L/usr/local/cuda/lib64", "-lcudart", "-ldl", "-lpthread", "-Wl,-rpath,$ORIGIN:/usr/local/cuda/lib64"])
    return exe


def build_c1_fixture():
    """BASELINE.json configs[0] literally: a scratch copy of the reference's data/configs directory entry scene_vorts1.json
    (copied at build time into the git-ignored build/ovr/c1/, never into the repository) next to a SYNTHESISED payload
    data/data/vorts1.data -- the reference ships the scene description but not the 8 MiB volume it names. The
    payload is a smooth analytic field of the file's dimensions (128^3 fp32) spanning the scene's scalarMappingRange."""
    src = os.path.join(REF, "data", "configs", "scene_vorts1.json")
    if not os.path.exists(src):
        return None
    import shutil
    import numpy as np
    root = os.path.join(OUT, "c1")
    os.makedirs(os.path.join(root, "data", "configs"), exist_ok=True)
    os.makedirs(os.path.join(root, "data", "data"), exist_ok=True)
    shutil.copyfile(src, os.path.join(root, "data", "configs", "scene_vorts1.json"))
    payload = os.path.join(root, "data", "data", "vorts1.data")
    if not os.path.exists(payload):
        n = 128
        g = np.linspace(0.0, 1.0, n, dtype=np.float64)
        z, y, x = np.meshgrid(g, g, g, indexing="ij")
        f = (np.sin(6.3 * x + 1.0) * np.cos(4.1 * y) + np.sin(5.2 * z + 2.0 * x) * np.cos(3.3 * y * z + 0.5) +
             0.7 * np.cos(9.0 * (x - 0.5) * (y - 0.5)) * np.sin(7.0 * z))
        f = (f - f.min()) / (f.max() - f.min())
        lo, hi = 0.0070086014457046986, 12.139389038085938   # scalarMappingRange of the scene file
        (lo + (hi - lo) * f ** 3).astype("<f4").tofile(payload)   # cubed: most of the volume stays in the transparent foot of the TF
    return root


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(build_c1_fixture())
