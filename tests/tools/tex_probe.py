"""GPU experiment runner: reads tests/tools/_probe_job.npz (keys <name>_vol [nz,ny,nx] f32 and <name>_pts [n,3] f32, made
locally), samples every job with the hardware texture unit and the software sampler through b200vr_probe and writes
gpurun_out/probe_result.npz (<name>_tex, <name>_man). Used to fit the oracle's texture-unit model to the real unit."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import ovr_b200  # noqa: E402

capi, S = ovr_b200.capi, ovr_b200.scenes
job = np.load(os.path.join(ROOT, "tests", "tools", "_probe_job.npz"))
out = {}
for key in job.files:
    if not key.endswith("_vol"):
        continue
    name = key[:-4]
    vol = np.ascontiguousarray(job[key], np.float32)
    pts = np.ascontiguousarray(job[name + "_pts"], np.float32)
    nz, ny, nx = vol.shape
    sc = S.Scene(S.Volume(vol), S.make_tfn("ramp", value_range=(0.0, 1.0)), S.camera_preset("front", (nx, ny, nz)))
    r = ovr_b200.create_renderer("b200")
    r.set_fbsize((16, 16))
    r.init(sc, sc.camera)
    for tag, smp in (("tex", capi.SAMPLER_TEX3D), ("man", capi.SAMPLER_MANUAL)):
        vals = np.empty(len(pts), np.float32)
        r._ck(capi.lib().b200vr_probe(r.handle, pts.ctypes.data_as(C.POINTER(C.c_float)), len(pts), smp,
                                      vals.ctypes.data_as(C.POINTER(C.c_float)), None))
        out[f"{name}_{tag}"] = vals
    r.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "probe_result.npz"), **out)
print("probed", sorted(out))
