"""Hardware tex3D fetches on LARGE 3-D textures (one axis of 256 ... 2048 texels): the coordinate model of DESIGN.md section 4 was
fitted on textures of at most 64 texels per axis. Writes gpurun_out/tex_capture3.npz."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import ovr_b200  # noqa: E402

capi = ovr_b200.capi
S = ovr_b200.scenes
L = capi.lib()
FP = C.POINTER(C.c_float)
rng = np.random.default_rng(31337)
out = {}
k = 0
for dims in ((1024, 6, 5), (6, 1024, 5), (6, 5, 1024), (2048, 4, 4), (512, 4, 6), (256, 256, 4), (1000, 7, 3), (300, 5, 700)):
    vol = rng.uniform(0, 1, (dims[2], dims[1], dims[0])).astype(np.float32)
    sc = S.Scene(S.Volume(vol), S.make_tfn("ramp", value_range=(0.0, 1.0)), S.camera_preset("oblique", dims))
    r = ovr_b200.create_renderer("b200")
    r.set_fbsize((16, 16))
    r.init(sc, sc.camera)
    r.commit()
    pts = rng.uniform(-0.002, 1.002, (30000, 3)).astype(np.float32)
    vals = np.empty(len(pts), np.float32)
    r._ck(L.b200vr_probe(r.handle, pts.ctypes.data_as(FP), len(pts), capi.SAMPLER_TEX3D, vals.ctypes.data_as(FP), None))
    man = np.empty(len(pts), np.float32)
    r._ck(L.b200vr_probe(r.handle, pts.ctypes.data_as(FP), len(pts), capi.SAMPLER_MANUAL, man.ctypes.data_as(FP), None))
    out[f"v{k}_vol"], out[f"v{k}_pts"], out[f"v{k}_hw"] = vol, pts, vals
    print(dims, "hardware == software sampler:", float(np.mean(vals.view(np.uint32) == man.view(np.uint32))), "max abs", float(np.abs(vals - man).max()))
    k += 1
    r.close()
out["n_vol"] = np.array(k)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "tex_capture3.npz"), **out)
