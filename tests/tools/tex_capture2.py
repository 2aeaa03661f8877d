"""Captures hardware tex3D / tex1D fetches on the GPU box for the texture-unit model fit: volumes and tables
with values of one sign and of both signs and with wide exponent spreads, random coordinates (a few out of range). Writes
gpurun_out/tex_capture2.npz; tests/golden/make_golden.py-style fixture material, checked offline against oracle/oracle_sample.cpp."""
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
rng = np.random.default_rng(20261105)
out = {}
vols = {
    "unit": lambda s: rng.uniform(0, 1, s),
    "signed": lambda s: rng.normal(0, 1, s),
    "wide": lambda s: rng.uniform(-1, 1, s) * np.exp2(rng.integers(-12, 6, s)),
    "small": lambda s: rng.uniform(0, 1, s) * 1e-3,
}
k = 0
for name, gen in vols.items():
    for dims in ((16, 12, 10), (37, 5, 64), (64, 64, 64)):
        vol = gen((dims[2], dims[1], dims[0])).astype(np.float32)
        sc = S.Scene(S.Volume(vol), S.make_tfn("ramp", value_range=(0.0, 1.0)), S.camera_preset("oblique", dims))
        r = ovr_b200.create_renderer("b200")
        r.set_fbsize((16, 16))
        r.init(sc, sc.camera)
        r.commit()
        pts = rng.uniform(-0.03, 1.03, (20000, 3)).astype(np.float32)
        vals = np.empty(len(pts), np.float32)
        r._ck(L.b200vr_probe(r.handle, pts.ctypes.data_as(FP), len(pts), capi.SAMPLER_TEX3D, vals.ctypes.data_as(FP), None))
        out[f"v{k}_vol"], out[f"v{k}_pts"], out[f"v{k}_hw"] = vol, pts, vals
        out[f"v{k}_name"] = np.array(name)
        k += 1
        r.close()
out["n_vol"] = np.array(k)
r = ovr_b200.create_renderer("b200")
r.set_fbsize((16, 16))
sc = S.noise_scene(16)
r.init(sc, sc.camera)
r.commit()
k = 0
for name in ("unit", "signed", "wide"):
    for n, ch in ((1024, 1), (1024, 4), (100, 4), (7, 1), (256, 1)):
        table = vols[name]((n, ch)).astype(np.float32)
        u = np.concatenate([rng.uniform(-0.01, 1.01, 30000), np.linspace(0, 1, 4097)]).astype(np.float32)
        hw = np.empty((len(u), ch), np.float32)
        sw = np.empty((len(u), ch), np.float32)
        r._ck(L.b200vr_probe_tex1d(r.handle, table.ctypes.data_as(FP), n, ch, u.ctypes.data_as(FP), len(u), hw.ctypes.data_as(FP), sw.ctypes.data_as(FP)))
        out[f"t{k}_table"], out[f"t{k}_u"], out[f"t{k}_hw"] = table, u, hw
        k += 1
out["n_tab"] = np.array(k)
r.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "tex_capture2.npz"), **out)
print("captured", int(out["n_vol"]), "volumes,", int(out["n_tab"]), "tables")
