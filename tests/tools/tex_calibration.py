"""GPU experiment: recover the texture unit's interpolation-weight staircase w(x) and compare it with the oracle's model.
A 2-voxel ramp volume (0, 1) along one axis makes tex3D return the weight itself. Run under gpurun; writes
gpurun_out/tex_calibration.npz + a text summary."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ovr_b200  # noqa: E402

capi = ovr_b200.capi
S = ovr_b200.scenes
out = {}
lines = []


def probe(r, pts, sampler):
    pts = np.ascontiguousarray(pts, np.float32)
    vals = np.empty(len(pts), np.float32)
    r._ck(capi.lib().b200vr_probe(r.handle, pts.ctypes.data_as(C.POINTER(C.c_float)), len(pts), sampler,
                                  vals.ctypes.data_as(C.POINTER(C.c_float)), None))
    return vals


for n in (2, 3, 40, 64, 512):
    for axis in (0, 1, 2):
        dims = [4, 4, 4]
        dims[axis] = n
        nx, ny, nz = dims
        vol = np.zeros((nz, ny, nx), np.float32)
        idx = np.arange(n, dtype=np.float32)
        shape = [1, 1, 1]
        shape[2 - axis] = n
        vol += idx.reshape(shape)
        sc = S.Scene(S.Volume(vol), S.make_tfn("ramp", value_range=(0.0, float(n))), S.camera_preset("front", (nx, ny, nz)))
        r = ovr_b200.create_renderer("b200")
        r.set_fbsize((16, 16))
        r.init(sc, sc.camera)
        # sweep one texel interval [k+0.5, k+1.5]/n finely, at texel k = n//2 - 1
        k = max(n // 2 - 1, 0)
        M = 1 << 20
        t = (np.arange(M, dtype=np.float64) + 0.5) / M
        x = ((k + 0.5 + t) / n).astype(np.float32)
        pts = np.full((M, 3), 0.5, np.float32)
        pts[:, axis] = x
        v = probe(r, pts, capi.SAMPLER_TEX3D)
        w_hw = v - np.float32(k)                       # ramp: value = k + w
        xb = x * np.float32(n) - np.float32(0.5)
        frac = xb - np.floor(xb)
        w_rn = np.floor(frac * np.float32(256) + np.float32(0.5)) / np.float32(256)
        q_hw = np.round(w_hw.astype(np.float64) * 256)
        q_rn = np.round(w_rn.astype(np.float64) * 256)
        mism = float(np.mean(q_hw != q_rn))
        # where does the hardware step? fractional position (in quanta) of each transition
        tr = np.nonzero(np.diff(q_hw) != 0)[0]
        pos = frac[tr + 1].astype(np.float64) * 256
        off = pos - np.round(pos)                       # offset of the step from the half-quantum lattice ... model steps at .5
        step_frac = pos - np.floor(pos)
        levels = np.unique(q_hw)
        lines.append(f"n={n} axis={axis}: levels {len(levels)} (min {levels.min()}, max {levels.max()}), mismatch vs RN {mism:.4f}, "
                     f"step position within quantum: mean {step_frac.mean():.4f} std {step_frac.std():.4f} (RN model: 0.5), "
                     f"non-integer hw weights {float(np.mean(np.abs(w_hw * 256 - q_hw) > 1e-3)):.4f}")
        out[f"n{n}_a{axis}_x"] = x[::64]
        out[f"n{n}_a{axis}_whw"] = w_hw[::64]
        out[f"n{n}_a{axis}_steps"] = step_frac
        r.close()

# random volume + random points: raw material for fitting the unit's lerp arithmetic offline
rng = np.random.default_rng(7)
for tag, dims in (("a", (16, 12, 10)), ("b", (37, 5, 64))):
    nx, ny, nz = dims
    vol = rng.uniform(0.0, 1.0, (nz, ny, nx)).astype(np.float32)
    sc = S.Scene(S.Volume(vol), S.make_tfn("ramp", value_range=(0.0, 1.0)), S.camera_preset("front", dims))
    r = ovr_b200.create_renderer("b200")
    r.set_fbsize((16, 16))
    r.init(sc, sc.camera)
    pts = rng.uniform(0.0, 1.0, (200000, 3)).astype(np.float32)
    out[f"rand_{tag}_vol"] = vol
    out[f"rand_{tag}_pts"] = pts
    out[f"rand_{tag}_tex"] = probe(r, pts, capi.SAMPLER_TEX3D)
    out[f"rand_{tag}_man"] = probe(r, pts, capi.SAMPLER_MANUAL)
    r.close()

os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "tex_calibration.npz"), **out)
open(os.path.join(ROOT, "gpurun_out", "tex_calibration.txt"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
