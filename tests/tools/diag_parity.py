"""GPU diagnostic: for each ray-march parity case print the worst pixels (GPU vs oracle) with their alpha, so that
outliers can be attributed (low-alpha un-premultiplied pixels, texture-unit rounding, flipped sample counts).
Run under gpurun: python tools/diag_parity.py [case-name ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import max_abs, oracle_desc, ovr_b200, psnr, pyoracle  # noqa: E402
import test_gpu_parity as T  # noqa: E402

S = ovr_b200.scenes
capi = ovr_b200.capi
want = set(sys.argv[1:])
for case in T.MARCH_CASES:
    name, n, tf, cam, rate, shading, sampler = case
    if want and name not in want:
        continue
    W, H = 200, 120
    sc = S.noise_scene(n, tf=tf, camera=cam, rate=rate)
    for smp in (capi.SAMPLER_TEX3D, capi.SAMPLER_MANUAL):
        r = T.make_renderer(sc, W, H, shading=shading, sampler=smp, skip=True)
        r.render()
        rgba, grad = r.download()
        d = oracle_desc(sc, W, H, shading=shading, volume=T.device_volume(r, (n, n, n)))
        o_rgba, o_grad, o_cnt = pyoracle.OracleScene(d, skip_empty=True).render_march()
        err = np.abs(rgba.astype(np.float64) - o_rgba).max(axis=-1)
        idx = np.argsort(err.ravel())[::-1][:6]
        print(f"== {name} sampler={'tex' if smp == 0 else 'manual'}: max-abs {max_abs(rgba, o_rgba):.3e} psnr {psnr(rgba, o_rgba):.1f} "
              f"n(err>1/255)={int((err > 1 / 255).sum())} n(err>1/510)={int((err > 1 / 510).sum())}")
        for i in idx:
            y, x = divmod(int(i), W)
            print(f"   px ({x},{y}) err {err[y, x]:.3e} gpu {rgba[y, x]} oracle {o_rgba[y, x]}")
        r.close()
