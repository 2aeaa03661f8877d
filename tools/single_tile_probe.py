"""Kernel time of every single super tile (64 x 32 pixels = 64 warp tiles) rendered alone on an otherwise idle GPU: the length of the
longest dependent chain in a frame, without any load from other tiles.  python tools/single_tile_probe.py C3 kernel=rq6 [kernel=qc6 ...]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ovr_b200  # noqa: E402

capi = ovr_b200.capi
L = capi.lib()
wl = dict(ovr_b200.scenes.WORKLOADS[sys.argv[1]])
sc = ovr_b200.scenes.noise_scene(wl["n"], tf=wl["tf"], camera=wl["camera"], rate=wl["rate"])
r = ovr_b200.create_renderer("b200")
r.shading = capi.SHADING_REF
r.set_fbsize((wl["width"], wl["height"]))
r.init(sc, sc.camera)
r.commit()
n_super = ((wl["width"] + 63) // 64) * ((wl["height"] + 31) // 32)
for spec in sys.argv[2:]:
    for kv in spec.split(","):
        k, v = kv.split("=")
        r.set_option(k, v)
    ms = []
    for rk in range(n_super):
        L.b200vr_set_tile_partition(r.handle, rk, n_super)
        r.render()
        r.render()
        ms.append(r.last_stats["ms"])
    ms = np.array(ms)
    top = np.argsort(ms)[::-1][:5]
    print(json.dumps({"workload": sys.argv[1], "set": spec, "kernel": r.last_kernel, "super_tiles": n_super, "max_ms": float(ms.max()), "median_ms": float(np.median(ms)),
                      "sum_ms": float(ms.sum()), "top5": [(int(i), round(float(ms[i]), 3)) for i in top]}), flush=True)
r.close()
