"""Kernel time of the C2 frame when only the tiles of (rank, nranks) are rendered, on ONE GPU: separates load imbalance /
tail effects of the sort-first partition from exchange costs."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ovr_b200  # noqa: E402

capi = ovr_b200.capi
L = capi.lib()
wl = dict(ovr_b200.scenes.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "C2"])
sc = ovr_b200.scenes.noise_scene(wl["n"], tf=wl["tf"], camera=wl["camera"], rate=wl["rate"])
r = ovr_b200.create_renderer("b200")
r.shading = capi.SHADING_REF if wl["shading"] == "ref" else capi.SHADING_NONE
r.set_fbsize((wl["width"], wl["height"]))
r.init(sc, sc.camera)
r.commit()
for nr in (1, 2, 4, 8):
    row = []
    for rk in range(nr):
        L.b200vr_set_tile_partition(r.handle, rk, nr)
        for _ in range(3):
            r.render()
        ms = []
        for _ in range(5):
            r.render()
            ms.append(r.last_stats["ms"])
        st = r.last_stats
        row.append((round(float(np.median(ms)), 3), round((st["primary_samples"] + st["shadow_samples"]) / 1e6, 1)))
    print(nr, row, flush=True)
