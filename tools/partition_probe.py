"""Kernel time of a frame when only the tiles of (rank, nranks) are rendered, on ONE GPU: separates load imbalance / tail effects of
the sort-first partition from exchange costs.  python tools/partition_probe.py C3 "kernel=qc6" "kernel=qa6" "kernel=ring" [--ranks 4,8]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ovr_b200  # noqa: E402

capi = ovr_b200.capi
L = capi.lib()
argv = [a for a in sys.argv[1:] if not a.startswith("--ranks")]
ranks = (1, 2, 4, 8)
for a in sys.argv[1:]:
    if a.startswith("--ranks"):
        ranks = tuple(int(v) for v in a.split("=")[1].split(","))
wl = dict(ovr_b200.scenes.WORKLOADS[argv[0] if argv else "C2"])
sets = argv[1:] or ["kernel=auto"]
sc = ovr_b200.scenes.noise_scene(wl["n"], tf=wl["tf"], camera=wl["camera"], rate=wl["rate"])
r = ovr_b200.create_renderer("b200")
r.shading = capi.SHADING_REF if wl["shading"] == "ref" else capi.SHADING_NONE
r.set_fbsize((wl["width"], wl["height"]))
r.init(sc, sc.camera)
r.commit()
for spec in sets:
    for kv in spec.split(","):
        k, v = kv.split("=")
        r.set_option(k, v)
    for nr in ranks:
        row = []
        for rk in range(nr):
            L.b200vr_set_tile_partition(r.handle, rk, nr)
            for _ in range(2):
                r.render()
            ms = []
            for _ in range(3):
                r.render()
                ms.append(r.last_stats["ms"])
            row.append(round(float(np.median(ms)), 3))
        print(json.dumps({"workload": argv[0] if argv else "C2", "set": spec, "nranks": nr, "kernel": r.last_kernel, "ms_by_rank": row, "max": max(row)}), flush=True)
r.close()
