#!/usr/bin/env python
"""Fetch-rate probe (b200vr_bench_fetch) on a 1024^3 fp32 volume: coherent 8x4 ray bundles stepping `stride` voxels per fetch through a
footprint of `fp` voxels -- what the texture path delivers when the samples of consecutive steps share (stride < 1) or do not share
(stride = 10: the shadow rays of SHADING_REF) their texels. Prints one JSON line per setting."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ovr_b200  # noqa: E402

capi = ovr_b200.capi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
sc = ovr_b200.scenes.noise_scene(n, tf="sparse", camera="oblique")
r = ovr_b200.create_renderer("b200", 0)
r.set_fbsize((64, 32))
r.init(sc, sc.camera)
r.commit()
L = capi.lib()
for fp, stride, iters, ctas in ((64, 1.0, 4096, 8), (256, 1.0, 2048, 8), (n, 1.0, 2048, 8), (n, 10.0, 2048, 8), (n, 10.0, 2048, 16), (n, 4.0, 2048, 8),
                                (256, 10.0, 2048, 8)):
    d = (C.c_float * 3)(0.35 * stride, 0.25 * stride, 0.9 * stride)
    ms, cnt = C.c_float(), C.c_uint64()
    rc = L.b200vr_bench_fetch(r.handle, capi.SAMPLER_TEX3D, fp, iters, d, ctas, C.byref(ms), C.byref(cnt))
    print(json.dumps({"volume": n, "footprint": fp, "stride_voxels": stride, "ctas_per_sm": ctas, "rc": rc, "ms": ms.value,
                      "gfetch_s": cnt.value / (ms.value * 1e-3) / 1e9 if ms.value > 0 else None}), flush=True)
r.close()
