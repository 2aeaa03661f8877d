#!/usr/bin/env python
"""A/B runs of one workload on one GPU: the same frame rendered under a list of tuning-option sets (b200vr_set_option), each timed
with CUDA events over a few frames after warm-up and compared with the first set's frame.

  python tools/ab_run.py --workload C3 --frames 5 --sets "kernel=q" "kernel=qb4,tile_order=7" ...

Prints one JSON line per set: ms per frame (median / min of the kernel's own event time), executed samples, fetches, the
fraction of the measured fetch roof, max-abs difference of the frame against the first set's frame. No oracle here (tests/ and
bench.py check parity); this is the knob-turning tool behind the numbers in DESIGN.md.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

DEFAULTS = {"kernel": "auto", "tile_order": -1, "sheet_px": 8, "max_ctas_per_sm": 0, "occ_smem": -1, "carveout": -1, "ring_high": 4, "rq_self": 1, "rq_cap": 128}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--frames", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--sets", nargs="+", default=["kernel=q"])
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import ovr_b200
    import bench
    capi = ovr_b200.capi
    wl, scene = bench.workload_scene(args.workload, 0)
    r = ovr_b200.create_renderer("b200", 0)
    r.shading = capi.SHADING_REF if wl["shading"] == "ref" else capi.SHADING_NONE
    r.set_fbsize((wl["width"], wl["height"]))
    r.set_path_tracing(wl["mode"] == "pathtrace")
    r.set_frame_accumulation(wl["mode"] == "pathtrace")
    r.init(scene, scene.camera)
    r.commit()
    p_fetch, _ = bench.measure_fetch_peak(r, capi, r.sampler)
    base = None
    lines = []
    for spec in args.sets:
        opts = dict(DEFAULTS)
        for kv in spec.split(","):
            if kv:
                k, v = kv.split("=")
                opts[k] = v
        sampler = opts.pop("sampler", "tex3d")   # not a tuning option: the sampler of b200vr_params (tex3d | manual)
        r.sampler = capi.SAMPLER_MANUAL if sampler == "manual" else capi.SAMPLER_TEX3D
        r.apply_device_options()
        for k, v in opts.items():
            r.set_option(k, v)
        ms = []
        st = None
        for i in range(args.warmup + args.frames):
            r.render()
            st = dict(r.last_stats)
            if i >= args.warmup:
                ms.append(st["ms"])
        rgba = r.download(want_grad=False)[0]
        if base is None:
            base = rgba.copy()
        samples = st["primary_samples"] + st["shadow_samples"] + st["delta_samples"]
        fetches = samples + st["gradient_fetches"]
        med = float(np.median(ms))
        line = {"set": spec, "ms_median": med, "ms_min": float(min(ms)), "samples": samples, "fetches": fetches,
                "primary": st["primary_samples"], "shadow": st["shadow_samples"], "shaded": st["gradient_fetches"] // 3,
                "msamples_s": samples / med / 1e3, "frac_fetch_roof": fetches / (med * 1e-3) / p_fetch if p_fetch else None,
                "max_abs_vs_first": float(np.abs(rgba - base).max()), "launches": st["kernel_launches"], "kernel_name": r.last_kernel}
        print(json.dumps(line), flush=True)
        lines.append(line)
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"workload": args.workload, "p_fetch_gfetch_s": p_fetch / 1e9, "runs": lines}, f, indent=1)
    r.close()


if __name__ == "__main__":
    main()
