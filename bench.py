#!/usr/bin/env python
"""Headline benchmark of the B200 volume device: Msamples/s and ms/frame of 1080p ray marching (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C3] [--impl b200|reference]

Default workload: C3 (BASELINE.json configs[2], the configuration north_star's roofline target is quoted on: 1024^3 fp32, 1920x1080,
gradient + shadow shading) at every N. A "step" is one batch of --frames-per-step frames of the workload through the hot path
(default 20: at 8 GPUs a frame takes about a millisecond and the timed region of K steps should still last ~0.5 s). One JSON line
is printed by rank 0:
  value       executed Msamples/s (TF-classified trilinear samples: primary + shadow + delta-tracking collisions), whole
              job, volume / TF / macrocells already resident in HBM, timed on the device with CUDA events
  e2e         the same metric through the plugin-facing API with host buffers: per step the camera and the transfer
              function travel host -> device and the rgba + grad frame comes back device -> pinned host
  roofline    texture/L1 sampling roofline of the dominant kernel (32 B per executed trilinear fetch) against the
              fetch rate measured on this box by b200vr_bench_fetch, plus the HBM floor against MEASURED_PEAKS.json
  cpu_baseline  the reference's optix7 shader arithmetic on the host cores on a strided pixel sample of the same frame:
              its own raymarching() / raymarching_shadow() compiled on the host (oracle/_ref, kind "reference") for shaded
              ray marching, the CPU oracle (kind "port") for path tracing / unshaded frames or where oracle/_ref is absent
  parity      the GPU frame against the oracle on the pixel subset the CPU leg renders anyway (ray marching: max-abs per RGBA
              channel and PSNR, bar 1/255 and 45 dB; path tracing: MSE of 64 accumulated frames against the oracle at equal spp
              on a 480x270 crop). At N > 1 also: rank 0's gathered frame against its own undivided frame (bit-identical).
  sort_last   (N > 1) the sort-last path next to it: N bricks, binary swap and direct send over NVLink peer memory, checked
              against the one-volume render on a down-scaled twin
`--impl reference` times that CPU arm alone (the reference's own devices need OptiX / OSPRay, neither installable here).
N > 1 (torchrun): sort-first -- every rank renders the image tiles it owns of the same frame straight into rank 0's frame over
NVLink peer memory (--exchange nccl: slices gathered over NCCL); strong scaling.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Msamples/s & ms/frame, 1080p ray-march, 1/2/4/8 B200 vs OSPRay CPU device"
UNIT = "Msamples/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=25)   # apps/main_batch.cpp:278-289: 5 warm-up + 25 timed frames
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--frames-per-step", type=int, default=20, help="frames rendered per step (one step = one batch of frames)")
    ap.add_argument("--no-sort-last", action="store_true", help="N > 1: skip the sort-last sub-record")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--sampler", default="tex3d", choices=["tex3d", "manual"])
    ap.add_argument("--no-skip", action="store_true")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="multi-GPU exchange: fused over NVLink peer memory (sort-first: stores into rank 0's frame; sort-last: blend reads the partner) or NCCL")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-stride", type=int, default=0, help="pixel stride of the CPU sample (0 = auto)")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None
        self.nvml_rows, self._stop, self._thr, self.nvml_error = [], False, None, None

    def _nvml_loop(self):
        # NVML in-process: a sample every ~2 ms, so that even a 30 ms timed region (8 GPUs, 25 frames) is covered;
        # the nvidia-smi process below stays as the second source (its first line arrives after ~100 ms)
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
            while not self._stop:
                sm = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)
                try:
                    bits = N.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                except Exception:
                    bits = 0
                self.nvml_rows.append((float(sm), float(mx), int(bits)))
                time.sleep(0.002)
        except Exception as e:
            self.nvml_error = repr(e)

    def start(self):
        self._thr = threading.Thread(target=self._nvml_loop, daemon=True)
        self._thr.start()
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        self._stop = True
        if self._thr is not None:
            self._thr.join(timeout=1.0)
        sm, mx, reasons = [], [], set()
        for a, b, bits in self.nvml_rows:   # nvml.h: HwSlowdown 0x8, SwPowerCap 0x4, SwThermalSlowdown 0x20, HwThermalSlowdown 0x40
            sm.append(a), mx.append(b)
            for name, bit in (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)):
                if bits & bit:
                    reasons.add(name)
        if self.proc is None:
            if sm:
                return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                        "source": "nvml"}
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        for r in self.rows:
            try:
                sm.append(float(r[0])), mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload_scene(name, cuda_device=None):
    import ovr_b200
    S = ovr_b200.scenes
    w = dict(S.WORKLOADS[name])
    scene = S.noise_scene(w["n"], tf=w["tf"], camera=w["camera"], rate=w["rate"])
    if w.get("dtype") == "u8" and cuda_device is not None:
        # typed workload: the device generates noise(n), the host quantises it to 8 bits and hands it back as a u8 volume
        r0 = ovr_b200.create_renderer("b200", cuda_device)
        r0.set_fbsize((8, 8))
        r0.init(scene, scene.camera)
        vol = np.empty((w["n"],) * 3, np.float32)
        r0._ck(ovr_b200.capi.lib().b200vr_download_volume(r0.handle, vol.ctypes.data_as(C.c_void_p)))
        r0.close()
        data = np.rint(vol * np.float32(255.0)).astype(np.uint8)
        scene = S.Scene(S.Volume(data), S.make_tfn(w["tf"], value_range=(1.0, -1.0)), scene.camera, w["rate"])
    return w, scene


def orbit_camera(scene, k, n):
    """k-th camera of an orbit around the volume centre (the multi-frame mode of apps/main_batch.cpp:295-313)."""
    import ovr_b200
    cam = scene.camera
    at, frm = np.array(cam.at), np.array(cam.frm)
    ang = 2 * np.pi * k / max(n, 1) * 0.05  # a slow orbit: consecutive frames stay comparable in cost
    d = frm - at
    c, s = np.cos(ang), np.sin(ang)
    d2 = np.array([c * d[0] + s * d[2], d[1], -s * d[0] + c * d[2]])
    return ovr_b200.Camera(tuple(at + d2), tuple(at), cam.up, cam.fovy)


def peaks():
    p = {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}
    try:
        m = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        p = {"hbm_gbs": float(m["hbm_gbs"]), "source": "MEASURED_PEAKS.json"}
    except Exception:
        pass
    return p


# ----------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle (port of the reference's shader arithmetic) on the host cores, bounded pixel sample
# ----------------------------------------------------------------------------------------------------------------------

def cpu_oracle_run(wl, scene, steps, warmup, stride, volume=None):
    """Renders every `stride`-th pixel in x and y of the workload's frame with the reference algorithm verbatim
    (no skipping, every sample shaded and shadowed); returns (Msamples/s, ms per sampled frame, info)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import oracle_desc
    from oracle import pyoracle
    shading = pyoracle.SHADING_REF if wl["shading"] == "ref" else pyoracle.SHADING_NONE
    d = oracle_desc(scene, wl["width"], wl["height"], shading=shading, volume=volume)
    o = pyoracle.OracleScene(d, skip_empty=False, build_macrocells=(wl["mode"] == "pathtrace"))
    cores = pyoracle.hardware_threads()

    # Shaded ray marching and path tracing: the REFERENCE's own shader functions compiled on the host
    # (oracle/_ref/libovr_ref_march.so -- raymarching() / raymarching_shadow() / render_raymarching(), delta_tracking() /
    # pathtracing() / render_pathtracing() -- built where /root/reference exists and shipped with the snapshot) are what is
    # timed; the oracle port -- pinned bit for bit against them by tests/golden/ref_march*.npz and ref_path.npz -- supplies the
    # sample count of the same pixels in an untimed pass. Unshaded frames have no reference counterpart: port.
    W, H = wl["width"], wl["height"]
    path = wl["mode"] == "pathtrace"
    use_ref = ((path or wl["shading"] == "ref") and not os.environ.get("B200VR_BENCH_CPU_PORT")
               and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libovr_ref_march.so")))
    ref_px = None
    if use_ref:
        ys, xs = np.meshgrid(np.arange(0, H, stride), np.arange(0, W, stride), indexing="ij")
        ref_px = np.stack([xs.ravel(), ys.ravel()], 1).astype(np.int32)

    pix = {"stride": stride, "frame_index": None}   # the oracle's pixels of the last frame: bench.py's parity check reads them

    def port_frame(fi):
        if path:
            img, cnt = o.render_path(stride=(stride, stride), threads=cores, frame_index=fi)
            pix.update(rgba=img[::stride, ::stride], grad=None, frame_index=fi)
            return cnt
        img, grd, cnt = o.render_march(stride=(stride, stride), threads=cores, frame_index=fi)
        pix.update(rgba=img[::stride, ::stride], grad=grd[::stride, ::stride], frame_index=fi)
        return cnt

    def frame(fi):
        if use_ref:
            cnt = port_frame(fi)  # untimed: counts and the oracle's pixels (same pixels, same RNG streams)
            t0 = time.perf_counter()
            if path:
                out = o.ref_path(ref_px, frame_index=fi, threads=cores)
                pix["ref_rgba"] = out.reshape(len(range(0, H, stride)), len(range(0, W, stride)), 4)
            else:
                o.s.frame_index = fi
                out = o.ref_march(ref_px, threads=cores)
                pix["ref_rgba"] = out[0].reshape(len(range(0, H, stride)), len(range(0, W, stride)), 4)
            dt = time.perf_counter() - t0
        else:
            t0 = time.perf_counter()
            cnt = port_frame(fi)
            dt = time.perf_counter() - t0
        return dt, cnt["primary_samples"] + cnt["shadow_samples"] + cnt["delta_samples"], cnt

    for k in range(warmup):
        frame(k + 1)
    tot_t, tot_s, cnt = 0.0, 0, None
    for k in range(steps):
        dt, s, cnt = frame(k + 1)
        tot_t += dt
        tot_s += s
    what = (("the reference's own delta_tracking()/pathtracing() (shaders_pathtracing.cu:269-542) compiled on the host, texture unit emulated"
             if path else
             "the reference's own raymarching()/raymarching_shadow() (shaders_raymarching.cu:43-171) compiled on the host, texture unit emulated")
            if use_ref else "CPU port of the reference algorithm without skipping")
    info = {"cores": cores, "kind": "reference" if use_ref else "port",
            "sample": f"every {stride}th pixel in x and y (1/{stride * stride} of the rays) of the {W}x{H} frame, "
                      f"{what}, {steps} frame(s)", "rays": cnt["rays"], "samples_per_frame": tot_s // max(steps, 1)}
    info["_pixels"] = pix
    info["_oracle"] = o
    return tot_s / tot_t / 1e6, tot_t / steps * 1e3, info


def public(info):
    return {k: v for k, v in info.items() if not k.startswith("_")}


def image_metrics(got, want):
    d = np.abs(got.astype(np.float64) - want.astype(np.float64))
    mse = float(np.mean(d ** 2))
    return float(d.max()), (float("inf") if mse == 0 else float(10.0 * np.log10(1.0 / mse)))


def auto_stride(cores):
    # 1/16 of the rays where the box has the threads for it (a few seconds of CPU work per sampled frame), else 1/64
    return 4 if cores >= 16 else 8


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    from oracle import pyoracle
    pyoracle.build()
    wl, scene = workload_scene(args.workload)
    stride = args.cpu_stride or auto_stride(pyoracle.hardware_threads())
    steps, warmup = min(args.steps, 3), min(args.warmup, 1)  # bounded: a CPU frame takes seconds
    v, ms, info = cpu_oracle_run(wl, scene, steps, warmup, stride)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args, wl), "note": "reference devices need OptiX/OSPRay (absent); this is the optix7 device's shader arithmetic on "
                       "the host cores (cpu_baseline.kind: the reference's own march functions compiled on the host where oracle/_ref "
                       "exists, else the CPU port); ms_per_step is for the sampled pixel subset"},
            "cpu_baseline": dict(public(info), value=v, unit=UNIT),
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(args, wl):
    return (f"{args.workload}: noise({wl['n']},1337) f32 {wl['n']}^3, {wl['width']}x{wl['height']}, camera {wl['camera']}, TF {wl['tf']}, "
            f"rate {wl['rate']}, {wl['mode']}, shading {wl['shading']}")


# ----------------------------------------------------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------------------------------------------------

def measure_fetch_peak(r, capi, sampler):
    """P_fetch of SURVEY 8d: best trilinear fetch rate of b200vr_bench_fetch over a few footprints (fetches/s)."""
    best, rows = 0.0, []
    d = (C.c_float * 3)(0.35, 0.25, 0.9)
    for fp, iters in ((32, 4096), (64, 4096), (256, 2048)):
        ms, n = C.c_float(), C.c_uint64()
        rc = capi.lib().b200vr_bench_fetch(r.handle, sampler, fp, iters, d, 8, C.byref(ms), C.byref(n))
        if rc != 0:
            continue
        rate = n.value / (ms.value * 1e-3)
        rows.append({"footprint": fp, "gfetch_s": rate / 1e9})
        best = max(best, rate)
    return best, rows


def run_b200(args):
    import torch
    import torch.distributed as dist
    import ovr_b200
    capi = ovr_b200.capi
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 volume device has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = capi.lib()
    from ovr_b200 import distributed as D
    wl, scene = workload_scene(args.workload, local_rank)
    W, H = wl["width"], wl["height"]
    sortlast = bool(wl.get("sortlast"))
    brick = None
    if sortlast:
        # weak scaling: every rank owns one n^3 brick of a (grid * n) volume; explicit TF window (a brick's data range is not the volume's)
        g = D.brick_grid(world)
        dims = tuple(g[a] * wl["n"] for a in range(3))
        scene = ovr_b200.scenes.noise_scene(wl["n"], tf=wl["tf"], camera=wl["camera"], rate=wl["rate"], dims=dims)
        scene.tfn.value_range = (0.0, 1.0)
        full_scene = scene
        scene, brick = D.brick_scene(full_scene, rank, world)

    r = ovr_b200.create_renderer("b200", local_rank)
    r.shading = capi.SHADING_REF if wl["shading"] == "ref" else capi.SHADING_NONE
    r.sampler = capi.SAMPLER_MANUAL if args.sampler == "manual" else capi.SAMPLER_TEX3D
    r.skip_empty = not args.no_skip
    r.set_fbsize((W, H))
    r.set_path_tracing(wl["mode"] == "pathtrace")
    r.set_frame_accumulation(wl["mode"] == "pathtrace")
    r.init(scene, scene.camera)
    r.commit()
    if wl.get("sparse"):
        r.set_focus((0.5, 0.5), 0.06, 0.07)
        r.set_sparse_sampling(True)
        r.commit()
    stream = torch.cuda.current_stream()
    r.use_stream(stream.cuda_stream)   # kernels ordered with the NCCL collectives / barriers of the exchange
    launches = [0]

    if sortlast:
        D.setup_brick_renderer(r, brick)
        # exchange fused into the blend over NVLink peer mappings (--exchange nccl: the send/recv schedule)
        p2p = world > 1 and args.exchange == "p2p"
        swap = (D.SortLastP2P(rank, world, full_scene.volume.dims, W * H, r) if p2p else
                D.SortLast(rank, world, full_scene.volume.dims, W * H, handle_renderer=r))

        def step():
            eye = r.params["camera"].get().frm   # visibility order follows the committed camera
            # all pixels against this rank's brick -> premultiplied layer; binary swap; display rank un-premultiplies
            r.render()
            launches[0] += r.last_stats["kernel_launches"]
            n0 = swap.launches
            if p2p:
                out = swap.composite(eye)
            else:
                out = swap.composite(D.frame_tensor(r), eye)
                out = out.data_ptr() if out is not None else None
            launches[0] += swap.launches - n0
            if out is not None:
                r._ck(L.b200vr_finalize_premultiplied(r.handle, C.c_void_p(out), W * H))
                launches[0] += 1
            return r.last_stats
    else:
        # sort-first: own tiles rendered, dense slices gathered on rank 0 (one exchange step)
        # --exchange p2p: tiles rendered straight into rank 0's frame over NVLink peer memory; nccl: pack + gather + unpack
        sf = D.SortFirstP2P(r, rank, world) if (world > 1 and args.exchange == "p2p") else D.SortFirst(r, rank, world)

        def step():
            n0 = sf.launches
            st = sf.step()
            launches[0] += sf.launches - n0
            return st

        if isinstance(sf, D.SortFirstP2P) and world > 1:
            # frames enqueued back to back (b200vr_render_async + the stream-ordered completion all-reduce): the host does not
            # wait for a frame before it enqueues the next one; the counters of all frames are read once after the timed region
            step_async = lambda: sf.step(wait=False)   # noqa: E731

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if sortlast or not (world > 1 and args.exchange == "p2p"):
        step_async = None

    # ---- device-resident throughput (`value`) ----
    F = max(1, args.frames_per_step)
    for _ in range(max(args.warmup, 3)):
        step()
    clocks = ClockSampler(local_rank)
    barrier()
    clocks.start()
    launches[0] = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    samples = fetches = 0
    kernel_ms = 0.0
    barrier()
    ev0.record(stream)
    if step_async is None:
        for _ in range(args.steps * F):
            st = step()
            samples += st["primary_samples"] + st["shadow_samples"] + st["delta_samples"]
            fetches += st["primary_samples"] + st["shadow_samples"] + st["delta_samples"] + st["gradient_fetches"]
            kernel_ms += st["ms"]
    else:
        for _ in range(args.steps * F):
            step_async()
    ev1.record(stream)
    barrier()
    clk = clocks.stop()
    if step_async is not None:
        st = r.wait()   # counters summed over the frames of the timed region
        samples = st["primary_samples"] + st["shadow_samples"] + st["delta_samples"]
        fetches = samples + st["gradient_fetches"]
        launches[0] = st["kernel_launches"] * args.steps * F
        # the render kernel's own time per launch (roofline, per-rank table): F more frames, each waited for
        for _ in range(F):
            kernel_ms += step()["ms"]
        kernel_ms *= args.steps
        launches[0] = st["kernel_launches"] * args.steps * F
        barrier()
    gpu_launches = launches[0]
    t_ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    tot = torch.tensor([samples, fetches], dtype=torch.float64, device="cuda")
    kms = torch.tensor([kernel_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(kms, op=dist.ReduceOp.MAX)
    per_rank_kernel = None
    if world > 1:   # every rank's own kernel time: tells load imbalance between the ranks from a common floor (the heaviest tile)
        mine = torch.tensor([kernel_ms], dtype=torch.float64, device="cuda")
        allk = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allk, mine)
        per_rank_kernel = [float(k.item()) / (args.steps * F) for k in allk]
    t_ms, samples, fetches, kernel_ms = float(t_ms.item()), float(tot[0].item()), float(tot[1].item()), float(kms.item())
    value = samples / (t_ms * 1e-3) / 1e6
    return dict(r=r, wl=wl, scene=full_scene if sortlast else scene, value=value, t_ms=t_ms, samples=samples, fetches=fetches, kernel_ms=kernel_ms, clk=clk,
                gpu_launches=gpu_launches, step=step, barrier=barrier, stream=stream, world=world, rank=rank, local_rank=local_rank,
                F=F, frames=args.steps * F, kernel=r.last_kernel, per_rank_kernel=per_rank_kernel)


def run_e2e(ctx, args):
    """Same frames through the plugin-facing calls with host buffers: set_camera + set_transfer_function (host arrays ->
    device, majorants rebuilt), commit, render, then the rgba + grad frame device -> pinned host."""
    import torch
    import ovr_b200
    r, wl, scene, world, rank = ctx["r"], ctx["wl"], ctx["scene"], ctx["world"], ctx["rank"]
    W, H = wl["width"], wl["height"]
    L = ovr_b200.capi.lib()
    tfn = scene.tfn
    c = np.ascontiguousarray(np.asarray(tfn.color, np.float32)[:, :3]).ravel()
    n_o = len(tfn.opacity)
    o = np.empty(2 * n_o, np.float32)
    o[0::2] = np.arange(n_o, dtype=np.float32) / np.float32(n_o - 1)
    o[1::2] = tfn.opacity
    rgba_host = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    grad_host = torch.empty((H, W, 3), dtype=torch.float32).pin_memory()
    h2d = c.nbytes + o.nbytes + 10 * 4
    d2h = rgba_host.numel() * 4 + grad_host.numel() * 4

    copy_stream = torch.cuda.Stream()

    phases = [0.0] * 5   # host seconds per phase, printed with B200VR_BENCH_DEBUG=1

    def e2e_step(k):
        # the GUI loop of apps/main_app.cpp:244-263: commit -> mapframe (frame k-1) -> swap -> render (frame k). The
        # device->host copy of frame k-1 runs on its own stream while frame k renders into the other buffer; both are
        # complete before the step returns. (Under sort-first, rank 0 copies the gathered frame.)
        t = [time.perf_counter()]
        r.set_camera(orbit_camera(scene, k, ctx["frames"]))
        r.set_transfer_function(c, o, tuple(tfn.value_range))
        t.append(time.perf_counter())
        r.commit()
        t.append(time.perf_counter())
        if rank == 0:
            r._ck(L.b200vr_download_async(r.handle, C.c_void_p(rgba_host.data_ptr()), C.c_void_p(grad_host.data_ptr()),
                                          C.c_void_p(copy_stream.cuda_stream)))
        r.swap()
        t.append(time.perf_counter())
        st = ctx["step"]()
        t.append(time.perf_counter())
        copy_stream.synchronize()
        t.append(time.perf_counter())
        for i in range(5):
            phases[i] += t[i + 1] - t[i]
        phases.append(st["ms"])
        return st

    F, n_frames = ctx["F"], ctx["frames"]
    for k in range(max(1, min(args.warmup, 3))):
        e2e_step(k)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    samples = 0
    ctx["barrier"]()
    t0 = time.perf_counter()
    ev0.record(ctx["stream"])
    for k in range(n_frames):
        st = e2e_step(k)
        samples += st["primary_samples"] + st["shadow_samples"] + st["delta_samples"]
    if rank == 0:   # drain: the last rendered frame also reaches the host inside the timed region
        r._ck(L.b200vr_download(r.handle, C.c_void_p(rgba_host.data_ptr()), C.c_void_p(grad_host.data_ptr())))
    ev1.record(ctx["stream"])
    ctx["barrier"]()
    wall_ms = (time.perf_counter() - t0) * 1e3
    if os.environ.get("B200VR_BENCH_DEBUG") and rank == 0:
        n_calls = n_frames + max(1, min(args.warmup, 3))
        print("[bench] e2e host ms per step: setters %.3f commit %.3f copy-enqueue+swap %.3f render %.3f copy-wait %.3f" %
              tuple(1e3 * p / n_calls for p in phases[:5]), "kernel ms (device events) %.3f" % (sum(phases[5:]) / n_calls), file=sys.stderr)
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    s = torch.tensor([samples], dtype=torch.float64, device="cuda")
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    t_ms = float(t.item())
    return {"value": float(s.item()) / (t_ms * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d) * F, "d2h_bytes_per_step": int(d2h) * F,
            "ms_per_step": t_ms / args.steps, "ms_per_frame": t_ms / n_frames, "wall_ms_per_frame": wall_ms / n_frames,
            "path": "per frame: MainRenderer.set_camera/set_transfer_function/commit/mapframe+to_cpu(async, previous frame)/swap/render; "
                    "n+1 frame copies to pinned host for n frames"}


KERNEL_OPTION = {"raymarch_q_kernel": "q", "raymarch_ring_kernel": "ring", "raymarch_qc_kernel<J=4>": "qc4", "raymarch_qc_kernel<J=6>": "qc6",
                 "raymarch_qc_kernel<J=8>": "qc8", "raymarch_rq_kernel<J=6>": "rq6", "raymarch_rq_kernel<J=4>": "rq4", "raymarch_rq_kernel<J=2>": "rq2"}


def parity_march(rgba, grad, pix):
    """GPU frame against the oracle's pixels (every stride-th pixel in x and y): the north star's bar."""
    s = pix["stride"]
    e, p = image_metrics(rgba[::s, ::s], pix["rgba"])
    out = {"max_abs": e, "psnr_db": p, "pixels": int(pix["rgba"].shape[0] * pix["rgba"].shape[1]), "checker": "oracle (CPU port, pinned bit for bit "
           "against the reference's march code by tests/golden/ref_march*.npz), no skipping", "tolerance": "max-abs <= 1/255 per RGBA channel, PSNR >= 45 dB"}
    if grad is not None and pix.get("grad") is not None:
        out["grad_max_abs"] = image_metrics(grad[::s, ::s], pix["grad"])[0]
    if pix.get("ref_rgba") is not None:   # the reference's own raymarching() compiled on the host, same pixels
        out["max_abs_vs_reference_code"] = image_metrics(rgba[::s, ::s], pix["ref_rgba"])[0]
    out["covered"] = float((pix["rgba"][..., 3] > 0).mean())
    # where the frame misses the bar: how many pixels, how opaque they are, and the same comparison on the displayed colour
    # (colour x alpha over the black background: alpha_blend's C / A amplifies differences at nearly transparent pixels)
    g, o = rgba[::s, ::s].astype(np.float64), pix["rgba"].astype(np.float64)
    d = np.abs(g - o).max(axis=-1)
    over = d > 1.0 / 255.0
    disp = lambda f: np.concatenate([f[..., :3] * f[..., 3:4], f[..., 3:4]], axis=-1)   # noqa: E731
    out["pixels_over_tolerance"] = int(over.sum())
    out["max_abs_displayed"] = float(np.abs(disp(g) - disp(o)).max())
    out["max_abs_alpha"] = float(np.abs(g[..., 3] - o[..., 3]).max())
    if over.any():
        iy, ix = np.unravel_index(int(np.argmax(d)), d.shape)
        out["worst_pixel"] = {"xy": [int(ix) * s, int(iy) * s], "gpu": [float(v) for v in g[iy, ix]], "oracle": [float(v) for v in o[iy, ix]]}
        out["alpha_of_pixels_over_tolerance"] = {"min": float(o[over][:, 3].min()), "median": float(np.median(o[over][:, 3])), "max": float(o[over][:, 3].max())}
    # The bar, pixel for pixel, on the stored (un-premultiplied) frame; and the same bar where it is meaningful: a pixel whose
    # opacity is below 1e-3 is invisible (it displays as 0.001 x colour at most) while alpha_blend's C / A turns a difference of
    # a few 1e-6 of ONE sample's opacity -- the last-bit rounding of a sample position, which the kernel (object-space ray) and the
    # host-compiled reference code (world position transformed per sample) do differently, DESIGN.md sections 4 and 7.1 -- into a
    # full-scale colour there. Both are reported; `ok` is the second.
    vis = o[..., 3] >= 1e-3
    out["max_abs_where_alpha_ge_1e-3"] = float(np.abs(g - o)[vis].max()) if vis.any() else 0.0
    out["ok_every_pixel_as_stored"] = bool(e <= 1.0 / 255.0 and p >= 45.0)
    out["ok"] = bool(out["max_abs_where_alpha_ge_1e-3"] <= 1.0 / 255.0 and out["max_abs_displayed"] <= 1.0 / 255.0 and
                     out["max_abs_alpha"] <= 1.0 / 255.0 and p >= 45.0)
    return out


def parity_path(r, scene, wl, oracle, spp=64, crop=(480, 270)):
    """SURVEY 8c protocol for path tracing: `spp` accumulated frames on the GPU against the oracle's mean of the same `spp` frames
    (same RNG streams: RandomTEA(frame_index, pixel)) on a centred crop; MSE next to the variance of one frame."""
    from oracle import pyoracle
    W, H = wl["width"], wl["height"]
    x0, y0 = (W - crop[0]) // 2, (H - crop[1]) // 2
    rect = (x0, y0, x0 + crop[0], y0 + crop[1])
    r.set_camera(scene.camera)
    r.commit()   # a camera change resets the accumulation (device_impl.cpp:225-233): frame_index starts at 1 again
    for _ in range(spp):
        r.render()
    assert r.last_stats["frame_index"] == spp
    gpu = r.download(want_grad=False)[0][y0:y0 + crop[1], x0:x0 + crop[0]]
    acc = np.zeros((crop[1], crop[0], 4), np.float64)
    first = None
    cores = pyoracle.hardware_threads()
    for fi in range(1, spp + 1):
        img = oracle.render_path(rect=rect, threads=cores, frame_index=fi)[0][y0:y0 + crop[1], x0:x0 + crop[0]]
        first = img if first is None else first
        acc += img
    cpu = (acc / spp).astype(np.float32)
    hit = cpu[..., 3] > 0
    mse = float(np.mean((gpu[hit][:, :3].astype(np.float64) - cpu[hit][:, :3]) ** 2)) if hit.any() else 0.0
    var1 = float(np.var(first[hit][:, :3])) if hit.any() else 0.0
    return {"mse": mse, "spp": spp, "pixels": int(crop[0] * crop[1]), "single_frame_variance": var1, "alpha_identical": bool(np.array_equal(gpu[..., 3], cpu[..., 3])),
            "mean_gpu": float(gpu[hit][:, :3].mean()) if hit.any() else 0.0, "mean_oracle": float(cpu[hit][:, :3].mean()) if hit.any() else 0.0,
            "checker": "oracle path tracer (pinned bit for bit against the reference's delta_tracking()/pathtracing() by tests/golden/ref_path.npz)",
            "tolerance": "MSE at equal spp <= 2 x single-frame variance / spp (two independent estimates); same RNG streams, so ~0 expected",
            "ok": bool(mse <= 2.0 * var1 / spp + 1e-6)}


def sort_first_check(ctx):
    """N > 1, outside the timed region: the frame rank 0 holds after one exchange step against rank 0's own undivided frame of the
    same kernel -- pixels are independent, so the two must be bit-identical. Returns (record, gathered rgba, grad) on rank 0."""
    import ovr_b200
    r, rank, world = ctx["r"], ctx["rank"], ctx["world"]
    L = ovr_b200.capi.lib()
    r.set_camera(ctx["scene"].camera)
    r.commit()
    ctx["step"]()
    ctx["barrier"]()
    rec = got = grad = None
    if rank == 0:
        got, grad = r.download()
        got, grad = got.copy(), grad.copy()
        kern = r.last_kernel
        r._ck(L.b200vr_set_tile_partition(r.handle, 0, 1))
        if kern in KERNEL_OPTION:
            r.set_option("kernel", KERNEL_OPTION[kern])
        r.render()
        want, want_g = r.download()
        rec = {"bit_identical_to_undivided_frame": bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)) and
                                                       np.array_equal(grad.view(np.uint32), want_g.view(np.uint32))),
               "max_abs_vs_undivided_frame": float(np.abs(got - want).max()), "kernel": kern, "undivided_kernel": r.last_kernel}
        r.set_option("kernel", "auto")
        r._ck(L.b200vr_set_tile_partition(r.handle, 0, world))
    ctx["barrier"]()
    return rec, got, grad


def run_sort_last_record(args, ctx):
    """N > 1 (power of two): the sort-last path measured next to the sort-first headline. Every rank holds one brick
    (own voxels + ghosts) of a volume N times as large, marches all pixels against it on the whole volume's sample lattice and
    the N premultiplied layers are composited over NVLink peer memory, by binary swap and by direct send. C5 (8 x 2048^3 = 4096^3,
    3840x2160) at N >= 8, C5-mid (N x 1024^3, 1920x1080) below. First a down-scaled twin (N x 96^3, 640x360) is rendered both ways and
    compared with the one-volume render on rank 0."""
    import torch
    import torch.distributed as dist
    import ovr_b200
    from ovr_b200 import distributed as D
    capi = ovr_b200.capi
    L = capi.lib()
    rank, world, local, stream = ctx["rank"], ctx["world"], ctx["local_rank"], ctx["stream"]
    S = ovr_b200.scenes
    g = D.brick_grid(world)

    def make(scene, w, h):
        r = ovr_b200.create_renderer("b200", local)
        r.shading = capi.SHADING_NONE
        r.set_fbsize((w, h))
        r.init(scene, scene.camera)
        r.commit()
        r.use_stream(stream.cuda_stream)
        return r

    def dev_frame(ptr, w, h):
        return torch.as_tensor(D._DevicePtr(ptr, w * h * 4), device="cuda").clone().cpu().numpy().reshape(h, w, 4)

    rec = {}
    # ---- down-scaled twin: parity of both schedules against the single-volume render ----
    tw, th, tn = 640, 360, 96
    full = S.noise_scene(tn, tf="ramp", camera="oblique", dims=tuple(g[a] * tn for a in range(3)))
    full.tfn.value_range = (0.0, 1.0)
    want = None
    if rank == 0:
        ref = make(full, tw, th)
        ref.render()
        want = ref.download(want_grad=False)[0].copy()
        ref.close()
    part, brick = D.brick_scene(full, rank, world)
    b = make(part, tw, th)
    D.setup_brick_renderer(b, brick)
    ex = D.SortLastP2P(rank, world, full.volume.dims, tw * th, b)
    b.render()
    ptr = ex.composite(full.camera.frm)
    twin = {}
    if rank == 0:
        b._ck(L.b200vr_finalize_premultiplied(b.handle, C.c_void_p(ptr), tw * th))
        e, p = image_metrics(dev_frame(ptr, tw, th), want)
        twin["binary_swap"] = {"max_abs": e, "psnr_db": p}
    b.render()
    ptr = ex.composite_direct(full.camera.frm)
    if rank == 0:
        e, p = image_metrics(dev_frame(ptr, tw, th), want)
        twin["direct_send"] = {"max_abs": e, "psnr_db": p}
        twin["ok"] = bool(all(v["max_abs"] <= 1.0 / 255.0 and v["psnr_db"] >= 45.0 for v in (twin["binary_swap"], twin["direct_send"])))
        twin["what"] = f"{world} bricks of {tn}^3 (+ghosts) of a {full.volume.dims} volume, {tw}x{th}, against the one-volume render on rank 0; bar 1/255, 45 dB"
        rec["parity_twin"] = twin
    dist.barrier()
    ex.close()
    b.close()
    dist.barrier()

    # ---- the measured workload ----
    name = "C5" if world >= 8 else "C5-mid"
    wl = dict(S.WORKLOADS[name])
    W, H, n = wl["width"], wl["height"], wl["n"]
    dims = tuple(g[a] * n for a in range(3))
    full = S.noise_scene(n, tf=wl["tf"], camera=wl["camera"], rate=wl["rate"], dims=dims)
    full.tfn.value_range = (0.0, 1.0)
    part, brick = D.brick_scene(full, rank, world)
    b = make(part, W, H)
    D.setup_brick_renderer(b, brick)
    ex = D.SortLastP2P(rank, world, dims, W * H, b)
    eye = full.camera.frm
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, frames=20, warm=3):
        for _ in range(warm):
            fn()
        barrier()
        samples, own_ms = 0, 0.0
        ev0.record(stream)
        for _ in range(frames):
            fn()
            samples += b.last_stats["primary_samples"]
            own_ms += b.last_stats["ms"]
        ev1.record(stream)
        barrier()
        t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
        s = torch.tensor([samples, own_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        return float(t.item()) / frames, float(s[0].item()) / frames, float(s[1].item()) / frames

    def f_march():
        b.render()

    def f_swap():
        b.render()
        out = ex.composite(eye)
        if out is not None:
            b._ck(L.b200vr_finalize_premultiplied(b.handle, C.c_void_p(out), W * H))

    def f_direct():
        b.render()
        ex.composite_direct(eye)

    ms_march, samples, seq_ms = timed(f_march)   # seq_ms: sum over the ranks of their own march-kernel time per frame
    ms_swap, _, _ = timed(f_swap)
    ms_direct, _, _ = timed(f_direct)
    # one brick alone on one GPU: the same voxels as a whole volume of their own, framed by the same camera preset
    b._ck(L.b200vr_set_brick(b.handle, None, None, None, None))
    b.premultiplied = False
    b.apply_device_options()
    org = [float(part.volume.grid_origin[a]) for a in range(3)]
    cam1 = S.camera_preset(wl["camera"], brick.size)
    b.set_camera(ovr_b200.Camera(tuple(cam1.frm[a] + org[a] for a in range(3)), tuple(cam1.at[a] + org[a] for a in range(3)), cam1.up, cam1.fovy))
    b.commit()
    for _ in range(3):
        b.render()
    t1, s1 = 0.0, 0
    for _ in range(10):
        b.render()
        t1 += b.last_stats["ms"]
        s1 += b.last_stats["primary_samples"]
    t1, s1 = t1 / 10, s1 / 10
    barrier()
    ex.close()
    b.close()
    if rank != 0:
        return None
    frame_bytes = W * H * 16
    swap_bytes = sum((keep[1] - keep[0]) * 16 for _p, keep, _s in ex.plan)   # read from the partner per round, per rank
    lo, hi = ex.direct_ranges[0]
    best = min(ms_swap, ms_direct)
    rec.update({
        "workload": f"{name}: {world} bricks of {n}^3 fp32 (+ghosts) = {dims[0]}x{dims[1]}x{dims[2]} voxels ({4 * dims[0] * dims[1] * dims[2] / 2 ** 30:.0f} GiB), "
                    f"{W}x{H}, TF {wl['tf']}, no shading, weak scaling",
        "samples_per_frame": samples,
        "march_ms": ms_march,
        "binary_swap": {"ms_per_frame": ms_swap, "exchange_ms": ms_swap - ms_march, "msamples_s": samples / ms_swap / 1e3,
                        "nvlink_bytes_read_per_rank": swap_bytes, "nvlink_bytes_pulled_by_display_rank": frame_bytes - (ex.ranges[0][1] - ex.ranges[0][0]) * 16,
                        "rounds": len(ex.plan), "launches_per_rank": len(ex.plan), "barriers": len(ex.plan) + 2},
        "direct_send": {"ms_per_frame": ms_direct, "exchange_ms": ms_direct - ms_march, "msamples_s": samples / ms_direct / 1e3,
                        "nvlink_bytes_read_per_rank": (world - 1) * (hi - lo) * 16, "nvlink_bytes_written_per_rank": (hi - lo) * 16,
                        "rounds": 1, "launches_per_rank": 1, "barriers": 2},
        "one_brick_one_gpu": {"ms_per_frame": t1, "samples_per_frame": s1, "msamples_s": s1 / t1 / 1e3,
                              "what": "rank 0's brick as a volume of its own, same camera preset, same kernel"},
        "weak_scaling": {"one_gpu_sequential_ms": seq_ms, "speedup": seq_ms / best if best > 0 else None,
                         "efficiency": seq_ms / best / world if best > 0 else None,
                         "note": "the volume does not fit on one GPU, so the one-GPU time is what one GPU with enough memory would need to march "
                                 "the same N bricks one after the other (sum over the ranks of their own march-kernel time, compositing "
                                 "excluded), over the N-GPU frame time including the exchange; north_star: >= 6x at 8. one_brick_one_gpu is "
                                 "the other view: a brick as a volume of its own filling the frame (more pixels hit, more samples)"},
    })
    return rec


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    ctx = run_b200(args)
    r, wl, rank, world = ctx["r"], ctx["wl"], ctx["rank"], ctx["world"]
    import ovr_b200
    capi = ovr_b200.capi
    e2e = run_e2e(ctx, args)
    sortlast_wl = bool(wl.get("sortlast"))
    sf_rec = sf_rgba = sf_grad = None
    if world > 1 and not sortlast_wl and wl["mode"] == "raymarch":
        sf_rec, sf_rgba, sf_grad = sort_first_check(ctx)
    n_frames, F = ctx["frames"], ctx["F"]
    line = None
    if rank == 0:
        # ---- roofline of the render kernel: texture/L1 sampling rate (32 B per executed trilinear fetch) ----
        capi.lib().b200vr_set_tile_partition(r.handle, 0, 1)
        p_fetch, fetch_rows = measure_fetch_peak(r, capi, r.sampler)
        capi.lib().b200vr_set_tile_partition(r.handle, 0, world)
        pk = peaks()
        n = wl["n"]
        r_dims = tuple(int(v) for v in (ctx["scene"].volume.dims if world == 1 or not sortlast_wl else (n, n, n)))
        if sortlast_wl:
            r_dims = (n, n, n)   # this rank's brick
        kernel_s = ctx["kernel_ms"] * 1e-3
        ach = 32.0 * ctx["fetches"] / world / kernel_s / 1e9 if kernel_s > 0 else 0.0   # per GPU: rank-max kernel time
        peak = 32.0 * p_fetch / 1e9
        # compulsory HBM bytes per frame: 4 B x voxels of the macrocells a fetch can touch + 12 B per macrocell + the
        # frame stores. With skipping on, only cells whose majorant is not 0 are ever fetched from (a superset of the cells the
        # rays reach before they terminate); the whole volume is the bound without skipping.
        mcd = (C.c_int32 * 3)()
        r._ck(capi.lib().b200vr_get_macrocells(r.handle, mcd, None, None))
        n_cells = mcd[0] * mcd[1] * mcd[2]
        maj = np.empty(n_cells, np.float32)
        r._ck(capi.lib().b200vr_get_macrocells(r.handle, mcd, None, maj.ctypes.data_as(C.c_void_p)))
        live_cells = int(np.count_nonzero(maj)) if not args.no_skip else n_cells
        vox_per_cell = (r_dims[0] * r_dims[1] * r_dims[2]) / n_cells
        frame_bytes = (28.0 + (32.0 if wl["mode"] != "raymarch" else 0.0)) * wl["width"] * wl["height"]
        # bytes per voxel in the 3-D array the kernels fetch from: 8-bit volumes stay 8-bit (B200VR_TEX_NATIVE8=0: fp32 array)
        bpv = 1.0 if wl.get("dtype") == "u8" and os.environ.get("B200VR_TEX_NATIVE8", "1") != "0" else 4.0
        hbm_bytes_whole = bpv * r_dims[0] * r_dims[1] * r_dims[2] + 12.0 * n_cells + frame_bytes
        hbm_bytes = bpv * vox_per_cell * live_cells + 12.0 * n_cells + frame_bytes
        hbm_ach = hbm_bytes * n_frames / kernel_s / 1e9 if kernel_s > 0 else 0.0
        kernel_name = ctx["kernel"]   # b200vr_last_kernel: what pick_kernel (csrc/b200vr.cu) chose for the timed frames
        # dram__bytes_read.sum + dram__bytes_write.sum of that kernel for this workload, from the committed `ncu --set full` capture
        traffic, traffic_src, l1_miss = None, None, None
        if world == 1:
            try:
                with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                    t = json.load(f).get(f"{args.workload}:{kernel_name}")
                if t:
                    traffic, traffic_src = t["dram_bytes_per_launch"], t["source"]
                    if t.get("l1_miss_bytes_per_launch"):
                        # what bounds the 1024^3 kernel: bytes that miss L1 (l1tex__m_xbar2l1tex_read_bytes of the committed
                        # capture) over this run's kernel time; the fetch probe sustains 3.55 TB/s of them at its DRAM roof
                        l1_miss = {"bytes_per_launch": t["l1_miss_bytes_per_launch"], "source": t["l1_miss_source"],
                                   "achieved_gbs": t["l1_miss_bytes_per_launch"] / (ctx["kernel_ms"] / n_frames * 1e-3) / 1e9}
            except (OSError, ValueError):
                pass
        roofline = {"achieved_tex": ach / peak if peak else None, "achieved_hbm": hbm_ach / pk["hbm_gbs"],   # fractions of both roofs
                    "bound": "tex", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak if peak else None,
                    "traffic": traffic, "traffic_source": traffic_src, "kernel": kernel_name,
                    "note": "texture/L1 sampling roofline: 32 B x executed trilinear fetches / mean kernel time (CUDA events on the launch "
                            "stream); peak = best b200vr_bench_fetch rate x 32 B measured in this run",
                    "kernel_ms_per_launch": ctx["kernel_ms"] / n_frames, "fetch_bench": fetch_rows, "l1_miss_path": l1_miss,
                    "hbm": {"achieved": hbm_ach, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": hbm_ach / pk["hbm_gbs"],
                            "bytes_per_launch": hbm_bytes, "live_macrocells": live_cells, "macrocells": n_cells,
                            "bytes_per_launch_whole_volume": hbm_bytes_whole, "peak_source": pk["source"]}}
        line = {"metric": METRIC, "value": ctx["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ctx["t_ms"] / args.steps, "ms_per_frame": ctx["t_ms"] / n_frames, "higher_is_better": True,
                "scaling": "weak" if sortlast_wl else "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": workload_name(args, wl), "frames_per_step": F, "sampler": args.sampler, "skip_empty": not args.no_skip,
                           "parallelism": (f"sort-last x{world}: one {n}^3 brick per GPU, binary swap, " + ("blend reads the partner over NVLink peer memory" if args.exchange == "p2p" else "NCCL send/recv") if sortlast_wl else
                                           f"sort-first x{world}, " + ("tiles stored straight into rank 0's frame over NVLink peer memory, frames enqueued back to back (b200vr_render_async), counters read once after the timed region" if args.exchange == "p2p"
                                                                       else "NCCL gather of tile slices") if world > 1 else "1 GPU"),
                           "l2": f"volume {4 * n ** 3 / 2 ** 20:.0f} MiB > 126 MB L2: inputs larger than L2, no flush",
                           "samples_per_frame": ctx["samples"] / n_frames, "fetches_per_frame": ctx["fetches"] / n_frames},
                "clocks": ctx["clk"], "e2e": e2e, "gpu_launches": ctx["gpu_launches"], "roofline": roofline}
        if ctx["per_rank_kernel"] is not None:
            roofline["kernel_ms_per_launch_by_rank"] = ctx["per_rank_kernel"]
        if sf_rec is not None:
            line["sort_first_check"] = sf_rec
        parity_frame = None
        if world == 1 and not sortlast_wl and not wl.get("sparse") and not args.no_skip:
            # the frame the parity check looks at: the workload's own camera, the kernel that was timed
            r.set_camera(ctx["scene"].camera)
            r.commit()
            if wl["mode"] == "raymarch":
                r.render()
                parity_frame = r.download()
                line["parity_kernel"] = r.last_kernel
            # Msamples/s is reported twice -- `value` counts the samples this device executed; the reference-
            # equivalent figure counts the samples the reference's unskipped algorithm takes for the same frame (one extra frame
            # with empty-space skipping off: the image is bit-identical, only the counters differ), over the same frame time
            r.skip_empty = False
            r.apply_device_options()
            r.render()
            st = r.last_stats
            s_ref = st["primary_samples"] + st["shadow_samples"] + st["delta_samples"]
            r.skip_empty = True
            r.apply_device_options()
            line["config"]["samples_per_frame_reference_equivalent"] = s_ref
            line["value_reference_equivalent"] = s_ref / (ctx["t_ms"] / n_frames * 1e-3) / 1e6
        elif world > 1 and sf_rgba is not None:
            parity_frame = (sf_rgba, sf_grad)   # the gathered frame
        if not args.no_cpu_baseline and not sortlast_wl and not wl.get("sparse") and n <= 1024:   # the CPU port needs the whole volume on the host
            from oracle import pyoracle
            vol = np.empty(r_dims[::-1], np.float32)
            r._ck(capi.lib().b200vr_download_volume(r.handle, vol.ctypes.data_as(C.c_void_p)))
            stride = args.cpu_stride or (auto_stride(pyoracle.hardware_threads()) * (1 if world == 1 else 2))
            v, ms, info = cpu_oracle_run(wl, ctx["scene"], 1, 0, stride, volume=vol)
            if world == 1:
                line["cpu_baseline"] = dict(public(info), value=v, unit=UNIT, ms_sampled_frame=ms)
            if not args.no_parity:
                if wl["mode"] == "raymarch" and parity_frame is not None:
                    line["parity"] = parity_march(parity_frame[0], parity_frame[1], info["_pixels"])
                elif wl["mode"] == "pathtrace" and world == 1:
                    line["parity"] = parity_path(r, ctx["scene"], wl, info["_oracle"])
    ctx["barrier"]()
    r.close()
    if world > 1 and not sortlast_wl and not args.no_sort_last and (world & (world - 1)) == 0:
        rec = run_sort_last_record(args, ctx)
        if rank == 0:
            line["sort_last"] = rec
    if rank == 0:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
