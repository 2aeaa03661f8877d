"""Generates the committed fixtures under tests/golden/ from the REFERENCE itself (run in the authoring container only,
where /root/reference exists):

  ref_kat.npz       outputs of the reference's own dda.h (dda3 + DDAIter), RandomTEA and gdt transforms compiled on the
                    host by oracle/Makefile (target `ref`, oracle/ref_shim.cpp)
  scene_vorts1.npz  the bundled scene data/configs/scene_vorts1.json (BASELINE config C1): camera, sample distance,
                    scalar mapping range, 1024-entry alpha array, colour controls

  sparse_mask.npz   sparse-sampling masks from the reference's own mask kernels run on the host (same shim)

  tfn_tables.npz    RGBA tables of every bundled scene's transfer function from the reference's tfn::loadTransferFunction

  delta_hashit.npz  free-path iteration of delta tracking from the reference's own DeltaTrackingIter<true>::hashit

  shader_helpers.npz  intersect_box and uniform_sample_sphere from the reference's shaders_common.h

  ref_march.npz     whole frames from the reference's own raymarching() / raymarching_shadow() run on the host
  host_derivations.npz  device camera and instance matrices from the reference's own host statements (device_impl.cpp)
  ref_march_spp.npz the same with 2-4 samples per pixel (RandomTEA pixel jitter of the raygen program)
  value_ranges.npz  macrocell value ranges from the reference's value_range_kernel run on the host
  scene_dump.json   synthetic VIDI3D scenes (all value types, endianness, offset, scales) as parsed by the reference's own loader
  ref_path.npz      whole frames from the reference's own delta_tracking() / pathtracing() run on the host (both scattered-ray clips)

Usage: python tests/golden/make_golden.py [sparse|tfn|delta|helpers|march|march_spp|path|host|ranges|scenes]
"""
import base64
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import sparse_noise_table  # noqa: E402

REF = "/root/reference"


def f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def i3(v):
    return (C.c_int32 * 3)(*[int(x) for x in v])


def main():
    pyoracle.build()
    L = pyoracle.ref_kat_lib()
    assert L is not None, "oracle/_ref/libovr_ref_kat.so missing (needs /root/reference)"
    rng = np.random.default_rng(1337)
    out = {}

    # ---- RandomTEA ----
    seeds = np.array([[1, 0], [1, 1], [2, 12345], [64, 1920 * 1080 - 1], [0xFFFFFFFF, 0xDEADBEEF], [7, 99]], np.uint32)
    draws = np.zeros((len(seeds), 8, 2), np.float32)
    for k, (a, b) in enumerate(seeds):
        L.ref_tea(C.c_uint32(int(a)), C.c_uint32(int(b)), 8, draws[k].ctypes.data_as(C.POINTER(C.c_float)))
    out["tea_seeds"], out["tea_draws"] = seeds, draws

    # ---- DDA: rays through grids, both the one-shot dda3 and the resumable DDAIter ----
    grids = [(8, 8, 8), (32, 16, 8), (4, 7, 64)]
    rays = []
    MAXC = 256
    for g in grids:
        for _ in range(24):
            gdim = np.array(g, np.float32)
            org = (rng.uniform(-0.5, 1.5, 3) * gdim).astype(np.float32)
            tgt = (rng.uniform(0.1, 0.9, 3) * gdim).astype(np.float32)
            d = (tgt - org).astype(np.float32)
            if rng.uniform() < 0.25:
                d[rng.integers(3)] = 0.0  # axis-parallel component
                org = np.clip(org, 0.25, gdim - 0.25).astype(np.float32)
            d = (d / np.float32(np.linalg.norm(d)) * np.float32(rng.uniform(0.5, 2.0))).astype(np.float32)
            ok, t0, t1 = pyoracle.clip(org, d, (0, 0, 0), g)  # entry/exit like the shaders do before the DDA
            if not ok:
                continue
            rec = {}
            for name, fn in (("dda3", L.ref_dda3), ("iter", L.ref_dda_iter)):
                cells = np.zeros((MAXC, 3), np.int32)
                tt = np.zeros((MAXC, 2), np.float32)
                n = fn(f3(org), f3(d), C.c_float(t0), C.c_float(t1), i3(g), MAXC, cells.ctypes.data_as(C.c_void_p),
                       tt.ctypes.data_as(C.POINTER(C.c_float)))
                rec[name] = (cells[:n].copy(), tt[:n].copy())
            rays.append((np.array(g, np.int32), org, d, np.float32(t0), np.float32(t1), rec))
    out["dda_n"] = np.int32(len(rays))
    for k, (g, org, d, t0, t1, rec) in enumerate(rays):
        out[f"dda{k}_grid"], out[f"dda{k}_org"], out[f"dda{k}_dir"] = g, org, d
        out[f"dda{k}_t"] = np.array([t0, t1], np.float32)
        for name in ("dda3", "iter"):
            out[f"dda{k}_{name}_cells"], out[f"dda{k}_{name}_t"] = rec[name]

    # ---- gdt transforms ----
    mats, vecs, res = [], [], []
    for k in range(32):
        if k % 2 == 0:  # translate * scale, what buildScene produces
            s = rng.uniform(0.5, 600, 3)
            o = rng.uniform(-10, 10, 3)
            m = np.array([s[0], 0, 0, o[0], 0, s[1], 0, o[1], 0, 0, s[2], o[2]], np.float32)
        else:
            m = rng.uniform(-2, 2, 12).astype(np.float32)
        v = rng.uniform(-3, 3, 3).astype(np.float32)
        r = np.zeros(24, np.float32)
        L.ref_xfm(m.ctypes.data_as(C.POINTER(C.c_float)), v.ctypes.data_as(C.POINTER(C.c_float)),
                  r.ctypes.data_as(C.POINTER(C.c_float)))
        mats.append(m), vecs.append(v), res.append(r)
    out["xfm_m"], out["xfm_v"], out["xfm_out"] = np.array(mats), np.array(vecs), np.array(res)
    np.savez_compressed(os.path.join(HERE, "ref_kat.npz"), **out)
    print("wrote ref_kat.npz:", len(rays), "dda rays")

    # ---- bundled scene C1 ----
    js = json.load(open(os.path.join(REF, "data/configs/scene_vorts1.json")))
    ds, view = js["dataSource"][0], js["view"]
    vol, cam = view["volume"], view["camera"]
    tf = vol["transferFunction"]
    alpha = np.frombuffer(base64.b64decode(tf["alphaArray"]["data"]), dtype="<f4").astype(np.float32)
    cc = np.array([[c["position"], c["color"]["r"], c["color"]["g"], c["color"]["b"]] for c in tf["colorControls"]], np.float32)
    np.savez_compressed(
        os.path.join(HERE, "scene_vorts1.npz"),
        dims=np.array([ds["dimensions"][k] for k in "xyz"], np.int32), type=np.array(ds["type"]),
        eye=np.array([cam["eye"][k] for k in "xyz"], np.float32), center=np.array([cam["center"][k] for k in "xyz"], np.float32),
        up=np.array([cam["up"][k] for k in "xyz"], np.float32), fovy=np.float32(cam["fovy"]),
        sample_distance=np.float32(vol["sampleDistance"]),
        mapping_range=np.array([vol["scalarMappingRange"]["minimum"], vol["scalarMappingRange"]["maximum"]], np.float32),
        alpha=alpha, color_controls=cc, resolution=np.int32(tf["resolution"]))
    print("wrote scene_vorts1.npz: alpha", alpha.shape, "max", alpha.max())


SPARSE_CASES = [  # (width, height, frame_index, tile_xy, focus centre, focus scale, base noise)
    (160, 96, 1, 16, (0.5, 0.5), 0.06, 0.07),     # the defaults of apps/main_app.cpp:122-124
    (257, 131, 7, 16, (0.3, 0.7), 0.15, 0.01),
    (320, 200, 64, 64, (0.9, 0.1), 0.5, 0.2),     # frame_index wraps the 64 time slices
    (64, 64, 129, 64, (0.5, 0.5), 1.0, 1.0),      # base noise 1: p = 1 everywhere -> the noise value 1.0 alone is rejected
    (96, 128, 3, 16, (1.5, -0.5), 0.05, 0.0),     # focus outside the image, no base noise: almost empty
]


def make_sparse():
    """sparse_mask.npz: masks produced by the reference's generate_blue_noise_kernel + generate_sparse_samples (cut out of
    ovr/common/random/blue_noise.h and ovr/common/generate_mask.cu and run on the host by oracle/ref_shim.cpp)."""
    pyoracle.build()
    assert pyoracle.ref_kat_lib() is not None, "oracle/_ref/libovr_ref_kat.so missing (needs /root/reference)"
    out = {"n": np.int32(len(SPARSE_CASES))}
    for k, (w, h, fi, tile, mean, sigma, base) in enumerate(SPARSE_CASES):
        xy = pyoracle.ref_sparse_mask(w, h, fi, sparse_noise_table(tile), tile, mean, sigma, base)
        keep = xy[:, 0] >= 0
        idx = np.nonzero(keep)[0]
        assert np.array_equal(xy[idx, 0], idx % w) and np.array_equal(xy[idx, 1], idx // w) and (xy[~keep] == -1).all()
        out[f"case{k}"] = np.array([w, h, fi, tile], np.int32)
        out[f"case{k}_f"] = np.array([mean[0], mean[1], sigma, base], np.float32)
        out[f"case{k}_keep"] = np.packbits(keep)
        print(f"sparse case {k}: {keep.sum()} of {w * h} pixels kept")
    np.savez_compressed(os.path.join(HERE, "sparse_mask.npz"), **out)


def make_tfn():
    """tfn_tables.npz: for every bundled data/configs/*.json its "transferFunction" object (text) and the RGBA table the
    reference's tfn::loadTransferFunction builds from it (alpha array, colour controls, Gaussian objects, opacity controls),
    plus two synthetic objects that exercise opacityControl and a missing alphaArray."""
    import glob
    from ovr_b200 import scene_io
    pyoracle.build()
    out, names = {}, []
    cases = [(os.path.basename(p)[:-5], scene_io.load_json_with_comments(p)["view"]["volume"]["transferFunction"])
             for p in sorted(glob.glob(os.path.join(REF, "data/configs/*.json")))]
    cases.append(("synthetic_opacity_controls", {
        "resolution": 256, "colorControls": [{"position": 0.2, "color": {"r": 1, "g": 0, "b": 0}},
                                             {"position": 0.9, "color": {"r": 0, "g": 0.5, "b": 1}}],
        "opacityControl": [{"position": {"x": 0.1, "y": 0.0}}, {"position": {"x": 0.5, "y": 0.8}}, {"position": {"x": 0.75, "y": 0.1}}],
        "gaussianObjects": [{"mean": 0.3, "sigma": 0.02, "heightFactor": 0.03}]}))
    cases.append(("synthetic_no_controls", {"resolution": 64, "gaussianObjects": [{"mean": 0.5, "sigma": 0.1, "heightFactor": 0.1}]}))
    for name, tf in cases:
        text = json.dumps(tf)
        table = pyoracle.ref_tfn_table(text)
        assert table is not None, "oracle/_ref/libovr_ref_tfn.so missing (needs /root/reference)"
        names.append(name)
        out[name + "_json"], out[name + "_table"] = np.array(text), table
        print(f"tfn {name}: resolution {len(table)}, alpha max {table[:, 3].max():.4f}")
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "tfn_tables.npz"), **out)


def delta_cases():
    """Seeded rays through random majorant grids with zero-majorant (skipped) cells; shared by the generator and the test."""
    rng = np.random.default_rng(4242)
    cases = []
    for grid in [(8, 8, 8), (16, 4, 9), (3, 64, 5)]:
        for k in range(12):
            maj = rng.uniform(0.0, 1.5, grid[::-1]).astype(np.float32)
            maj[rng.uniform(size=maj.shape) < (0.15 + 0.05 * k)] = 0.0          # empty cells: majorant <= epsilon
            g = np.array(grid, np.float32)
            spac_rcp = g                                                          # object space = unit cube
            org = rng.uniform(-0.6, 1.6, 3).astype(np.float32)
            tgt = rng.uniform(0.1, 0.9, 3).astype(np.float32)
            d = tgt - org
            if k % 4 == 3:
                d[rng.integers(3)] = 0.0
                org = np.clip(org, 0.05, 0.95).astype(np.float32)
            d = (d / np.float32(np.linalg.norm(d))).astype(np.float32)
            ok, t0, t1 = pyoracle.clip(org, d, (0, 0, 0), (1, 1, 1))
            if not ok:
                continue
            taus = (-np.log(1.0 - rng.uniform(0.0, 1.0, 40)) * rng.choice([0.02, 0.2, 1.0])).astype(np.float32)
            cases.append((maj, spac_rcp, org, d, np.float32(t0), np.float32(t1), taus))
    return cases


def shader_helper_cases():
    """Seeded inputs of intersect_box and uniform_sample_sphere; shared by the generator and the test."""
    rng = np.random.default_rng(777)
    rays = []
    for k in range(600):
        lo = rng.uniform(-0.2, 0.4, 3).astype(np.float32) if k % 3 else np.zeros(3, np.float32)
        hi = (lo + rng.uniform(0.1, 1.0, 3)).astype(np.float32) if k % 3 else np.ones(3, np.float32)
        org = rng.uniform(-2.0, 3.0, 3).astype(np.float32)
        d = rng.normal(size=3).astype(np.float32)
        if k % 2 == 0:                                     # aimed at the box: most of these hit
            d = ((lo + (hi - lo) * rng.uniform(0.05, 0.95, 3)).astype(np.float32) - org).astype(np.float32)
        if k % 5 == 0:
            d[rng.integers(3)] = 0.0                       # |dir| < FLT_MIN on one axis
        if k % 7 == 0:
            d[rng.integers(3)] = np.float32(1e-39)         # denormal component: also "small"
        if k % 11 == 0:
            org = (lo + (hi - lo) * rng.uniform(0.1, 0.9, 3)).astype(np.float32)   # origin inside the box
        rays.append((org, d, lo, hi))
    s = rng.uniform(0.0, 1.0, (1500, 2)).astype(np.float32)
    s[:8] = [[0, 0], [1, 1], [0.5, 0.5], [0.25, 0], [0, 1], [1, 0], [0.75, 0.999999], [0.999999, 0.5]]
    return rays, s


def make_shader_helpers():
    """shader_helpers.npz: the reference's own intersect_box and uniform_sample_sphere (shaders_common.h:156-184,339-354)."""
    pyoracle.build()
    assert pyoracle.ref_kat_lib() is not None, "oracle/_ref/libovr_ref_kat.so missing (needs /root/reference)"
    rays, s = shader_helper_cases()
    box = np.array([[float(h), t0, t1] for h, t0, t1 in (pyoracle.ref_clip(*r) for r in rays)], np.float32)
    sph = pyoracle.uniform_sphere(s, use_ref=True)
    np.savez_compressed(os.path.join(HERE, "shader_helpers.npz"), box=box, sphere=sph)
    print(f"wrote shader_helpers.npz: {int(box[:, 0].sum())} of {len(rays)} rays hit, {len(sph)} sphere samples")


MARCH_REF_CASES = [  # n, TF preset, camera preset, sampling rate, (W, H)
    (32, "sparse", "oblique", 1.0, (48, 32)),   # step 1: opacity correction only where the running sum drifts / last step
    (24, "ramp", "inside", 2.0, (40, 24)),      # camera inside the volume, rate 2: correction on every sample
    (20, "ramp", "oblique", 0.7, (32, 20)),     # step > 1 voxel
    (28, "dense", "front", 1.0, (36, 24)),      # no empty space, early ray termination
]


def march_ref_scene(case):
    """OracleScene of one MARCH_REF_CASES entry (shared by the generator and the test)."""
    import ovr_b200
    from helpers import oracle_desc
    n, tf, cam, rate, (w, h) = case
    sc = ovr_b200.scenes.noise_scene(n, tf=tf, camera=cam, rate=rate)
    return pyoracle.OracleScene(oracle_desc(sc, w, h)), w, h


def make_march_ref():
    """ref_march.npz: frames computed by the REFERENCE's own raymarching() / raymarching_shadow() and their helpers, cut out of
    shaders_raymarching.cu / shaders_common.h and executed on the host (oracle/ref_march_shim.cpp), with texture fetches
    through the oracle's model of the unit."""
    pyoracle.build()
    out = {"n": np.int32(len(MARCH_REF_CASES))}
    for k, case in enumerate(MARCH_REF_CASES):
        o, w, h = march_ref_scene(case)
        ys, xs = np.mgrid[0:h, 0:w]
        r = o.ref_march(np.stack([xs.ravel(), ys.ravel()], 1))
        assert r is not None, "oracle/_ref/libovr_ref_march.so missing (needs /root/reference)"
        out[f"rgba{k}"], out[f"grad{k}"] = r[0].reshape(h, w, 4), r[1].reshape(h, w, 3)
        print(f"march case {k} {case}: {(out[f'rgba{k}'][..., 3] > 0).mean():.2f} of the pixels hit, max alpha {out[f'rgba{k}'][..., 3].max():.4f}")
    np.savez_compressed(os.path.join(HERE, "ref_march.npz"), **out)


# n, TF preset, camera preset, sampling rate, spp, frame index, (W, H): pixel jitter + mean over spp of __raygen__render_frame
MARCH_SPP_CASES = [(32, "sparse", "oblique", 1.0, 4, 1, (48, 32)), (24, "ramp", "inside", 2.0, 2, 5, (40, 24)),
                   (28, "dense", "front", 0.7, 3, 2, (36, 24))]


def march_spp_scene(case):
    import ovr_b200
    from helpers import oracle_desc
    n, tf, cam, rate, spp, fi, (w, h) = case
    sc = ovr_b200.scenes.noise_scene(n, tf=tf, camera=cam, rate=rate)
    return pyoracle.OracleScene(oracle_desc(sc, w, h, spp=spp)), w, h, fi


def make_march_spp():
    """ref_march_spp.npz: ray-marched frames with several samples per pixel -- the reference's render_raymarching() (two traces,
    background blend) and raymarching() under the raygen loop of __raygen__render_frame (RandomTEA pixel jitter, mean)."""
    pyoracle.build()
    out = {"n": np.int32(len(MARCH_SPP_CASES))}
    for k, case in enumerate(MARCH_SPP_CASES):
        o, w, h, fi = march_spp_scene(case)
        o.s.frame_index = fi
        ys, xs = np.mgrid[0:h, 0:w]
        r = o.ref_march(np.stack([xs.ravel(), ys.ravel()], 1), threads=0)
        assert r is not None, "oracle/_ref/libovr_ref_march.so missing (needs /root/reference)"
        out[f"rgba{k}"], out[f"grad{k}"] = r[0].reshape(h, w, 4), r[1].reshape(h, w, 3)
    np.savez_compressed(os.path.join(HERE, "ref_march_spp.npz"), **out)
    print(f"wrote ref_march_spp.npz: {len(MARCH_SPP_CASES)} frames")


# (volume size, TF, camera, spp, frame index, (width, height)): two sizes that are multiples of the 16-voxel macrocell and two that
# are not (there the reference's scattered rays run on into the overhang of the last macrocells)
PATH_REF_CASES = [(32, "sparse", "oblique", 1, 1, (48, 32)), (32, "ramp", "inside", 4, 7, (40, 24)),
                  (20, "ramp", "oblique", 2, 1, (32, 20)), (28, "dense", "oblique", 1, 3, (36, 24))]


def path_ref_scene(case, child_instance_xfm):
    """OracleScene of one PATH_REF_CASES entry (shared by the generator and the test)."""
    import ovr_b200
    from helpers import oracle_desc
    n, tf, cam, spp, fi, (w, h) = case
    sc = ovr_b200.scenes.noise_scene(n, tf=tf, camera=cam, rate=1.0)
    o = pyoracle.OracleScene(oracle_desc(sc, w, h, spp=spp))
    o.s.pt_child_instance_xfm = int(child_instance_xfm)
    return o, w, h, fi


def make_path_ref():
    """ref_path.npz: frames computed by the REFERENCE's own delta_tracking() / pathtracing() / render_pathtracing(), its macrocell
    iterator and RandomTEA, cut out of shaders_pathtracing.cu / spatial_partition.h / random.h and executed on the host
    (oracle/ref_march_shim.cpp, namespace pt). Two frames per case: x1 = scattered rays carried through the instance transform
    again before the clip, as OptiX does (the reference's behaviour); x0 = clipped as the object-space rays they are."""
    pyoracle.build()
    out = {"n": np.int32(len(PATH_REF_CASES))}
    for k, case in enumerate(PATH_REF_CASES):
        for x in (0, 1):
            o, w, h, fi = path_ref_scene(case, x)
            ys, xs = np.mgrid[0:h, 0:w]
            r = o.ref_path(np.stack([xs.ravel(), ys.ravel()], 1), frame_index=fi)
            assert r is not None, "oracle/_ref/libovr_ref_march.so missing (needs /root/reference)"
            out[f"rgba{k}_x{x}"] = r.reshape(h, w, 4)
        d = (out[f"rgba{k}_x0"].view(np.uint32) != out[f"rgba{k}_x1"].view(np.uint32)).any(-1)
        print(f"path case {k} {case}: mean radiance {out[f'rgba{k}_x1'][..., :3].mean():.4f} (x1) {out[f'rgba{k}_x0'][..., :3].mean():.4f} (x0), "
              f"{int(d.sum())} of {w * h} pixels differ between the two clips")
    np.savez_compressed(os.path.join(HERE, "ref_path.npz"), **out)


def scene_dump_cases(tmpdir):
    """Synthetic VIDI3D configs + payloads written into tmpdir: every value type, big-endian payloads, a file offset, `scales`,
    an unnormalised mapping range, alpha ends below 0.01. -> [(name, json path)] (shared by the generator and the test)."""
    import json
    import ovr_b200
    from ovr_b200 import scene_io
    S = ovr_b200.scenes
    rng = np.random.default_rng(77)
    cases = []
    kinds = [("u8", np.uint8, 255), ("i8", np.int8, 127), ("u16", np.uint16, 60000), ("i16", np.int16, 30000), ("u32", np.uint32, 4e9),
             ("i32", np.int32, 2e9), ("f32", np.float32, 1.0), ("f64", np.float64, 1.0)]
    for k, (name, dt, top) in enumerate(kinds):
        dims = (5 + k, 4 + (k % 3), 3 + (k % 2))  # nz, ny, nx
        vol = (rng.random(dims) * top).astype(dt)
        cam = S.Camera((1.5 + k, 2.5, -3.0), (0.1, 0.2 * k, 0.3), (0.0, 1.0, 0.0), 30.0 + 5 * k)
        alpha = rng.random(8 + 3 * k).astype(np.float32)
        alpha[0], alpha[-1] = (0.005, 0.5) if k % 2 else (0.3, 0.009)
        cc = [(0.0, (0, 0, 1)), (0.35, (0, 1, 0.25)), (1.0, (1, 0, 0))]
        path = scene_io.write_vidi3d_scene(os.path.join(tmpdir, f"case_{name}.json"), vol, alpha, cc, cam, sample_distance=0.25 * (k + 1),
                                           scalar_range=(0.1, 0.9))
        root = json.load(open(path))
        ds = root["dataSource"][0]
        raw = ds["fileName"][0]
        if k % 3 == 1:   # big-endian payload behind a 12-byte header
            with open(raw, "wb") as f:
                f.write(b"x" * 12)
                f.write(vol.astype(vol.dtype.newbyteorder(">")).tobytes())
            ds["endian"], ds["offset"] = "BIG_ENDIAN", 12
        if k % 3 == 2:
            ds["scales"] = {"x": 1.0, "y": 0.5 + k, "z": 2.25}
        if k % 4 == 3:
            root["view"]["volume"]["scalarMappingRangeUnnormalized"] = {"minimum": 3.0 + k, "maximum": 40.5 + k}
        json.dump(root, open(path, "w"), indent=1)
        cases.append((name, path))
    return cases


def make_scene_dump():
    """scene_dump.json: what the reference's own scene loader (create_scene: scene.cpp + serializer_vidi3d.cpp, compiled from
    /root/reference into build/ovr/scenedump behind oracle/ref_scene_dump.cpp) makes of scene_dump_cases()."""
    import json
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "build", "ovr", "scenedump")
    assert os.path.exists(exe), "build/ovr/scenedump missing: python open-volume-renderer_b200/device/build.py (needs /root/reference)"
    with tempfile.TemporaryDirectory() as d:
        cases = scene_dump_cases(d)
        r = subprocess.run([exe] + [p for _, p in cases], capture_output=True, text=True, check=True)
        rows = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(rows) == len(cases) and not any("error" in x for x in rows), rows
    for (name, _), row in zip(cases, rows):
        row["file"] = name
    json.dump(rows, open(os.path.join(HERE, "scene_dump.json"), "w"), separators=(",", ":"))
    print(f"wrote scene_dump.json: {len(rows)} scenes")


VALUE_RANGE_DIMS = [(32, 32, 32), (70, 45, 33), (17, 16, 15), (100, 1, 37), (48, 31, 64)]


def value_range_volume(dims):
    rng = np.random.default_rng(dims[0] * 1000003 + dims[1] * 1009 + dims[2])
    return rng.random((dims[2], dims[1], dims[0]), dtype=np.float32)


def make_value_ranges():
    """value_ranges.npz: macrocell value ranges from the reference's value_range_kernel (accel/sp_singlemc.cu:10-54) run on the host."""
    pyoracle.build()
    out = {f"r{k}": pyoracle.ref_value_ranges(value_range_volume(d)) for k, d in enumerate(VALUE_RANGE_DIMS)}
    np.savez_compressed(os.path.join(HERE, "value_ranges.npz"), **out)
    print(f"wrote value_ranges.npz: {len(VALUE_RANGE_DIMS)} volumes")


def host_derivation_inputs():
    """Seeded look-at cameras / frame sizes and volume placements (shared by the generator and the test)."""
    rng = np.random.default_rng(20261105)
    cams = []
    for k in range(400):
        frm = (rng.normal(size=3) * rng.choice([1, 100, 1000])).astype(np.float32)
        at = (rng.normal(size=3) * 10).astype(np.float32)
        up = rng.normal(size=3).astype(np.float32)
        cams.append((frm, at, up, np.float32(rng.uniform(5, 120)), int(rng.integers(16, 4000)), int(rng.integers(16, 3000))))
    cams.append((np.array([0, 0, -5], np.float32), np.zeros(3, np.float32), np.array([0, 1, 0], np.float32), np.float32(60), 1920, 1080))
    vols = []
    for k in range(200):
        org = (rng.normal(size=3) * rng.choice([0, 1, 10])).astype(np.float32)
        vols.append((np.abs(org) * np.sign(org + 1e-30).astype(np.float32) if k % 2 else np.zeros(3, np.float32),
                     rng.uniform(0.1, 3, size=3).astype(np.float32), rng.integers(2, 4097, size=3).astype(np.int32)))
    return cams, vols


# value type of the texture the reference creates for each scene value type (CreateArray3DScalarOptix7, array.cpp:322-351):
# 16-bit and double arrays are converted to float; indices follow B200VR_TYPE_* (u8 i8 u16 i16 u32 i32 f32 f64)
TEXTURE_TYPE = {0: 0, 1: 1, 2: 6, 3: 6, 4: 4, 5: 5, 6: 6, 7: 6}


def integer_normalize_inputs():
    rng = np.random.default_rng(9)
    rows = []
    for t in range(8):
        top = {0: 255, 1: 127, 2: 65535, 3: 32767, 4: 4294967295, 5: 2147483647, 6: 1.0, 7: 1.0}[t]
        lo = {1: -128, 3: -32768, 5: -2147483648}.get(t, 0)
        vals = np.concatenate([rng.uniform(lo, top, 200), [lo, top, 0, 1, top / 2]]).astype(np.float32)
        if t in (4, 5):  # stay inside the integer range: the cast of a larger float is undefined
            vals = np.clip(vals, lo, np.float32(top) - 256).astype(np.float32)
        rows += [(t, v) for v in vals]
    return rows


def make_host_derivations():
    """host_derivations.npz: the device camera of DeviceOptix7::Impl::commit (device_impl.cpp:128-136) and the instance matrix of
    buildScene (:288-296) with its gdt inverse, computed by the reference's own statements (cut out at build time, ref_shim.cpp)."""
    pyoracle.build()
    cams, vols = host_derivation_inputs()
    import ctypes as C
    L = pyoracle.ref_kat_lib()
    L.ref_integer_normalize.restype = C.c_float
    out = {"camera": np.array([pyoracle.ref_camera(*c) for c in cams]),
           "matrix": np.array([pyoracle.ref_volume_matrix(*v) for v in vols]),
           # integer_normalize(float, ValueType) (array.h:64-106) for the texture's element type
           "norm": np.array([L.ref_integer_normalize(C.c_float(v), TEXTURE_TYPE[t]) for t, v in integer_normalize_inputs()], np.float32)}
    np.savez_compressed(os.path.join(HERE, "host_derivations.npz"), **out)
    print(f"wrote host_derivations.npz: {len(cams)} cameras, {len(vols)} volume placements")


def make_delta():
    """delta_hashit.npz: outputs of the reference's own DeltaTrackingIter<true>::hashit (oracle/ref_shim.cpp) for delta_cases()."""
    pyoracle.build()
    out = {}
    cases = delta_cases()
    n_found = 0
    for k, (maj, sr, org, d, t0, t1, taus) in enumerate(cases):
        rows = pyoracle.delta_hashit(maj, sr, org, d, t0, t1, taus, use_ref=True)
        assert rows is not None, "oracle/_ref/libovr_ref_kat.so missing (needs /root/reference)"
        out[f"case{k}"] = rows
        n_found += int(rows[:, 0].sum())
    out["n"] = np.int32(len(cases))
    np.savez_compressed(os.path.join(HERE, "delta_hashit.npz"), **out)
    print(f"wrote delta_hashit.npz: {len(cases)} rays, {n_found} tentative collisions")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "delta":
        make_delta()
    elif len(sys.argv) > 1 and sys.argv[1] == "helpers":
        make_shader_helpers()
    elif len(sys.argv) > 1 and sys.argv[1] == "march":
        make_march_ref()
    elif len(sys.argv) > 1 and sys.argv[1] == "path":
        make_path_ref()
    elif len(sys.argv) > 1 and sys.argv[1] == "march_spp":
        make_march_spp()
    elif len(sys.argv) > 1 and sys.argv[1] == "ranges":
        make_value_ranges()
    elif len(sys.argv) > 1 and sys.argv[1] == "scenes":
        make_scene_dump()
    elif len(sys.argv) > 1 and sys.argv[1] == "host":
        make_host_derivations()
    elif len(sys.argv) > 1 and sys.argv[1] == "sparse":
        make_sparse()
    elif len(sys.argv) > 1 and sys.argv[1] == "tfn":
        make_tfn()
    else:
        main()
        make_sparse()
        make_tfn()
        make_delta()
        make_shader_helpers()
        make_march_ref()
        make_path_ref()
        make_host_derivations()
        make_march_spp()
        make_value_ranges()
        make_scene_dump()
