"""Pins the CPU oracle against fixtures produced by the reference's own code compiled on the host
(tests/golden/ref_kat.npz from dda.h, RandomTEA and gdt; generator: tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from helpers import ovr_b200, pyoracle

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_kat.npz"))


def test_tea_matches_reference_bit_exact():
    for (v0, v1), want in zip(G["tea_seeds"], G["tea_draws"]):
        got, _ = pyoracle.tea(int(v0), int(v1), len(want))
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_tea_state_advances_in_place():
    a, st = pyoracle.tea(3, 5, 4)
    b, _ = pyoracle.tea(int(st[1, 0]), int(st[1, 1]), 2)  # restart from the state after draw 2
    assert np.array_equal(a[2:], b)
    assert (a >= 0).all() and (a <= 1).all()


@pytest.mark.parametrize("variant", ["dda3", "iter"])
def test_dda_cell_sequence_matches_reference(variant):
    n = int(G["dda_n"])
    assert n >= 60
    for k in range(n):
        cells, tt = pyoracle.dda(G[f"dda{k}_org"], G[f"dda{k}_dir"], float(G[f"dda{k}_t"][0]), float(G[f"dda{k}_t"][1]),
                                 G[f"dda{k}_grid"])
        want_c, want_t = G[f"dda{k}_{variant}_cells"], G[f"dda{k}_{variant}_t"]
        assert np.array_equal(cells, want_c), f"ray {k}"
        assert np.array_equal(tt.view(np.uint32), want_t.view(np.uint32)), f"ray {k}"
        # traversal sanity: contiguous segments, unit steps
        assert (np.abs(np.diff(cells, axis=0)).sum(axis=1) >= 1).all()
        assert np.allclose(tt[1:, 0], tt[:-1, 1])


def test_gdt_transforms_match_reference():
    """xfmPoint / xfmVector / xfmNormal / normalize / affine inverse vs gdt compiled from the reference tree."""
    for m, v, want in zip(G["xfm_m"], G["xfm_v"], G["xfm_out"]):
        got = pyoracle.xfm(m, v)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        M = m.reshape(3, 4).astype(np.float64)  # and the fixture itself is sane
        assert np.allclose(want[0:3], M[:, :3] @ v + M[:, 3], rtol=1e-5, atol=1e-5)


def test_scene_vorts1_fixture():
    S = np.load(os.path.join(os.path.dirname(__file__), "golden", "scene_vorts1.npz"))
    assert tuple(S["dims"]) == (128, 128, 128) and str(S["type"]) == "FLOAT"
    assert S["alpha"].shape == (1024,) and abs(float(S["alpha"].max()) - 0.49999) < 1e-4
    assert float(S["sample_distance"]) == 0.25 and len(S["color_controls"]) == 7


# ---- texture unit: outputs of the real B200 tex3D captured on the GPU box (tests/tools/tex_calibration.py, tests/tools/tex_probe.py)
HW = np.load(os.path.join(os.path.dirname(__file__), "golden", "tex_b200_hw.npz"))


@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_texture_model_reproduces_captured_b200_tex3d(tag):
    """The reference samples through hardware tex3D; the oracle's FILTER_HW model must return what the unit returned."""
    vol, pts, hw = HW[f"{tag}_vol"], HW[f"{tag}_pts"], HW[f"{tag}_tex3d"]
    got = pyoracle.sample_volume(vol, pts, pyoracle.FILTER_HW)
    ulp = np.abs(got.view(np.int32).astype(np.int64) - hw.view(np.int32).astype(np.int64))
    assert ulp.max() == 0, f"{int((ulp > 0).sum())} of {len(ulp)} fetches differ from the unit"   # bit for bit, every fetch
    # the CUDA-guide model (separable 8-bit weights) misses the unit by up to ~1.5 weight quanta on this random volume
    old = pyoracle.sample_volume(vol, pts, pyoracle.FILTER_FIXED8_RN)
    assert np.abs(old - hw).max() > 1e-3 > 1e3 * np.abs(got - hw).max()


HW2 = np.load(os.path.join(os.path.dirname(__file__), "golden", "tex_b200_hw2.npz"))


def _same_bits(a, b):
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    return (a.view(np.int32) == b.view(np.int32)) | ((a == 0) & (b == 0))


def test_oracle_texture_model_is_bit_exact_on_the_second_capture():
    """tests/tools/tex_capture2.py on the B200: tex3D on volumes of 16x12x10 and 37x5x64 voxels with values in [0, 1), around 1e-3,
    of both signs, and of both signs with exponents spread over 18 binades; tex1D on 15 tables (7 to 1024 entries, 1 and 4
    channels, the same four kinds of values), coordinates from -0.03 to 1.03. The fixed-point adder of oracle_sample.cpp
    (filter_group / filter_add / filter_result) reproduces EVERY fetch of non-negative data and every 1-D fetch bit for bit;
    3-D fetches that mix signs are exact in > 99.5 % of the cases and a few ulp off otherwise (cancellation)."""
    n3 = 0
    for k in range(int(HW2["n_vol"])):
        vol, pts, hw, name = HW2[f"v{k}_vol"], HW2[f"v{k}_pts"], HW2[f"v{k}_hw"], str(HW2[f"v{k}_name"])
        got = pyoracle.sample_volume(vol, pts, pyoracle.FILTER_HW)
        ok = _same_bits(got, hw)
        n3 += len(hw)
        if name in ("unit", "small"):
            assert ok.all(), (name, vol.shape, int((~ok).sum()))
        else:
            assert ok.mean() > (0.9995 if name == "signed" else 0.995), (name, vol.shape, ok.mean())
            assert np.abs(got - hw).max() <= 4e-6 * np.abs(vol).max()
    n1 = 0
    for k in range(int(HW2["n_tab"])):
        tab, u, hw = HW2[f"t{k}_table"], HW2[f"t{k}_u"], HW2[f"t{k}_hw"]
        ch = tab.shape[1]
        got = np.asarray(pyoracle.tex1d(tab[:, 0].copy() if ch == 1 else tab, u), np.float32).reshape(len(u), ch)
        assert _same_bits(got, hw).all(), (k, tab.shape)
        n1 += hw.size
    assert n3 >= 40000 and n1 >= 70000


def test_oracle_tap_weights_match_captured_b200_weights():
    """Eight one-hot 2x2x2 textures expose the unit's per-tap weights; fractions sit at quantum centres (k + 0.25) / 256."""
    frac, want = HW["tap_frac"].astype(np.float64), HW["tap_weights"].astype(np.int64)
    pts = (((frac + 0.25) / 256 + 0.5) / 2).astype(np.float32)
    for t in range(8):
        vol = np.zeros((2, 2, 2), np.float32)
        vol[t >> 2, (t >> 1) & 1, t & 1] = 1
        got = np.round(pyoracle.sample_volume(vol, pts, pyoracle.FILTER_HW).astype(np.float64) * 256).astype(np.int64)
        assert np.array_equal(got, want[:, t]), f"tap {t}"
    assert (want.sum(axis=1) == 256).all()


def test_sparse_mask_matches_reference_kernels():
    """ovr_oracle_sparse_mask against masks made by the reference's own generate_blue_noise_kernel + generate_sparse_samples
    (tests/golden/sparse_mask.npz, generator: make_golden.py sparse). Both run exp() on the host: bit-exact."""
    from helpers import sparse_noise_table
    M = np.load(os.path.join(os.path.dirname(__file__), "golden", "sparse_mask.npz"))
    assert int(M["n"]) >= 5
    kept = []
    for k in range(int(M["n"])):
        w, h, fi, tile = [int(v) for v in M[f"case{k}"]]
        cx, cy, sigma, base = [float(v) for v in M[f"case{k}_f"]]
        want = np.unpackbits(M[f"case{k}_keep"])[: w * h].astype(bool).reshape(h, w)
        keep, p = pyoracle.sparse_mask(w, h, fi, sparse_noise_table(tile), tile, 64, (cx, cy), sigma, base)
        assert np.array_equal(keep, want), f"case {k}"
        assert (p >= min(base, 1.0) - 1e-6).all() and (p <= 1.0 + 1e-6).all()
        kept.append(int(keep.sum()))
    assert kept[4] == 0 and 0 < kept[0] < 160 * 96 // 4  # empty list edge case; default focus keeps a small fraction


def test_scene_file_transfer_functions_match_the_reference_loader():
    """scene_io.transfer_function_table against tests/golden/tfn_tables.npz: the RGBA tables the reference's own
    tfn::loadTransferFunction (extern/tfn/core.h, compiled on the host by oracle/ref_tfn_shim.cpp) builds for every bundled
    scene config plus two synthetic objects (opacity controls, no alpha array). Alpha arrays and control colours are
    bit-exact; interpolated colours and Gaussians (exp / sqrt through numpy instead of libm's float versions) to 2 ulp."""
    import json

    from ovr_b200 import scene_io
    T = np.load(os.path.join(os.path.dirname(__file__), "golden", "tfn_tables.npz"))
    names = [str(n) for n in T["names"]]
    assert len(names) >= 20
    n_gauss = 0
    for name in names:
        jt = json.loads(str(T[name + "_json"]))
        want = T[name + "_table"]
        color, alpha = scene_io.transfer_function_table(jt)
        assert color.shape == (len(want), 4) and alpha.shape == (len(want),)
        assert np.abs(color[:, :3] - want[:, :3]).max() <= 3e-7, name
        assert np.abs(alpha - want[:, 3]).max() <= 3e-7, name
        if not jt.get("gaussianObjects") and not jt.get("opacityControl"):
            assert np.array_equal(alpha, want[:, 3]), name
        n_gauss += bool(jt.get("gaussianObjects"))
    assert n_gauss >= 10  # the bundled scenes that the alpha array alone does not describe


def test_delta_tracking_free_path_iteration_matches_the_reference_iterator():
    """The oracle's hashit (oracle_path.cpp) against tests/golden/delta_hashit.npz: (found, t, majorant, tau) of every call of
    the reference's own DeltaTrackingIter<true>::hashit (accel/spatial_partition.h:56-96, cut out and compiled on the host by
    oracle/ref_shim.cpp) along 36 seeded rays through grids with empty cells -- incl. its stale-t behaviour across skipped
    cells. Bit-exact."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(__file__), "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    D = np.load(os.path.join(os.path.dirname(__file__), "golden", "delta_hashit.npz"))
    cases = mg.delta_cases()
    assert len(cases) == int(D["n"]) >= 30
    found = skipped_cells = 0
    for k, (maj, sr, org, d, t0, t1, taus) in enumerate(cases):
        want = D[f"case{k}"]
        got = pyoracle.delta_hashit(maj, sr, org, d, float(t0), float(t1), taus)
        assert got.shape == want.shape, f"ray {k}"
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"ray {k}"
        found += int(want[:, 0].sum())
        skipped_cells += int((maj == 0).sum())
        hits = want[want[:, 0] == 1]
        assert (np.diff(hits[:, 1]) >= 0).all() and (hits[:, 1] >= t0).all() and (hits[:, 1] <= t1 + 1e-5).all()
    assert found > 200 and skipped_cells > 1000


def test_clip_and_sphere_sampling_match_the_reference_functions():
    """ovr_oracle_clip / the oracle's uniform_sphere against the reference's own intersect_box and uniform_sample_sphere
    (shaders_common.h:156-184, 339-354; cut out and compiled on the host) on seeded inputs incl. zero / denormal direction
    components and origins inside the box. Bit-exact."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(__file__), "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    H = np.load(os.path.join(os.path.dirname(__file__), "golden", "shader_helpers.npz"))
    rays, s = mg.shader_helper_cases()
    got = np.array([[float(h), t0, t1] for h, t0, t1 in (pyoracle.clip(*r) for r in rays)], np.float32)
    assert np.array_equal(got.view(np.uint32), H["box"].view(np.uint32))
    assert 100 < H["box"][:, 0].sum() < len(rays) - 100
    sph = pyoracle.uniform_sphere(s)
    assert np.array_equal(sph.view(np.uint32), H["sphere"].view(np.uint32))
    assert np.allclose(np.linalg.norm(sph.astype(np.float64), axis=1), 1.0, atol=1e-6)


def test_oracle_march_equals_the_reference_shader_code_bit_for_bit():
    """The oracle's ray march (oracle_march.cpp: primary march, gradient, shadow march, shading, compositing, final blend)
    against tests/golden/ref_march.npz -- whole frames computed by the REFERENCE's own raymarching() / raymarching_shadow()
    and their helpers, cut out of shaders_raymarching.cu / shaders_common.h and executed on the host by
    oracle/ref_march_shim.cpp (texture fetches through the oracle's model of the texture unit, which is itself pinned against
    B200 hardware fetches). Four scenes: sparse / ramp / dense TFs, camera outside and inside, rates 1, 2 and 0.7 (opacity
    correction active). Every pixel, colour, alpha and gradient layer: bit-exact."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(__file__), "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_march.npz"))
    assert int(G["n"]) == len(mg.MARCH_REF_CASES) >= 4
    for k, case in enumerate(mg.MARCH_REF_CASES):
        o, w, h = mg.march_ref_scene(case)
        rgba, grad, cnt = o.render_march()
        assert np.array_equal(rgba.view(np.uint32), G[f"rgba{k}"].view(np.uint32)), f"case {k} {case}"
        assert np.array_equal(grad.view(np.uint32), G[f"grad{k}"].view(np.uint32)), f"case {k} {case}"
        assert (rgba[..., 3] > 0).mean() > 0.1 and cnt["shadow_samples"] > cnt["primary_samples"] > 0
        # the bit-exact empty-space skipping of the oracle (what the GPU kernels are compared with) changes nothing
        o2, _, _ = mg.march_ref_scene(case)
        o2.s.skip_empty = 1
        rgba2, grad2, _ = o2.render_march()
        assert np.array_equal(rgba2.view(np.uint32), rgba.view(np.uint32)) and np.array_equal(grad2.view(np.uint32), grad.view(np.uint32))


def test_reference_march_shim_live_and_threaded_equals_the_golden_frames():
    """Where oracle/_ref exists (this container; it ships prebuilt to the GPU box): the reference's own march functions run
    live -- single loop and ref_march_rays_mt on all host threads, the timed CPU arm of `bench.py --impl reference` --
    reproduce tests/golden/ref_march.npz bit for bit."""
    import importlib.util
    if not os.path.exists(os.path.join(os.path.dirname(pyoracle.__file__), "_ref", "libovr_ref_march.so")):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(__file__), "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_march.npz"))
    for k, case in enumerate(mg.MARCH_REF_CASES[:2]):
        o, w, h = mg.march_ref_scene(case)
        ys, xs = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
        px = np.stack([xs.ravel(), ys.ravel()], 1)
        for threads in (1, 0):
            rgba, grad = o.ref_march(px, threads=threads)
            assert np.array_equal(rgba.reshape(h, w, 4).view(np.uint32), G[f"rgba{k}"].view(np.uint32)), (k, threads)
            assert np.array_equal(grad.reshape(h, w, 3).view(np.uint32), G[f"grad{k}"].view(np.uint32)), (k, threads)


def _make_golden_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(os.path.dirname(__file__), "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    return mg


def test_oracle_path_tracer_equals_the_reference_shader_code_bit_for_bit():
    """The oracle's delta-tracking path tracer (oracle_path.cpp: TEA streams, pixel jitter, macrocell free-path iteration,
    TF-classified tentative collisions, uniform sphere scattering, scatter_index bookkeeping, ambient term, albedo product)
    against tests/golden/ref_path.npz -- whole frames computed by the REFERENCE's own delta_tracking() / pathtracing() /
    render_pathtracing(), DeltaTrackingIter<true> and RandomTEA, cut out of its sources and executed on the host by
    oracle/ref_march_shim.cpp. Both clips of the scattered rays: x1 = through the instance transform a second
    time, as OptiX does with the object-space ray pathtracing() hands it (the reference's behaviour), x0 = clipped against the
    volume (the B200 kernels' default). spp 1, 2 and 4 (pixel jitter on), frame indices 1, 3 and 7, volume sizes that are and
    are not multiples of the 16-voxel macrocell. Every pixel bit-exact (host logf / sincosf on both sides)."""
    mg = _make_golden_module()
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_path.npz"))
    assert int(G["n"]) == len(mg.PATH_REF_CASES) >= 4
    differ = []
    for k, case in enumerate(mg.PATH_REF_CASES):
        for x in (0, 1):
            o, w, h, fi = mg.path_ref_scene(case, x)
            rgba, cnt = o.render_path(frame_index=fi)
            assert np.array_equal(rgba.view(np.uint32), G[f"rgba{k}_x{x}"].view(np.uint32)), f"case {k} {case} x{x}"
            assert cnt["delta_samples"] > cnt["rays"] > 0 and rgba[..., :3].max() > 0
        differ.append(int((G[f"rgba{k}_x0"].view(np.uint32) != G[f"rgba{k}_x1"].view(np.uint32)).any(-1).sum()))
    # q14 shows exactly where the last macrocells overhang the volume: 32^3 no, 20^3 and 28^3 yes
    assert differ[0] == 0 and differ[1] == 0 and differ[2] > 50 and differ[3] > 50


def test_reference_path_shim_live_equals_the_golden_frames():
    """Where oracle/_ref exists: the reference's path-tracing functions run live reproduce tests/golden/ref_path.npz."""
    if not os.path.exists(os.path.join(os.path.dirname(pyoracle.__file__), "_ref", "libovr_ref_march.so")):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    mg = _make_golden_module()
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_path.npz"))
    for k in (1, 2):
        for x in (0, 1):
            o, w, h, fi = mg.path_ref_scene(mg.PATH_REF_CASES[k], x)
            ys, xs = np.mgrid[0:h, 0:w]
            for threads in (1, 0):  # 0: ref_path_rays_mt on all host threads, the timed CPU arm of bench.py's path-tracing workloads
                r = o.ref_path(np.stack([xs.ravel(), ys.ravel()], 1), frame_index=fi, threads=threads).reshape(h, w, 4)
                assert np.array_equal(r.view(np.uint32), G[f"rgba{k}_x{x}"].view(np.uint32)), (k, x, threads)


def test_host_derivations_equal_the_reference_statements():
    """Camera set-up of commit() (device_impl.cpp:128-136; its tan(float) is the float overload) and the instance matrix of
    buildScene() (:288-296) with the gdt inverse the shaders rebuild: the oracle's ovr_oracle_camera (same expression as
    b200vr.cu derive_camera) bit-exact on 401 cameras; Volume.obj_to_world + the oracle's affine inverse equal on 200 placements
    (numerically: the reference's translate * scale product turns a -0 origin into +0)."""
    mg = _make_golden_module()
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "host_derivations.npz"))
    cams, vols = mg.host_derivation_inputs()
    assert len(cams) == len(G["camera"]) and len(vols) == len(G["matrix"])
    for c, want in zip(cams, G["camera"]):
        got = pyoracle.camera(*c)
        got = np.concatenate([np.asarray(got[k], np.float32) for k in ("position", "direction", "horizontal", "vertical")])
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), c
    for (org, sp, dims), want in zip(vols, G["matrix"]):
        v = ovr_b200.scenes.Volume(noise={"seed": 1, "org": (0, 0, 0), "dims": tuple(int(d) for d in dims)},
                                   grid_origin=tuple(org), grid_spacing=tuple(sp))
        m = np.asarray(v.obj_to_world(), np.float32)
        inv = pyoracle.xfm(m, np.zeros(3, np.float32))[12:24]
        assert np.array_equal(m, want[:12]) and np.array_equal(inv, want[12:]), (org, sp, dims)


def test_oracle_march_with_pixel_jitter_equals_the_reference_shader_code():
    """spp 2 / 3 / 4, frame indices 1 / 5 / 2: the oracle's raygen loop (TEA(frame_index, pixel_index), jitter, mean over spp,
    background blend) against frames from the reference's render_raymarching() + raymarching() under the same loop
    (tests/golden/ref_march_spp.npz); every pixel and layer bit-exact."""
    mg = _make_golden_module()
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_march_spp.npz"))
    assert int(G["n"]) == len(mg.MARCH_SPP_CASES)
    for k, case in enumerate(mg.MARCH_SPP_CASES):
        o, w, h, fi = mg.march_spp_scene(case)
        rgba, grad, cnt = o.render_march(frame_index=fi)
        assert cnt["rays"] == w * h * case[4]
        assert np.array_equal(rgba.view(np.uint32), G[f"rgba{k}"].view(np.uint32)), case
        assert np.array_equal(grad.view(np.uint32), G[f"grad{k}"].view(np.uint32)), case


def test_macrocell_value_ranges_against_the_reference_kernel():
    """tests/golden/value_ranges.npz holds the output of the reference's value_range_kernel (accel/sp_singlemc.cu:10-54) run on the
    host for five volumes (ragged, one-voxel-thin, smaller than a cell). The oracle's MC_REFERENCE mode reproduces it exactly;
    the footprint the B200 device builds (MC_CONSERVATIVE, bit-exact with the device in the GPU tests) contains it in every
    cell, so its majorants are never below the reference's."""
    mg = _make_golden_module()
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "value_ranges.npz"))
    for k, dims in enumerate(mg.VALUE_RANGE_DIMS):
        vol = mg.value_range_volume(dims)
        want = G[f"r{k}"]
        assert np.array_equal(pyoracle.build_ranges(vol, pyoracle.MC_REFERENCE), want), dims
        cons = pyoracle.build_ranges(vol, pyoracle.MC_CONSERVATIVE)
        assert (cons[..., 0] <= want[..., 0]).all() and (cons[..., 1] >= want[..., 1]).all(), dims


def test_scene_io_reads_what_the_reference_loader_reads(tmp_path):
    """tests/golden/scene_dump.json = the reference's own scene loader (ovr::scene::create_scene, compiled from /root/reference
    into build/ovr/scenedump behind oracle/ref_scene_dump.cpp) on eight synthetic VIDI3D scenes: every value type, big-endian
    payloads behind a file offset, `scales`, normalised and unnormalised mapping ranges, alpha ends below 0.01. scene_io, the
    Python mirror of that loader, must hand over the same voxels (FNV-1a of the bytes, type, dims, min / max), grid spacing,
    transfer-function tables and value range, camera and sampling rate."""
    import json
    mg = _make_golden_module()
    rows = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "scene_dump.json")))
    cases = mg.scene_dump_cases(str(tmp_path))
    assert [r["file"] for r in rows] == [n for n, _ in cases] and len(rows) == 8
    f32 = lambda x: np.asarray(x, np.float32)
    for (name, path), want in zip(cases, rows):
        sc = ovr_b200.scene_io.read_vidi3d_scene(path)
        data = sc.volume.data
        assert [data.shape[2], data.shape[1], data.shape[0]] == want["dims"] and data.dtype.itemsize == want["type_bytes"], name
        h = 1469598103934665603
        for b in data.tobytes():
            h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        assert f"{h:016x}" == want["fnv"], name
        assert float(data.min()) == want["min"] and float(data.max()) == want["max"], name
        assert np.array_equal(f32(sc.volume.grid_spacing), f32(want["spacing"])) and np.array_equal(f32(sc.volume.grid_origin), f32(want["origin"])), name
        assert np.array_equal(f32(sc.tfn.value_range), f32(want["tf_range"])), (name, sc.tfn.value_range, want["tf_range"])
        assert np.float32(sc.volume_sampling_rate) == np.float32(want["rate"]), name
        assert np.array_equal(f32(sc.camera.frm), f32(want["eye"])) and np.array_equal(f32(sc.camera.at), f32(want["center"])), name
        assert np.array_equal(f32(sc.camera.up), f32(want["up"])) and np.float32(sc.camera.fovy) == np.float32(want["fovy"]), name
        assert np.array_equal(f32(sc.tfn.opacity), f32(want["opacity"])), name
        assert np.allclose(f32(sc.tfn.color).ravel(), f32(want["color"]), rtol=0, atol=1e-6), name
