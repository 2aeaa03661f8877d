"""CPU-side checks of the drop-in boundary: libb200vr.so loads, exports exactly what include/b200vr.h declares, refuses
to work without a GPU (no CPU fallback), and the host-side MainRenderer mirror keeps the reference's call semantics."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from helpers import ROOT, ovr_b200

capi = ovr_b200.capi


def header_symbols():
    text = open(os.path.join(ROOT, "include", "b200vr.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200vr_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = C.CDLL(capi.LIB_PATH)
    names = header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b200vr.h but not exported"
    assert sorted(capi.SYMBOLS) == names, "ctypes table and header drifted apart"
    assert capi.lib().b200vr_abi_version() == 4


def test_struct_layouts_match_header():
    assert C.sizeof(capi.Params) == 60
    assert C.sizeof(capi.Stats) == 56
    p = capi.Params()
    capi.lib().b200vr_default_params(C.byref(p))
    assert (p.sampling_rate, p.base, p.spp, p.mode, p.shading, p.skip_empty, p.max_num_scatters) == (1.0, 1.0, 1, 0, 1, 1, 24)
    assert abs(p.light_pos[1] - 2205.875) < 1e-3 and p.light_ambient == 1.0 and p.pt_child_instance_xfm == 1


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback_without_a_gpu():
    h = C.c_void_p()
    rc = capi.lib().b200vr_create(C.byref(h), 0)
    assert rc == capi.ENODEV and not h.value
    assert b"no CPU fallback" in capi.lib().b200vr_last_error(None)
    r = ovr_b200.create_renderer("b200")
    with pytest.raises(RuntimeError):
        r.init(ovr_b200.scenes.noise_scene(16))


def test_create_renderer_unknown_name_raises():
    with pytest.raises(RuntimeError):
        ovr_b200.create_renderer("optix7")


def test_transactional_setters_queue_until_commit():
    r = ovr_b200.create_renderer("b200")
    assert r.unsafe_get_fbsize() is None and r.unsafe_get_variance() == float("inf")
    r.set_fbsize((64, 32))
    assert r.unsafe_get_fbsize() is None          # queued, not committed
    assert r.params["fbsize"].update() and r.unsafe_get_fbsize() == (64, 32)
    assert not r.params["fbsize"].update()        # consumed
    r.set_transfer_function([0, 0, 1, 1, 0, 0], [0, 0.1, 1, 0.9], (0.0, 2.0))
    assert r.params["tfn"].update()
    c, o, rng = r.unsafe_get_tfn()
    assert c.shape == (6,) and o.shape == (4,) and rng == (0.0, 2.0)
    with pytest.raises(Exception):
        r.render()  # device not initialised: must raise, never silently fall back


def test_host_arithmetic_of_the_device_equals_the_reference_statements():
    """b200vr_host_camera / b200vr_host_integer_normalize run the very functions b200vr_set_camera_lookat and
    b200vr_set_transfer_function use (no GPU needed). tests/golden/host_derivations.npz holds the outputs of the reference's own
    statements compiled on the host (device_impl.cpp:128-136; array.h:64-106 for the element type the reference's texture has):
    bit-exact on 401 cameras and 1640 range bounds of every value type."""
    import importlib.util
    import numpy as np
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    G = np.load(os.path.join(ROOT, "tests", "golden", "host_derivations.npz"))
    L = capi.lib()
    fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
    cams, _ = mg.host_derivation_inputs()
    assert len(cams) == len(G["camera"]) == 401
    for c, want in zip(cams, G["camera"]):
        out = np.zeros(12, np.float32)
        assert L.b200vr_host_camera(fp(c[0]), fp(c[1]), fp(c[2]), C.c_float(c[3]), c[4], c[5], out.ctypes.data_as(C.POINTER(C.c_float))) == capi.OK
        assert np.array_equal(out.view(np.uint32), want.view(np.uint32)), c
    rows = mg.integer_normalize_inputs()
    assert len(rows) == len(G["norm"]) == 1640
    got = np.array([L.b200vr_host_integer_normalize(C.c_float(v), t) for t, v in rows], np.float32)
    assert np.array_equal(got.view(np.uint32), G["norm"].view(np.uint32))
    assert L.b200vr_host_camera(None, None, None, C.c_float(45), 0, 0, None) == capi.EINVAL
