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
    assert capi.lib().b200vr_abi_version() == 2


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
