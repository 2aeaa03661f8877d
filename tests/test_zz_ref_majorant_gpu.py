"""SURVEY 8 a16 pinned on the GPU: the reference's own majorant_kernel (ovr/devices/optix7/accel/sp_singlemc.cu:56-97), cut out of
the reference tree at build time and compiled unchanged by nvcc (oracle/ref_majorant_shim.cu -> oracle/_ref/libovr_ref_majorant.so,
test infrastructure, built where /root/reference exists and shipped prebuilt), runs on the B200 beside the product's majorant
kernel (csrc/build.cuh) and the CPU oracle's restatement on the same macrocell ranges and the same opacity tables. All three must
agree bit for bit. The kernel cannot run on the host: it converts a negative float to uint32_t, undefined in C++ and resolved
differently by x86 and by the GPU. Grids are chosen so that the cell count is a multiple of the table length: in a
partly filled last block the reference kernel reads table entries it never loaded (the other half of q7)."""
import ctypes as C

import numpy as np
import pytest

from helpers import ovr_b200, pyoracle

pytestmark = pytest.mark.gpu
S = ovr_b200.scenes
capi = ovr_b200.capi

CASES = [
    # dims (macrocells = dims / 16), opacity table length, TF preset, explicit TF window (None: the volume's own range)
    ((64, 64, 256), 256, "sparse", None),      # 4 x 4 x 16 = 256 cells, one block
    ((128, 128, 256), 1024, "sparse", None),   # 8 x 8 x 16 = 1024 cells, one block of 1024 threads
    ((128, 128, 256), 256, "ramp", None),      # four blocks
    ((64, 128, 128), 64, "dense", None),       # eight blocks; every cell non-empty
    ((128, 128, 256), 1024, "sparse", (0.3, 0.7)),   # window inside the data range: both clamps of the kernel are exercised
    ((64, 64, 256), 256, "ramp", (-2.0, 3.0)),       # window wider than the data
]


@pytest.mark.parametrize("dims,n_tf,tf,window", CASES, ids=[f"{c[0][0]}x{c[0][1]}x{c[0][2]}-tf{c[1]}-{c[2]}-{'win' if c[3] else 'auto'}" for c in CASES])
def test_majorants_equal_the_reference_kernel_run_on_the_gpu(dims, n_tf, tf, window):
    import torch
    ref = pyoracle.ref_majorant_lib()
    if ref is None:
        pytest.skip("oracle/_ref/libovr_ref_majorant.so not built (needs the reference tree at build time)")
    sc = S.noise_scene(max(dims), dims=dims, tf=tf, camera="oblique")
    sc.tfn = S.make_tfn(tf, n=n_tf, value_range=window if window else (1.0, -1.0))
    r = ovr_b200.create_renderer("b200")
    r.set_fbsize((32, 16))
    r.init(sc, sc.camera)
    r.commit()
    mc = (C.c_int32 * 3)()
    r._ck(capi.lib().b200vr_get_macrocells(r.handle, mc, None, None))
    n_cells = mc[0] * mc[1] * mc[2]
    assert n_cells % n_tf == 0, "q7: the reference kernel is only defined when the last block is full"
    ranges = np.empty((n_cells, 2), np.float32)
    maj = np.empty(n_cells, np.float32)
    r._ck(capi.lib().b200vr_get_macrocells(r.handle, mc, ranges.ctypes.data_as(C.c_void_p), maj.ctypes.data_as(C.c_void_p)))
    lo, hi = window if window else r.value_range
    alphas = np.ascontiguousarray(sc.tfn.opacity, np.float32)
    assert len(alphas) == n_tf
    d_ranges, d_alpha = torch.from_numpy(ranges).cuda(), torch.from_numpy(alphas).cuda()
    d_out = torch.full((n_cells,), -1.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    rc = ref.ref_majorant_kernel_run(d_ranges.data_ptr(), n_cells, d_alpha.data_ptr(), n_tf, float(lo), float(hi), d_out.data_ptr())
    assert rc == 0, f"reference majorant_kernel: CUDA error {rc}"
    want = d_out.cpu().numpy()
    assert (want >= 0).all()
    # product kernel == reference kernel, bit for bit, on the same ranges
    assert np.array_equal(maj.view(np.uint32), want.view(np.uint32)), f"{int((maj != want).sum())} of {n_cells} cells differ"
    # and the oracle's restatement (what the CPU tests and the march oracle use) == reference kernel
    o_maj = pyoracle.build_majorants(ranges.reshape(mc[2], mc[1], mc[0], 2), alphas, float(lo), float(hi)).ravel()
    assert np.array_equal(o_maj.view(np.uint32), want.view(np.uint32))
    if tf == "sparse":
        assert (want == 0).any()
    assert (want > 0).any()
    print(f"majorants {dims} tf {n_tf} {tf}: {n_cells} cells, {int((want == 0).sum())} empty, product == reference kernel == oracle")
    r.close()
