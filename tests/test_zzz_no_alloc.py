"""GPU test of b200vr_sparse_samples_noalloc (include/b200vr.h): the allocation-free variant of the sparse-sample compaction,
whose per-block scratch lives in a static device array instead of cudaMalloc'ed memory of the handle. It has no counterpart in
the reference; what it must equal is the existing entry point -- the pixel list
of a sparse b200vr_render as b200vr_get_sparse_samples returns it -- pair for pair (index work: bit-exact), on two calls in a
row that share the scratch array. Collected last (a failure here hides nothing after it)."""
import ctypes as C

import numpy as np
import pytest

from helpers import ovr_b200, sparse_noise_table

pytestmark = pytest.mark.gpu
S = ovr_b200.scenes
capi = ovr_b200.capi


def _sparse_renderer(W, H, table, tile, focus, scale, base):
    r = ovr_b200.create_renderer("b200")
    r.shading = capi.SHADING_NONE
    r.set_fbsize((W, H))
    r.init(S.noise_scene(32, tf="sparse", camera="oblique"))
    r.set_sparse_noise(table, tile, 64)
    r.set_focus(focus, scale, base)
    r.set_sparse_sampling(True)
    r.commit()
    return r


def _noalloc(r, noise_dev, tile, W, H, frame_index, focus, scale, base, xy_dev, count_dev):
    fc = (C.c_float * 2)(float(focus[0]), float(focus[1]))
    return capi.lib().b200vr_sparse_samples_noalloc(r.handle, C.c_void_p(noise_dev.data_ptr()), tile, 64, W, H, frame_index, fc,
                                                    C.c_float(scale), C.c_float(base), C.c_void_p(xy_dev.data_ptr()),
                                                    xy_dev.shape[0], C.c_void_p(count_dev.data_ptr()))


def test_no_alloc_sparse_list_equals_the_render_path_on_two_calls_in_a_row():
    import torch
    tile, focus, scale, base = 16, (0.45, 0.6), 0.2, 0.1
    table = sparse_noise_table(tile)
    noise_dev = torch.from_numpy(table).cuda()
    # two frame sizes, the larger one first: the second call finds the first call's block counts in the shared scratch
    for W, H in ((333, 190), (203, 117)):
        r = _sparse_renderer(W, H, table, tile, focus, scale, base)
        xy_dev = torch.full((W * H, 2), -1, dtype=torch.int32, device="cuda")
        count_dev = torch.zeros(1, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        for call in range(2):   # consecutive frames: another time slice of the noise table, another pixel set
            r.render()
            want = r.sparse_samples()
            fi = r.last_stats["frame_index"]
            assert 0.05 * W * H < len(want) < 0.9 * W * H
            rc = _noalloc(r, noise_dev, tile, W, H, fi, focus, scale, base, xy_dev, count_dev)
            assert rc == capi.OK, capi.lib().b200vr_last_error(r.handle)
            r._ck(capi.lib().b200vr_wait(r.handle, None))   # the call itself does not wait: drain the handle's stream
            n = int(count_dev.cpu().numpy().view(np.uint32)[0])
            got = xy_dev[:n].cpu().numpy()
            assert n == len(want), (W, H, call)
            assert np.array_equal(got, want), (W, H, call)     # same pixels in the same (tile) order
            assert (xy_dev[n:] == -1).all()                    # nothing written beyond the count
            xy_dev.fill_(-1)
            torch.cuda.synchronize()
        r.close()


def test_no_alloc_rejects_what_does_not_fit_the_static_scratch():
    import torch
    r = _sparse_renderer(64, 32, sparse_noise_table(16), 16, (0.5, 0.5), 0.06, 0.07)
    noise_dev = torch.zeros(16 * 16 * 64, dtype=torch.float32, device="cuda")
    xy_dev = torch.zeros((64 * 32, 2), dtype=torch.int32, device="cuda")
    count_dev = torch.zeros(1, dtype=torch.int32, device="cuda")
    # 8192 x 4096 = 33.5 M pixels > B200VR_NOALLOC_MAX_ENTRIES (16.7 M): refused with a message that names the limit
    assert _noalloc(r, noise_dev, 16, 8192, 4096, 1, (0.5, 0.5), 0.06, 0.07, xy_dev, count_dev) == capi.EINVAL
    assert b"B200VR_NOALLOC_MAX_ENTRIES" in capi.lib().b200vr_last_error(r.handle)
    # output buffer smaller than the frame
    assert _noalloc(r, noise_dev, 16, 128, 64, 1, (0.5, 0.5), 0.06, 0.07, xy_dev, count_dev) == capi.EINVAL
    # and the handle still works afterwards
    assert _noalloc(r, noise_dev, 16, 64, 32, 1, (0.5, 0.5), 0.06, 0.07, xy_dev, count_dev) == capi.OK
    r._ck(capi.lib().b200vr_wait(r.handle, None))
    assert int(count_dev.cpu()[0]) == 64 * 32   # an all-zero noise table keeps every pixel (0 < p everywhere)
    r.close()
