"""Host-side logic that needs no GPU: tile ownership, synthetic scene definitions, TF rasterisation."""
import numpy as np

from helpers import ovr_b200

S = ovr_b200.scenes


def super_tiles(w, h):
    return (w + 63) // 64, (h + 31) // 32


def owned(rank, nranks, n):
    return len(range(rank, n, nranks))


def test_tile_partition_covers_every_pixel_once():
    for (w, h) in [(1920, 1080), (100, 70), (64, 32)]:
        sw, sh = super_tiles(w, h)
        for nranks in (1, 2, 3, 8):
            cover = np.zeros((sh * 32, sw * 64), np.int32)
            for rank in range(nranks):
                n_owned = owned(rank, nranks, sw * sh)
                for k in range(n_owned * 64):  # warp tiles: same arithmetic as warp_tile_origin (launch_params.cuh)
                    m, r = divmod(k, 64)
                    s = rank + m * nranks
                    b, i = divmod(r, 16)  # 2 x 2 blocks of 4 x 4 tiles
                    px = (s % sw) * 64 + ((b & 1) * 4 + (i & 3)) * 8
                    py = (s // sw) * 32 + ((b >> 1) * 4 + (i >> 2)) * 4
                    cover[py:py + 4, px:px + 8] += 1
            assert (cover == 1).all()


def test_rainbow_rasterisation_and_presets():
    c = S.rasterize_color_controls(n=1024)
    assert c.shape == (1024, 4) and np.all(c[:, 3] == 1)
    # updateColorMap evaluates entry i at the cell centre (i + 0.5) / n (extern/tfn/core.h:606)
    w = (0.5 / 1024) / (1 / 6)
    assert np.allclose(c[0, :3], [0, 0.364706 + w * (1 - 0.364706), 1 + w * (0.976471 - 1)], atol=1e-6)
    assert np.allclose(c[-1, :3], [1 + (1 - w) * (0.662745 - 1), 0, (1 - w) * 1], atol=1e-5)
    a = S.alpha_preset("sparse")
    assert a[0] == 0 and a[613] == 0 and a[-1] == np.float32(0.6) and (np.diff(a) >= 0).all()
    assert S.alpha_preset("dense").min() > 0


def test_scene_geometry():
    sc = S.noise_scene(512)
    assert sc.volume.dims == (512, 512, 512)
    assert np.array_equal(sc.volume.obj_to_world(), np.array([512, 0, 0, 0, 0, 512, 0, 0, 0, 0, 512, 0], np.float32))
    cam = S.camera_preset("front", (128, 128, 128))
    assert np.allclose(cam.frm, (64, 64, 64 + 2.63 * 128)) and cam.fovy == 45.0
    v = S.Volume(np.zeros((4, 6, 8), np.uint8), grid_origin=(1, 2, 3), grid_spacing=(2, 2, 2))
    assert v.dims == (8, 6, 4)
    assert np.array_equal(v.obj_to_world(), np.array([16, 0, 0, 1, 0, 12, 0, 2, 0, 0, 8, 3], np.float32))


def test_sparse_enumeration_visits_every_pixel_once():
    """sparse.cuh lists the kept pixels in the order the render kernels walk the image (super tile -> 2 x 2 blocks of 4 x 4 warp
    tiles -> 8 x 4 pixels), not row-major: same arithmetic as sparse_entry_pixel, checked to be a bijection onto the image plus
    padding beyond the right / top edge, with the 32 entries of a warp tile forming one 8 x 4 block."""
    for (w, h) in [(203, 117), (1920, 1080), (64, 32), (7, 3)]:
        sw, sh = super_tiles(w, h)
        e = np.arange(sw * sh * 2048)
        s, r = e // 2048, e % 2048
        t, lane = r >> 5, r & 31
        b, i = t // 16, t % 16
        x = (s % sw) * 64 + ((b & 1) * 4 + (i & 3)) * 8 + (lane & 7)
        y = (s // sw) * 32 + ((b >> 1) * 4 + (i >> 2)) * 4 + lane // 8
        inside = (x < w) & (y < h)
        flat = (x + y * w)[inside]
        assert len(flat) == w * h and len(np.unique(flat)) == w * h
        xt, yt = x.reshape(-1, 32), y.reshape(-1, 32)
        assert (xt.max(1) - xt.min(1) == 7).all() and (yt.max(1) - yt.min(1) == 3).all()


def test_every_bundled_scene_config_parses():
    """the reference's data/configs/*.json (u8 / u16 / f32 volumes, optional `scales`, BASE64 alpha arrays) go
    through read_vidi3d_scene with the serializer's derived values (serializer_vidi3d.cpp:203-412). The raw payloads are not
    shipped, so a zero volume of the declared type and size stands in (configs above 64 M voxels: header fields only).
    Runs where the reference tree is mounted; the GPU box has no copy of it."""
    import glob

    import pytest
    cfgs = sorted(glob.glob("/root/reference/data/configs/scene_*.json"))
    if not cfgs:
        pytest.skip("reference tree not present")
    from ovr_b200 import scene_io
    seen = set()
    for path in cfgs:
        ds = scene_io.load_json_with_comments(path)["dataSource"][0]
        d = ds["dimensions"]
        n = int(d["x"]) * int(d["y"]) * int(d["z"])
        dtype = np.dtype(scene_io._TYPES[ds["type"]])
        seen.add(ds["type"])
        if n > 64 * 2 ** 20:
            continue
        sc = scene_io.read_vidi3d_scene(path, payload=np.zeros(n, dtype))
        assert sc.volume.dims == (int(d["x"]), int(d["y"]), int(d["z"])) and sc.volume.data.dtype == dtype
        assert sc.tfn.color.shape[1] == 4 and len(sc.tfn.opacity) >= 2 and sc.tfn.opacity[0] >= 0
        lo, hi = sc.tfn.value_range
        assert hi > lo and sc.volume_sampling_rate > 0 and sc.camera.fovy > 0
        if dtype.kind != "f":
            assert hi <= np.iinfo(dtype).max * 1.0001  # scalarMappingRange x type maximum (serializer_vidi3d.cpp:235-269)
    assert {"UNSIGNED_BYTE", "UNSIGNED_SHORT", "FLOAT"} <= seen


def test_scene_file_payload_offset_endianness_and_errors(tmp_path):
    """CreateArray3DScalarFromFile (ovr/scene.cpp:181-245): `offset` bytes are skipped, BIG_ENDIAN payloads are byte-swapped,
    a file shorter than offset + data raises; `scales` becomes the grid spacing (serializer_vidi3d.cpp:298-301)."""
    import json

    import pytest
    from ovr_b200 import scene_io
    data = (np.arange(4 * 5 * 6, dtype=np.uint16) * 517 % 65521).astype(np.uint16).reshape(6, 5, 4)
    cam = S.camera_preset("front", (4, 5, 6))
    path = scene_io.write_vidi3d_scene(str(tmp_path / "s.json"), data, S.alpha_preset("ramp", 16), S.RAINBOW_CONTROLS, cam,
                                       sample_distance=0.5)
    root = json.load(open(path))
    raw = root["dataSource"][0]["fileName"][0]
    with open(raw, "wb") as f:
        f.write(b"\x01" * 7 + data.astype(">u2").tobytes())
    root["dataSource"][0].update(offset=7, endian="BIG_ENDIAN", scales={"x": 1.0, "y": 2.0, "z": 0.5})
    json.dump(root, open(path, "w"))
    sc = scene_io.read_vidi3d_scene(path)
    assert np.array_equal(sc.volume.data, data) and sc.volume.data.dtype == np.uint16
    assert tuple(sc.volume.grid_spacing) == (1.0, 2.0, 0.5) and sc.volume_sampling_rate == 2.0
    # scalarMappingRange x type maximum, in fp32 like the serializer (serializer_vidi3d.cpp:235-269)
    assert sc.tfn.value_range == tuple(float(np.float32(65535) * np.float32(v)) for v in (data.min(), data.max()))
    root["dataSource"][0]["offset"] = 9
    json.dump(root, open(path, "w"))
    with pytest.raises(RuntimeError, match="File size does not match data size"):
        scene_io.read_vidi3d_scene(path)
    root["dataSource"][0]["fileName"] = ["/nonexistent/volume.raw"]
    json.dump(root, open(path, "w"))
    with pytest.raises(RuntimeError, match="Cannot find volume file"):
        scene_io.read_vidi3d_scene(path)
