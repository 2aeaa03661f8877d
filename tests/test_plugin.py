"""The drop-in boundary: libdevice_b200.so (DeviceB200 : ovr::MainRenderer, compiled against the reference's unchanged
headers) driven by the reference's own apps/main_batch.cpp (build/ovr/renderbatch, compiled from /root/reference by
open-volume-renderer_b200/device/build.py; the binaries travel to the GPU box prebuilt)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from helpers import ROOT, max_abs, oracle_desc, ovr_b200, psnr, pyoracle

OVR = os.path.join(ROOT, "build", "ovr")
PLUGIN = os.path.join(OVR, "libdevice_b200.so")
BATCH = os.path.join(OVR, "renderbatch")
S = ovr_b200.scenes
needs_build = pytest.mark.skipif(not (os.path.exists(PLUGIN) and os.path.exists(BATCH)),
                                 reason="build/ovr missing (needs the reference tree at build time)")


def write_scene(tmp_path, n=48, tf="ramp", cam="oblique", sample_distance=1.0):
    vol = pyoracle.noise(n)
    path = ovr_b200.scene_io.write_vidi3d_scene(str(tmp_path / "scene.json"), vol, S.alpha_preset(tf), S.RAINBOW_CONTROLS,
                                                S.camera_preset(cam, (n, n, n)), sample_distance)
    return path, vol


def run_batch(scene, exp, size, extra=(), env_extra=None):
    env = dict(os.environ, LD_LIBRARY_PATH=OVR + os.pathsep + os.environ.get("LD_LIBRARY_PATH", ""))
    env.update(env_extra or {})
    cmd = [BATCH, "--scene", scene, "--device", "b200", "--num-frames", "1", "--fbsize", f"({size[0]},{size[1]})", "--exp", exp]
    return subprocess.run(cmd + list(extra), capture_output=True, text=True, env=env, timeout=600)


def test_scene_file_round_trip(tmp_path):
    path, vol = write_scene(tmp_path, n=20)
    sc = ovr_b200.scene_io.read_vidi3d_scene(path)
    assert sc.volume.dims == (20, 20, 20) and np.array_equal(sc.volume.data, vol)
    assert sc.tfn.value_range == (float(vol.min()), float(vol.max())) and sc.volume_sampling_rate == 1.0
    assert sc.tfn.opacity[0] == 0.0 and sc.tfn.color.shape == (1024, 4) and sc.camera.fovy == 45.0
    # colour rasterisation = updateColorMap: cell-centred, clamped outside the controls
    c = sc.tfn.color
    assert np.allclose(c[0, :3], [0.0, 0.36470601 + (0.5 / 1024) / 0.16666667 * (1 - 0.36470601), 1.0 - (0.5 / 1024) / 0.16666667 * (1 - 0.97647101)], atol=1e-6)
    assert (c[:, 3] == 1).all()


@needs_build
def test_plugin_exports_the_factory_symbol():
    """create_renderer("b200") resolves `ovr_create_renderer__b200` (ovr/common/dylink/ObjectFactory.h:51-64,77-85)."""
    lib = C.CDLL(PLUGIN, mode=C.RTLD_LOCAL)
    assert hasattr(lib, "ovr_create_renderer__b200")


@needs_build
@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="GPU present")
def test_renderbatch_without_a_gpu_fails_loudly(tmp_path):
    path, _ = write_scene(tmp_path, n=16)
    r = run_batch(path, str(tmp_path / "out"), (32, 32))
    assert r.returncode != 0
    assert "no CUDA device" in r.stderr and "no CPU fallback" in r.stderr


@needs_build
@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["raymarch", "raymarch-rate2"])
def test_renderbatch_drives_the_b200_device(tmp_path, mode):
    """apps/main_batch.cpp: 5 warm-up + 25 timed frames with frame accumulation, then a PNG. The PNG must be what the
    Python host mirror and the oracle produce for the same scene file."""
    from PIL import Image
    W, H, n = 200, 120, 48
    path, vol = write_scene(tmp_path, n=n, sample_distance=0.5 if mode.endswith("rate2") else 1.0)
    exp = str(tmp_path / "out")
    # main_batch always calls set_volume_sampling_rate(--sampling-rate, default 1), which overrides the scene file's
    # 1 / sampleDistance in commit (main_batch.cpp:260, device_impl.cpp:189-195)
    r = run_batch(path, exp, (W, H), extra=("--sampling-rate", "2") if mode.endswith("rate2") else ())
    assert r.returncode == 0, r.stderr[-2000:]
    assert "fps = " in r.stdout
    png = np.asarray(Image.open(exp + "000000.png").convert("RGBA"), np.uint8)
    assert png.shape == (H, W, 4)
    sc = ovr_b200.scene_io.read_vidi3d_scene(path)
    # main_batch passes Camera{from, at, up}: fovy falls back to the struct default 60 (ovr/scene.h:219, main_batch.cpp:268)
    sc.camera = S.Camera(sc.camera.frm, sc.camera.at, sc.camera.up, 60.0)
    o_rgba = pyoracle.OracleScene(oracle_desc(sc, W, H), skip_empty=False).render_march()[0]
    want = (np.clip(o_rgba, 0.0, 1.0) * 255.0).astype(np.uint8)[::-1]     # save_image: truncation + vertical flip
    diff = np.abs(png.astype(np.int32) - want.astype(np.int32))
    print(f"renderbatch {mode}: png vs oracle max level diff {diff.max()}, differing pixels {(diff.max(axis=-1) > 0).mean():.4f}")
    assert (o_rgba[..., 3] > 0).mean() > 0.05
    assert diff.max() <= 1 and (diff.max(axis=-1) > 0).mean() < 0.02
