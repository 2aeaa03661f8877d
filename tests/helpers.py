"""Shared test helpers: build oracle scene descriptions from ovr_b200 scenes, image metrics."""
import os
import sys
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import ovr_b200  # noqa: E402
from oracle import pyoracle  # noqa: E402


def oracle_desc(scene, width, height, spp=1, shading=pyoracle.SHADING_REF, volume=None, tf_range=None):
    """SimpleNamespace understood by pyoracle.OracleScene for an ovr_b200.Scene."""
    vol = scene.volume
    if volume is None:
        if vol.data is not None:
            volume = np.ascontiguousarray(vol.data, np.float32)
        else:
            nz = vol.noise
            volume = pyoracle.noise(max(nz["dims"]), nz["seed"], nz["org"], nz["dims"])
    if tf_range is None:
        tf_range = scene.tfn.value_range
        if tf_range[1] < tf_range[0]:
            tf_range = (float(volume.min()), float(volume.max()))
    cam = scene.camera
    return SimpleNamespace(
        volume=volume, obj_to_world=vol.obj_to_world(), tf_color=scene.tfn.color, tf_alpha=scene.tfn.opacity,
        tf_range=tf_range, step=np.float32(1.0) / np.float32(scene.volume_sampling_rate), base=1.0,
        cam=pyoracle.camera(cam.frm, cam.at, cam.up, cam.fovy, width, height), width=width, height=height, spp=spp,
        shading=shading)


def psnr(a, b, peak=1.0):
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)


def max_abs(a, b):
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))))


def sparse_noise_table(tile_xy, seed=1337):
    """Synthetic stand-in for the reference's embedded noise tables (blue_noise.h:60-78): values on the same 8-bit grid
    k / 255, laid out [y][x][t] with 64 time slices. tests/golden/make_golden.py feeds the same tables to the reference's
    mask kernels."""
    k = np.random.default_rng(seed + tile_xy).integers(0, 256, size=tile_xy * tile_xy * 64, dtype=np.uint32)
    return (k.astype(np.float32) / np.float32(255.0)).astype(np.float32)
