"""Closed-form and property checks of the CPU oracle (no reference golden exists for whole frames: the reference has no
tests)."""
from types import SimpleNamespace

import numpy as np
import pytest

from helpers import oracle_desc, ovr_b200, pyoracle

S = ovr_b200.scenes


def homogeneous_desc(n=32, value=0.5, alpha=0.1, w=48, h=32, rate=1.0, shading=pyoracle.SHADING_NONE):
    vol = np.full((n, n, n), value, np.float32)
    col = np.ones((16, 4), np.float32) * np.array([0.2, 0.5, 0.8, 1.0], np.float32)
    sc = S.Scene(S.Volume(vol), S.TransferFunction(col, np.full(16, alpha, np.float32), (0.0, 1.0)),
                 S.camera_preset("front", (n, n, n)), rate)
    return oracle_desc(sc, w, h, shading=shading)


def test_homogeneous_alpha_closed_form():
    """alpha = 1 - (1-a)^(k + (r)^..): k unit steps give 1-(1-a)^k; a short last step uses the corrected opacity."""
    a = 0.1
    d = homogeneous_desc(alpha=a)
    o = pyoracle.OracleScene(d)
    rgba, grad, cnt = o.render_march()
    cy, cx = d.height // 2, d.width // 2
    # central ray travels ~n world units through the box: opacity correction makes alpha = 1-(1-a)^L for any step split
    org, dr = d.cam["position"], d.cam["direction"]
    hit, t0, t1 = pyoracle.clip(org / 32.0, dr / 32.0)
    L = t1 - t0
    # the central pixel is half a pixel off axis, so its chord differs from the axial one by ~1e-3 relative
    assert hit and abs(rgba[cy, cx, 3] - (1 - (1 - a) ** L)) < 1e-3
    # un-premultiplied colour of a constant-colour medium is the colour itself
    assert np.allclose(rgba[cy, cx, :3], [0.2, 0.5, 0.8], atol=1e-5)
    assert cnt["shadow_samples"] == 0 and cnt["gradient_fetches"] == 0


def test_sampling_rate_invariance_of_corrected_opacity():
    a = 0.2
    img = [pyoracle.OracleScene(homogeneous_desc(alpha=a, rate=r)).render_march()[0] for r in (1.0, 2.0, 4.0)]
    m = img[0][..., 3] > 0
    assert m.any()
    assert np.abs(img[0][..., 3] - img[1][..., 3])[m].max() < 2e-3
    assert np.abs(img[0][..., 3] - img[2][..., 3])[m].max() < 2e-3


def test_flat_region_shades_black_q5():
    """Zero gradient -> NaN normal -> clamp01(NaN) = 0: homogeneous media render black with REF shading (A.7-q5)."""
    d = homogeneous_desc(shading=pyoracle.SHADING_REF)
    rgba, grad, cnt = pyoracle.OracleScene(d).render_march()
    m = rgba[..., 3] > 0
    assert m.any() and np.all(rgba[m][:, :3] == 0) and np.all(grad[m] == 0)
    assert cnt["shadow_samples"] > 0


def test_miss_pixels_are_zero_and_early_termination():
    sc = S.noise_scene(32, tf="dense", camera="front")
    d = oracle_desc(sc, 64, 40)
    d.tf_alpha = np.full(1024, 0.9, np.float32)
    rgba, _, cnt = pyoracle.OracleScene(d).render_march()
    assert np.all(rgba[0, 0] == 0)                       # corner ray misses the box
    hit = rgba[..., 3] > 0
    assert rgba[..., 3].max() >= 0.9999                  # ERT threshold reached
    assert cnt["primary_samples"] < hit.sum() * 8        # ~4 samples/ray at alpha 0.9 instead of 32


@pytest.mark.parametrize("shading", [pyoracle.SHADING_NONE, pyoracle.SHADING_REF])
@pytest.mark.parametrize("camera", ["oblique", "inside"])
def test_skipping_is_bit_exact(shading, camera):
    sc = S.noise_scene(48, tf="sparse", camera=camera)
    d = oracle_desc(sc, 72, 40, shading=shading)
    a0, g0, c0 = pyoracle.OracleScene(d, skip_empty=False).render_march()
    a1, g1, c1 = pyoracle.OracleScene(d, skip_empty=True).render_march()
    assert np.array_equal(a0.view(np.uint32), a1.view(np.uint32))
    assert np.array_equal(g0.view(np.uint32), g1.view(np.uint32))
    assert c1["primary_samples"] < c0["primary_samples"]


def test_majorants_bound_the_classified_opacity():
    vol = pyoracle.noise(40, 7, (0, 0, 0), (40, 33, 37))
    tf = S.make_tfn("sparse", value_range=(float(vol.min()), float(vol.max())))
    ranges = pyoracle.build_ranges(vol)
    maj = pyoracle.build_majorants(ranges, tf.opacity, *tf.value_range)
    assert ranges.shape == (3, 3, 3, 2) and (maj == 0).any() and (maj > 0).any()
    sc = S.Scene(S.Volume(vol), tf, S.camera_preset("front", (40, 33, 37)))
    o = pyoracle.OracleScene(oracle_desc(sc, 8, 8))
    rng = np.random.default_rng(3)
    pts = rng.uniform(0, 1, (20000, 3)).astype(np.float32)
    vals = pyoracle.sample_volume(vol, pts)
    alpha = o.sample_tf(vals)[:, 3]
    cx = np.minimum((pts[:, 0] * np.float32(40)).astype(int) // 16, 2)
    cy = np.minimum((pts[:, 1] * np.float32(33)).astype(int) // 16, 2)
    cz = np.minimum((pts[:, 2] * np.float32(37)).astype(int) // 16, 2)
    assert np.all(alpha <= maj[cz, cy, cx] + 1e-7)
    assert np.all(vals >= ranges[cz, cy, cx, 0]) and np.all(vals <= ranges[cz, cy, cx, 1])
    # the reference footprint (upper apron missing) is a subset of the conservative one
    ref = pyoracle.build_ranges(vol, pyoracle.MC_REFERENCE)
    assert np.all(ref[..., 0] >= ranges[..., 0]) and np.all(ref[..., 1] <= ranges[..., 1])


def test_texture_model_taps():
    vol = np.arange(4 * 3 * 2, dtype=np.float32).reshape(2, 3, 4)
    centres = np.array([[(x + .5) / 4, (y + .5) / 3, (z + .5) / 2] for z in range(2) for y in range(3) for x in range(4)], np.float32)
    assert np.array_equal(pyoracle.sample_volume(vol, centres), vol.ravel())      # voxel centres are exact
    edge = np.array([[0, 0, 0], [1, 1, 1]], np.float32)
    assert np.array_equal(pyoracle.sample_volume(vol, edge), [0.0, 23.0])        # clamp addressing
    mid = np.array([[0.5, 0.5 / 3, 0.25]], np.float32)                            # halfway between x=1 and x=2
    assert pyoracle.sample_volume(vol, mid)[0] == 1.5
    q = np.array([[(1.5 + 1 / 512) / 4, 0.5 / 3, 0.25]], np.float32)             # weight 1/512 rounds to 1/256 (RN) or 0
    assert pyoracle.sample_volume(vol, q, pyoracle.FILTER_FIXED8_RN)[0] == 1 + 1 / 256
    assert pyoracle.sample_volume(vol, q, pyoracle.FILTER_FIXED8_TRUNC)[0] == 1.0


def test_noise_is_deterministic_and_brick_consistent():
    full = pyoracle.noise(32, 1337)
    brick = pyoracle.noise(32, 1337, (8, 4, 16), (16, 20, 10))
    assert np.array_equal(full[16:26, 4:24, 8:24], brick)
    assert 0.0 <= full.min() and full.max() < 1.0 and full.std() > 0.05


def test_path_tracer_statistics():
    """Homogeneous medium, albedo (.2,.5,.8): alpha is 1 on the box; a ray with optical depth tau escapes unscattered
    with probability exp(-tau) and then carries radiance 0 (ambient only after >= 1 scatter)."""
    d = homogeneous_desc(n=16, alpha=0.05, w=24, h=16)
    o = pyoracle.OracleScene(d)
    acc = np.zeros((16, 24, 4))
    N = 64
    for f in range(1, N + 1):
        acc += o.render_path(frame_index=f)[0]
    acc /= N
    hit = acc[..., 3] > 0
    assert hit.any() and np.all(acc[hit][:, 3] == 1.0) and np.all(acc[~hit] == 0)
    c = acc[8, 12, :3]
    assert c[0] < c[1] < c[2] < 1.0 and c[2] > 0.15   # tinted by the albedo, bounded by the ambient radiance
    # chord ~16 world units at sigma 0.05: P(no collision) = exp(-0.8) ~ 0.45 -> E[blue] <= 0.55*0.8
    assert c[2] < 0.55 * 0.8 + 0.08
