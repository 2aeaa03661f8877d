"""Host-side scene types (the subset of ovr::scene the single-volume devices accept, ovr/scene.h:201-371,413-426) and
the synthetic workloads of SURVEY.md section 8(d). Pure numpy; no device code and no oracle here."""
from dataclasses import dataclass, field

import numpy as np

NOISE_SEED = 1337  # the reference's own RNG seed (ovr/common/generate_mask.cu:35)


@dataclass
class Camera:  # ovr::scene::Camera (ovr/scene.h:201-222)
    frm: tuple = (0.0, 0.0, -1.0)
    at: tuple = (0.0, 0.0, 0.0)
    up: tuple = (0.0, 1.0, 0.0)
    fovy: float = 60.0


@dataclass
class TransferFunction:  # ovr::scene::TransferFunction: float4 colours, float opacities, value range
    color: np.ndarray = None      # [Nc, 4] float32 (w = 1)
    opacity: np.ndarray = None    # [Na] float32
    value_range: tuple = (1.0, -1.0)


@dataclass
class Volume:  # scene::Volume::structured_regular (ovr/scene.h:262-282)
    data: np.ndarray = None       # [nz, ny, nx], any ValueType dtype; None with `noise` set = generate on device
    grid_origin: tuple = (0.0, 0.0, 0.0)
    grid_spacing: tuple = (1.0, 1.0, 1.0)
    noise: dict = None            # {"seed": int, "org": (x,y,z), "dims": (nx,ny,nz)}

    @property
    def dims(self):
        if self.data is not None:
            nz, ny, nx = self.data.shape
            return (nx, ny, nz)
        return tuple(self.noise["dims"])

    def obj_to_world(self):
        """translate(grid_origin) * scale(grid_spacing * dims) as 3x4 row-major (device_impl.cpp:288-296)."""
        d = self.dims
        s = [np.float32(self.grid_spacing[i]) * np.float32(d[i]) for i in range(3)]
        o = self.grid_origin
        return np.array([s[0], 0, 0, o[0], 0, s[1], 0, o[1], 0, 0, s[2], o[2]], np.float32)


@dataclass
class Scene:  # ovr::scene::Scene reduced to what parse_single_volume_scene accepts
    volume: Volume = None
    tfn: TransferFunction = None
    camera: Camera = field(default_factory=Camera)
    volume_sampling_rate: float = 1.0


# ---- transfer functions -------------------------------------------------------------------------------------------

RAINBOW_CONTROLS = [  # tfn::TransferFunctionCore::fromRainbowMap / the bundled scenes' colorControls
    (0.0, (0.0, 0.36470600962638855, 1.0)),
    (0.1666666716337204, (0.0, 1.0, 0.97647100687026978)),
    (0.3333333432674408, (0.0, 1.0, 0.10588199645280838)),
    (0.5, (0.96862697601318359, 1.0, 0.0)),
    (0.66666668653488159, (1.0, 0.49019598960876465, 0.0)),
    (0.83333331346511841, (1.0, 0.0, 0.0)),
    (1.0, (0.66274499893188477, 0.0, 1.0)),
]


def rasterize_color_controls(controls=RAINBOW_CONTROLS, n=1024):
    """Colour control points -> n float4 entries (w = 1): tfn::TransferFunctionCore::updateColorMap
    (extern/tfn/core.h:598-634) restated -- entry i is evaluated at the cell centre (i + 0.5) / n in fp32, between the
    controls found by upper_bound, clamped to the end controls outside their span."""
    f32 = np.float32
    ctl = sorted(controls, key=lambda c: c[0])
    pos = np.array([c[0] for c in ctl], f32)
    rgb = np.array([c[1] for c in ctl], f32)
    value = (np.arange(n, dtype=f32) + f32(0.5)) / f32(n)
    ub = np.searchsorted(pos, value, side="right")          # first control with position > value
    lo = np.clip(ub - 1, 0, len(ctl) - 1)
    hi = np.clip(ub, 0, len(ctl) - 1)
    span = np.abs(pos[hi] - pos[lo])
    w = np.where((ub <= 0) | (ub >= len(ctl)), f32(0), np.abs(value - pos[lo]) / np.where(span > 0, span, f32(1))).astype(f32)
    out = np.ones((n, 4), f32)
    out[:, :3] = rgb[lo] + w[:, None] * (rgb[hi] - rgb[lo])          # mix(x, y, a) = x + a * (y - x), core.h:77-81
    out[ub <= 0, :3] = rgb[0]
    out[ub >= len(ctl), :3] = rgb[-1]
    return out


def alpha_preset(name, n=1024):
    u = np.arange(n, dtype=np.float64) / (n - 1)
    if name == "sparse":
        a = 0.6 * np.clip((u - 0.60) / 0.25, 0.0, 1.0)
    elif name == "dense":
        a = 0.05 + 0.25 * u
    elif name == "pt":
        a = 0.02 + 0.3 * u
    elif name == "ramp":
        a = np.clip((u - 0.3) / 0.7, 0.0, 1.0) * 0.5
    else:
        raise ValueError(name)
    return a.astype(np.float32)


def make_tfn(preset="sparse", n=1024, value_range=(0.0, 1.0)):
    return TransferFunction(rasterize_color_controls(n=n), alpha_preset(preset, n), tuple(value_range))


# ---- cameras --------------------------------------------------------------------------------------------------------

def camera_preset(name, dims):
    n = np.array(dims, np.float64)
    c = n / 2.0
    big = float(max(dims))
    if name == "front":  # same framing as scene_vorts1.json (eye z = 400 for N = 128)
        frm = c + np.array([0.0, 0.0, 2.63 * big])
    elif name == "oblique":
        frm = c + big * np.array([1.1, 0.8, 1.5])
    elif name == "inside":
        frm = c + big * np.array([0.05, 0.1, 0.2])
    else:
        raise ValueError(name)
    return Camera(tuple(float(v) for v in frm), tuple(float(v) for v in c), (0.0, 1.0, 0.0), 45.0)


# ---- synthetic workloads (BASELINE.json configs) ---------------------------------------------------------

def noise_scene(n, tf="sparse", camera="oblique", rate=1.0, seed=NOISE_SEED, dims=None, org=(0, 0, 0)):
    """Scene over the on-device synthetic volume noise(n, seed). The TF value range must be set to the volume's exact
    [min, max] after the device has generated it (MainRenderer.init does this when value_range is empty)."""
    dims = tuple(dims) if dims is not None else (n, n, n)
    vol = Volume(data=None, noise={"seed": seed, "org": tuple(org), "dims": dims})
    return Scene(vol, make_tfn(tf, value_range=(1.0, -1.0)), camera_preset(camera, dims), rate)


WORKLOADS = {
    # name: (volume edge, width, height, tf preset, camera, sampling rate, mode, shading)
    "C2": dict(n=512, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="ref"),
    "C2-noshade": dict(n=512, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="none"),
    # C2 with sparse (foveated) sampling at the GUI's default focus (apps/main_app.cpp:122-124): ~9 % of the pixels per frame
    "C2-sparse": dict(n=512, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="ref", sparse=True),
    # C2 on 8-bit data (the bundled engine / teapot / bonsai scenes are UNSIGNED_BYTE): noise(512) quantised to u8, kept as an
    # 8-bit 3-D array like the reference does (array.cpp:300-306); B200VR_TEX_NATIVE8=0 renders the same voxels from an fp32 array
    "C2-u8": dict(n=512, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="ref", dtype="u8"),
    "C3": dict(n=1024, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="ref"),
    "C3-u8": dict(n=1024, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="ref", dtype="u8"),
    "C3-noshade": dict(n=1024, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="none"),
    "C4": dict(n=1024, width=1920, height=1080, tf="pt", camera="front", rate=1.0, mode="pathtrace", shading="ref"),
    # sort-last: `n` is the edge of ONE brick; N ranks render a (grid * n)^3-ish volume, 4K frame, no shading
    "C5": dict(n=2048, width=3840, height=2160, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="none", sortlast=True),
    "C5-mid": dict(n=1024, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="none", sortlast=True),
    "C5-small": dict(n=256, width=1920, height=1080, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="none", sortlast=True),
    "tiny": dict(n=64, width=160, height=96, tf="sparse", camera="oblique", rate=1.0, mode="raymarch", shading="ref"),
}
