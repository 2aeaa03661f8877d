"""Python host mirror of the reference's device-plugin interface for this path: `ovr::MainRenderer`
(ovr/renderer.h:82-288) and the factory `create_renderer(name)` (ovr/renderer.cpp:42-61), driving the b200vr C ABI.

Same method names, argument meaning and error behaviour as the C++ interface: thread-safe setters queue values
(TransactionalValue, ovr/common/vidi_transactional_value.h:75-146), `commit()` applies only what changed and resets
accumulation, `render()` is synchronous and adds milliseconds to `render_time`, `mapframe()` hands out non-owning
device pointers, errors are RuntimeError (the C++ device throws std::runtime_error). The compiled C++ plugin
(device/device_b200.cpp -> libdevice_b200.so) is the same thing for OVR's own apps.
"""
import ctypes as C
import threading
import time

import numpy as np

from . import _capi as capi
from .scenes import Camera, Scene, TransferFunction  # noqa: F401  (re-exported)

_DTYPE_TO_TYPE = {np.dtype(np.uint8): capi.TYPE_UINT8, np.dtype(np.int8): capi.TYPE_INT8,
                  np.dtype(np.uint16): capi.TYPE_UINT16, np.dtype(np.int16): capi.TYPE_INT16,
                  np.dtype(np.uint32): capi.TYPE_UINT32, np.dtype(np.int32): capi.TYPE_INT32,
                  np.dtype(np.float32): capi.TYPE_FLOAT, np.dtype(np.float64): capi.TYPE_DOUBLE}


class _Transactional:
    """TransactionalValue<T>: producer assigns under a lock, consumer update() swaps the queued value in."""

    def __init__(self, value=None):
        self._lock = threading.Lock()
        self._queued = None
        self._has = False
        self._value = value

    def assign(self, v):
        with self._lock:
            self._queued, self._has = v, True

    def update(self):
        with self._lock:
            if not self._has:
                return False
            self._value, self._queued, self._has = self._queued, None, False
            return True

    def ref(self):
        return self._value

    get = ref


class FrameBufferData:
    """MainRenderer::FrameBufferData (ovr/renderer.h:89-97): two CrossDeviceBuffers; here (device pointer, nbytes)."""

    def __init__(self):
        self.rgba = (None, 0)
        self.grad = (None, 0)
        self.size = (0, 0)


def _f3(v):
    return (C.c_float * 3)(float(v[0]), float(v[1]), float(v[2]))


class MainRenderer:
    """DeviceB200: the B200-native OVR device."""

    _PARAM_NAMES = ("tfn", "fbsize", "sample_per_pixel", "volume_sampling_rate", "volume_density_scale", "ambient", "diffuse",
                    "specular", "shininess", "radius", "phi", "theta", "intensity", "focus_center", "focus_scale", "base_noise",
                    "add_lights", "sparse_sampling", "path_tracing", "photonmapping", "frame_accumulation", "camera")

    def __init__(self, cuda_device=0):
        self.render_time = 0.0           # ms, cumulative (ovr/renderer.h:87; device.cpp:34-43)
        self.variance = float("inf")     # optix7 sets 0 after the first frame (device_impl.cpp:266)
        self.current_scene = None
        self.params = {k: _Transactional() for k in self._PARAM_NAMES}
        self.last_stats = {}
        self._cuda_device = cuda_device
        self._h = None
        self._initialized = False
        # device-specific switches (the C++ plugin reads them from the environment: OVR_B200_SHADING, OVR_B200_SAMPLER, ...)
        self._options = {}               # tuning options set before init (b200vr_set_option)
        self.shading = capi.SHADING_REF
        self.sampler = capi.SAMPLER_TEX3D
        self.skip_empty = True
        self.pt_child_instance_xfm = True  # b200vr_params::pt_child_instance_xfm (the reference's scattered-ray clip, q14)
        self.premultiplied = False
        # sparse sampling state as the device holds it (defaults of apps/main_app.cpp:122-126)
        self._focus_center, self._focus_scale, self._base_noise, self._sparse = (0.5, 0.5), 0.06, 0.07, False

    # ---- thread-safe setters (ovr/renderer.h:135-248) ----
    def set_fbsize(self, fbsize):
        self.params["fbsize"].assign((int(fbsize[0]), int(fbsize[1])))

    def set_camera(self, camera_or_from, at=None, up=None):
        cam = camera_or_from if at is None else Camera(tuple(camera_or_from), tuple(at), tuple(up))
        self.params["camera"].assign(cam)

    def set_transfer_function(self, c, o, r):
        """c: 3*Nc floats RGB, o: 2*No floats (position, alpha), r: (lo, hi) -- MainRenderer::set_transfer_function."""
        self.params["tfn"].assign((np.asarray(c, np.float32).ravel().copy(), np.asarray(o, np.float32).ravel().copy(),
                                   (float(r[0]), float(r[1]))))

    def set_focus(self, center, scale, base_noise):
        self.params["focus_center"].assign(center)
        self.params["focus_scale"].assign(scale)
        self.params["base_noise"].assign(base_noise)

    def set_sample_per_pixel(self, spp):
        self.params["sample_per_pixel"].assign(int(spp))

    def set_sparse_sampling(self, v):
        self.params["sparse_sampling"].assign(bool(v))

    def set_path_tracing(self, v):
        self.params["path_tracing"].assign(bool(v))

    def set_frame_accumulation(self, v):
        self.params["frame_accumulation"].assign(bool(v))

    def set_volume_sampling_rate(self, v):
        self.params["volume_sampling_rate"].assign(float(v))

    # accepted and ignored, exactly like the optix7 device (no .update() for them in device_impl.cpp:113-197)
    def set_add_lights(self, v): self.params["add_lights"].assign(v)
    def set_photonmapping(self, v): self.params["photonmapping"].assign(v)
    def set_volume_density_scale(self, v): self.params["volume_density_scale"].assign(v)
    def set_mat_ambient(self, v): self.params["ambient"].assign(v)
    def set_mat_diffuse(self, v): self.params["diffuse"].assign(v)
    def set_mat_specular(self, v): self.params["specular"].assign(v)
    def set_mat_shininess(self, v): self.params["shininess"].assign(v)
    def set_light_phi(self, v): self.params["phi"].assign(v)
    def set_light_theta(self, v): self.params["theta"].assign(v)
    def set_light_radius(self, v): self.params["radius"].assign(v)
    def set_light_intensity(self, v): self.params["intensity"].assign(v)

    def set_option(self, name, value):
        """b200vr_set_option: tuning switches that choose between kernels / launch orders rendering the same frame
        (include/b200vr.h). Before init the option is kept and applied when the device is created."""
        if self._h is None:
            self._options[str(name)] = str(value)
        else:
            self._ck(capi.lib().b200vr_set_option(self._h, str(name).encode(), str(value).encode()))

    # ---- unsafe getters ----
    def unsafe_get_fbsize(self):
        return self.params["fbsize"].get()

    def unsafe_get_variance(self):
        return self.variance

    def unsafe_get_tfn(self):
        return self.params["tfn"].ref()

    # ---- helpers ----
    def _ck(self, rc):
        if rc != capi.OK:
            msg = capi.lib().b200vr_last_error(self._h)
            raise RuntimeError(f"[b200] {msg.decode() if msg else 'error'} (code {rc})")

    def _push_params(self):
        L = capi.lib()
        p = capi.Params()
        self._ck(L.b200vr_get_params(self._h, C.byref(p)))
        p.sampling_rate = self._rate
        p.spp = self._spp
        p.mode = capi.MODE_PATHTRACE if self._path_tracing else capi.MODE_RAYMARCH
        p.accumulate = int(self._accumulate)
        p.shading, p.sampler, p.skip_empty = int(self.shading), int(self.sampler), int(bool(self.skip_empty))
        p.premultiplied = int(bool(self.premultiplied))
        p.pt_child_instance_xfm = int(bool(self.pt_child_instance_xfm))
        self._ck(L.b200vr_set_params(self._h, C.byref(p)))

    # ---- MainRenderer::init (ovr/renderer.h:290-297) + DeviceOptix7::init (device.cpp:16-20) ----
    def init(self, scene, camera=None, argv=()):
        if self._initialized:
            raise RuntimeError("[b200] device already initialized!")  # device_impl.cpp:73-75
        L = capi.lib()
        h = C.c_void_p()
        rc = L.b200vr_create(C.byref(h), self._cuda_device)
        if rc != capi.OK:
            raise RuntimeError("[b200] " + L.b200vr_last_error(None).decode())
        self._h = h
        for k, v in self._options.items():
            self._ck(L.b200vr_set_option(self._h, k.encode(), v.encode()))
        self._rate, self._spp, self._path_tracing, self._accumulate = float(scene.volume_sampling_rate), 1, False, False
        # set_scene (ovr/renderer.h:299-341): flatten the TF into (rgb...) and (pos, alpha)... float arrays
        tfn = scene.tfn
        if tfn is not None:
            c = np.asarray(tfn.color, np.float32).reshape(-1, 4)[:, :3].ravel()
            n_o = len(tfn.opacity)
            o = np.empty(2 * n_o, np.float32)
            o[0::2] = np.arange(n_o, dtype=np.float32) / np.float32(max(n_o - 1, 1))
            o[1::2] = np.asarray(tfn.opacity, np.float32)
            self.set_transfer_function(c, o, tfn.value_range)
        self.current_scene = scene
        self.set_camera(camera if camera is not None else scene.camera)
        # buildScene (device_impl.cpp:283-302)
        vol = scene.volume
        if vol is None:
            raise RuntimeError("[b200] scene has no structured regular volume")  # parse_single_volume_scene throws
        otw = np.ascontiguousarray(vol.obj_to_world(), np.float32)
        otw_p = otw.ctypes.data_as(C.POINTER(C.c_float))
        if vol.data is not None:
            data = np.ascontiguousarray(vol.data)
            if data.dtype not in _DTYPE_TO_TYPE:
                raise RuntimeError("[b200] unexpected volume type ...")
            nz, ny, nx = data.shape
            self._ck(L.b200vr_set_volume(self._h, data.ctypes.data_as(C.c_void_p), _DTYPE_TO_TYPE[data.dtype], nx, ny, nz, otw_p, 0))
        else:
            nx, ny, nz = vol.noise["dims"]
            org = (C.c_int32 * 3)(*[int(v) for v in vol.noise["org"]])
            self._ck(L.b200vr_set_volume_noise(self._h, C.c_uint32(int(vol.noise["seed"])), org, nx, ny, nz, otw_p))
        lo, hi = C.c_float(), C.c_float()
        self._ck(L.b200vr_get_value_range(self._h, C.byref(lo), C.byref(hi)))
        self.value_range = (lo.value, hi.value)
        print(f"[b200] volume range = {lo.value} {hi.value}")
        self._push_params()
        self._initialized = True
        self.commit()
        print("[b200] setup done")

    # ---- Impl::commit (device_impl.cpp:113-197) ----
    def commit(self):
        L = capi.lib()
        P = self.params
        if P["fbsize"].update():
            w, hgt = P["fbsize"].ref()
            self._ck(L.b200vr_resize(self._h, w, hgt))
        if P["camera"].update():
            cam = P["camera"].ref()
            self._ck(L.b200vr_set_camera_lookat(self._h, _f3(cam.frm), _f3(cam.at), _f3(cam.up), C.c_float(cam.fovy)))
        if P["tfn"].update():
            c, o, r = P["tfn"].ref()
            self._ck(L.b200vr_set_transfer_function(self._h, c.ctypes.data_as(C.POINTER(C.c_float)), len(c) // 3,
                                                    o.ctypes.data_as(C.POINTER(C.c_float)), len(o) // 2, r[0], r[1]))
        dirty = False
        sparse_dirty = False
        for name, attr in (("focus_center", "_focus_center"), ("focus_scale", "_focus_scale"), ("base_noise", "_base_noise"),
                           ("sparse_sampling", "_sparse")):
            if P[name].update():  # device_impl.cpp:154-183
                setattr(self, attr, P[name].ref())
                sparse_dirty = True
        if sparse_dirty:
            fc = (C.c_float * 2)(float(self._focus_center[0]), float(self._focus_center[1]))
            self._ck(L.b200vr_set_sparse_sampling(self._h, 1 if self._sparse else 0, fc, C.c_float(self._focus_scale),
                                                  C.c_float(self._base_noise)))
        if P["sample_per_pixel"].update():
            self._spp, dirty = P["sample_per_pixel"].ref(), True
        if P["path_tracing"].update():
            self._path_tracing, dirty = P["path_tracing"].ref(), True
        if P["frame_accumulation"].update():
            self._accumulate, dirty = P["frame_accumulation"].ref(), True
        if P["volume_sampling_rate"].update():
            self._rate, dirty = P["volume_sampling_rate"].ref(), True
        if dirty:
            self._push_params()

    def apply_device_options(self):
        """Push shading / sampler / skip_empty / premultiplied after changing them on the Python object."""
        self._push_params()

    # ---- DeviceOptix7::render (device.cpp:34-43) ----
    def render(self):
        if self.params["fbsize"].get() is None:
            return  # launch only after the first resize (device_impl.cpp:215-216)
        t0 = time.perf_counter()
        st = capi.Stats()
        self._ck(capi.lib().b200vr_render(self._h, C.byref(st)))
        self.render_time += (time.perf_counter() - t0) * 1e3
        self.last_stats = st.as_dict()
        self.variance = 0.0

    def render_async(self):
        """b200vr_render_async: the frame is enqueued on the device's stream; wait() collects its counters."""
        if self.params["fbsize"].get() is None:
            return
        self._ck(capi.lib().b200vr_render_async(self._h))
        self.variance = 0.0

    def wait(self):
        st = capi.Stats()
        self._ck(capi.lib().b200vr_wait(self._h, C.byref(st)))
        self.last_stats = st.as_dict()
        return self.last_stats

    @property
    def last_kernel(self):
        """b200vr_last_kernel: name of the render kernel the last frame was enqueued with."""
        return capi.lib().b200vr_last_kernel(self._h).decode()

    def use_stream(self, cuda_stream_handle):
        """Run every kernel / copy of this device on the given CUDA stream (e.g. torch.cuda.current_stream().cuda_stream), so
        that they are ordered with the caller's own work (NCCL collectives, peer copies). Handle 0 is the legacy default
        stream, which the C ABI wants spelled cudaStreamLegacy (0x1): NULL means "the device's own stream" there."""
        h = int(cuda_stream_handle)
        self._ck(capi.lib().b200vr_set_stream(self._h, C.c_void_p(h if h != 0 else 1)))

    def swap(self):
        self._ck(capi.lib().b200vr_swap(self._h))

    def mapframe(self, fb=None):
        fb = fb if fb is not None else FrameBufferData()
        rgba, grad = C.c_void_p(), C.c_void_p()
        self._ck(capi.lib().b200vr_map(self._h, C.byref(rgba), C.byref(grad)))
        w, hgt = self.params["fbsize"].get()
        fb.rgba = (rgba.value, w * hgt * 16)
        fb.grad = (grad.value, w * hgt * 12)
        fb.size = (w, hgt)
        return fb

    # CrossDeviceBuffer::to_cpu() for both layers
    def download(self, want_grad=True):
        w, hgt = self.params["fbsize"].get()
        rgba = np.empty((hgt, w, 4), np.float32)
        grad = np.empty((hgt, w, 3), np.float32) if want_grad else None
        self._ck(capi.lib().b200vr_download(self._h, rgba.ctypes.data_as(C.c_void_p),
                                            grad.ctypes.data_as(C.c_void_p) if want_grad else None))
        return rgba, grad

    def set_sparse_noise(self, table, tile_xy, tile_t):
        """Device option: the [y][x][t] noise table the sparse-sampling mask is thresholded against (the reference embeds
        one at build time, blue_noise.h:60-78); None restores the built-in table."""
        if table is None:
            self._ck(capi.lib().b200vr_set_sparse_noise(self._h, None, 0, 0))
            return
        t = np.ascontiguousarray(table, np.float32).ravel()
        if t.size != tile_xy * tile_xy * tile_t:
            raise RuntimeError("[b200] noise table size does not match tile_xy^2 * tile_t")
        self._ck(capi.lib().b200vr_set_sparse_noise(self._h, t.ctypes.data_as(C.c_void_p), tile_xy, tile_t))

    def sparse_samples(self):
        """(n, 2) int32 array of the pixels (x, y) the last sparse frame rendered (xs_and_ys, device_impl.cpp:332-340)."""
        L = capi.lib()
        n = int(L.b200vr_get_sparse_samples(self._h, None, 0))
        if n < 0:
            self._ck(n)
        out = np.empty((n, 2), np.int32)
        if n and L.b200vr_get_sparse_samples(self._h, out.ctypes.data_as(C.c_void_p), n) != n:
            self._ck(capi.ECUDA)
        return out

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h is not None:
            capi.lib().b200vr_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def create_renderer(name="b200", cuda_device=0):
    """create_renderer (ovr/renderer.cpp:42-61): only the B200 device lives here; unknown names raise like
    UnknownDynamicModuleError does."""
    if name != "b200":
        raise RuntimeError(f"unknown device '{name}' (this package provides 'b200')")
    return MainRenderer(cuda_device)
