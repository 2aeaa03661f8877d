"""TEST INFRASTRUCTURE ONLY -- ctypes binding of the CPU oracle (oracle/liboracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may import this.
The product path (open-volume-renderer_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

SHADING_NONE, SHADING_REF = 0, 1
FILTER_FIXED8_RN, FILTER_FIXED8_TRUNC, FILTER_FP32, FILTER_HW = 0, 1, 2, 3
MC_REFERENCE, MC_CONSERVATIVE = 0, 1

_f3 = C.c_float * 3
_i3 = C.c_int32 * 3


class Scene(C.Structure):
    _fields_ = [
        ("dims", _i3), ("voxels", C.c_void_p), ("obj_to_world", C.c_float * 12),
        ("tf_color", C.c_void_p), ("n_color", C.c_int32), ("tf_alpha", C.c_void_p), ("n_alpha", C.c_int32),
        ("tf_lower", C.c_float), ("tf_upper", C.c_float), ("step", C.c_float), ("base", C.c_float),
        ("cam_position", _f3), ("cam_direction", _f3), ("cam_horizontal", _f3), ("cam_vertical", _f3),
        ("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("frame_index", C.c_int32),
        ("shading", C.c_int32), ("skip_empty", C.c_int32), ("filter_mode", C.c_int32),
        ("max_num_scatters", C.c_int32), ("light_pos", _f3), ("light_ambient", C.c_float),
        ("clip_lo", _f3), ("clip_hi", _f3), ("mc_majorants", C.c_void_p), ("mc_dims", _i3),
        ("pt_child_instance_xfm", C.c_int32),
    ]


class Counters(C.Structure):
    _fields_ = [("primary_samples", C.c_uint64), ("shadow_samples", C.c_uint64), ("gradient_fetches", C.c_uint64),
                ("delta_samples", C.c_uint64), ("rays", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def _stale(target, sources):
    """True when `target` is missing or older than any of `sources` (files under oracle/)."""
    if not os.path.exists(target):
        return True
    mt = os.path.getmtime(target)
    return any(os.path.getmtime(os.path.join(_HERE, s)) > mt for s in sources if os.path.exists(os.path.join(_HERE, s)))


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present); rebuilt when a source is newer."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = ("oracle_sample.cpp", "oracle_grid.cpp", "oracle_march.cpp", "oracle_path.cpp", "oracle_math.h", "oracle_internal.h",
            "ovr_oracle.h")
    if force or _stale(so, srcs):
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, capture_output=True)
    refs = ("libovr_ref_kat.so", "libovr_ref_tfn.so", "libovr_ref_march.so", "libovr_ref_majorant.so")
    ref_srcs = ("ref_shim.cpp", "ref_tfn_shim.cpp", "ref_march_shim.cpp", "ref_majorant_shim.cu", "extract_march.py", "Makefile", "liboracle.so")
    if os.path.isdir("/root/reference/ovr") and (force or any(_stale(os.path.join(_HERE, "_ref", r), ref_srcs) for r in refs)):
        subprocess.run(["make", "-C", _HERE, "ref"], check=True, capture_output=True)
    build_scenedump(force)
    return so


def build_scenedump(force=False):
    """build/ovr/scenedump: oracle/ref_scene_dump.cpp linked with the reference's own host objects (build/ovr/libovr_host.a, made by
    open-volume-renderer_b200/device/build.py from /root/reference) -- the reference's scene loader as a command-line dump. Skipped
    where the reference tree or the archive is absent."""
    root = os.path.dirname(_HERE)
    ref, hostlib = "/root/reference", os.path.join(root, "build", "ovr", "libovr_host.a")
    exe, src = os.path.join(root, "build", "ovr", "scenedump"), os.path.join(_HERE, "ref_scene_dump.cpp")
    if not (os.path.isdir(os.path.join(ref, "ovr")) and os.path.exists(hostlib)):
        return None
    if not force and os.path.exists(exe) and os.path.getmtime(exe) >= max(os.path.getmtime(src), os.path.getmtime(hostlib)):
        return exe
    inc = [ref, f"{ref}/ovr", f"{ref}/ovr/common", f"{ref}/ovr/common/cuda", f"{ref}/extern", f"{ref}/extern/gdt", f"{ref}/extern/tfn",
           "/usr/local/cuda/include"]
    subprocess.run(["g++", "-std=c++17", "-O1", "-w", "-DOVR_BUILD_CUDA_DEVICES", "-rdynamic", "-o", exe, src] + [f"-I{i}" for i in inc] +
                   ["-Wl,--whole-archive", hostlib, "-Wl,--no-whole-archive", "-L/usr/local/cuda/lib64", "-lcudart", "-ldl", "-lpthread",
                    "-Wl,-rpath,/usr/local/cuda/lib64"], check=True, capture_output=True)
    return exe


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.ovr_oracle_hardware_threads.restype = C.c_int
        L.ovr_oracle_dda.restype = C.c_int
        L.ovr_oracle_clip.restype = C.c_int
        _LIB = L
    return _LIB


def ref_majorant_lib():
    """The reference's majorant_kernel compiled for the GPU (oracle/ref_majorant_shim.cu); None where oracle/_ref was not built."""
    so = os.path.join(_HERE, "_ref", "libovr_ref_majorant.so")
    if not os.path.exists(so):
        return None
    L = C.CDLL(so)
    L.ref_majorant_kernel_run.restype = C.c_int
    L.ref_majorant_kernel_run.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_float, C.c_float, C.c_void_p]
    return L


def ref_kat_lib():
    """The reference's own dda.h / RandomTEA / gdt compiled on the host (exists only where /root/reference does)."""
    so = os.path.join(_HERE, "_ref", "libovr_ref_kat.so")
    if not os.path.exists(so):
        return None
    L = C.CDLL(so)
    L.ref_dda3.restype = C.c_int
    L.ref_dda_iter.restype = C.c_int
    return L


def ref_tfn_table(tfn_json_text, capacity=8192):
    """The reference's own tfn::loadTransferFunction (oracle/_ref/libovr_ref_tfn.so, built where /root/reference exists):
    (N, 4) float32 RGBA table of a VIDI3D "transferFunction" object; None when the library is not there."""
    so = os.path.join(_HERE, "_ref", "libovr_ref_tfn.so")
    if not os.path.exists(so):
        return None
    L = C.CDLL(so)
    L.ref_tfn_table.restype = C.c_int
    out = np.zeros((capacity, 4), np.float32)
    n = L.ref_tfn_table(tfn_json_text.encode(), out.ctypes.data_as(C.POINTER(C.c_float)), capacity)
    assert n > 0
    return out[:n].copy()


def delta_hashit(majorants, spac_rcp, org, dr, tnear, tfar, taus, use_ref=False):
    """Free-path iteration of delta tracking along one ray (spatial_partition.h:56-96): rows (found, rayt, majorant, tau)
    per hashit call. use_ref: the reference's own iterator from oracle/_ref (None when not built)."""
    maj = np.ascontiguousarray(majorants, np.float32)
    grid = _i3(maj.shape[2], maj.shape[1], maj.shape[0])
    taus = np.ascontiguousarray(taus, np.float32)
    out = np.zeros((len(taus), 4), np.float32)
    if use_ref:
        L = ref_kat_lib()
        if L is None:
            return None
        fn = L.ref_delta_hashit
    else:
        fn = lib().ovr_oracle_delta_hashit
    fn.restype = C.c_int
    n = fn(_fp(maj), grid, _f3(*spac_rcp), _f3(*org), _f3(*dr), C.c_float(tnear), C.c_float(tfar), _fp(taus), len(taus), _fp(out))
    return out[:n].copy()


def uniform_sphere(s, use_ref=False):
    """uniform_sample_sphere(1, s) for rows s = (sx, sy); use_ref: the reference's own function from oracle/_ref."""
    s = np.ascontiguousarray(s, np.float32)
    L = ref_kat_lib() if use_ref else lib()
    if L is None:
        return None
    fn = L.ref_uniform_sphere if use_ref else L.ovr_oracle_uniform_sphere
    out = np.zeros((len(s), 3), np.float32)
    for k in range(len(s)):
        fn(C.c_float(float(s[k, 0])), C.c_float(float(s[k, 1])), _fp(out[k]))
    return out


def ref_clip(org, dr, lo, hi, t0=0.0, t1=np.finfo(np.float32).max):
    """The reference's own intersect_box (oracle/_ref): (hit, t0, t1)."""
    L = ref_kat_lib()
    a, b = C.c_float(t0), C.c_float(t1)
    L.ref_intersect_box.restype = C.c_int
    ok = L.ref_intersect_box(_f3(*org), _f3(*dr), _f3(*lo), _f3(*hi), C.byref(a), C.byref(b))
    return bool(ok), a.value, b.value


def ref_camera(frm, at, up, fovy, width, height):
    """Device camera from the look-at camera by the reference's own statements (device_impl.cpp:128-136, oracle/ref_shim.cpp):
    float32 [12] = position, direction, horizontal, vertical."""
    out = np.zeros(12, np.float32)
    ref_kat_lib().ref_camera(_fp(np.ascontiguousarray(frm, np.float32)), _fp(np.ascontiguousarray(at, np.float32)),
                             _fp(np.ascontiguousarray(up, np.float32)), C.c_float(fovy), int(width), int(height), _fp(out))
    return out


def ref_value_ranges(vol):
    """Macrocell value ranges [mz, my, mx, 2] by the reference's own value_range_kernel run on the host (oracle/ref_shim.cpp)."""
    vol = np.ascontiguousarray(vol, np.float32)
    dims = (vol.shape[2], vol.shape[1], vol.shape[0])
    mc = macrocell_dims(dims)
    r = np.empty((mc[2], mc[1], mc[0], 2), np.float32)
    ref_kat_lib().ref_value_ranges(_fp(vol), _i3(*dims), _fp(r))
    return r


def ref_volume_matrix(grid_origin, grid_spacing, dims):
    """Object-to-world matrix of the volume instance and its inverse by the reference's own statements
    (device_impl.cpp:288-296, gdt rcp): float32 [24] = otw 3x4 row major, wto 3x4 row major."""
    out = np.zeros(24, np.float32)
    ref_kat_lib().ref_volume_matrix(_fp(np.ascontiguousarray(grid_origin, np.float32)), _fp(np.ascontiguousarray(grid_spacing, np.float32)),
                                    _p(np.ascontiguousarray(dims, np.int32)), _fp(out))
    return out


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def hardware_threads():
    return int(lib().ovr_oracle_hardware_threads())


def camera(frm, at, up, fovy, width, height):
    out = [np.zeros(3, np.float32) for _ in range(4)]
    lib().ovr_oracle_camera(_f3(*frm), _f3(*at), _f3(*up), C.c_float(fovy), width, height, *[_fp(o) for o in out])
    return dict(position=out[0], direction=out[1], horizontal=out[2], vertical=out[3])


def noise(n, seed=1337, org=(0, 0, 0), bdims=None, threads=0):
    bdims = tuple(bdims) if bdims is not None else (n, n, n)
    out = np.empty((bdims[2], bdims[1], bdims[0]), np.float32)
    lib().ovr_oracle_noise(n, C.c_uint32(seed), _i3(*org), _i3(*bdims), _fp(out), threads)
    return out


def sparse_mask(width, height, frame_index, noise_table, tile_xy, tile_t, mean, sigma, base_noise):
    """(keep, p): boolean (H, W) mask of the pixels a sparse frame renders and their acceptance probability."""
    t = np.ascontiguousarray(noise_table, np.float32).ravel()
    assert t.size == tile_xy * tile_xy * tile_t
    keep = np.zeros((height, width), np.uint8)
    p = np.zeros((height, width), np.float32)
    m = (C.c_float * 2)(float(mean[0]), float(mean[1]))
    lib().ovr_oracle_sparse_mask(width, height, int(frame_index), _fp(t), tile_xy, tile_t, m, C.c_float(sigma),
                                 C.c_float(base_noise), _p(keep), _fp(p))
    return keep.astype(bool), p


def ref_sparse_mask(width, height, frame_index, noise_table, tile_xy, mean, sigma, base_noise):
    """The reference's own two kernels run on the host (oracle/_ref; the table's time extent is the reference's 64)."""
    L = ref_kat_lib()
    t = np.ascontiguousarray(noise_table, np.float32).ravel().copy()
    assert t.size == tile_xy * tile_xy * 64
    xy = np.zeros((height * width, 2), np.int32)
    m = (C.c_float * 2)(float(mean[0]), float(mean[1]))
    L.ref_sparse_mask(width, height, int(frame_index), _fp(t), tile_xy, m, C.c_float(sigma), C.c_float(base_noise), _p(xy))
    return xy


def macrocell_dims(dims):
    mc = _i3()
    lib().ovr_oracle_macrocell_dims(_i3(*dims), mc)
    return tuple(mc)


def build_ranges(vol, mode=MC_CONSERVATIVE):
    """vol: float32 [nz, ny, nx]; returns ranges [mz, my, mx, 2]."""
    vol = np.ascontiguousarray(vol, np.float32)
    dims = (vol.shape[2], vol.shape[1], vol.shape[0])
    mc = macrocell_dims(dims)
    r = np.empty((mc[2], mc[1], mc[0], 2), np.float32)
    lib().ovr_oracle_build_ranges(_fp(vol), _i3(*dims), mode, _fp(r))
    return r


def build_majorants(ranges, tf_alpha, win_lo, win_hi):
    ranges = np.ascontiguousarray(ranges, np.float32)
    tf_alpha = np.ascontiguousarray(tf_alpha, np.float32)
    m = np.empty(ranges.shape[:-1], np.float32)
    lib().ovr_oracle_build_majorants(_fp(ranges), C.c_int64(m.size), _fp(tf_alpha), len(tf_alpha),
                                     C.c_float(win_lo), C.c_float(win_hi), _fp(m))
    return m


def tea(v0, v1, n):
    out = np.empty((n, 2), np.float32)
    st = np.empty((n, 2), np.uint32)
    lib().ovr_oracle_tea(C.c_uint32(v0), C.c_uint32(v1), n, _fp(out), st.ctypes.data_as(C.c_void_p))
    return out, st


def dda(org, dr, tmin, tmax, grid, max_cells=4096):
    cells = np.empty((max_cells, 3), np.int32)
    tt = np.empty((max_cells, 2), np.float32)
    n = lib().ovr_oracle_dda(_f3(*org), _f3(*dr), C.c_float(tmin), C.c_float(tmax), _i3(*grid), max_cells,
                             cells.ctypes.data_as(C.c_void_p), _fp(tt))
    return cells[:n].copy(), tt[:n].copy()


def clip(org, dr, lo=(0, 0, 0), hi=(1, 1, 1)):
    t0, t1 = C.c_float(), C.c_float()
    hit = lib().ovr_oracle_clip(_f3(*org), _f3(*dr), _f3(*lo), _f3(*hi), C.byref(t0), C.byref(t1))
    return bool(hit), t0.value, t1.value


def sample_volume(vol, pts, filter_mode=FILTER_HW):
    vol = np.ascontiguousarray(vol, np.float32)
    pts = np.ascontiguousarray(pts, np.float32)
    out = np.empty(len(pts), np.float32)
    dims = (vol.shape[2], vol.shape[1], vol.shape[0])
    lib().ovr_oracle_sample_volume(_fp(vol), _i3(*dims), filter_mode, _fp(pts), C.c_int64(len(pts)), _fp(out))
    return out


def tex1d(table, u, filter_mode=FILTER_HW):
    """table [n] or [n, 4] float32; u [m] float32 normalised coordinates -> [m] or [m, 4]."""
    table = np.ascontiguousarray(table, np.float32)
    u = np.ascontiguousarray(u, np.float32)
    ch = 1 if table.ndim == 1 else table.shape[1]
    out = np.empty(len(u) if ch == 1 else (len(u), ch), np.float32)
    lib().ovr_oracle_tex1d(_fp(table), len(table), ch, filter_mode, _fp(u), C.c_int64(len(u)), _fp(out))
    return out


def xfm(m, v):
    m = np.ascontiguousarray(m, np.float32)
    v = np.ascontiguousarray(v, np.float32)
    out = np.empty(24, np.float32)
    lib().ovr_oracle_xfm(_fp(m), _fp(v), _fp(out))
    return out


class OracleScene:
    """Owns the numpy buffers behind an ovr_oracle_scene. `desc` is a scenes.SceneDesc-like object with
    attributes: volume [nz,ny,nx] f32, obj_to_world (12,), tf_color [Nc,4], tf_alpha [Na], tf_range (lo,hi),
    step, base, cam (dict from camera()), width, height, spp, shading."""

    def __init__(self, desc, skip_empty=False, filter_mode=FILTER_HW, mc_mode=MC_CONSERVATIVE,
                 build_macrocells=True):
        self.vol = np.ascontiguousarray(desc.volume, np.float32)
        self.tf_color = np.ascontiguousarray(desc.tf_color, np.float32).reshape(-1, 4)
        self.tf_alpha = np.ascontiguousarray(desc.tf_alpha, np.float32)
        s = Scene()
        lib().ovr_oracle_default_scene(C.byref(s))
        nz, ny, nx = self.vol.shape
        s.dims = _i3(nx, ny, nz)
        s.voxels = _p(self.vol)
        s.obj_to_world = (C.c_float * 12)(*[float(x) for x in desc.obj_to_world])
        s.tf_color, s.n_color = _p(self.tf_color), len(self.tf_color)
        s.tf_alpha, s.n_alpha = _p(self.tf_alpha), len(self.tf_alpha)
        s.tf_lower, s.tf_upper = float(desc.tf_range[0]), float(desc.tf_range[1])
        s.step, s.base = float(desc.step), float(desc.base)
        cam = desc.cam
        s.cam_position, s.cam_direction = _f3(*cam["position"]), _f3(*cam["direction"])
        s.cam_horizontal, s.cam_vertical = _f3(*cam["horizontal"]), _f3(*cam["vertical"])
        s.width, s.height, s.spp = int(desc.width), int(desc.height), int(desc.spp)
        s.shading = int(desc.shading)
        s.skip_empty = int(bool(skip_empty))
        s.filter_mode = int(filter_mode)
        clip = getattr(desc, "clip", None)
        if clip is not None:
            s.clip_lo, s.clip_hi = _f3(*clip[0]), _f3(*clip[1])
        self.s = s
        self.ranges = self.majorants = None
        if build_macrocells:
            self.ranges = build_ranges(self.vol, mc_mode)
            self.majorants = build_majorants(self.ranges, self.tf_alpha, s.tf_lower, s.tf_upper)
            s.mc_majorants = _p(self.majorants)
            s.mc_dims = _i3(*macrocell_dims((nx, ny, nz)))

    def render_march(self, rect=None, stride=(1, 1), premultiplied=False, threads=0, frame_index=1):
        s = self.s
        s.frame_index = frame_index
        W, H = s.width, s.height
        x0, y0, x1, y1 = rect if rect is not None else (0, 0, W, H)
        rgba = np.zeros((H, W, 4), np.float32)
        grad = np.zeros((H, W, 3), np.float32)
        cnt = Counters()
        lib().ovr_oracle_render_march(C.byref(s), x0, y0, x1, y1, stride[0], stride[1], _fp(rgba), _fp(grad),
                                      int(premultiplied), C.byref(cnt), threads)
        return rgba, grad, cnt.as_dict()

    def ref_march(self, pixels_xy, threads=1):
        """The REFERENCE's own raymarching() / raymarching_shadow() (cut out of its shader sources and compiled on the host by
        oracle/ref_march_shim.cpp, texture fetches through this oracle's model of the unit) for the listed pixels:
        (rgba [n, 4], grad [n, 3]); None where oracle/_ref was not built (needs /root/reference). threads: 1 = the plain loop,
        0 = all hardware threads (ref_march_rays_mt, same pixels)."""
        so = os.path.join(_HERE, "_ref", "libovr_ref_march.so")
        if not os.path.exists(so):
            return None
        lib()  # liboracle.so first: the shim links against it
        L = C.CDLL(so)
        px = np.ascontiguousarray(pixels_xy, np.int32).reshape(-1, 2)
        out = np.zeros((len(px), 7), np.float32)
        if threads == 1:
            L.ref_march_rays(C.byref(self.s), _p(px), C.c_int64(len(px)), _fp(out))
        else:
            L.ref_march_rays_mt(C.byref(self.s), _p(px), C.c_int64(len(px)), _fp(out), C.c_int(threads))
        return out[:, :4].copy(), out[:, 4:].copy()

    def ref_path(self, pixels_xy, frame_index=1, threads=1):
        """The REFERENCE's own delta_tracking() / pathtracing() / render_pathtracing() (oracle/ref_march_shim.cpp, namespace pt) for
        the listed pixels, one frame: rgba [n, 4]; None where oracle/_ref was not built. self.s.pt_child_instance_xfm = 1:
        scattered rays go through the instance transform a second time before they are clipped, as OptiX does; 0: clipped as
        the object-space rays they are."""
        so = os.path.join(_HERE, "_ref", "libovr_ref_march.so")
        if not os.path.exists(so):
            return None
        lib()
        L = C.CDLL(so)
        self.s.frame_index = frame_index
        px = np.ascontiguousarray(pixels_xy, np.int32).reshape(-1, 2)
        out = np.zeros((len(px), 4), np.float32)
        if threads == 1:
            L.ref_path_rays(C.byref(self.s), _p(px), C.c_int64(len(px)), _fp(out))
        else:
            L.ref_path_rays_mt(C.byref(self.s), _p(px), C.c_int64(len(px)), _fp(out), C.c_int(threads))
        return out

    def render_path(self, rect=None, stride=(1, 1), threads=0, frame_index=1):
        s = self.s
        s.frame_index = frame_index
        W, H = s.width, s.height
        x0, y0, x1, y1 = rect if rect is not None else (0, 0, W, H)
        rgba = np.zeros((H, W, 4), np.float32)
        cnt = Counters()
        lib().ovr_oracle_render_path(C.byref(s), x0, y0, x1, y1, stride[0], stride[1], _fp(rgba), C.byref(cnt),
                                     threads)
        return rgba, cnt.as_dict()

    def sample_tf(self, values):
        values = np.ascontiguousarray(values, np.float32)
        out = np.empty((len(values), 4), np.float32)
        lib().ovr_oracle_sample_tf(C.byref(self.s), _fp(values), C.c_int64(len(values)), _fp(out))
        return out
