"""ctypes binding of the b200vr C ABI (include/b200vr.h -> libb200vr.so). Fails loudly when the CUDA library is missing:
there is no Python / CPU fallback for any entry point."""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
# B200VR_LIB_PATH: another build of the same library (A/B runs of compile-time tuning constants); the product is the in-tree one
LIB_PATH = os.environ.get("B200VR_LIB_PATH") or os.path.join(_HERE, "libb200vr.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC"]

OK, EINVAL, ECUDA, ENODEV, ESTATE = 0, -1, -2, -3, -4
TYPE_UINT8, TYPE_INT8, TYPE_UINT16, TYPE_INT16, TYPE_UINT32, TYPE_INT32, TYPE_FLOAT, TYPE_DOUBLE = range(8)
MODE_RAYMARCH, MODE_PATHTRACE = 0, 1
SHADING_NONE, SHADING_REF = 0, 1
SAMPLER_TEX3D, SAMPLER_MANUAL = 0, 1


class Params(C.Structure):
    _fields_ = [("sampling_rate", C.c_float), ("base", C.c_float), ("spp", C.c_int32), ("mode", C.c_int32),
                ("shading", C.c_int32), ("sampler", C.c_int32), ("skip_empty", C.c_int32), ("accumulate", C.c_int32),
                ("max_num_scatters", C.c_int32), ("premultiplied", C.c_int32), ("light_pos", C.c_float * 3),
                ("light_ambient", C.c_float), ("pt_child_instance_xfm", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("ms", C.c_float), ("primary_samples", C.c_uint64), ("shadow_samples", C.c_uint64),
                ("gradient_fetches", C.c_uint64), ("delta_samples", C.c_uint64), ("rays", C.c_uint64),
                ("frame_index", C.c_int32), ("kernel_launches", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# every symbol include/b200vr.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_FP = C.POINTER(C.c_float)
_IP = C.POINTER(C.c_int32)
SYMBOLS = {
    "b200vr_abi_version": (C.c_int, []),
    "b200vr_host_camera": (C.c_int, [_FP, _FP, _FP, C.c_float, C.c_int, C.c_int, _FP]),
    "b200vr_host_integer_normalize": (C.c_float, [C.c_float, C.c_int]),
    "b200vr_create": (C.c_int, [C.POINTER(_P), C.c_int]),
    "b200vr_destroy": (None, [_P]),
    "b200vr_last_error": (C.c_char_p, [_P]),
    "b200vr_default_params": (None, [C.POINTER(Params)]),
    "b200vr_set_stream": (C.c_int, [_P, _P]),
    "b200vr_set_volume": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, _FP, C.c_int]),
    "b200vr_set_volume_noise": (C.c_int, [_P, C.c_uint32, _IP, C.c_int, C.c_int, C.c_int, _FP]),
    "b200vr_get_value_range": (C.c_int, [_P, _FP, _FP]),
    "b200vr_get_macrocells": (C.c_int, [_P, _IP, _P, _P]),
    "b200vr_download_volume": (C.c_int, [_P, _P]),
    "b200vr_probe": (C.c_int, [_P, _FP, C.c_int64, C.c_int, _FP, _FP]),
    "b200vr_probe_tex1d": (C.c_int, [_P, _FP, C.c_int, C.c_int, _FP, C.c_int64, _FP, _FP]),
    "b200vr_bench_fetch": (C.c_int, [_P, C.c_int, C.c_int, C.c_int, _FP, C.c_int, _FP, C.POINTER(C.c_uint64)]),
    "b200vr_set_transfer_function": (C.c_int, [_P, _FP, C.c_int, _FP, C.c_int, C.c_float, C.c_float]),
    "b200vr_set_camera": (C.c_int, [_P, _FP, _FP, _FP, _FP]),
    "b200vr_set_camera_lookat": (C.c_int, [_P, _FP, _FP, _FP, C.c_float]),
    "b200vr_get_camera": (C.c_int, [_P, _FP, _FP, _FP, _FP]),
    "b200vr_set_params": (C.c_int, [_P, C.POINTER(Params)]),
    "b200vr_get_params": (C.c_int, [_P, C.POINTER(Params)]),
    "b200vr_resize": (C.c_int, [_P, C.c_int, C.c_int]),
    "b200vr_reset_accumulation": (C.c_int, [_P]),
    "b200vr_set_clip_box": (C.c_int, [_P, _FP, _FP]),
    "b200vr_set_tile_partition": (C.c_int, [_P, C.c_int, C.c_int]),
    "b200vr_render": (C.c_int, [_P, C.POINTER(Stats)]),
    "b200vr_render_async": (C.c_int, [_P]),
    "b200vr_wait": (C.c_int, [_P, C.POINTER(Stats)]),
    "b200vr_last_kernel": (C.c_char_p, [_P]),
    "b200vr_set_option": (C.c_int, [_P, C.c_char_p, C.c_char_p]),
    "b200vr_map": (C.c_int, [_P, C.POINTER(_P), C.POINTER(_P)]),
    "b200vr_swap": (C.c_int, [_P]),
    "b200vr_map_index": (C.c_int, [_P, C.c_int, C.POINTER(_P), C.POINTER(_P)]),
    "b200vr_download": (C.c_int, [_P, _P, _P]),
    "b200vr_download_async": (C.c_int, [_P, _P, _P, _P]),
    "b200vr_set_brick": (C.c_int, [_P, _P, _P, _P, _P]),
    "b200vr_ipc_export_frame": (C.c_int, [_P, C.c_int, _P]),
    "b200vr_ipc_open": (C.c_int, [_P, _P, _P]),
    "b200vr_ipc_close": (C.c_int, [_P, _P]),
    "b200vr_current_buffer": (C.c_int, [_P]),
    "b200vr_set_frame_target": (C.c_int, [_P, _P, _P]),
    "b200vr_group_create": (C.c_int, [C.POINTER(_P), _IP, C.c_int]),
    "b200vr_group_destroy": (None, [_P]),
    "b200vr_group_size": (C.c_int, [_P]),
    "b200vr_group_member": (_P, [_P, C.c_int]),
    "b200vr_group_render": (C.c_int, [_P, C.POINTER(Stats)]),
    "b200vr_group_last_error": (C.c_char_p, [_P]),
    "b200vr_copy_device": (C.c_int, [_P, _P, _P, C.c_int64]),
    "b200vr_tile_slice_floats": (C.c_int64, [_P, C.c_int, C.c_int]),
    "b200vr_pack_tiles": (C.c_int, [_P, _P]),
    "b200vr_unpack_tiles": (C.c_int, [_P, _P, C.c_int, C.c_int]),
    "b200vr_composite_over": (C.c_int, [_P, _P, _P, _P, C.c_int64]),
    "b200vr_finalize_premultiplied": (C.c_int, [_P, _P, C.c_int64]),
    "b200vr_composite_layers": (C.c_int, [_P, C.POINTER(_P), C.c_int, _P, C.c_int64, C.c_int]),
    "b200vr_set_sparse_sampling": (C.c_int, [_P, C.c_int, _FP, C.c_float, C.c_float]),
    "b200vr_set_sparse_noise": (C.c_int, [_P, _P, C.c_int, C.c_int]),
    "b200vr_get_sparse_samples": (C.c_int64, [_P, _P, C.c_int64]),
    "b200vr_sparse_samples_noalloc": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _FP, C.c_float, C.c_float,
                                              _P, C.c_int64, _P]),
}

_lib = None


def build(force=False, verbose=False):
    """Compile libb200vr.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    src = os.path.join(_HERE, "csrc", "b200vr.cu")
    deps = [os.path.join(_HERE, "csrc", f) for f in os.listdir(os.path.join(_HERE, "csrc"))]
    deps.append(os.path.join(_ROOT, "include", "b200vr.h"))
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    extra = os.environ.get("B200VR_NVCC_EXTRA", "").split()  # tuning runs: e.g. -DB200VR_Q_MIN_CTAS=3
    cmd = ["nvcc"] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH, src]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
                               "the B200 volume device has no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the ABI and the header drifted apart
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib
