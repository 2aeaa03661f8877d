"""VIDI3D-JSON scene files, the format of the reference's bundled `data/configs/*.json` read by
ovr::scene::create_json_scene_vidi3d (ovr/serializer/serializer_vidi3d.cpp:203-412).

`read_vidi3d_scene` mirrors that serializer for the Python host mirror (same keys, same derived values: TF arrays,
value-range scaling by the type maximum, alpha end snapping, sampling rate = 1 / sampleDistance, grid spacing from
`scales`). `write_vidi3d_scene` emits a file of the same schema so that the reference's own `renderbatch`
(apps/main_batch.cpp) can be driven with synthetic volumes on a box that has no copy of the reference's data."""
import base64
import json
import os

import numpy as np

from .scenes import Camera, Scene, TransferFunction, Volume, rasterize_color_controls

_TYPES = {"UNSIGNED_BYTE": np.uint8, "BYTE": np.int8, "UNSIGNED_SHORT": np.uint16, "SHORT": np.int16,
          "UNSIGNED_INT": np.uint32, "INT": np.int32, "FLOAT": np.float32, "DOUBLE": np.float64}
_TYPE_NAMES = {np.dtype(v): k for k, v in _TYPES.items()}


def _xyz(d):
    return (float(d["x"]), float(d["y"]), float(d["z"]))


def write_vidi3d_scene(json_path, data, alpha, color_controls, camera, sample_distance=1.0, scalar_range=None,
                       raw_name=None):
    """data: [nz, ny, nx] numpy array (any supported dtype); alpha: [N] float32; color_controls: [(pos, (r, g, b)), ...].
    Writes <json_path> and the raw payload next to it (absolute fileName: valid_filename returns the bare name when it
    only finds the file relative to the config directory, serializer_vidi3d.cpp:160-176)."""
    data = np.ascontiguousarray(data)
    alpha = np.ascontiguousarray(alpha, np.float32)
    nz, ny, nx = data.shape
    json_path = os.path.abspath(json_path)
    raw_path = os.path.join(os.path.dirname(json_path), raw_name or (os.path.splitext(os.path.basename(json_path))[0] + ".raw"))
    data.astype(data.dtype.newbyteorder("<")).tofile(raw_path)
    if scalar_range is None:
        scalar_range = (float(data.min()), float(data.max()))
    root = {
        "dataSource": [{
            "dimensions": {"x": nx, "y": ny, "z": nz}, "endian": "LITTLE_ENDIAN", "fileName": [raw_path],
            "fileUpperLeft": False, "format": "REGULAR_GRID_RAW_BINARY", "id": 1, "name": os.path.basename(raw_path),
            "offset": 0, "type": _TYPE_NAMES[data.dtype]}],
        "view": {
            "camera": {"center": dict(zip("xyz", map(float, camera.at))), "eye": dict(zip("xyz", map(float, camera.frm))),
                       "fovy": float(camera.fovy), "projectionMode": "PERSPECTIVE", "up": dict(zip("xyz", map(float, camera.up)))},
            "volume": {
                "dataId": 1, "sampleDistance": float(sample_distance),
                "scalarMappingRange": {"minimum": float(scalar_range[0]), "maximum": float(scalar_range[1])},
                "transferFunction": {
                    "alphaArray": {"data": base64.b64encode(alpha.astype("<f4").tobytes()).decode(), "encoding": "BASE64"},
                    "colorControls": [{"color": {"r": float(c[0]), "g": float(c[1]), "b": float(c[2])}, "position": float(p)}
                                      for p, c in color_controls],
                    "resolution": int(len(alpha))},
                "transferFunctionType": "TRANSFER_FUNCTION", "visible": True}}}
    with open(json_path, "w") as f:
        json.dump(root, f, indent=1)
    return json_path


def _strip_json_comments(text):
    """The reference parses with nlohmann's ignore_comments (json::parse(text, nullptr, true, true),
    serializer_vidi3d.cpp:417): // and /* */ comments outside strings are skipped. Two bundled configs rely on it."""
    out, i, n, in_str = [], 0, len(text), False
    while i < n:
        c = text[i]
        if in_str:
            out.append(c)
            if c == "\\" and i + 1 < n:
                out.append(text[i + 1])
                i += 1
            elif c == '"':
                in_str = False
        elif c == '"':
            in_str = True
            out.append(c)
        elif c == "/" and i + 1 < n and text[i + 1] == "/":
            while i < n and text[i] != "\n":
                i += 1
            continue
        elif c == "/" and i + 1 < n and text[i + 1] == "*":
            j = text.find("*/", i + 2)
            i = n if j < 0 else j + 2
            out.append(" ")
            continue
        else:
            out.append(c)
        i += 1
    return "".join(out)


def load_json_with_comments(json_path):
    with open(json_path) as f:
        return json.loads(_strip_json_comments(f.read()))


def transfer_function_table(jt):
    """tfn::loadTransferFunction + updateColorMap (extern/tfn/core.h:711-760, 598-634, 344-378, 652-685) restated: the RGBA
    table create_scene_tfn reads (serializer_vidi3d.cpp:210-221). `jt` is the "transferFunction" object. Entry i:
    colour from the colour controls at (i + 0.5) / n; alpha = max(alphaArray[i], every Gaussian object at (i + 0.5) / n
    clamped to [0, 1], the opacity-control polyline at i / (n - 1)). Returns (color [n, 4] with w = 1, alpha [n])."""
    f32 = np.float32
    n = int(jt.get("resolution", 1024))
    alpha = None
    aa = jt.get("alphaArray")
    if aa and "data" in aa and aa.get("encoding") == "BASE64":
        alpha = np.frombuffer(base64.b64decode(aa["data"]), "<f4").astype(f32).copy()
        n = len(alpha)
    if alpha is None:
        alpha = np.zeros(n, f32)
    controls = [(float(c["position"]), (float(c["color"]["r"]), float(c["color"]["g"]), float(c["color"]["b"])))
                for c in jt.get("colorControls", []) if "position" in c and "color" in c]
    color = rasterize_color_controls(controls or [(0.0, (0.0, 0.0, 0.0))], n)
    x = (np.arange(n, dtype=f32) + f32(0.5)) * (f32(1.0) / f32(n))
    for g in jt.get("gaussianObjects", []):
        if not all(k in g for k in ("mean", "sigma", "heightFactor")):
            continue
        mean, sigma, hf = f32(g["mean"]), f32(g["sigma"]), f32(g["heightFactor"])
        diff = x - mean
        val = hf / (sigma * np.sqrt(f32(2.0) * f32(np.pi))) * np.exp(-(diff * diff) / (f32(2.0) * sigma * sigma))
        alpha = np.maximum(alpha, np.clip(val.astype(f32), f32(0), f32(1)))
    oc = sorted(((float(c["position"]["x"]), float(c["position"]["y"])) for c in jt.get("opacityControl", []) if "position" in c),
                key=lambda c: c[0])
    if oc:
        px, py = np.array([c[0] for c in oc], f32), np.array([c[1] for c in oc], f32)
        value = np.arange(n, dtype=np.float64) / np.float64(n - 1)
        ub = np.searchsorted(px, value.astype(f32), side="right")
        lo, hi = np.clip(ub - 1, 0, len(oc) - 1), np.clip(ub, 0, len(oc) - 1)
        span = np.abs(px[hi].astype(np.float64) - px[lo].astype(np.float64))
        w = (np.abs(value - px[lo].astype(np.float64)) / np.where(span > 0, span, 1.0)).astype(f32)
        a = (py[lo] + w * (py[hi] - py[lo])).astype(f32)
        a[ub <= 0] = py[0]
        a[ub >= len(oc)] = py[-1]
        alpha = np.maximum(alpha, a)
    return color, alpha.astype(f32)


def read_vidi3d_scene(json_path, payload=None):
    """-> scenes.Scene. `payload` overrides the raw file (bundled configs reference blobs that are not shipped)."""
    root = load_json_with_comments(json_path)
    ds, view = root["dataSource"][0], root["view"]
    dims = ds["dimensions"]
    nx, ny, nz = int(dims["x"]), int(dims["y"]), int(dims["z"])
    dtype = np.dtype(_TYPES[ds["type"]])
    if payload is None:
        names = ds["fileName"] if isinstance(ds["fileName"], list) else [ds["fileName"]]
        base = os.path.dirname(os.path.abspath(json_path))
        found = next((p for n in names for p in (n, os.path.join(base, n)) if os.path.exists(p)), None)
        if found is None:
            raise RuntimeError("Cannot find volume file.")  # serializer_vidi3d.cpp:188
        offset = int(ds.get("offset", 0))
        if os.path.getsize(found) < offset + nx * ny * nz * dtype.itemsize:
            raise RuntimeError("File size does not match data size")  # CreateArray3DScalarFromFile, ovr/scene.cpp:207-212
        order = ">" if ds.get("endian", "LITTLE_ENDIAN") == "BIG_ENDIAN" else "<"
        payload = np.fromfile(found, dtype.newbyteorder(order), nx * ny * nz, offset=int(ds.get("offset", 0))).astype(dtype)
    data = np.ascontiguousarray(payload, dtype).reshape(nz, ny, nx)
    vol = Volume(data=data)
    if "scales" in ds:
        vol.grid_spacing = _xyz(ds["scales"])
    jv = view["volume"]
    jt = jv["transferFunction"]
    color, alpha = transfer_function_table(jt)
    if alpha[0] < 0.01:      # serializer_vidi3d.cpp:222-223
        alpha[0] = 0.0
    if alpha[-1] < 0.01:
        alpha[-1] = 0.0
    if "scalarMappingRangeUnnormalized" in jv:
        r = jv["scalarMappingRangeUnnormalized"]
        rng = (float(r["minimum"]), float(r["maximum"]))
    else:
        r = jv["scalarMappingRange"]
        scale = 1.0 if dtype.kind == "f" else float(np.iinfo(dtype).max)   # serializer_vidi3d.cpp:235-269
        rng = (np.float32(scale) * np.float32(r["minimum"]), np.float32(scale) * np.float32(r["maximum"]))
        rng = (float(rng[0]), float(rng[1]))
    jc = view["camera"]
    cam = Camera(_xyz(jc["eye"]), _xyz(jc["center"]), _xyz(jc["up"]), float(jc["fovy"]))
    rate = float(np.float32(1.0) / np.float32(float(jv["sampleDistance"])))
    return Scene(vol, TransferFunction(color, alpha, rng), cam, rate)
