"""Scene and camera file formats either side of the rasteriser: what a user needs to point the B200
path at a trained 3DGRUT scene or a COLMAP reconstruction.  Host-side numpy only -- these are byte formats, not kernels.

  read_ply / write_ply          the reference's Gaussian PLY layout: MixtureOfGaussians.init_from_ply
                                (threedgrut/model/model.py:671-719) and PLYExporter.export
                                (threedgrut/export/ply_exporter.py:34-84); no `plyfile` dependency
  parameters_from_checkpoint    the checkpoint dict of get_model_parameters / init_from_checkpoint (model.py:108-140, 417-436)
  read_colmap_cameras / _images cameras.{bin,txt} and images.{bin,txt} (threedgrut/datasets/utils.py:336-565)
  colmap_views                  pose and intrinsics per image as ColmapDataset builds them
                                (threedgrut/datasets/dataset_colmap.py:130-283): C2W = inv([R(qvec) | tvec]),
                                pinhole from SIMPLE_PINHOLE / PINHOLE with the principal point forced to the image centre,
                                OPENCV_FISHEYE with max_angle from the largest distance to an image corner

All particle tensors are the model's RAW (pre-activation) parameters, the form Tracer.render's fused path takes.
"""
import struct

import numpy as np

PARAMETER_NAMES = ("positions", "rotation", "scale", "density", "features_albedo", "features_specular")

# ---------------------------------------------------------------------------------------------------
# PLY
# ---------------------------------------------------------------------------------------------------
_PLY_TYPES = {"char": "i1", "uchar": "u1", "short": "i2", "ushort": "u2", "int": "i4", "uint": "u4", "float": "f4", "double": "f8",
              "int8": "i1", "uint8": "u1", "int16": "i2", "uint16": "u2", "int32": "i4", "uint32": "u4", "float32": "f4", "float64": "f8"}


def ply_attribute_names(n_albedo=3, n_specular=45, n_scale=3, n_rotation=4):
    """PLYExporter._construct_list_of_attributes (ply_exporter.py:34-47)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(n_albedo)]
    names += [f"f_rest_{i}" for i in range(n_specular)]
    names.append("opacity")
    names += [f"scale_{i}" for i in range(n_scale)]
    names += [f"rot_{i}" for i in range(n_rotation)]
    return names


def _read_ply_vertices(path):
    """The first element of a PLY file as a numpy structured array (ascii, binary_little_endian, binary_big_endian)."""
    with open(path, "rb") as fh:
        if fh.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, props, in_first, seen_element = None, None, [], False, False
        while True:
            line = fh.readline()
            if not line:
                raise ValueError(f"{path}: PLY header has no end_header")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] == "comment" or tok[0] == "obj_info":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_first = not seen_element
                if in_first:
                    count = int(tok[2])
                seen_element = True
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError(f"{path}: list properties in the vertex element are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt is None or count is None:
            raise ValueError(f"{path}: incomplete PLY header")
        if fmt == "ascii":
            rows = [fh.readline().split() for _ in range(count)]
            data = np.empty(count, dtype=[(n, t) for n, t in props])
            for i, (n, _t) in enumerate(props):
                data[n] = [float(r[i]) for r in rows]
            return data
        order = "<" if fmt == "binary_little_endian" else ">"
        dtype = np.dtype([(n, order + t) for n, t in props])
        raw = fh.read(count * dtype.itemsize)
        if len(raw) != count * dtype.itemsize:
            raise ValueError(f"{path}: truncated vertex data")
        return np.frombuffer(raw, dtype=dtype, count=count)


def _numbered(names, prefix):
    sel = [n for n in names if n.startswith(prefix)]
    return sorted(sel, key=lambda n: int(n.split("_")[-1]))


def read_ply(path, max_n_features=3):
    """MixtureOfGaussians.init_from_ply (model.py:671-713): -> dict of raw float32 parameters.
    f_rest_* is stored channel-major ([N,3,S]) in the file and coefficient-major ([N,S,3]) in the model."""
    v = _read_ply_vertices(path)
    names = v.dtype.names
    n = v.shape[0]
    out = {"positions": np.stack([v["x"], v["y"], v["z"]], axis=1).astype(np.float32),
           "density": np.asarray(v["opacity"], dtype=np.float32)[:, None],
           "features_albedo": np.stack([v["f_dc_0"], v["f_dc_1"], v["f_dc_2"]], axis=1).astype(np.float32)}
    rest = _numbered(names, "f_rest_")
    n_spec = (max_n_features + 1) ** 2 - 1
    if len(rest) != 3 * n_spec:
        raise ValueError(f"{path}: {len(rest)} f_rest_* properties, expected {3 * n_spec} for SH degree {max_n_features}")
    spec = np.stack([v[k] for k in rest], axis=1).astype(np.float32) if rest else np.zeros((n, 0), np.float32)
    out["features_specular"] = np.ascontiguousarray(spec.reshape(n, 3, n_spec).transpose(0, 2, 1).reshape(n, 3 * n_spec))
    out["scale"] = np.stack([v[k] for k in _numbered(names, "scale_")], axis=1).astype(np.float32)
    out["rotation"] = np.stack([v[k] for k in _numbered(names, "rot")], axis=1).astype(np.float32)
    return out


def write_ply(path, params):
    """PLYExporter.export (ply_exporter.py:50-84): binary little-endian, float32, the reference's property order;
    normals are the constant (0, 0, 1), density / scale / rotation are written pre-activation."""
    p = {k: np.asarray(params[k], dtype=np.float32) for k in PARAMETER_NAMES}
    n = p["positions"].shape[0]
    n_spec = p["features_specular"].shape[1] // 3
    spec = p["features_specular"].reshape(n, n_spec, 3).transpose(0, 2, 1).reshape(n, 3 * n_spec)
    normals = np.repeat(np.array([[0, 0, 1]], dtype=np.float32), n, axis=0)
    table = np.concatenate([p["positions"], normals, p["features_albedo"], spec, p["density"].reshape(n, 1), p["scale"], p["rotation"]],
                           axis=1).astype("<f4")
    names = ply_attribute_names(p["features_albedo"].shape[1], spec.shape[1], p["scale"].shape[1], p["rotation"].shape[1])
    assert table.shape[1] == len(names)
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {n}\n" + "".join(f"property float {k}\n" for k in names) + "end_header\n"
    with open(path, "wb") as fh:
        fh.write(header.encode("ascii"))
        fh.write(np.ascontiguousarray(table).tobytes())


def parameters_from_checkpoint(checkpoint):
    """The particle part of a reference checkpoint dict (model.py:108-140; consumed by init_from_checkpoint, :417-426)
    -> (dict of raw float32 numpy parameters, dict of the scalar fields)."""
    def arr(t):
        t = t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)
        return np.ascontiguousarray(t, dtype=np.float32)
    params = {k: arr(checkpoint[k]) for k in PARAMETER_NAMES}
    meta = {k: checkpoint[k] for k in ("n_active_features", "max_n_features", "scene_extent") if k in checkpoint}
    return params, meta


# ---------------------------------------------------------------------------------------------------
# COLMAP
# ---------------------------------------------------------------------------------------------------
# model id -> (name, number of parameters): the public COLMAP camera model table (datasets/utils.py:237-249)
COLMAP_CAMERA_MODELS = {0: ("SIMPLE_PINHOLE", 3), 1: ("PINHOLE", 4), 2: ("SIMPLE_RADIAL", 4), 3: ("RADIAL", 5), 4: ("OPENCV", 8),
                        5: ("OPENCV_FISHEYE", 8), 6: ("FULL_OPENCV", 12), 7: ("FOV", 5), 8: ("SIMPLE_RADIAL_FISHEYE", 4),
                        9: ("RADIAL_FISHEYE", 5), 10: ("THIN_PRISM_FISHEYE", 12)}


def read_colmap_cameras(path):
    """cameras.bin / cameras.txt -> {camera_id: dict(id, model, width, height, params float64[...])}
    (read_colmap_intrinsics_binary / _text, datasets/utils.py:336-409)."""
    cams = {}
    if str(path).endswith(".txt"):
        with open(path, "r") as fh:
            for line in fh:
                line = line.strip()
                if not line or line.startswith("#"):
                    continue
                cid, model, w, h, *prm = line.split()
                cams[int(cid)] = {"id": int(cid), "model": model, "width": int(w), "height": int(h), "params": np.array([float(x) for x in prm])}
        return cams
    with open(path, "rb") as fh:
        (count,) = struct.unpack("<Q", fh.read(8))
        for _ in range(count):
            cid, mid, w, h = struct.unpack("<iiQQ", fh.read(24))
            if mid not in COLMAP_CAMERA_MODELS:
                raise KeyError(f"Invalid camera model ID: {mid}")
            name, k = COLMAP_CAMERA_MODELS[mid]
            prm = struct.unpack("<" + "d" * k, fh.read(8 * k))
            assert cid not in cams, f"Camera ID {cid} already exists"
            cams[cid] = {"id": cid, "model": name, "width": int(w), "height": int(h), "params": np.array(prm)}
    return cams


def read_colmap_images(path):
    """images.bin / images.txt -> list of dict(id, qvec (w,x,y,z), tvec, camera_id, name), sorted by name
    (read_colmap_extrinsics_binary / _text, datasets/utils.py:439-565).  The 2D observations are skipped."""
    images = []
    if str(path).endswith(".txt"):
        # like the reference: blank and comment lines are dropped first, then lines are taken in (image, 2D points) pairs --
        # so an image without observations (an empty points line) is not representable in the text format
        with open(path, "r") as fh:
            lines = [ln.strip() for ln in fh]
        lines = [ln for ln in lines if ln and not ln.startswith("#")]
        if len(lines) % 2:
            raise ValueError(f"Error parsing extrinsics file: missing points data for image {lines[-1].split()[-1]}")
        for line in lines[0::2]:
            tok = line.split()
            if len(tok) < 10:
                raise ValueError(f"Error parsing extrinsics file: invalid image line format: {line}")
            images.append({"id": int(tok[0]), "qvec": np.array([float(x) for x in tok[1:5]]), "tvec": np.array([float(x) for x in tok[5:8]]),
                           "camera_id": int(tok[8]), "name": tok[9]})
        return sorted(images, key=lambda im: im["name"])
    with open(path, "rb") as fh:
        (count,) = struct.unpack("<Q", fh.read(8))
        for _ in range(count):
            props = struct.unpack("<idddddddi", fh.read(64))
            name = bytearray()
            while True:
                c = fh.read(1)
                if c == b"\x00" or c == b"":
                    break
                name += c
            (n2d,) = struct.unpack("<Q", fh.read(8))
            fh.seek(24 * n2d, 1)
            images.append({"id": props[0], "qvec": np.array(props[1:5]), "tvec": np.array(props[5:8]), "camera_id": props[8],
                           "name": name.decode("utf-8")})
    return sorted(images, key=lambda im: im["name"])


def qvec_to_so3(q):
    """datasets/utils.py:412-431 (q = w, x, y, z)."""
    w, x, y, z = q
    return np.array([[1 - 2 * y * y - 2 * z * z, 2 * x * y - 2 * w * z, 2 * z * x + 2 * w * y],
                     [2 * x * y + 2 * w * z, 1 - 2 * x * x - 2 * z * z, 2 * y * z - 2 * w * x],
                     [2 * z * x - 2 * w * y, 2 * y * z + 2 * w * x, 1 - 2 * x * x - 2 * y * y]])


def colmap_pose(image):
    """dataset_colmap.py:261-267: W2C = [R(qvec) | tvec] in float32, C2W = inv(W2C)."""
    w2c = np.zeros((4, 4), dtype=np.float32)
    w2c[:3, :3] = qvec_to_so3(image["qvec"])
    w2c[:3, 3] = np.array(image["tvec"])
    w2c[3, 3] = 1.0
    return np.linalg.inv(w2c)


def colmap_intrinsics(camera, downsample=1):
    """dataset_colmap.py:133-250 -> (kind, dict) with kind one of the Batch field names
    'intrinsics_OpenCVPinholeCameraModelParameters' / 'intrinsics_OpenCVFisheyeCameraModelParameters'.
    `downsample` is the integer factor between the calibrated and the loaded image size (images_2, images_4 ...)."""
    w, h = int(round(camera["width"] / downsample)), int(round(camera["height"] / downsample))
    prm = np.array(camera["params"], dtype=np.float64)
    if camera["model"] in ("SIMPLE_PINHOLE", "PINHOLE"):
        fx = prm[0] / downsample
        fy = (prm[0] if camera["model"] == "SIMPLE_PINHOLE" else prm[1]) / downsample
        return "intrinsics_OpenCVPinholeCameraModelParameters", {
            "resolution": np.array([w, h], dtype=np.int64), "shutter_type": "GLOBAL",
            "principal_point": np.array([w, h], dtype=np.float32) / 2, "focal_length": np.array([fx, fy], dtype=np.float32),
            "radial_coeffs": np.zeros((6,), np.float32), "tangential_coeffs": np.zeros((2,), np.float32),
            "thin_prism_coeffs": np.zeros((4,), np.float32)}
    if camera["model"] == "OPENCV_FISHEYE":
        prm = prm.copy()
        prm[:4] = prm[:4] / downsample
        principal, focal = prm[2:4].astype(np.float32), prm[0:2].astype(np.float32)
        res = np.array([w, h], dtype=np.float64)
        # compute_max_radius (datasets/utils.py:170-191): largest distance from the principal point to an image corner
        corners = np.array([[0.0, 0.0], [res[0], 0.0], [0.0, res[1]], [res[0], res[1]]])
        max_radius = float(np.max(np.linalg.norm(corners - principal[None, :].astype(np.float64), axis=1)))
        max_angle = max(2.0 * max_radius / focal[0], 2.0 * max_radius / focal[1]) / 2.0
        return "intrinsics_OpenCVFisheyeCameraModelParameters", {
            "resolution": res.astype(np.int64), "shutter_type": "GLOBAL", "principal_point": principal, "focal_length": focal,
            "radial_coeffs": prm[4:].astype(np.float32), "max_angle": float(max_angle)}
    raise ValueError(f"Colmap camera model '{camera['model']}' not handled: only undistorted datasets (PINHOLE, SIMPLE_PINHOLE or "
                     "OPENCV_FISHEYE cameras) are supported")


def colmap_views(cameras, images, downsample=1):
    """One dict per image (sorted by name): name, T_to_world [4,4] float32, and the intrinsics field of the Batch."""
    views = []
    for im in images:
        kind, intr = colmap_intrinsics(cameras[im["camera_id"]], downsample)
        views.append({"name": im["name"], "T_to_world": colmap_pose(im).astype(np.float32), kind: intr})
    return views
