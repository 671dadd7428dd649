#!/usr/bin/env python
"""Generates tests/golden/camera/expected_from_reference.json: what the REFERENCE's own Python
(`Tracer.__create_camera_parameters` and `SensorPose3DModel`, /root/reference/threedgut_tracer/tracer.py:60-151,362-431)
returns for seeded poses and the three intrinsics kinds.  Run in the build container only (it imports the reference file by
path, with stand-ins for the modules that are not installed: omegaconf, dataclasses_json, and the native plugin, whose factory
functions are replaced by recorders -- the code under test is the host-side conversion, which runs unchanged).

    python tests/golden/make_camera_golden.py
"""
import enum
import importlib.util
import json
import math
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "camera", "expected_from_reference.json")
REF = "/root/reference"


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


class _PluginShutter(enum.IntEnum):   # lib3dgut_cc.ShutterType (bindings.cpp:87-92), native order
    ROLLING_TOP_TO_BOTTOM = 0
    ROLLING_LEFT_TO_RIGHT = 1
    ROLLING_BOTTOM_TO_TOP = 2
    ROLLING_RIGHT_TO_LEFT = 3
    GLOBAL = 4


def load_reference_tracer():
    sys.modules.setdefault("omegaconf", types.SimpleNamespace(OmegaConf=object))
    dj = types.ModuleType("dataclasses_json")
    dj.DataClassJsonMixin = type("DataClassJsonMixin", (), {})
    dj.config = lambda **kw: {}
    sys.modules.setdefault("dataclasses_json", dj)
    for pkg in ("threedgrut", "threedgrut.datasets"):
        sys.modules.setdefault(pkg, types.ModuleType(pkg))
    _load("threedgrut.datasets.protocols", os.path.join(REF, "threedgrut/datasets/protocols.py"))
    cm = _load("threedgrut.datasets.camera_models", os.path.join(REF, "threedgrut/datasets/camera_models.py"))
    tr = _load("ref_threedgut_tracer", os.path.join(REF, "threedgut_tracer/tracer.py"))
    rec = types.SimpleNamespace(ShutterType=_PluginShutter)
    rec.fromOpenCVPinholeCameraModelParameters = lambda **kw: ("pinhole", kw)
    rec.fromOpenCVFisheyeCameraModelParameters = lambda **kw: ("fisheye", kw)
    tr._3dgut_plugin = rec
    return tr, cm


def random_c2w(rng, k):
    """A camera-to-world matrix: random rotation (QR of a Gaussian matrix, det +1) and translation; every fifth one is a
    near-degenerate rotation (180 degrees about an axis: exercises the non-trace branches of the matrix -> quaternion code)."""
    if k % 5 == 4:
        axis = np.eye(3)[k % 3]
        R = 2 * np.outer(axis, axis) - np.eye(3)
        R = R @ _small_rotation(rng, 1e-3)
    else:
        Q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        R = Q * np.sign(np.linalg.det(Q))
    m = np.eye(4, dtype=np.float32)
    m[:3, :3] = R.astype(np.float32)
    m[:3, 3] = rng.uniform(-6, 6, size=3).astype(np.float32)
    return m


def _small_rotation(rng, s):
    w = rng.normal(size=3) * s
    K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    return np.eye(3) + K + 0.5 * K @ K


def jsonable(v):
    if isinstance(v, enum.Enum):
        return v.name
    if isinstance(v, np.ndarray):
        return {"dtype": str(v.dtype), "values": [float(x) for x in v.reshape(-1)]}
    if isinstance(v, (np.floating, float)):
        return float(v)
    if isinstance(v, (np.integer, int)):
        return int(v)
    raise TypeError(type(v))


def main():
    tr, cm = load_reference_tracer()
    Batch = sys.modules["threedgrut.datasets.protocols"].Batch
    create = getattr(tr.Tracer, "_Tracer__create_camera_parameters")
    rng = np.random.default_rng(2024)
    cases = []
    shutters = list(cm.ShutterType)
    for k in range(24):
        c2w = random_c2w(rng, k)
        rays = torch.zeros((1, 2, 2, 3))
        T = torch.from_numpy(c2w[None].copy())
        kind = k % 3
        w, h = int(rng.integers(64, 2000)), int(rng.integers(64, 1200))
        kw = {}
        if kind == 0:      # [fx, fy, cx, cy] with a fractional principal point (the reference forces it to int(2c)/2)
            kw["intrinsics"] = [float(rng.uniform(0.5, 2.0) * w), float(rng.uniform(0.5, 2.0) * w), w / 2.0 + float(rng.uniform(0, 0.49)),
                                h / 2.0 + float(rng.uniform(0, 0.49))]
        elif kind == 1:
            kw["intrinsics_OpenCVPinholeCameraModelParameters"] = {
                "resolution": np.array([w, h], dtype=np.int64), "shutter_type": shutters[k % 5],
                "principal_point": rng.uniform(0.4, 0.6, size=2).astype(np.float32) * np.array([w, h], np.float32),
                "focal_length": rng.uniform(0.5, 2.0, size=2).astype(np.float32) * np.float32(w),
                "radial_coeffs": (rng.normal(size=6) * 0.05).astype(np.float32), "tangential_coeffs": (rng.normal(size=2) * 1e-3).astype(np.float32),
                "thin_prism_coeffs": (rng.normal(size=4) * 1e-3).astype(np.float32)}
        else:
            kw["intrinsics_OpenCVFisheyeCameraModelParameters"] = {
                "resolution": np.array([w, h], dtype=np.int64), "shutter_type": shutters[(k + 2) % 5],
                "principal_point": rng.uniform(0.4, 0.6, size=2).astype(np.float32) * np.array([w, h], np.float32),
                "focal_length": rng.uniform(0.3, 0.8, size=2).astype(np.float32) * np.float32(w),
                "radial_coeffs": (rng.normal(size=4) * 0.01).astype(np.float32), "max_angle": float(rng.uniform(0.8, 1.5))}
        batch = Batch(rays_ori=rays, rays_dir=rays, T_to_world=T, **kw)
        (model, sensor_kw), pose = create(batch)
        key = next(iter(kw))
        inp = kw[key] if kind == 0 else {a: jsonable(b) for a, b in kw[key].items()}
        cases.append({
            "T_to_world": [float(x) for x in c2w.reshape(-1)], "intrinsics_key": key, "intrinsics": inp,
            "expected_model": model, "expected_sensor": {a: jsonable(b) for a, b in sensor_kw.items()},
            "expected_poses": [[float(x) for x in p.numpy().astype(np.float32)] for p in pose.T_world_sensors],
            "expected_timestamps_us": [int(t) for t in pose.timestamps_us]})
    # the unsupported-intrinsics error path (tracer.py:429-431)
    # NB: with the reference's own Batch dataclass the f-string calls gpu_batch.keys(), which Batch does not have, so the
    # reference actually raises AttributeError there; recorded as found
    try:
        create(Batch(rays_ori=rays, rays_dir=rays, T_to_world=T))
        err = None
    except Exception as e:
        err = f"{type(e).__name__}: {e}"
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as fh:
        json.dump({"source": "threedgut_tracer/tracer.py:60-151,362-431 executed by tests/golden/make_camera_golden.py",
                   "cases": cases, "no_intrinsics_error": err}, fh, indent=1)
    print("wrote", OUT, len(cases), "cases; error text:", err)


if __name__ == "__main__":
    main()
