"""camera.py against what the REFERENCE's own Python returns: tests/golden/camera/expected_from_reference.json
was produced by executing /root/reference/threedgut_tracer/tracer.py (`Tracer.__create_camera_parameters`, `SensorPose3DModel`)
on seeded poses and the three intrinsics kinds (tests/golden/make_camera_golden.py).  Poses and intrinsics must come out
bit-identical in fp32."""
import importlib
import json
import os
import types

import numpy as np
import pytest
import torch

camera = importlib.import_module("3dgrut_b200.camera")
abi = importlib.import_module("3dgrut_b200.abi")
protocols = importlib.import_module("3dgrut_b200.protocols")

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "camera", "expected_from_reference.json")
DATA = json.load(open(PATH))


def _arr(d):
    return np.array(d["values"], dtype=d["dtype"])


def _batch(case):
    T = torch.tensor(case["T_to_world"], dtype=torch.float32).reshape(1, 4, 4)
    rays = torch.zeros((1, 2, 2, 3))
    key = case["intrinsics_key"]
    if key == "intrinsics":
        val = list(case["intrinsics"])
    else:
        val = {k: (v if isinstance(v, (str, float, int)) else _arr(v)) for k, v in case["intrinsics"].items()}
    return protocols.Batch(rays_ori=rays, rays_dir=rays, T_to_world=T, **{key: val})


@pytest.mark.parametrize("i", range(len(DATA["cases"])))
def test_camera_parameters_equal_the_reference_python(i):
    case = DATA["cases"][i]
    cam = camera.camera_from_batch(_batch(case))
    exp = case["expected_sensor"]
    assert cam.model == (0 if case["expected_model"] == "pinhole" else 1)
    assert abi.ShutterType(cam.shutter).name == exp["shutter_type"]
    f32 = lambda a: np.array(list(a), dtype=np.float32)  # noqa: E731
    assert np.array_equal(f32(cam.principal_point), _arr(exp["principal_point"]).astype(np.float32))
    assert np.array_equal(f32(cam.focal_length), _arr(exp["focal_length"]).astype(np.float32))
    if case["expected_model"] == "pinhole":
        assert np.array_equal(f32(cam.radial_coeffs), _arr(exp["radial_coeffs"]).astype(np.float32))
        assert np.array_equal(f32(cam.tangential_coeffs), _arr(exp["tangential_coeffs"]).astype(np.float32))
        assert np.array_equal(f32(cam.thin_prism_coeffs), _arr(exp["thin_prism_coeffs"]).astype(np.float32))
    else:
        assert np.array_equal(f32(cam.radial_coeffs)[:4], _arr(exp["radial_coeffs"]).astype(np.float32))
        assert np.float32(cam.max_angle) == np.float32(exp["max_angle"])
    # poses: [t3, q xyzw], the same pose twice, timestamps 0 / 1 (tracer.py:138-151)
    want0, want1 = (np.array(p, dtype=np.float32) for p in case["expected_poses"])
    assert np.array_equal(f32(cam.start_pose), want0), (f32(cam.start_pose), want0)
    assert np.array_equal(f32(cam.end_pose), want1)
    assert [cam.start_timestamp_us, cam.end_timestamp_us] == case["expected_timestamps_us"]


def test_intrinsics_path_forces_the_principal_point_to_the_image_centre():
    case = next(c for c in DATA["cases"] if c["intrinsics_key"] == "intrinsics")
    fx, fy, cx, cy = case["intrinsics"]
    pp = _arr(case["expected_sensor"]["principal_point"])
    assert pp[0] == int(2 * cx) / 2 and pp[1] == int(2 * cy) / 2 and (pp[0] != cx or pp[1] != cy)


def test_missing_intrinsics_raise():
    """The reference raises here too -- an AttributeError in practice, because its f-string calls Batch.keys(), which its own
    Batch dataclass lacks (recorded in the fixture); this build raises the ValueError the reference's code intends."""
    assert DATA["no_intrinsics_error"].startswith("AttributeError")
    b = types.SimpleNamespace(T_to_world=torch.eye(4)[None], intrinsics=None, intrinsics_OpenCVPinholeCameraModelParameters=None,
                              intrinsics_OpenCVFisheyeCameraModelParameters=None)
    with pytest.raises(ValueError):
        camera.camera_from_batch(b)
