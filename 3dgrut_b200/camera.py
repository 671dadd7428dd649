"""Host-side camera / pose conversion of the 3DGUT plugin.

Mirrors ``Tracer.__create_camera_parameters`` and ``SensorPose3DModel`` of the
reference (threedgut_tracer/tracer.py:54-151, 362-431): a ``Batch`` carries a
camera-to-world matrix and one of three intrinsics descriptions; the native
layer wants a ``GutCamera`` (intrinsics + two world->sensor ``[t, q_xyzw]``
poses).  Host-side numpy (+ one CPU torch call), no CUDA, so it is testable on a CPU box; pinned on the reference's own
Python by tests/test_camera_golden.py.
"""
import math

import numpy as np
import torch

from .abi import GutCamera, ShutterType, fisheye_camera, pinhole_camera, set_poses


def so3_to_quat_xyzw(R: np.ndarray) -> np.ndarray:
    """Rotation matrix -> unit quaternion (x,y,z,w).

    Same branch selection as the reference's ``__so3_matrix_to_quat``
    (tracer.py:88-136): pick the largest of (R00, R11, R22, trace), fp32.
    """
    R = np.asarray(R, dtype=np.float32).reshape(3, 3)
    diag = np.array([R[0, 0], R[1, 1], R[2, 2]], dtype=np.float32)
    trace = np.float32(diag.sum(dtype=np.float32))
    decision = np.concatenate([diag, [trace]]).astype(np.float32)
    choice = int(np.argmax(decision))
    q = np.empty(4, dtype=np.float32)
    if choice != 3:
        i = choice
        j = (i + 1) % 3
        k = (j + 1) % 3
        q[i] = np.float32(1) - trace + np.float32(2) * R[i, i]
        q[j] = R[j, i] + R[i, j]
        q[k] = R[k, i] + R[i, k]
        q[3] = R[k, j] - R[j, k]
    else:
        q[0] = R[2, 1] - R[1, 2]
        q[1] = R[0, 2] - R[2, 0]
        q[2] = R[1, 0] - R[0, 1]
        q[3] = np.float32(1) + trace
    # the reference normalises with torch (`quat / torch.norm(quat, dim=1)[:, None]`, tracer.py:131); torch's fp32 norm and
    # numpy's differ in the last bit for some inputs, and the pose must come out bit-identical (tests/test_camera_golden.py)
    qt = torch.from_numpy(q.reshape(1, 4))
    return (qt / torch.norm(qt, dim=1)[:, None]).reshape(4).numpy().astype(np.float32)


def c2w_to_sensor_pose(T_to_world: np.ndarray) -> np.ndarray:
    """4x4 (or 3x4) camera-to-world -> world->sensor ``[tx,ty,tz,qx,qy,qz,qw]`` fp32.

    reference: tracer.py:373-383 (fp64 inverse of the C2W matrix), then
    SensorPose3DModel.__getWorld2View2 (tracer.py:75-85: two more fp64
    inverses, narrowed to fp32) and get_sensor_pose (tracer.py:138-151).
    """
    pose = np.asarray(T_to_world)
    pose = pose.reshape(pose.shape[-2], pose.shape[-1])
    C2W = np.concatenate((pose[:3, :4], np.zeros((1, 4))))  # promotes to fp64 like the reference
    C2W[3, 3] = 1.0
    W2C = np.linalg.inv(C2W)
    # __getWorld2View2(R = W2C[:3,:3]^T, t): Rt[:3,:3] = R^T, translate = 0, scale = 1
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = W2C[:3, :3]
    Rt[:3, 3] = W2C[:3, 3]
    Rt[3, 3] = 1.0
    Rt = np.float32(np.linalg.inv(np.linalg.inv(Rt)))
    t = Rt[:3, 3]
    q = so3_to_quat_xyzw(Rt[:3, :3])
    return np.concatenate([t, q]).astype(np.float32)


def _focal2fov(focal: float, pixels: int) -> float:
    return 2 * math.atan(pixels / (2 * focal))


def camera_from_batch(batch, end_T_to_world=None) -> GutCamera:
    """Build the native camera description from a ``Batch``-like object.

    reference: Tracer.__create_camera_parameters (tracer.py:362-431).  The
    reference always passes the same pose for shutter start and end
    (tracer.py:148-150); ``end_T_to_world`` (or ``batch.T_to_world_end``) is
    this build's extension for real rolling-shutter input and defaults to the
    reference behaviour.
    """
    T = batch.T_to_world
    T = T.detach().cpu().numpy() if hasattr(T, "detach") else np.asarray(T)
    T = np.squeeze(T)
    assert T.ndim == 2, "3DGUT renders one view per call (tracer.py:374-375)"
    start = c2w_to_sensor_pose(T)
    if end_T_to_world is None:
        end_T_to_world = getattr(batch, "T_to_world_end", None)
    if end_T_to_world is not None:
        E = end_T_to_world.detach().cpu().numpy() if hasattr(end_T_to_world, "detach") else np.asarray(end_T_to_world)
        end = c2w_to_sensor_pose(np.squeeze(E))
    else:
        end = start

    K = getattr(batch, "intrinsics", None)
    if K is not None:
        fx, fy, cx, cy = (float(K[0]), float(K[1]), float(K[2]), float(K[3]))
        w = int(2 * cx)
        h = int(2 * cy)
        fov_x = _focal2fov(fx, w)
        fov_y = _focal2fov(fy, h)
        cam = pinhole_camera(
            principal_point=np.array([w, h], dtype=np.float32) / 2,
            focal_length=np.array([w / (2.0 * math.tan(fov_x * 0.5)), h / (2.0 * math.tan(fov_y * 0.5))], dtype=np.float32),
            shutter_type=ShutterType.GLOBAL,
        )
        return set_poses(cam, start, end)

    K = getattr(batch, "intrinsics_OpenCVPinholeCameraModelParameters", None)
    if K is not None:
        cam = pinhole_camera(
            principal_point=K["principal_point"], focal_length=K["focal_length"], radial_coeffs=K["radial_coeffs"],
            tangential_coeffs=K["tangential_coeffs"], thin_prism_coeffs=K["thin_prism_coeffs"],
            shutter_type=_shutter(K["shutter_type"]),
        )
        return set_poses(cam, start, end)

    K = getattr(batch, "intrinsics_OpenCVFisheyeCameraModelParameters", None)
    if K is not None:
        cam = fisheye_camera(
            principal_point=K["principal_point"], focal_length=K["focal_length"], radial_coeffs=K["radial_coeffs"],
            max_angle=K["max_angle"], shutter_type=_shutter(K["shutter_type"]),
        )
        return set_poses(cam, start, end)

    keys = list(vars(batch).keys()) if hasattr(batch, "__dict__") else []
    raise ValueError(f"Camera intrinsics unavailable or unsupported, input keys are [{', '.join(keys)}]")


def _shutter(value) -> ShutterType:
    """Accept this package's enum, the reference's ``datasets.camera_models.ShutterType``
    (an ``auto()`` Enum whose *names* match), an int in native order, or a name string."""
    if isinstance(value, ShutterType):
        return value
    name = getattr(value, "name", None)
    if name is not None:
        return ShutterType[name]
    if isinstance(value, str):
        return ShutterType[value]
    return ShutterType(int(value))
