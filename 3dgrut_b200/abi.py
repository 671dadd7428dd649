"""ctypes mirror of ``include/gut_b200.h`` (the C ABI of libgut_b200.so).

Field order and types must match the header exactly; ``tests/test_abi.py``
checks the struct sizes against ``gut_abi_sizeof_*`` exported by the library.
"""
import ctypes as C
import enum

GUT_ABI_VERSION = 1
GUT_COLOUR_DENSE, GUT_COLOUR_COMPACT = 0, 1     # gut_set_colour_layout


class GutStatus(enum.IntEnum):
    OK = 0
    BAD_INPUT = 1
    CUDA = 2
    NO_FORWARD = 3
    OUT_OF_MEMORY = 4
    UNSUPPORTED = 5


class ShutterType(enum.IntEnum):
    """reference: ``lib3dgut_cc.ShutterType`` (threedgut_tracer/bindings.cpp:87-92)."""

    ROLLING_TOP_TO_BOTTOM = 0
    ROLLING_LEFT_TO_RIGHT = 1
    ROLLING_BOTTOM_TO_TOP = 2
    ROLLING_RIGHT_TO_LEFT = 3
    GLOBAL = 4


class CameraModel(enum.IntEnum):
    OPENCV_PINHOLE = 0
    OPENCV_FISHEYE = 1


class GutConfig(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("kernel_degree", C.c_int32),
        ("min_kernel_response", C.c_float),
        ("min_alpha", C.c_float),
        ("max_alpha", C.c_float),
        ("min_transmittance", C.c_float),
        ("sph_max_degree", C.c_int32),
        ("enable_hitcounts", C.c_int32),
        ("rect_bounding", C.c_int32),
        ("tight_opacity_bounding", C.c_int32),
        ("tile_based_culling", C.c_int32),
        ("n_rolling_shutter_iterations", C.c_int32),
        ("ut_alpha", C.c_float),
        ("ut_beta", C.c_float),
        ("ut_kappa", C.c_float),
        ("ut_in_image_margin_factor", C.c_float),
        ("ut_require_all_sigma_points_valid", C.c_int32),
        ("k_buffer_size", C.c_int32),
        ("global_z_order", C.c_int32),
        ("enable_kernel_timings", C.c_int32),
    ]


def default_config(**overrides) -> GutConfig:
    """Reference defaults: configs/render/3dgut.yaml + configs/render/3dgrt.yaml."""
    cfg = GutConfig(
        abi_version=GUT_ABI_VERSION,
        kernel_degree=2,
        min_kernel_response=0.0113,
        min_alpha=1.0 / 255.0,
        max_alpha=0.99,
        min_transmittance=0.0001,
        sph_max_degree=3,
        enable_hitcounts=1,
        rect_bounding=1,
        tight_opacity_bounding=1,
        tile_based_culling=1,
        n_rolling_shutter_iterations=5,
        ut_alpha=1.0,
        ut_beta=2.0,
        ut_kappa=0.0,
        ut_in_image_margin_factor=0.1,
        ut_require_all_sigma_points_valid=0,
        k_buffer_size=0,
        global_z_order=1,
        enable_kernel_timings=0,
    )
    for k, v in overrides.items():
        if not hasattr(cfg, k):
            raise AttributeError(f"GutConfig has no field {k!r}")
        setattr(cfg, k, v)
    return cfg


class GutCamera(C.Structure):
    _fields_ = [
        ("model", C.c_int32),
        ("shutter", C.c_int32),
        ("principal_point", C.c_float * 2),
        ("focal_length", C.c_float * 2),
        ("radial_coeffs", C.c_float * 6),
        ("tangential_coeffs", C.c_float * 2),
        ("thin_prism_coeffs", C.c_float * 4),
        ("max_angle", C.c_float),
        ("start_pose", C.c_float * 7),
        ("end_pose", C.c_float * 7),
        ("start_timestamp_us", C.c_int64),
        ("end_timestamp_us", C.c_int64),
    ]


class GutStageOutputs(C.Structure):
    _fields_ = [
        ("d_tiles_count", C.c_void_p),
        ("d_proj_position", C.c_void_p),
        ("d_proj_conic_opacity", C.c_void_p),
        ("d_proj_extent", C.c_void_p),
        ("d_global_depth", C.c_void_p),
        ("d_features", C.c_void_p),
        ("d_sorted_keys", C.c_void_p),
        ("d_sorted_particle_idx", C.c_void_p),
        ("d_tile_ranges", C.c_void_p),
        ("capacity_intersections", C.c_uint64),
        ("num_intersections", C.c_uint64),
    ]


class GutFrameStats(C.Structure):
    _fields_ = [
        ("num_intersections", C.c_uint64),
        ("num_tiles", C.c_uint32),
        ("max_tile_list_length", C.c_uint32),
        ("kernel_launches", C.c_uint64),
    ]


class Activation(enum.IntEnum):
    """GutActivation (include/gut_b200.h): the model's activation table, threedgrut/utils/misc.py:48-53."""
    NONE = 0
    EXP = 1
    SIGMOID = 2


GUT_ADAM_MAX_GROUPS = 8


class GutAdamGroup(C.Structure):
    """One parameter group of gut_selective_adam_step: contiguous fp32 [N, width] tensors and its learning rate."""
    _fields_ = [
        ("param", C.c_void_p),
        ("grad", C.c_void_p),
        ("exp_avg", C.c_void_p),
        ("exp_avg_sq", C.c_void_p),
        ("width", C.c_uint32),
        ("lr", C.c_float),
    ]


class GutAdamTensors(C.Structure):
    """One parameter of the fused optimiser epilogue (gut_density_adam / gut_sh_expand_adam): raw parameter and its two
    moments, contiguous fp32, updated in place, and the learning rate."""
    _fields_ = [
        ("param", C.c_void_p),
        ("exp_avg", C.c_void_p),
        ("exp_avg_sq", C.c_void_p),
        ("lr", C.c_float),
    ]


def _fill(arr, values, n):
    vals = [float(v) for v in values]
    if len(vals) != n:
        raise ValueError(f"expected {n} values, got {len(vals)}")
    for i, v in enumerate(vals):
        arr[i] = v


def pinhole_camera(principal_point, focal_length, radial_coeffs=(0,) * 6, tangential_coeffs=(0, 0),
                   thin_prism_coeffs=(0,) * 4, shutter_type=ShutterType.GLOBAL) -> GutCamera:
    """reference: fromOpenCVPinholeCameraModelParameters (bindings.cpp:33-57)."""
    cam = GutCamera(model=int(CameraModel.OPENCV_PINHOLE), shutter=int(shutter_type))
    _fill(cam.principal_point, principal_point, 2)
    _fill(cam.focal_length, focal_length, 2)
    _fill(cam.radial_coeffs, radial_coeffs, 6)
    _fill(cam.tangential_coeffs, tangential_coeffs, 2)
    _fill(cam.thin_prism_coeffs, thin_prism_coeffs, 4)
    return cam


def fisheye_camera(principal_point, focal_length, radial_coeffs, max_angle,
                   shutter_type=ShutterType.GLOBAL) -> GutCamera:
    """reference: fromOpenCVFisheyeCameraModelParameters (bindings.cpp:59-77)."""
    cam = GutCamera(model=int(CameraModel.OPENCV_FISHEYE), shutter=int(shutter_type))
    _fill(cam.principal_point, principal_point, 2)
    _fill(cam.focal_length, focal_length, 2)
    rc = list(radial_coeffs) + [0.0, 0.0]
    _fill(cam.radial_coeffs, rc, 6)
    cam.max_angle = float(max_angle)
    return cam


def set_poses(cam: GutCamera, start_pose, end_pose=None, timestamps_us=(0, 1)) -> GutCamera:
    """Poses are world->sensor ``[tx,ty,tz,qx,qy,qz,qw]`` (sensors.h:33)."""
    if end_pose is None:
        end_pose = start_pose
    _fill(cam.start_pose, start_pose, 7)
    _fill(cam.end_pose, end_pose, 7)
    cam.start_timestamp_us = int(timestamps_us[0])
    cam.end_timestamp_us = int(timestamps_us[1])
    return cam
