"""Seeded synthetic scenes and cameras of the shapes BASELINE.json names
(SURVEY.md section 8d).  CPU torch/numpy only; callers move tensors to the GPU.

Scene "S-init" mirrors ``MixtureOfGaussians.init_from_random_point_cloud``
(threedgrut/model/model.py:368-415, configs/initialization/random.yaml):
positions U[-1.5,1.5]^3, raw rotation U[0,1]^4 with w:=1 then normalised,
isotropic scale ~ nearest-neighbour distance, density 0.1, SH DC U[0,1]/255.
For large N the k-d tree is replaced by the closed form of the expected
nearest-neighbour distance of a uniform cloud, 0.554 (V/N)^(1/3), times a
LogNormal(0, 0.25) spread.  Rays follow ``pinhole_camera_rays``
(threedgrut/datasets/utils.py:39-59).
"""
import math

import numpy as np
import torch


def make_scene(n, sph_max_degree=3, seed=1000, kind="init", higher_order_std=0.05):
    """Returns dict of post-activation particle tensors (CPU fp32):
    positions [N,3], rotation [N,4] (wxyz, unit), scale [N,3], density [N,1], features [N,(d+1)^2*3]."""
    g = torch.Generator().manual_seed(int(seed))
    pos = torch.rand((n, 3), generator=g) * 3.0 - 1.5
    rot = torch.rand((n, 4), generator=g)
    rot[:, 0] = 1.0
    rot = torch.nn.functional.normalize(rot, dim=1)
    ncoef = (sph_max_degree + 1) ** 2
    feats = torch.zeros((n, ncoef, 3))
    feats[:, 0, :] = torch.rand((n, 3), generator=g) / 255.0
    if ncoef > 1:
        feats[:, 1:, :] = torch.randn((n, ncoef - 1, 3), generator=g) * higher_order_std
    if kind == "init":
        nn_dist = 0.554 * (27.0 / n) ** (1.0 / 3.0)
        spread = torch.exp(torch.randn((n, 1), generator=g) * 0.25)
        scale = (nn_dist * spread).clamp_min(1e-3).repeat(1, 3)
        density = torch.full((n, 1), 0.1)
    elif kind == "dense":
        scale = torch.exp(math.log(0.02) + 0.6 * torch.randn((n, 3), generator=g))
        density = torch.sigmoid(torch.randn((n, 1), generator=g) * 2.0)
    else:
        raise ValueError(kind)
    return {
        "positions": pos.contiguous(), "rotation": rot.contiguous(), "scale": scale.contiguous(),
        "density": density.contiguous(), "features": feats.reshape(n, ncoef * 3).contiguous(),
    }


def pack_density(scene):
    """[pos, density, rot(wxyz), scale, 0] -> [N,12]  (threedgut_tracer/tracer.py:176-178)."""
    n = scene["positions"].shape[0]
    return torch.cat([scene["positions"], scene["density"], scene["rotation"], scene["scale"],
                      torch.zeros((n, 1))], dim=1).contiguous()


def look_at_c2w(eye, target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0)):
    """Camera-to-world 4x4, OpenCV convention (+x right, +y down, +z forward)."""
    eye = np.asarray(eye, np.float64); target = np.asarray(target, np.float64); up = np.asarray(up, np.float64)
    fwd = target - eye; fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, up); right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = right, down, fwd, eye
    return m.astype(np.float32)


def orbit_c2w(k, n_views, radius=4.0, elevation_deg=20.0):
    az = 2.0 * math.pi * k / max(n_views, 1)
    el = math.radians(elevation_deg)
    eye = (radius * math.cos(el) * math.cos(az), radius * math.cos(el) * math.sin(az), radius * math.sin(el))
    return look_at_c2w(eye)


def pinhole_rays(w, h, fx, fy):
    """[1,H,W,3] origins (0) and unit directions; pixel centre +0.5 (datasets/utils.py:39-59)."""
    xs = ((np.arange(w, dtype=np.float32) + 0.5) - 0.5 * w) / fx
    ys = ((np.arange(h, dtype=np.float32) + 0.5) - 0.5 * h) / fy
    gx, gy = np.meshgrid(xs, ys)
    d = np.stack((gx, gy, np.ones_like(gx)), axis=-1).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    return torch.zeros((1, h, w, 3)), torch.from_numpy(d[None].astype(np.float32)).contiguous()


def fisheye_max_angle(w, h, fx, fy, cx, cy, k):
    """Largest valid ray angle: invert the forward polynomial at the farthest image corner
    (same idea as compute_max_radius, threedgrut/datasets/dataset_colmap.py:166-172), capped below pi/2."""
    r = max(math.hypot((u - cx) / fx, (v - cy) / fy) for u in (0.0, float(w)) for v in (0.0, float(h)))
    return float(min(_fisheye_theta_from_radius(np.array([r]), k)[0], 0.5 * math.pi - 1e-3))


def _fisheye_theta_from_radius(r, k, iters=20):
    theta = np.minimum(r.astype(np.float64), 1.5)
    for _ in range(iters):
        t2 = theta * theta
        f = theta * (1 + t2 * (k[0] + t2 * (k[1] + t2 * (k[2] + t2 * k[3])))) - r
        df = 1 + t2 * (3 * k[0] + t2 * (5 * k[1] + t2 * (7 * k[2] + t2 * 9 * k[3])))
        theta = theta - f / df
    return theta


def fisheye_rays(w, h, fx, fy, cx, cy, k):
    """Unit rays of an OpenCV-fisheye camera through pixel centres: the inverse of projectPoint
    (threedgut_tracer/include/3dgut/kernels/cuda/sensors/cameraProjections.cuh:105-128), by Newton
    iteration like image_points_to_camera_rays (threedgrut/datasets/camera_models.py:156-236)."""
    u = (np.arange(w, dtype=np.float64) + 0.5 - cx) / fx
    v = (np.arange(h, dtype=np.float64) + 0.5 - cy) / fy
    gu, gv = np.meshgrid(u, v)
    r = np.sqrt(gu * gu + gv * gv)
    theta = _fisheye_theta_from_radius(r, k)
    s = np.where(r > 1e-12, np.sin(theta) / np.maximum(r, 1e-12), 1.0)
    d = np.stack((gu * s, gv * s, np.cos(theta)), axis=-1)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    return torch.zeros((1, h, w, 3)), torch.from_numpy(d[None].astype(np.float32)).contiguous()


# distortion set of the "opencv_pinhole" synthetic camera: rational k1..k6, tangential p1/p2, thin-prism s1..s4, all
# non-zero and large enough that image corners leave the 0.8 < icD < 1.2 trust region of the reference's projection
# (cameraProjections.cuh:57-103)
OPENCV_PINHOLE_DISTORTION = {
    "radial_coeffs": (-0.28, 0.07, -0.008, 0.02, -0.01, 0.002),
    "tangential_coeffs": (0.0015, -0.0011),
    "thin_prism_coeffs": (0.0009, -0.0004, -0.0007, 0.0003),
}


def _opencv_distort(u, v, rc, tc, pc):
    """Normalised image point -> distorted normalised point (the forward model of cameraProjections.cuh:66-90)."""
    r2 = u * u + v * v
    icd = (1 + r2 * (rc[0] + r2 * (rc[1] + r2 * rc[2]))) / (1 + r2 * (rc[3] + r2 * (rc[4] + r2 * rc[5])))
    dx = 2 * tc[0] * u * v + tc[1] * (r2 + 2 * u * u) + r2 * (pc[0] + r2 * pc[1])
    dy = tc[0] * (r2 + 2 * v * v) + 2 * tc[1] * u * v + r2 * (pc[2] + r2 * pc[3])
    return u * icd + dx, v * icd + dy


def opencv_pinhole_rays(w, h, fx, fy, cx, cy, rc, tc, pc, iters=30):
    """Unit rays of a distorted OpenCV pinhole camera through pixel centres: fixed-point inversion of the distortion
    (as image_points_to_camera_rays does, threedgrut/datasets/camera_models.py:156-236)."""
    xd = (np.arange(w, dtype=np.float64) + 0.5 - cx) / fx
    yd = (np.arange(h, dtype=np.float64) + 0.5 - cy) / fy
    gxd, gyd = np.meshgrid(xd, yd)
    u, v = gxd.copy(), gyd.copy()
    for _ in range(iters):
        du, dv = _opencv_distort(u, v, rc, tc, pc)
        u = u - 0.7 * (du - gxd)
        v = v - 0.7 * (dv - gyd)
    d = np.stack((u, v, np.ones_like(u)), axis=-1)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    return torch.zeros((1, h, w, 3)), torch.from_numpy(d[None].astype(np.float32)).contiguous()


def make_view(w, h, k=0, n_views=1, model="pinhole", rolling=False, shutter=None):
    """One synthetic view: dict(rays_ori, rays_dir, T_to_world [1,4,4], intrinsics kwargs for Batch, T_to_world_end).
    ``shutter``: a ShutterType name for the OpenCV models; default ROLLING_TOP_TO_BOTTOM when ``rolling`` else GLOBAL."""
    c2w = orbit_c2w(k, n_views)
    out = {"T_to_world": torch.from_numpy(c2w[None].copy())}
    if shutter is None:
        shutter = "ROLLING_TOP_TO_BOTTOM" if rolling else "GLOBAL"
    rolling = shutter != "GLOBAL"
    if model == "pinhole":
        fx = fy = 1.0 * w
        out["rays_ori"], out["rays_dir"] = pinhole_rays(w, h, fx, fy)
        out["intrinsics"] = [fx, fy, w / 2.0, h / 2.0]
    elif model == "fisheye":
        fx = fy = 0.45 * w
        cx, cy = w / 2.0, h / 2.0
        kk = (-0.02, 0.003, 0.0, 0.0)
        out["rays_ori"], out["rays_dir"] = fisheye_rays(w, h, fx, fy, cx, cy, kk)
        out["intrinsics_OpenCVFisheyeCameraModelParameters"] = {
            "resolution": np.array([w, h], dtype=np.int64),
            "shutter_type": shutter,
            "principal_point": np.array([cx, cy], dtype=np.float32),
            "focal_length": np.array([fx, fy], dtype=np.float32),
            "radial_coeffs": np.array(kk, dtype=np.float32),
            "max_angle": fisheye_max_angle(w, h, fx, fy, cx, cy, kk),
        }
    elif model == "opencv_pinhole":
        fx, fy = 0.62 * w, 0.60 * w          # wide enough that r^2 reaches ~1 in the corners
        cx, cy = 0.49 * w, 0.52 * h          # principal point off the image centre
        dd = OPENCV_PINHOLE_DISTORTION
        out["rays_ori"], out["rays_dir"] = opencv_pinhole_rays(w, h, fx, fy, cx, cy, dd["radial_coeffs"], dd["tangential_coeffs"],
                                                               dd["thin_prism_coeffs"])
        out["intrinsics_OpenCVPinholeCameraModelParameters"] = {
            "resolution": np.array([w, h], dtype=np.int64),
            "shutter_type": shutter,
            "principal_point": np.array([cx, cy], dtype=np.float32),
            "focal_length": np.array([fx, fy], dtype=np.float32),
            "radial_coeffs": np.array(dd["radial_coeffs"], dtype=np.float32),
            "tangential_coeffs": np.array(dd["tangential_coeffs"], dtype=np.float32),
            "thin_prism_coeffs": np.array(dd["thin_prism_coeffs"], dtype=np.float32),
        }
    else:
        raise ValueError(model)
    if rolling:
        # end pose = start pose composed with a 2 degree yaw and a 0.05 translation
        a = math.radians(2.0)
        yaw = np.array([[math.cos(a), 0, math.sin(a), 0.05], [0, 1, 0, 0], [-math.sin(a), 0, math.cos(a), 0], [0, 0, 0, 1]], np.float32)
        out["T_to_world_end"] = torch.from_numpy((c2w @ yaw)[None].astype(np.float32).copy())
    return out
