"""Shared test helpers: a duck-typed Gaussian model, Batch construction, oracle glue and comparison metrics."""
import importlib
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

abi = importlib.import_module("3dgrut_b200.abi")
syn = importlib.import_module("3dgrut_b200.synthetic")
camera = importlib.import_module("3dgrut_b200.camera")
protocols = importlib.import_module("3dgrut_b200.protocols")

# relative distance to an accept/kill threshold below which one fp32 implementation may take the other branch
BORDERLINE_MARGIN = 2e-5


def make_conf(**render_overrides):
    """A plain-dict stand-in for the Hydra config (configs/render/3dgut.yaml + 3dgrt.yaml defaults)."""
    splat = {
        "rect_bounding": True, "tight_opacity_bounding": True, "tile_based_culling": True, "n_rolling_shutter_iterations": 5,
        "ut_alpha": 1.0, "ut_beta": 2.0, "ut_kappa": 0.0, "ut_in_image_margin_factor": 0.1,
        "ut_require_all_sigma_points_valid": False, "k_buffer_size": 0, "global_z_order": True,
    }
    render = {
        "method": "3dgut", "particle_kernel_degree": 2, "particle_kernel_min_response": 0.0113, "min_transmittance": 0.0001,
        "particle_kernel_min_alpha": 1.0 / 255.0, "particle_kernel_max_alpha": 0.99, "particle_radiance_sph_degree": 3,
        "enable_hitcounts": True, "enable_kernel_timings": False, "splat": splat,
    }
    for k, v in render_overrides.items():
        if k in splat:
            splat[k] = v
        else:
            render[k] = v
    return {"render": render}


def cfg_from_conf(conf):
    tracer_mod = importlib.import_module("3dgrut_b200.tracer")
    return tracer_mod.config_from_conf(conf)


class SyntheticGaussians:
    """The attributes of MixtureOfGaussians the Tracer touches (threedgut_tracer/tracer.py:310-339), post-activation."""

    def __init__(self, scene, device, n_active_features=3, requires_grad=False):
        def p(t):
            t = t.to(device).clone()
            return t.requires_grad_(requires_grad)

        self.positions = p(scene["positions"])
        self._rotation = p(scene["rotation"])
        self._scale = p(scene["scale"])
        self._density = p(scene["density"])
        self._features = p(scene["features"])
        self.n_active_features = n_active_features
        self.num_gaussians = self.positions.shape[0]

    def get_rotation(self):
        return self._rotation

    def get_scale(self):
        return self._scale

    def get_density(self):
        return self._density

    def get_features(self):
        return self._features

    def background(self, T_to_world, rays_d, rgb, opacity, train):
        return rgb, opacity  # black background (threedgrut/model/background.py:81-95)


def make_batch(view, device, pose_device=None):
    """``pose_device="cpu"`` keeps T_to_world a host tensor: the tracer consumes the pose on the host, and a device copy is read
    straight back with a stream-wide synchronisation (what the reference does, tracer.py:376)."""
    pose_device = device if pose_device is None else pose_device
    kw = {k: view[k] for k in ("intrinsics", "intrinsics_OpenCVPinholeCameraModelParameters",
                               "intrinsics_OpenCVFisheyeCameraModelParameters") if k in view}
    end = view.get("T_to_world_end")
    return protocols.Batch(rays_ori=view["rays_ori"].to(device), rays_dir=view["rays_dir"].to(device),
                           T_to_world=view["T_to_world"].to(pose_device),
                           T_to_world_end=None if end is None else end.to(pose_device), **kw)


def camera_of(view):
    b = types.SimpleNamespace(T_to_world=view["T_to_world"], T_to_world_end=view.get("T_to_world_end"),
                              intrinsics=view.get("intrinsics"),
                              intrinsics_OpenCVPinholeCameraModelParameters=view.get("intrinsics_OpenCVPinholeCameraModelParameters"),
                              intrinsics_OpenCVFisheyeCameraModelParameters=view.get("intrinsics_OpenCVFisheyeCameraModelParameters"))
    return camera.camera_from_batch(b)


def oracle_reference(scene, view, w, h, sph_degree=3, cfg=None, rgba_grad=None, dist_grad=None):
    """Run the CPU oracle on a synthetic scene/view; optionally also its backward."""
    from oracle import gut_oracle as O

    cfg = cfg if cfg is not None else abi.default_config()
    cam = camera_of(view)
    d12 = syn.pack_density(scene).numpy()
    sph = scene["features"].numpy()
    ro, rd = view["rays_ori"].numpy(), view["rays_dir"].numpy()
    fwd = O.forward(cfg, cam, d12, sph, ro, rd, w, h, sph_degree)
    fwd["borderline"] = fwd["margin"] < BORDERLINE_MARGIN
    if rgba_grad is not None:
        g12, gsph, gfeat = O.backward(cfg, cam, d12, sph, ro, rd, w, h, sph_degree, fwd, rgba_grad, dist_grad)
        fwd["density_grad"], fwd["radiance_grad"], fwd["feature_grad"] = g12, gsph, gfeat
    return fwd


def compare_images(test_rgba, ref_rgba, borderline):
    """max-abs over pixels whose accept decisions are not borderline in the oracle, plus the borderline statistics."""
    err = np.abs(test_rgba.reshape(ref_rgba.shape).astype(np.float64) - ref_rgba.astype(np.float64)).max(axis=-1)
    clean = ~borderline
    return {
        "max_abs_clean": float(err[clean].max()) if clean.any() else 0.0,
        "max_abs_borderline": float(err[borderline].max()) if borderline.any() else 0.0,
        "borderline_fraction": float(borderline.mean()),
        "max_abs_all": float(err.max()),
    }


def rel_l2(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / den) if den > 0 else float(np.linalg.norm(a))
