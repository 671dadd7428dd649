"""CPU checks of the oracle itself (no GPU): the scalar C restatement (oracle/gut_oracle.c, hand-derived backward)
against the independent PyTorch transcription (oracle/gut_torch_ref.py, autograd) and against closed-form cases.
PARITY UNPINNED: the reference has no vectors for this path, so agreement of two differently written restatements
plus analytic cases is what anchors the oracle."""
import math

import numpy as np
import pytest
import torch

from oracle import gut_oracle as O
from oracle import gut_torch_ref as T
from tests.helpers import BORDERLINE_MARGIN, abi, camera_of, rel_l2, syn


def _inputs(n, w, h, seed, kind="init", model="pinhole", rolling=False):
    scene = syn.make_scene(n, 3, seed=seed, kind=kind)
    view = syn.make_view(w, h, model=model, rolling=rolling)
    cam = camera_of(view)
    return scene, view, cam, syn.pack_density(scene).numpy(), scene["features"].numpy()


@pytest.mark.parametrize("model,rolling,deg", [("pinhole", False, 3), ("fisheye", False, 2), ("fisheye", True, 1), ("pinhole", True, 0)])
def test_projection_two_restatements_agree(model, rolling, deg):
    n, w, h = 3000, 160, 96
    cfg = abi.default_config()
    scene, view, cam, d12, sph = _inputs(n, w, h, seed=11, model=model, rolling=rolling)
    if model == "pinhole" and rolling:  # the [fx,fy,cx,cy] path is always global shutter: use the OpenCV-pinhole dict with distortion
        view.pop("intrinsics")
        view["intrinsics_OpenCVPinholeCameraModelParameters"] = {
            "resolution": np.array([w, h]), "shutter_type": "ROLLING_LEFT_TO_RIGHT", "principal_point": np.array([w / 2, h / 2], np.float32),
            "focal_length": np.array([w, w], np.float32), "radial_coeffs": np.array([0.05, -0.01, 0.002, 0.01, 0.0, 0.0], np.float32),
            "tangential_coeffs": np.array([1e-3, -5e-4], np.float32), "thin_prism_coeffs": np.array([1e-3, 0, -1e-3, 0], np.float32)}
        cam = camera_of(view)
    a = O.project(cfg, cam, d12, sph, w, h, deg)
    b = T.project(cfg, cam, d12, sph, w, h, deg)
    same = a["tiles_count"] == b["tiles_count"]
    assert same.mean() > 0.997, f"tiles_count differs on {(~same).sum()} of {n} particles"
    m = same & (a["tiles_count"] > 0)
    assert m.sum() > 200
    for k in ("proj_pos", "conic_opacity", "extent", "depth", "features"):
        np.testing.assert_allclose(a[k][m], b[k][m], rtol=2e-4, atol=2e-4, err_msg=k)


def test_binning_bit_exact_against_numpy_sort():
    n, w, h = 5000, 200, 120
    cfg = abi.default_config()
    scene, view, cam, d12, sph = _inputs(n, w, h, seed=21)
    proj = O.project(cfg, cam, d12, sph, w, h, 0)
    bins = O.bin_and_sort(cfg, w, h, proj)
    tiles, lists = T.tile_lists_from_projection(cfg, w, h, proj["proj_pos"], proj["conic_opacity"], proj["extent"])
    assert np.array_equal(tiles, proj["tiles_count"])          # K2's re-walk reproduces K1's counts
    ref = T.bin_and_sort(w, h, tiles, lists, proj["depth"])
    assert ref["M"] == bins["M"] > 0
    assert np.array_equal(ref["keys"], bins["keys"]) and np.array_equal(ref["values"], bins["values"])
    assert np.array_equal(ref["sorted_keys"], bins["sorted_keys"])
    assert np.array_equal(ref["sorted_values"], bins["sorted_values"])
    assert np.array_equal(ref["tile_ranges"], bins["tile_ranges"])
    # sortedness + stability: within a tile depth ascends, ties keep ascending particle index
    k = bins["sorted_keys"]; v = bins["sorted_values"]
    assert np.all(k[1:] >= k[:-1])
    tie = k[1:] == k[:-1]
    assert np.all(v[1:][tie] > v[:-1][tie])


@pytest.mark.parametrize("kind,deg,seed", [("init", 3, 5), ("dense", 2, 9)])
def test_compositing_and_gradients_two_restatements_agree(kind, deg, seed):
    n, w, h = 1500, 64, 48
    cfg = abi.default_config()
    scene, view, cam, d12, sph = _inputs(n, w, h, seed=seed, kind=kind)
    ro, rd = view["rays_ori"].numpy(), view["rays_dir"].numpy()
    fwd = O.forward(cfg, cam, d12, sph, ro, rd, w, h, deg)
    border = fwd["margin"] < BORDERLINE_MARGIN
    rng = np.random.default_rng(seed)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    rgba_grad[border] = 0; dist_grad[border] = 0   # borderline pixels may legitimately take another branch: keep them out of the gradient
    g12, gsph, _ = O.backward(cfg, cam, d12, sph, ro, rd, w, h, deg, fwd, rgba_grad, dist_grad)
    ref = T.forward_backward(cfg, cam, d12, sph, ro, rd, w, h, deg, fwd["bins"]["tile_ranges"], fwd["bins"]["sorted_values"],
                             fwd["proj"]["tiles_count"], rgba_grad, dist_grad)
    clean = ~border
    assert clean.mean() > 0.98
    assert np.abs(ref["rgba"] - fwd["rgba"])[clean].max() < 5e-5   # fp32 C vs fp64 torch
    assert np.abs(ref["dist"] - fwd["dist"])[clean].max() < 5e-5 * max(1.0, float(fwd["dist"].max()))
    assert np.array_equal(ref["hits"][clean], fwd["hits"][clean])
    assert fwd["hits"].max() > 3
    for lo, hi, what in ((0, 3, "position"), (3, 4, "density"), (4, 8, "rotation"), (8, 11, "scale")):
        e = rel_l2(g12[:, lo:hi], ref["density_grad"][:, lo:hi])
        assert e < 2e-4, (what, e)
    assert np.all(g12[:, 11] == 0)
    assert rel_l2(gsph, ref["radiance_grad"]) < 2e-4


@pytest.mark.parametrize("kind,deg,seed,K", [("init", 3, 5, 16), ("dense", 2, 9, 4), ("init", 1, 11, 1)])
def test_kbuffer_compositing_and_gradients_two_restatements_agree(kind, deg, seed, K):
    """k_buffer_size > 0 (gutKBufferRenderer.cuh:28-85, 217-292): the C oracle's hand-derived true-derivative backward against
    torch autograd of alpha / hit distance driven by the reference's un-blend recurrences."""
    n, w, h = 900, 48, 32
    cfg = abi.default_config()
    cfg.k_buffer_size = K
    scene, view, cam, d12, sph = _inputs(n, w, h, seed=seed, kind=kind)
    if kind == "dense":   # some negative colours: the K>0 backward un-blends with the UNclamped feature (reference quirk)
        sph = sph.copy(); sph[::7, 0:3] -= 2.5
    ro, rd = view["rays_ori"].numpy(), view["rays_dir"].numpy()
    fwd = O.forward(cfg, cam, d12, sph, ro, rd, w, h, deg)
    border = fwd["margin"] < BORDERLINE_MARGIN
    rng = np.random.default_rng(seed)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    rgba_grad[border] = 0; dist_grad[border] = 0
    g12, gsph, gfeat = O.backward(cfg, cam, d12, sph, ro, rd, w, h, deg, fwd, rgba_grad, dist_grad)
    ref = T.forward_backward_kbuffer(cfg, cam, d12, sph, ro, rd, w, h, deg, fwd["bins"]["tile_ranges"], fwd["bins"]["sorted_values"],
                                     fwd["proj"]["tiles_count"], rgba_grad, dist_grad)
    clean = ~border
    assert clean.mean() > 0.97
    assert np.abs(ref["rgba"] - fwd["rgba"])[clean].max() < 5e-5
    assert np.abs(ref["dist"] - fwd["dist"])[clean].max() < 5e-5 * max(1.0, float(fwd["dist"].max()))
    assert np.array_equal(ref["hits"][clean], fwd["hits"][clean])
    assert fwd["hits"].max() > min(K, 3)
    if K > 1:   # the sorted path really reorders: it differs from the unsorted K = 0 composite somewhere
        cfg0 = abi.default_config()
        unsorted = O.forward(cfg0, cam, d12, sph, ro, rd, w, h, deg)
        assert np.abs(unsorted["rgba"] - fwd["rgba"]).max() > 1e-4
    assert rel_l2(gfeat, ref["feature_grad"]) < 2e-4
    for lo, hi, what in ((0, 3, "position"), (3, 4, "density"), (4, 8, "rotation"), (8, 11, "scale")):
        e = rel_l2(g12[:, lo:hi], ref["density_grad"][:, lo:hi])
        assert e < 5e-4, (what, e)
    assert np.all(g12[:, 11] == 0)
    assert rel_l2(gsph, ref["radiance_grad"]) < 2e-4


def _single_particle(density, scale, pos=(0.0, 0.0, 0.0), w=33, h=33):
    """One isotropic Gaussian at `pos`, camera on the -y axis looking at the origin from distance 4."""
    view = syn.make_view(w, h)
    view["T_to_world"] = torch.from_numpy(syn.look_at_c2w((0.0, -4.0, 0.0))[None].copy())
    cam = camera_of(view)
    d12 = np.zeros((1, 12), np.float32)
    d12[0, 0:3] = pos; d12[0, 3] = density; d12[0, 4] = 1.0; d12[0, 8:11] = scale
    sph = np.zeros((1, 48), np.float32); sph[0, 0:3] = (0.3 / T.SH_C0, 0.1 / T.SH_C0, -0.9 / T.SH_C0)  # DC -> 0.5+(0.3,0.1,-0.9)
    return view, cam, d12, sph


def test_closed_form_single_gaussian_on_axis():
    cfg = abi.default_config()
    w = h = 33
    view, cam, d12, sph = _single_particle(0.6, 0.05)
    out = O.forward(cfg, cam, d12, sph, view["rays_ori"].numpy(), view["rays_dir"].numpy(), w, h, 0)
    c = out["rgba"][16, 16]
    # centre ray passes through the mean: response 1 -> alpha = density, colour = alpha * max(0.5 + DC, 0), dist = alpha * distance
    assert abs(c[3] - 0.6) < 1e-6
    np.testing.assert_allclose(c[:3], 0.6 * np.array([0.8, 0.6, 0.0]), atol=1e-6)   # the negative channel is clamped at 0
    assert abs(out["dist"][16, 16] - 0.6 * 4.0) < 1e-5
    assert out["hits"][16, 16] == 1
    # off-centre pixel: alpha = density * exp(-0.5 d^2 / s^2), d = distance of the ray to the mean
    rd = view["rays_dir"].numpy()[0, 16, 18]
    ang = math.acos(float(rd[2]))
    dmin = 4.0 * math.sin(ang)
    assert abs(out["rgba"][16, 18, 3] - 0.6 * math.exp(-0.5 * (dmin / 0.05) ** 2)) < 1e-5
    # alpha is capped at max_alpha
    view, cam, d12, sph = _single_particle(1.0, 0.05)
    out = O.forward(cfg, cam, d12, sph, view["rays_ori"].numpy(), view["rays_dir"].numpy(), w, h, 0)
    assert abs(out["rgba"][16, 16, 3] - 0.99) < 1e-6


def test_closed_form_culling_rules():
    cfg = abi.default_config()
    w = h = 33
    for density, pos, why in ((1.0 / 300.0, (0, 0, 0), "opacity below 1/255"), (0.6, (0.0, -3.9, 0.0), "closer than z = 0.2"),
                              (0.6, (0.0, -8.0, 0.0), "behind the camera")):
        view, cam, d12, sph = _single_particle(density, 0.05, pos)
        out = O.forward(cfg, cam, d12, sph, view["rays_ori"].numpy(), view["rays_dir"].numpy(), w, h, 0)
        assert out["bins"]["M"] == 0, why
        assert out["rgba"].max() == 0 and out["dist"].min() == 1e6   # M == 0: outputs keep their initial values


def test_two_overlapping_gaussians_order_dependence():
    cfg = abi.default_config()
    w = h = 33
    view, cam, d12, sph = _single_particle(0.5, 0.05)
    d12 = np.concatenate([d12, d12]); sph = np.concatenate([sph, sph])
    d12[1, 1] = 0.5           # second particle half a unit further along the view axis (+y)
    sph[1, 0:3] = 0.0         # colour 0.5 grey
    out = O.forward(cfg, cam, d12, sph, view["rays_ori"].numpy(), view["rays_dir"].numpy(), w, h, 0)
    c = out["rgba"][16, 16]
    near, far = 0.5 * np.array([0.8, 0.6, 0.0]), 0.5 * 0.5 * np.array([0.5, 0.5, 0.5])
    np.testing.assert_allclose(c[:3], near + far, atol=1e-6)
    assert abs(c[3] - 0.75) < 1e-6
    assert abs(out["dist"][16, 16] - (0.5 * 4.0 + 0.25 * 4.5)) < 1e-5
    assert out["hits"][16, 16] == 2


def test_higher_msb_and_sort_bits():
    # sort end bit = 32 + higherMsb(T) (src/gutRenderer.cu:79-94): C2 4056 tiles -> 44, C4/C5 8160 tiles -> 45, C1 64 tiles -> 39
    assert 32 + O.higher_msb(64) == 39 and 32 + O.higher_msb(4056) == 44 and 32 + O.higher_msb(8160) == 45
