"""GPU parity: libgut_b200 (through the Tracer / SplatRaster API and the C ABI) against the CPU oracle.

Bars (BASELINE.json north_star): tile lists / sorted keys / ranges bit-exact; RGB/alpha/depth <= 1e-4 max-abs;
particle gradients <= 1e-3 relative L2.  Pixels whose accept/kill decisions sit within BORDERLINE_MARGIN of a
threshold in the oracle are compared separately (two fp32 implementations may legitimately take different
branches there and differ by one whole hit); their fraction is bounded.
"""
import importlib

import numpy as np
import pytest
import torch

from tests.helpers import (SyntheticGaussians, abi, camera_of, cfg_from_conf, compare_images, make_batch, make_conf,
                           oracle_reference, rel_l2, syn)

pytestmark = pytest.mark.gpu

TOL_IMG = 1e-4
TOL_GRAD = 1e-3


@pytest.fixture(scope="module")
def pkg():
    return importlib.import_module("3dgrut_b200")


def _raster(pkg, **over):
    native = importlib.import_module("3dgrut_b200.native")
    return native.SplatRaster(cfg_from_conf(make_conf(**over)))


def _run_forward(raster, scene, view, sph_degree, stage=True):
    dev = "cuda"
    d12 = syn.pack_density(scene).to(dev)
    sph = scene["features"].to(dev)
    cam = camera_of(view)
    out = raster.trace(0, sph_degree, d12, sph, view["rays_ori"].to(dev), view["rays_dir"].to(dev), None, cam, stage_outputs=stage)
    torch.cuda.synchronize()
    return out, (d12, sph, cam)


CASES = [
    # name, N, W, H, sph_degree, kind, seed
    ("c1_oracle_config", 10000, 128, 128, 0, "init", 1001),
    ("sh3_ragged_image", 6000, 150, 70, 3, "init", 5),
    ("dense_anisotropic", 3000, 96, 96, 2, "dense", 9),
]


@pytest.mark.parametrize("name,n,w,h,deg,kind,seed", CASES)
def test_forward_stages_and_image(pkg, name, n, w, h, deg, kind, seed):
    scene = syn.make_scene(n, 3, seed=seed, kind=kind)
    view = syn.make_view(w, h)
    ref = oracle_reference(scene, view, w, h, sph_degree=deg)
    raster = _raster(pkg)
    (rgba, dist, hits, vis, st), _ = _run_forward(raster, scene, view, deg)

    # K1: integer result exact, floats to 1e-5 relative (pinhole / global shutter is bit-reproducible)
    tc = st["tiles_count"].cpu().numpy().astype(np.uint32)
    assert np.array_equal(tc, ref["proj"]["tiles_count"]), f"{name}: tiles_count differs on {(tc != ref['proj']['tiles_count']).sum()} particles"
    m = tc > 0
    for key_gpu, key_ref in (("proj_position", "proj_pos"), ("proj_conic_opacity", "conic_opacity"), ("proj_extent", "extent"),
                             ("global_depth", "depth"), ("features", "features")):
        a = st[key_gpu].cpu().numpy()[m]; b = ref["proj"][key_ref][m]
        np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6, err_msg=f"{name}: {key_gpu}")
    assert np.array_equal(vis.cpu().numpy().reshape(-1)[m] > 0, ref["visibility"][m] > 0)

    # binning: bit-exact
    bins = ref["bins"]
    assert st["num_intersections"] == bins["M"]
    assert np.array_equal(st["sorted_keys"].cpu().numpy().view(np.uint64), bins["sorted_keys"])
    assert np.array_equal(st["sorted_particle_idx"].cpu().numpy().view(np.uint32), bins["sorted_values"])
    assert np.array_equal(st["tile_ranges"].cpu().numpy().view(np.uint32), bins["tile_ranges"])

    # image
    stats = compare_images(rgba.cpu().numpy(), ref["rgba"], ref["borderline"])
    assert stats["max_abs_clean"] <= TOL_IMG, (name, stats)
    assert stats["borderline_fraction"] < 0.02, (name, stats)
    assert stats["max_abs_borderline"] < 2e-2, (name, stats)
    clean = ~ref["borderline"]
    d_err = np.abs(dist.cpu().numpy().reshape(h, w) - ref["dist"])[clean]
    assert d_err.max() <= TOL_IMG * max(1.0, float(np.abs(ref["dist"]).max())), (name, d_err.max())
    assert np.array_equal(hits.cpu().numpy().reshape(h, w)[clean], ref["hits"][clean])


@pytest.mark.parametrize("name,n,w,h,deg,kind,seed", CASES)
def test_backward_gradients(pkg, name, n, w, h, deg, kind, seed):
    scene = syn.make_scene(n, 3, seed=seed, kind=kind)
    view = syn.make_view(w, h)
    rng = np.random.default_rng(seed)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    ref = oracle_reference(scene, view, w, h, sph_degree=deg, rgba_grad=rgba_grad, dist_grad=dist_grad)
    raster = _raster(pkg)
    (rgba, dist, hits, vis), (d12, sph, cam) = _run_forward(raster, scene, view, deg, stage=False)
    g12, gsph = raster.trace_bwd(0, deg, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                                 rgba, torch.from_numpy(rgba_grad).cuda(), dist, torch.from_numpy(dist_grad).cuda().reshape(h, w, 1))
    torch.cuda.synchronize()
    g12 = g12.cpu().numpy(); gsph = gsph.cpu().numpy()
    assert np.all(g12[:, 11] == 0)
    for lo, hi, what in ((0, 3, "position"), (3, 4, "density"), (4, 8, "rotation"), (8, 11, "scale")):
        e = rel_l2(g12[:, lo:hi], ref["density_grad"][:, lo:hi])
        assert e <= TOL_GRAD, (name, what, e)
    e = rel_l2(gsph, ref["radiance_grad"])
    assert e <= TOL_GRAD, (name, "sph", e)


def test_non_default_thresholds_take_the_run_time_path(pkg):
    """The reference's default thresholds are immediates in the degree-2 kernels (gut_render.cu, STD = true); any other value
    of particle_kernel_min_response / min_alpha / max_alpha / min_transmittance runs the instantiation that reads GutConfig."""
    n, w, h, deg = 3000, 80, 64, 2
    scene = syn.make_scene(n, 3, seed=31, kind="dense")
    view = syn.make_view(w, h)
    over = dict(particle_kernel_min_response=0.05, particle_kernel_min_alpha=0.01, particle_kernel_max_alpha=0.9, min_transmittance=0.02)
    cfg = cfg_from_conf(make_conf(**over))
    assert abs(cfg.max_alpha - 0.9) < 1e-6 and abs(cfg.min_transmittance - 0.02) < 1e-6
    rng = np.random.default_rng(31)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    ref0 = oracle_reference(scene, view, w, h, sph_degree=deg, cfg=cfg)
    border = ref0["borderline"]
    rgba_grad[border] = 0; dist_grad[border] = 0
    ref = oracle_reference(scene, view, w, h, sph_degree=deg, cfg=cfg, rgba_grad=rgba_grad, dist_grad=dist_grad)
    ref_std = oracle_reference(scene, view, w, h, sph_degree=deg)
    assert np.abs(ref["rgba"] - ref_std["rgba"]).max() > 1e-3       # the thresholds matter on this scene
    raster = _raster(pkg, **over)
    (rgba, dist, hits, vis), (d12, sph, cam) = _run_forward(raster, scene, view, deg, stage=False)
    stats = compare_images(rgba.cpu().numpy(), ref["rgba"], border)
    assert stats["max_abs_clean"] <= TOL_IMG, stats
    rgba_o = torch.from_numpy(ref["rgba"]).cuda(); dist_o = torch.from_numpy(ref["dist"]).cuda().reshape(h, w, 1)
    g12, gsph = raster.trace_bwd(0, deg, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                                 rgba_o, torch.from_numpy(rgba_grad).cuda(), dist_o, torch.from_numpy(dist_grad).cuda().reshape(h, w, 1))
    torch.cuda.synchronize()
    assert rel_l2(g12.cpu().numpy(), ref["density_grad"]) <= TOL_GRAD
    assert rel_l2(gsph.cpu().numpy(), ref["radiance_grad"]) <= TOL_GRAD


@pytest.mark.parametrize("K", [0, 8])
def test_rays_not_starting_at_the_frame_origin(pkg, K):
    """Ray bundles whose origins differ per pixel (e.g. an orthographic or multi-sensor rig) take the kernels' non-uniform
    path: the canonical origin is corrected per ray (gut_render.cu particle_test<.., UNIFORM=false>).  Forward and backward
    against the oracle, for the unsorted and the sorted compositing."""
    n, w, h, deg, seed = 4000, 80, 56, 2, 33
    scene = syn.make_scene(n, 3, seed=seed, kind="init")
    view = syn.make_view(w, h)
    g = torch.Generator().manual_seed(seed)
    view["rays_ori"] = view["rays_ori"] + (torch.rand(view["rays_ori"].shape, generator=g) - 0.5) * 0.2   # camera-space offsets
    cfg = abi.default_config(); cfg.k_buffer_size = K
    rng = np.random.default_rng(seed)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    ref0 = oracle_reference(scene, view, w, h, sph_degree=deg, cfg=cfg)
    border = ref0["borderline"]
    rgba_grad[border] = 0; dist_grad[border] = 0
    ref = oracle_reference(scene, view, w, h, sph_degree=deg, cfg=cfg, rgba_grad=rgba_grad, dist_grad=dist_grad)
    raster = _raster(pkg, k_buffer_size=K)
    (rgba, dist, hits, vis), (d12, sph, cam) = _run_forward(raster, scene, view, deg, stage=False)
    stats = compare_images(rgba.cpu().numpy(), ref["rgba"], border)
    assert stats["max_abs_clean"] <= TOL_IMG, stats
    assert stats["borderline_fraction"] < 0.08, stats
    clean = ~border
    assert np.abs(dist.cpu().numpy().reshape(h, w) - ref["dist"])[clean].max() <= TOL_IMG * max(1.0, float(np.abs(ref["dist"]).max()))
    assert np.array_equal(hits.cpu().numpy().reshape(h, w)[clean], ref["hits"][clean])
    rgba_o = torch.from_numpy(ref["rgba"]).cuda(); dist_o = torch.from_numpy(ref["dist"]).cuda().reshape(h, w, 1)
    g12, gsph = raster.trace_bwd(0, deg, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                                 rgba_o, torch.from_numpy(rgba_grad).cuda(), dist_o, torch.from_numpy(dist_grad).cuda().reshape(h, w, 1))
    torch.cuda.synchronize()
    g12 = g12.cpu().numpy()
    for lo, hi, what in ((0, 3, "position"), (3, 4, "density"), (4, 8, "rotation"), (8, 11, "scale")):
        e = rel_l2(g12[:, lo:hi], ref["density_grad"][:, lo:hi])
        assert e <= TOL_GRAD, (K, what, e)
    assert rel_l2(gsph.cpu().numpy(), ref["radiance_grad"]) <= TOL_GRAD


KB_CASES = [
    # name, N, W, H, sph_degree, kind, seed, K
    ("sorted_k16_init", 6000, 150, 70, 3, "init", 5, 16),
    ("sorted_k4_dense", 3000, 96, 96, 2, "dense", 9, 4),
    ("sorted_k8_init", 4000, 64, 80, 1, "init", 21, 8),
    ("sorted_k1_dense", 3000, 96, 96, 2, "dense", 10, 1),
    ("sorted_k2_init", 4000, 64, 80, 3, "init", 22, 2),
    ("sorted_k12_dense", 3000, 80, 64, 0, "dense", 11, 12),
]


@pytest.mark.parametrize("name,n,w,h,deg,kind,seed,K", KB_CASES)
def test_kbuffer_forward_and_backward(pkg, name, n, w, h, deg, kind, seed, K):
    """k_buffer_size > 0 (sorted compositing, gutKBufferRenderer.cuh:28-292): image and gradients against the oracle."""
    scene = syn.make_scene(n, 3, seed=seed, kind=kind)
    if kind == "dense":   # negative colours exercise the unclamped un-blend of the K>0 backward
        scene["features"][::7, 0:3] -= 2.5
    view = syn.make_view(w, h)
    cfg = abi.default_config(); cfg.k_buffer_size = K
    rng = np.random.default_rng(seed)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    ref0 = oracle_reference(scene, view, w, h, sph_degree=deg, cfg=cfg)
    border = ref0["borderline"]
    rgba_grad[border] = 0; dist_grad[border] = 0      # a ray that may take another branch is kept out of the gradient comparison
    ref = oracle_reference(scene, view, w, h, sph_degree=deg, cfg=cfg, rgba_grad=rgba_grad, dist_grad=dist_grad)
    raster = _raster(pkg, k_buffer_size=K)
    (rgba, dist, hits, vis), (d12, sph, cam) = _run_forward(raster, scene, view, deg, stage=False)
    stats = compare_images(rgba.cpu().numpy(), ref["rgba"], border)
    assert stats["max_abs_clean"] <= TOL_IMG, (name, stats)
    assert stats["borderline_fraction"] < 0.08, (name, stats)   # includes rays with two hit distances within 2e-6 (order may flip)
    assert stats["max_abs_borderline"] < 2e-2, (name, stats)
    clean = ~border
    d_err = np.abs(dist.cpu().numpy().reshape(h, w) - ref["dist"])[clean]
    assert d_err.max() <= TOL_IMG * max(1.0, float(np.abs(ref["dist"]).max())), (name, d_err.max())
    assert np.array_equal(hits.cpu().numpy().reshape(h, w)[clean], ref["hits"][clean])
    # the sorted composite differs from the unsorted one (otherwise this test would not exercise the buffer)
    ref_unsorted = oracle_reference(scene, view, w, h, sph_degree=deg)
    if K > 1:   # a buffer of one entry releases every hit as the next arrives: list order, the K = 0 image
        assert np.abs(ref_unsorted["rgba"] - ref["rgba"]).max() > 1e-4
    # gradients: feed the ORACLE's forward outputs so that both backward passes un-blend the same image
    rgba_o = torch.from_numpy(ref["rgba"]).cuda(); dist_o = torch.from_numpy(ref["dist"]).cuda().reshape(h, w, 1)
    g12, gsph = raster.trace_bwd(0, deg, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                                 rgba_o, torch.from_numpy(rgba_grad).cuda(), dist_o, torch.from_numpy(dist_grad).cuda().reshape(h, w, 1))
    torch.cuda.synchronize()
    g12 = g12.cpu().numpy(); gsph = gsph.cpu().numpy()
    assert np.all(g12[:, 11] == 0)
    for lo, hi, what in ((0, 3, "position"), (3, 4, "density"), (4, 8, "rotation"), (8, 11, "scale")):
        e = rel_l2(g12[:, lo:hi], ref["density_grad"][:, lo:hi])
        assert e <= TOL_GRAD, (name, what, e)
    e = rel_l2(gsph, ref["radiance_grad"])
    assert e <= TOL_GRAD, (name, "sph", e)


def test_tracer_autograd_matches_native(pkg):
    """The Tracer.render autograd path returns the same gradients as the raw trace/trace_bwd calls."""
    n, w, h, deg = 3000, 64, 48, 3
    scene = syn.make_scene(n, 3, seed=3)
    view = syn.make_view(w, h)
    tracer = pkg.Tracer(make_conf())
    g = SyntheticGaussians(scene, "cuda", n_active_features=deg, requires_grad=True)
    out = tracer.render(g, make_batch(view, "cuda"), train=True)
    assert out["pred_rgb"].shape == (1, h, w, 3) and out["pred_opacity"].shape == (1, h, w, 1)
    assert out["pred_dist"].shape == (1, h, w, 1) and out["hits_count"].shape == (1, h, w, 1)
    assert out["mog_visibility"].shape == (n, 1)
    assert out["pred_normals"].shape == (1, h, w, 3)     # normalize(ones_like(pred_rgb)), tracer.py:347
    assert torch.allclose(out["pred_normals"], torch.full((1, h, w, 3), 3.0 ** -0.5, device="cuda"))
    rng = np.random.default_rng(0)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    rg = torch.from_numpy(rgba_grad).cuda()
    loss = (out["pred_rgb"][0] * rg[..., :3]).sum() + (out["pred_opacity"][0] * rg[..., 3:]).sum() + \
        (out["pred_dist"][0, ..., 0] * torch.from_numpy(dist_grad).cuda()).sum()
    loss.backward()
    ref = oracle_reference(scene, view, w, h, sph_degree=deg, rgba_grad=rgba_grad, dist_grad=dist_grad)
    assert rel_l2(g.positions.grad.cpu().numpy(), ref["density_grad"][:, 0:3]) <= TOL_GRAD
    assert rel_l2(g._density.grad.cpu().numpy(), ref["density_grad"][:, 3:4]) <= TOL_GRAD
    assert rel_l2(g._rotation.grad.cpu().numpy(), ref["density_grad"][:, 4:8]) <= TOL_GRAD
    assert rel_l2(g._scale.grad.cpu().numpy(), ref["density_grad"][:, 8:11]) <= TOL_GRAD
    assert rel_l2(g._features.grad.cpu().numpy(), ref["radiance_grad"]) <= TOL_GRAD


@pytest.mark.parametrize("max_deg", [3, 2, 1, 0])
def test_deferred_sh_gradient_equals_per_view_accumulation(pkg, max_deg):
    """gut_backward_colour + gut_sh_gradient_expand over three views == gut_backward in accumulate mode over the same views,
    for every particle_radiance_sph_degree the reference accepts (16, 9, 4, 1 SH coefficients per particle)."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    n, w, h, deg = 5000, 96, 64, max_deg
    scene = syn.make_scene(n, max_deg, seed=17)
    d12 = syn.pack_density(scene).cuda(); sph = scene["features"].cuda()
    views, ups = [], []
    g = torch.Generator().manual_seed(4)
    for k in range(3):
        hv = syn.make_view(w, h, k=k, n_views=3)
        views.append({"rays_ori": hv["rays_ori"].cuda(), "rays_dir": hv["rays_dir"].cuda(), "camera": camera_of(hv)})
        ups.append(((torch.rand((h, w, 4), generator=g) * 2 - 1).cuda(), ((torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01).cuda()))
    raster = _raster(pkg, particle_radiance_sph_degree=max_deg)
    nsph = 3 * (max_deg + 1) ** 2
    assert sph.shape == (n, nsph)
    ex = gdist.GradientExchange(n, nsph, "cuda", world=1)
    gdist.training_step(raster, ex, views, deg, d12, sph, ups)
    dx = gdist.DeferredExchange(n, nsph, 3, "cuda", world=1)
    gdist.training_step_deferred(raster, dx, views, deg, d12, sph, ups)
    torch.cuda.synchronize()
    assert float(ex.radiance_grad.abs().max()) > 0
    assert rel_l2(dx.density_grad.cpu().numpy(), ex.density_grad.cpu().numpy()) <= 1e-5
    assert rel_l2(dx.radiance_grad.cpu().numpy(), ex.radiance_grad.cpu().numpy()) <= 1e-5
    # a second step overwrites (no stale accumulation)
    gdist.training_step_deferred(raster, dx, views, deg, d12, sph, ups)
    torch.cuda.synchronize()
    assert rel_l2(dx.radiance_grad.cpu().numpy(), ex.radiance_grad.cpu().numpy()) <= 1e-5


def test_two_lane_step_equals_one_lane_step(pkg):
    """dist.training_step_deferred with TWO handles ("lanes": consecutive views on two CUDA streams, each lane summing its own
    accumulators, converted one after the other into the same [N,12] buffer) == the one-lane step; five views, so the lanes hold
    three and two; twice in a row."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    n, w, h, deg, nv = 5000, 96, 64, 3, 5
    scene = syn.make_scene(n, 3, seed=19)
    d12 = syn.pack_density(scene).cuda(); sph = scene["features"].cuda()
    views, ups = [], []
    g = torch.Generator().manual_seed(7)
    for k in range(nv):
        hv = syn.make_view(w, h, k=k, n_views=nv)
        views.append({"rays_ori": hv["rays_ori"].cuda(), "rays_dir": hv["rays_dir"].cuda(), "camera": camera_of(hv)})
        ups.append(((torch.rand((h, w, 4), generator=g) * 2 - 1).cuda(), ((torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01).cuda()))
    one = gdist.DeferredExchange(n, 48, nv, "cuda", world=1, track_visibility=True)
    outs1 = gdist.training_step_deferred(_raster(pkg), one, views, deg, d12, sph, ups)
    two = gdist.DeferredExchange(n, 48, nv, "cuda", world=1, track_visibility=True)
    lanes = [_raster(pkg), _raster(pkg)]
    for _ in range(2):
        outs2 = gdist.training_step_deferred(lanes, two, views, deg, d12, sph, ups)
        torch.cuda.synchronize()
        for (a, b), (c, d) in zip(outs1, outs2):
            assert torch.equal(a, c) and torch.equal(b, d)                      # the images are the same bits
        assert float(one.radiance_grad.abs().max()) > 0
        assert rel_l2(two.density_grad.cpu().numpy(), one.density_grad.cpu().numpy()) <= 1e-5
        assert rel_l2(two.radiance_grad.cpu().numpy(), one.radiance_grad.cpu().numpy()) <= 1e-5
        assert torch.equal(two.visibility, one.visibility)


def test_one_view_per_rank_step_writes_the_density_gradient_directly(pkg):
    """views_per_rank == 1 (what a rank of an 8-GPU step holds): the view's backward converts its accumulators itself
    (gut_backward_colour_density) -- same [N,12] and [N,48] as gut_backward, no pass left pending, and a step in which the
    rank has no view leaves zeros, not the previous step's gradient."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    n, w, h, deg = 5000, 96, 64, 3
    scene = syn.make_scene(n, 3, seed=18)
    d12 = syn.pack_density(scene).cuda(); sph = scene["features"].cuda()
    hv = syn.make_view(w, h, k=1, n_views=3)
    views = [{"rays_ori": hv["rays_ori"].cuda(), "rays_dir": hv["rays_dir"].cuda(), "camera": camera_of(hv)}]
    g = torch.Generator().manual_seed(5)
    ups = [((torch.rand((h, w, 4), generator=g) * 2 - 1).cuda(), ((torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01).cuda())]
    raster = _raster(pkg)
    ex = gdist.GradientExchange(n, 48, "cuda", world=1)
    gdist.training_step(raster, ex, views, deg, d12, sph, ups)
    dx = gdist.DeferredExchange(n, 48, 1, "cuda", world=1)
    assert dx.density_out(0) is dx.density_grad
    for _ in range(2):
        gdist.training_step_deferred(raster, dx, views, deg, d12, sph, ups)
        torch.cuda.synchronize()
        assert float(ex.density_grad.abs().max()) > 0
        assert rel_l2(dx.density_grad.cpu().numpy(), ex.density_grad.cpu().numpy()) <= 1e-5
        assert rel_l2(dx.radiance_grad.cpu().numpy(), ex.radiance_grad.cpu().numpy()) <= 1e-5
    # nothing pending: the reference-shaped backward is accepted right away
    out = raster.trace(0, deg, d12, sph, views[0]["rays_ori"], views[0]["rays_dir"], None, views[0]["camera"])
    raster.trace_bwd(0, deg, d12, sph, views[0]["rays_ori"], views[0]["rays_dir"], None, views[0]["camera"], 0, 1, None, None,
                     out[0], ups[0][0], out[1], ups[0][1])
    # a step without a view for this rank
    gdist.training_step_deferred(raster, dx, [], deg, d12, sph, [])
    torch.cuda.synchronize()
    assert not dx.density_grad.any() and not dx.radiance_grad.any()
    assert gdist.DeferredExchange(n, 48, 2, "cuda", world=1).density_out(0) is None


def test_expand_gather_with_peer_buffer_equals_contiguous_expand(pkg):
    """gut_sh_gradient_expand_gather (one pointer per view) over slots of a gut_peer_buffer_create allocation == gut_sh_gradient_expand
    over a contiguous [V,N,3] buffer; 19 views exercise the chunking (16 per launch) and the accumulate flag.  Single process: the
    cross-rank mapping (gut_peer_buffer_open) is covered by scripts/multi_gpu_peer_check.py on 2+ GPUs."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    native = importlib.import_module("3dgrut_b200.native")
    raster = _raster(pkg)
    n, nv = 3001, 19
    gen = torch.Generator(device="cuda").manual_seed(9)
    d12 = syn.pack_density(syn.make_scene(n, 3, seed=2)).cuda()
    colour = torch.randn((nv, n, 3), device="cuda", generator=gen)
    colour[:, ::4] = 0.0                                   # untouched particles
    centres = torch.randn((nv, 3), device="cuda", generator=gen) * 4
    want = torch.empty((n, 48), device="cuda")
    raster.sh_gradient_expand(3, d12, centres, colour, want)
    ex = gdist.PeerExchange(raster, n, 48, nv, "cuda", rank=0, world=1)
    for i in range(nv):
        ex.slot_tensor(1, i).copy_(colour[i])
    ptrs = [ex.slot_ptr(0, 1, i) for i in range(nv)]
    assert all(p % 16 == 0 for p in ptrs)
    got = torch.full((n, 48), 7.0, device="cuda")
    raster.sh_gradient_expand_gather(3, d12, centres, ptrs, got)
    assert torch.equal(got, want)
    raster.sh_gradient_expand_gather(3, d12, centres, ptrs, got, accumulate=True)
    assert rel_l2(got.cpu().numpy(), 2 * want.cpu().numpy()) <= 1e-6     # (want + chunk 1) + chunk 2: another summation order
    raster.sh_gradient_expand_gather(3, d12, centres[:0], [], got)      # no views: zeros
    assert not got.any()
    with pytest.raises(native.GutError):
        raster.sh_gradient_expand_gather(3, d12, centres[:2], [ptrs[0], 0], got)
    # one-rank step through the peer formulation == the deferred formulation
    w, h = 64, 48
    hv = syn.make_view(w, h)
    views = [{"rays_ori": hv["rays_ori"].cuda(), "rays_dir": hv["rays_dir"].cuda(), "camera": camera_of(hv)}]
    ups = [(torch.rand((h, w, 4), device="cuda", generator=gen) * 2 - 1, torch.rand((h, w, 1), device="cuda", generator=gen) * 0.01)]
    sph = syn.make_scene(n, 3, seed=2)["features"].cuda()
    px = gdist.PeerExchange(raster, n, 48, 1, "cuda", rank=0, world=1)
    dx = gdist.DeferredExchange(n, 48, 1, "cuda", world=1)
    for _ in range(2):   # both buffer sets
        gdist.training_step_peer(raster, px, views, 3, d12, sph, ups)
        gdist.training_step_deferred(raster, dx, views, 3, d12, sph, ups)
        torch.cuda.synchronize()
        assert float(dx.radiance_grad.abs().max()) > 0
        assert rel_l2(px.radiance_grad.cpu().numpy(), dx.radiance_grad.cpu().numpy()) <= 1e-5
        assert rel_l2(px.density_grad.cpu().numpy(), dx.density_grad.cpu().numpy()) <= 1e-5
    px.close(); ex.close()


def _decode_compact_slot(flat, n):
    """A compact colour-gradient slot (include/gut_b200.h, GUT_COLOUR_COMPACT) as the dense [N,3] it stands for, plus its structure."""
    words = flat.view(torch.int32)
    count, groups = int(words[0]), (n + 31) // 32
    table = words[4:4 + 2 * groups].view(groups, 2)
    header = 4 + (2 * groups + 3) // 4 * 4
    bits = ((table[:, :1] >> torch.arange(32, device=flat.device, dtype=torch.int32)) & 1).bool()       # [groups, 32]
    row = table[:, 1:].long() + torch.cumsum(bits, dim=1) - bits.long()
    rows = flat[header:header + 3 * count].view(count, 3)
    dense = torch.zeros((groups * 32, 3), device=flat.device)
    dense[bits.reshape(-1)] = rows[row[bits]]
    return dense[:n].contiguous(), count, bits, row[bits]


def test_compact_colour_slots(pkg):
    """GUT_COLOUR_COMPACT: a view hands out only the rows it touched.  The slot decodes to a dense [N,3] with consistent structure
    (every claimed row used once), the expansion reads it to the bit like the dense one it stands for, stale memory behind the
    rows and in unused slots is never read, and the step's gradients are those of the dense layout."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    native = importlib.import_module("3dgrut_b200.native")
    raster = _raster(pkg)
    n, w, h, nv = 30011, 96, 64, 3
    scene = syn.make_scene(n, 3, seed=4)
    d12, sph = syn.pack_density(scene).cuda(), scene["features"].cuda()
    gen = torch.Generator(device="cuda").manual_seed(11)
    views, ups = [], []
    for k in range(nv):
        hv = syn.make_view(w, h, k=k, n_views=nv)
        views.append({"rays_ori": hv["rays_ori"].cuda(), "rays_dir": hv["rays_dir"].cuda(), "camera": camera_of(hv)})
        ups.append((torch.rand((h, w, 4), device="cuda", generator=gen) * 2 - 1, torch.rand((h, w, 1), device="cuda", generator=gen) * 0.01))
    assert raster.colour_slot_floats(n, 0) == (3 * n + 3) // 4 * 4
    assert raster.colour_slot_floats(n, 1) == 4 + (2 * ((n + 31) // 32) + 3) // 4 * 4 + (3 * n + 3) // 4 * 4
    cx = gdist.PeerExchange(raster, n, 48, nv + 1, "cuda", rank=0, world=1, compact=True)     # one slot more than views: stays empty
    dx = gdist.PeerExchange(raster, n, 48, nv + 1, "cuda", rank=0, world=1, compact=False)
    try:
        for lanes in (1, 2, 1):    # both buffer sets; one step with two handles writing slots
            rasters = [raster, _raster(pkg)] if lanes == 2 else raster
            cx.mine.fill_(float("nan"))                                      # nothing but what this step writes may be read
            gdist.training_step_deferred(rasters, cx, views, 3, d12, sph, ups)
            torch.cuda.synchronize()
            assert raster.colour_layout == 1
            dense = []
            for i in range(nv):
                dn, count, bits, rows_used = _decode_compact_slot(cx.slot(i), n)
                assert count == int(bits.sum()) == cx.moved_rows(i) and 0 < count < n
                assert torch.equal(torch.sort(rows_used)[0], torch.arange(count, device="cuda"))   # every claimed row, once
                assert bool(torch.isfinite(dn).all()) and bool((dn != 0).any(dim=1).sum() == count)
                dense.append(dn)
            assert cx.moved_rows(nv) == 0
            got = cx.radiance_grad.clone()
            assert bool(torch.isfinite(got).all()) and bool(torch.isfinite(cx.density_grad).all())
            # the same expansion from the decoded dense slots (plus an all-zero one for the unused slot): identical bits
            raster.set_colour_layout(0)
            dense.append(torch.zeros((n, 3), device="cuda"))
            want = torch.empty((n, 48), device="cuda")
            raster.sh_gradient_expand_gather(3, d12, cx.centres_all, [t.data_ptr() for t in dense], want)
            assert torch.equal(got, want)
            # and the step in the dense layout (another backward pass: equal up to the order of its atomics)
            gdist.training_step_deferred(rasters, dx, views, 3, d12, sph, ups)
            torch.cuda.synchronize()
            assert raster.colour_layout == 0
            for i in range(nv):
                assert rel_l2(dense[i].cpu().numpy(), dx.slot(i).cpu().numpy()) <= 1e-5
            assert rel_l2(got.cpu().numpy(), dx.radiance_grad.cpu().numpy()) <= 1e-5
            assert rel_l2(cx.density_grad.cpu().numpy(), dx.density_grad.cpu().numpy()) <= 1e-5
        # a handle in compact mode refuses dense arguments
        raster.set_colour_layout(1)
        with pytest.raises(native.GutError):
            raster.sh_gradient_expand(3, d12, cx.centres_all[:nv], torch.zeros((nv, n, 3), device="cuda"), torch.empty((n, 48), device="cuda"))
        v, (rg, dg) = views[0], ups[0]
        rgba, dist_, _hits, _vis = raster.trace(0, 3, d12, sph, v["rays_ori"], v["rays_dir"], None, v["camera"])
        with pytest.raises(native.GutError):
            raster.trace_bwd_colour(3, d12, sph, v["rays_ori"], v["rays_dir"], v["camera"], rgba, rg, dist_, dg, torch.zeros((n, 3), device="cuda"))
        with pytest.raises(native.GutError):
            raster.set_colour_layout(7)
    finally:
        raster.set_colour_layout(0)
        cx.close(); dx.close()


def test_empty_and_error_paths(pkg):
    native = importlib.import_module("3dgrut_b200.native")
    raster = _raster(pkg)
    w, h = 40, 24
    view = syn.make_view(w, h)
    cam = camera_of(view)
    # all particles behind the camera -> M == 0 -> outputs keep their initial values (gutRenderer.cu:323-325)
    scene = syn.make_scene(100, 3, seed=1)
    scene["positions"] = scene["positions"] * 0.01 + torch.tensor([20.0, 0.0, 0.0])
    d12 = syn.pack_density(scene).cuda(); sph = scene["features"].cuda()
    rgba, dist, hits, vis = raster.trace(0, 3, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam)
    torch.cuda.synchronize()
    assert raster.frame_stats()["num_intersections"] == 0
    assert float(rgba.abs().max()) == 0.0 and float(dist.min()) == 1e6 and float(hits.max()) == 0.0
    g12, gsph = raster.trace_bwd(0, 3, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                                 rgba, torch.ones_like(rgba), dist, torch.ones_like(dist))
    assert float(g12.abs().max()) == 0.0 and float(gsph.abs().max()) == 0.0
    # backward without forward on a fresh handle raises (the reference only logs, gutRenderer.cu:413-417)
    fresh = _raster(pkg)
    with pytest.raises(native.GutError):
        fresh.trace_bwd(0, 3, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                        rgba, torch.ones_like(rgba), dist, torch.ones_like(dist))
    with pytest.raises(native.GutError):
        raster.trace(0, 3, d12[:, :11].contiguous(), sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam)
    # a forward in inference mode records no hit masks: the backward refuses instead of walking stale ones
    scene2 = syn.make_scene(500, 3, seed=2)
    d12b = syn.pack_density(scene2).cuda(); sphb = scene2["features"].cuda()
    raster.set_training(False)
    rgba_i, dist_i, _, _ = raster.trace(0, 3, d12b, sphb, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam)
    assert raster.frame_stats()["num_intersections"] > 0
    with pytest.raises(native.GutError):
        raster.trace_bwd(0, 3, d12b, sphb, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                         rgba_i, torch.ones_like(rgba_i), dist_i, torch.ones_like(dist_i))
    raster.set_training(True)
    rgba_t, dist_t, _, _ = raster.trace(0, 3, d12b, sphb, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam)
    assert torch.equal(rgba_t, rgba_i) and torch.equal(dist_t, dist_i)   # recording does not change the image
    raster.trace_bwd(0, 3, d12b, sphb, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                     rgba_t, torch.ones_like(rgba_t), dist_t, torch.ones_like(dist_t))


def test_pack_and_unpack_kernels_equal_cat_and_split():
    """gut_pack_particles / gut_unpack_particle_gradient against the torch.cat / torch.split they replace (tracer.py:176-178)."""
    native = importlib.import_module("3dgrut_b200.native")
    raster = native.SplatRaster(cfg_from_conf(make_conf()))
    for n in (1, 33, 100_003):
        g = torch.Generator(device="cuda").manual_seed(n)
        pos, dns = torch.randn((n, 3), device="cuda", generator=g), torch.rand((n, 1), device="cuda", generator=g)
        rot, scl = torch.randn((n, 4), device="cuda", generator=g), torch.rand((n, 3), device="cuda", generator=g)
        packed = raster.pack_particles(pos, dns, rot, scl)
        assert torch.equal(packed, torch.cat([pos, dns, rot, scl, torch.zeros_like(dns)], dim=1))
        grad = torch.randn((n, 12), device="cuda", generator=g)
        parts = raster.unpack_particle_gradient(grad)
        for a, b in zip(parts, torch.split(grad, [3, 1, 4, 3, 1], dim=1)):
            assert a.is_contiguous() and torch.equal(a, b)
    with pytest.raises(native.GutError):
        raster.pack_particles(torch.zeros((4, 3), device="cuda"), torch.zeros((4, 1), device="cuda"), torch.zeros((4, 3), device="cuda"),
                              torch.zeros((4, 3), device="cuda"))


def test_host_buffer_entry_points_equal_the_device_pointer_calls(pkg):
    """gut_forward_host / gut_backward_host (include/gut_b200.h): rays and upstream gradients in pinned HOST memory, image /
    distance / hit counts back to host memory, copies enqueued on the call's stream -- same bits as gut_forward / gut_backward
    on device tensors (the backward's float atomics land in another order: gradients to 1e-6)."""
    n, w, h, deg = 6000, 150, 70, 3
    scene = syn.make_scene(n, 3, seed=5)
    view = syn.make_view(w, h)
    raster = _raster(pkg)
    (rgba, dist, hits, vis), (d12, sph, cam) = _run_forward(raster, scene, view, deg, stage=False)
    rng = np.random.default_rng(5)
    rg = torch.from_numpy(rng.uniform(-1, 1, (h, w, 4)).astype(np.float32))
    dg = torch.from_numpy(rng.uniform(-0.01, 0.01, (h, w, 1)).astype(np.float32))
    g12, gsph = raster.trace_bwd(0, deg, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, 0, 1, None, None,
                                 rgba, rg.cuda(), dist, dg.cuda())
    torch.cuda.synchronize()
    h_ro, h_rd = view["rays_ori"].pin_memory(), view["rays_dir"].pin_memory()
    h_rgba = torch.empty((h, w, 4)).pin_memory(); h_dist = torch.empty((h, w, 1)).pin_memory(); h_hits = torch.empty((h, w, 1)).pin_memory()
    vis2 = torch.empty((n, 1), device="cuda")
    raster.trace_host(deg, d12, sph, h_ro, h_rd, cam, h_rgba, h_dist, h_hits, vis2)
    g12h = torch.empty((n, 12), device="cuda"); gsphh = torch.empty((n, 48), device="cuda")
    raster.trace_bwd_host(deg, d12, sph, w, h, cam, rg.pin_memory(), dg.pin_memory(), g12h, gsphh)
    torch.cuda.synchronize()
    assert torch.equal(h_rgba, rgba.cpu()) and torch.equal(h_dist, dist.cpu()) and torch.equal(h_hits, hits.cpu())
    assert torch.equal(vis2, vis)
    assert rel_l2(g12h.cpu().numpy(), g12.cpu().numpy()) <= 1e-5 and rel_l2(gsphh.cpu().numpy(), gsph.cpu().numpy()) <= 1e-5
    # a backward without its host forward is refused
    other = _raster(pkg)
    with pytest.raises(Exception):
        other.trace_bwd_host(deg, d12, sph, w, h, cam, rg.pin_memory(), dg.pin_memory(), g12h, gsphh)
