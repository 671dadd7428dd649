"""GPU parity of the model-side kernels through the C ABI, against oracle/model_oracle.py and
against the torch functions / literal reference code they replace."""
import importlib

import numpy as np
import pytest
import torch

from oracle import model_oracle as mo
from tests.helpers import SyntheticGaussians, cfg_from_conf, make_batch, make_conf, rel_l2, syn

pytestmark = pytest.mark.gpu

TOL_ACT = 2e-6     # fp32 activations: CUDA expf vs libm / torch (relative)


@pytest.fixture(scope="module")
def native():
    return importlib.import_module("3dgrut_b200.native")


@pytest.fixture(scope="module")
def raster(native):
    return native.SplatRaster(cfg_from_conf(make_conf()))


def _raw(n, seed, device="cuda"):
    g = torch.Generator().manual_seed(seed)
    raw = {
        "positions": torch.randn(n, 3, generator=g),
        "rotation": torch.rand(n, 4, generator=g) + 0.05,
        "scale": torch.randn(n, 3, generator=g) * 0.7 - 3.0,
        "density": torch.randn(n, 1, generator=g) * 2.0,
        "albedo": torch.rand(n, 3, generator=g),
        "specular": torch.randn(n, 45, generator=g) * 0.05,
    }
    return {k: v.to(device) for k, v in raw.items()}


@pytest.mark.parametrize("n", [1, 31, 100_003])
def test_activate_pack_forward_and_backward(raster, n):
    raw = _raw(n, n)
    cpu = {k: v.cpu().numpy() for k, v in raw.items()}
    packed = raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"])
    want = mo.activate_pack(cpu["positions"], cpu["rotation"], cpu["scale"], cpu["density"])
    np.testing.assert_allclose(packed.cpu().numpy(), want, rtol=TOL_ACT, atol=1e-7)
    assert torch.equal(packed[:, 0:3], raw["positions"]) and (packed[:, 11] == 0).all()
    # the torch chain of the reference (model.py:77-93, tracer.py:176-178)
    dns = torch.sigmoid(raw["density"])
    ref = torch.cat([raw["positions"], dns, torch.nn.functional.normalize(raw["rotation"]), torch.exp(raw["scale"]), torch.zeros_like(dns)], dim=1)
    np.testing.assert_allclose(packed.cpu().numpy(), ref.cpu().numpy(), rtol=TOL_ACT, atol=1e-7)

    g12 = torch.randn((n, 12), device="cuda", generator=torch.Generator(device="cuda").manual_seed(n + 1))
    got = raster.activate_pack_backward(raw["rotation"], packed, g12)
    exp = mo.activate_pack_backward(cpu["rotation"], packed.cpu().numpy(), g12.cpu().numpy())
    for name, a, b in zip(("positions", "rotation", "scale", "density"), got, exp):
        assert a.is_contiguous()
        assert rel_l2(a.cpu().numpy(), b) <= TOL_ACT, name
    # and against autograd of the torch chain
    leaves = {k: raw[k].clone().requires_grad_(True) for k in ("positions", "rotation", "scale", "density")}
    d2 = torch.sigmoid(leaves["density"])
    chain = torch.cat([leaves["positions"], d2, torch.nn.functional.normalize(leaves["rotation"]), torch.exp(leaves["scale"]),
                       torch.zeros_like(d2)], dim=1)
    (chain * g12).sum().backward()
    for name, a in zip(("positions", "rotation", "scale", "density"), got):
        assert rel_l2(a.cpu().numpy(), leaves[name].grad.cpu().numpy()) <= 1e-5, name


def test_activation_kinds_and_degenerate_quaternion(raster, native):
    raw = _raw(257, 4)
    raw["rotation"][0] = 0.0          # |q| below the normalize eps: torch clamps the norm
    packed = raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"], mo.ACT_NONE, mo.ACT_NONE)
    assert torch.equal(packed[:, 8:11], raw["scale"]) and torch.equal(packed[:, 3:4], raw["density"])
    assert (packed[0, 4:8] == 0).all()
    g12 = torch.ones((257, 12), device="cuda")
    _, rot_g, scl_g, dns_g = raster.activate_pack_backward(raw["rotation"], packed, g12, mo.ACT_NONE, mo.ACT_NONE)
    assert torch.equal(scl_g, g12[:, 8:11]) and torch.equal(dns_g, g12[:, 3:4])
    assert torch.allclose(rot_g[0], torch.full((4,), 1e12, device="cuda"), rtol=1e-6)
    with pytest.raises(native.GutError):
        raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"], 7, mo.ACT_SIGMOID)
    with pytest.raises(native.GutError):
        raster.activate_pack(raw["positions"], raw["rotation"][:, :3], raw["scale"], raw["density"])


@pytest.mark.parametrize("n", [1, 5, 65_537])
def test_feature_cat_and_split_are_exact(raster, n):
    raw = _raw(n, 11 + n)
    feats = raster.pack_features(raw["albedo"], raw["specular"])
    assert torch.equal(feats, torch.cat((raw["albedo"], raw["specular"]), dim=1))
    g48 = torch.randn((n, 48), device="cuda", generator=torch.Generator(device="cuda").manual_seed(n))
    alb_g, spc_g = raster.unpack_feature_gradient(g48)
    assert alb_g.is_contiguous() and spc_g.is_contiguous()
    assert torch.equal(alb_g, g48[:, :3]) and torch.equal(spc_g, g48[:, 3:])


class RawGaussians:
    """The raw-parameter attributes of MixtureOfGaussians (model.py:163-167, 480-485) plus its activated getters."""

    def __init__(self, raw, n_active_features=3):
        for k in ("positions", "rotation", "scale", "density"):
            setattr(self, k, raw[k].clone().requires_grad_(True))
        self.features_albedo = raw["albedo"].clone().requires_grad_(True)
        self.features_specular = raw["specular"].clone().requires_grad_(True)
        self.scale_activation, self.density_activation = torch.exp, torch.sigmoid
        self.rotation_activation = torch.nn.functional.normalize
        self.n_active_features = n_active_features
        self.num_gaussians = self.positions.shape[0]

    def get_rotation(self):
        return self.rotation_activation(self.rotation)

    def get_scale(self):
        return self.scale_activation(self.scale)

    def get_density(self):
        return self.density_activation(self.density)

    def get_features(self):
        return torch.cat((self.features_albedo, self.features_specular), dim=1)

    def background(self, T_to_world, rays_d, rgb, opacity, train):
        return rgb, opacity

    def parameters(self):
        return [self.positions, self.rotation, self.scale, self.density, self.features_albedo, self.features_specular]


def _raw_from_scene(scene):
    """Invert the activations of a synthetic (post-activation) scene: what a MixtureOfGaussians would hold."""
    raw = {
        "positions": scene["positions"],
        "rotation": scene["rotation"] * 1.7,                       # any positive multiple normalises to the same quaternion
        "scale": torch.log(scene["scale"]),
        "density": torch.log(scene["density"] / (1 - scene["density"])),
        "albedo": scene["features"][:, :3].contiguous(),
        "specular": scene["features"][:, 3:].contiguous(),
    }
    return {k: v.cuda() for k, v in raw.items()}


def test_tracer_fused_raw_path_equals_torch_activation_path():
    """Tracer.render with raw parameters (one activation kernel each way) against the reference-shaped path where torch
    applies the activations and autograd differentiates them."""
    pkg = importlib.import_module("3dgrut_b200")
    n, w, h = 3000, 64, 48
    scene = syn.make_scene(n, 3, seed=21, kind="dense")
    view = syn.make_view(w, h)
    raw = _raw_from_scene(scene)
    outs, grads = [], []
    for fused in (True, False):
        conf = make_conf()
        conf["render"]["fused_activations"] = fused
        tracer = pkg.Tracer(conf)
        g = RawGaussians(raw)
        out = tracer.render(g, make_batch(view, "cuda"), train=True)
        gen = torch.Generator(device="cuda").manual_seed(1)
        loss = (out["pred_rgb"] * torch.rand(out["pred_rgb"].shape, device="cuda", generator=gen)).sum() + \
            (out["pred_opacity"] * torch.rand(out["pred_opacity"].shape, device="cuda", generator=gen)).sum() + \
            0.01 * (out["pred_dist"] * torch.rand(out["pred_dist"].shape, device="cuda", generator=gen)).sum()
        loss.backward()
        outs.append(out)
        grads.append([p.grad.clone() for p in g.parameters()])
    assert (outs[0]["pred_rgb"] - outs[1]["pred_rgb"]).abs().max() <= 1e-5
    assert (outs[0]["pred_opacity"] - outs[1]["pred_opacity"]).abs().max() <= 1e-5
    for name, a, b in zip(("positions", "rotation", "scale", "density", "albedo", "specular"), grads[0], grads[1]):
        assert a.shape == b.shape
        assert rel_l2(a.cpu().numpy(), b.cpu().numpy()) <= 1e-4, name


def _adam_groups(n, seed, widths=(3, 4, 3, 1, 3, 45)):
    g = torch.Generator(device="cuda").manual_seed(seed)
    groups = []
    for i, w in enumerate(widths):
        p = torch.randn((n, w), device="cuda", generator=g)
        m = torch.randn((n, w), device="cuda", generator=g) * 0.01
        v = torch.rand((n, w), device="cuda", generator=g) * 1e-4
        groups.append([p, torch.randn((n, w), device="cuda", generator=g), m, v, 10.0 ** (-2 - 0.5 * i)])
    return groups


@pytest.mark.parametrize("n", [1, 7, 10_001])
@pytest.mark.parametrize("masked", [False, True])
def test_selective_adam_step_is_bit_exact(raster, n, masked):
    """One launch over six parameter groups (N x {3,4,3,1,3,45}: aligned and unaligned rows) against the oracle's fp32
    restatement of selective_adam_update_kernel -- integer-exact comparison of the float bit patterns."""
    groups = _adam_groups(n, 100 + n)
    vis = None
    if masked:
        vis = (torch.rand((n, 1), device="cuda", generator=torch.Generator(device="cuda").manual_seed(n)) < 0.6).float()
    b1, b2, eps = 0.9, 0.999, 1e-15
    want = [mo.selective_adam_update(p.cpu().numpy(), g.cpu().numpy(), m.cpu().numpy(), v.cpu().numpy(),
                                     None if vis is None else vis.cpu().numpy(), lr, b1, b2, eps) for p, g, m, v, lr in groups]
    grads_before = [g.clone() for _, g, _, _, _ in groups]
    raster.selective_adam_step([tuple(x) for x in groups], vis, b1, b2, eps)
    for (p, g, m, v, _), (pw, mw, vw), g0 in zip(groups, want, grads_before):
        assert np.array_equal(p.cpu().numpy().view(np.uint32), pw.view(np.uint32))
        assert np.array_equal(m.cpu().numpy().view(np.uint32), mw.view(np.uint32))
        assert np.array_equal(v.cpu().numpy().view(np.uint32), vw.view(np.uint32))
        assert torch.equal(g, g0)   # gradients untouched without zero_grad
    raster.selective_adam_step([tuple(x) for x in groups], vis, b1, b2, eps, zero_grad=True)
    for _, g, _, _, _ in groups:
        assert (g == 0).all()


@pytest.mark.parametrize("n,masked", [(10_001, True), (100_000, False), (5, True)])
def test_selective_adam_bit_exact_against_the_reference_kernel(raster, n, masked):
    """gut_selective_adam_step (one launch, six groups) against the reference's own compiled selective_adam_update_kernel
    (threedgrut/optimizers/optimizers.cu, built unchanged by oracle/ref_build into oracle/_ref/lib_optimizers_cc.so), called the
    way SelectiveAdam.step calls it (optimizers/__init__.py:119-133), three steps in a row: every bit of param, exp_avg and
    exp_avg_sq.  The numpy/C oracle is compared with both."""
    from tests import refshim

    if not refshim.optimizers_available():
        pytest.skip("oracle/_ref/lib_optimizers_cc.so not built")
    ref = refshim.optimizers_module()
    ours = _adam_groups(n, 300 + n)
    theirs = [[t.clone() if torch.is_tensor(t) else t for t in g] for g in ours]
    orc = [[t.cpu().numpy().copy() if torch.is_tensor(t) else t for t in g] for g in ours]
    gen = torch.Generator(device="cuda").manual_seed(n)
    b1, b2, eps = 0.9, 0.999, 1e-15
    for step in range(3):
        vis = (torch.rand((n, 1), device="cuda", generator=gen) < 0.6).float() if masked else torch.ones((n, 1), device="cuda")
        for go, gt, gc in zip(ours, theirs, orc):
            grad = torch.randn(go[0].shape, device="cuda", generator=gen) * 10.0 ** (-step)
            go[1].copy_(grad); gt[1].copy_(grad); gc[1] = grad.cpu().numpy()
        for p, g, m, v, lr in theirs:
            ref.selective_adam_update(p, g, m, v, vis.bool().squeeze(1), lr, b1, b2, eps)
        raster.selective_adam_step([tuple(x) for x in ours], vis, b1, b2, eps)
        for gc in orc:
            gc[0], gc[2], gc[3] = mo.selective_adam_update(gc[0], gc[1], gc[2], gc[3], vis.cpu().numpy(), gc[4], b1, b2, eps)
        torch.cuda.synchronize()
        for i, (go, gt, gc) in enumerate(zip(ours, theirs, orc)):
            for k, what in ((0, "param"), (2, "exp_avg"), (3, "exp_avg_sq")):
                a, b = go[k].cpu().numpy().view(np.uint32), gt[k].cpu().numpy().view(np.uint32)
                assert np.array_equal(a, b), (step, i, what, int((a != b).sum()))
                assert np.array_equal(gc[k].view(np.uint32), b), ("oracle", step, i, what)


def test_selective_adam_class_mirrors_reference_api():
    """3dgrut_b200.SelectiveAdam: the reference's constructor / step(visibility) / state keys (optimizers/__init__.py:45-133)."""
    pkg = importlib.import_module("3dgrut_b200")
    n = 5000
    gen = torch.Generator(device="cuda").manual_seed(0)
    params = {name: torch.nn.Parameter(torch.randn((n, w), device="cuda", generator=gen)) for name, w in (("positions", 3), ("density", 1), ("features_specular", 45))}
    opt = pkg.SelectiveAdam([{"params": [p], "name": k, "lr": lr} for (k, p), lr in zip(params.items(), (1e-3, 5e-2, 2e-4))], lr=1e-3, eps=1e-15)
    state = {k: (p.detach().cpu().numpy().copy(), np.zeros((n, p.shape[1]), np.float32), np.zeros((n, p.shape[1]), np.float32)) for k, p in params.items()}
    for step in range(3):
        vis = (torch.rand((n, 1), device="cuda", generator=gen) < 0.5).float()
        for p in params.values():
            p.grad = torch.randn(p.shape, device="cuda", generator=gen)
        for (k, p), lr in zip(params.items(), (1e-3, 5e-2, 2e-4)):
            state[k] = mo.selective_adam_update(state[k][0], p.grad.cpu().numpy(), state[k][1], state[k][2], vis.cpu().numpy(), lr, 0.9, 0.999, 1e-15)
        opt.step(vis)
    for k, p in params.items():
        assert np.array_equal(p.detach().cpu().numpy(), state[k][0]), k
        st = opt.state[p]
        assert set(st) == {"step", "exp_avg", "exp_avg_sq"}
        assert np.array_equal(st["exp_avg"].cpu().numpy(), state[k][1]) and np.array_equal(st["exp_avg_sq"].cpu().numpy(), state[k][2])


def test_densify_stats_equal_literal_reference_code(raster):
    pkg = importlib.import_module("3dgrut_b200")
    n = 20_011
    gen = torch.Generator(device="cuda").manual_seed(2)
    pos = torch.nn.Parameter(torch.randn((n, 3), device="cuda", generator=gen))
    pos.grad = torch.randn((n, 3), device="cuda", generator=gen) * 1e-3
    pos.grad[::3] = 0.0
    sensor = torch.tensor([0.3, -4.0, 1.5], device="cuda")
    accum = torch.rand((n, 1), device="cuda", generator=gen)
    denom = torch.randint(0, 5, (n, 1), device="cuda", generator=gen).float()
    a_o, d_o = mo.update_gradient_buffer(pos.detach().cpu().numpy(), pos.grad.cpu().numpy(), sensor.cpu().numpy(), accum.cpu().numpy(),
                                         denom.cpu().numpy())
    # strategy/gs.py:104-115, literally, on clones
    a_t, d_t = accum.clone(), denom.clone()
    mask = (pos.grad != 0).max(dim=1)[0]
    dist = (pos.detach()[mask] - sensor).norm(dim=1, keepdim=True)
    a_t[mask] += torch.norm(pos.grad[mask] * dist, dim=-1, keepdim=True) / 2
    d_t[mask] += 1
    pkg.update_gradient_buffer(pos, accum, denom, sensor)
    np.testing.assert_allclose(accum.cpu().numpy(), a_o, rtol=1e-6)
    np.testing.assert_allclose(accum.cpu().numpy(), a_t.cpu().numpy(), rtol=1e-6)
    assert np.array_equal(denom.cpu().numpy(), d_o) and torch.equal(denom, d_t)


def test_ply_scene_renders_identically(tmp_path):
    """write_ply -> read_ply -> Tracer.render (fused raw-parameter path): the file format loses nothing."""
    pkg = importlib.import_module("3dgrut_b200")
    n, w, h = 2000, 48, 32
    raw = _raw_from_scene(syn.make_scene(n, 3, seed=5, kind="dense"))
    names = {"positions": "positions", "rotation": "rotation", "scale": "scale", "density": "density", "albedo": "features_albedo",
             "specular": "features_specular"}
    path = str(tmp_path / "scene.ply")
    pkg.write_ply(path, {names[k]: v.cpu().numpy() for k, v in raw.items()})
    back = pkg.read_ply(path)
    raw2 = {k: torch.from_numpy(back[names[k]]).cuda() for k in raw}
    view = syn.make_view(w, h)
    tracer = pkg.Tracer(make_conf())
    with torch.no_grad():
        a = tracer.render(RawGaussians(raw), make_batch(view, "cuda"))
        b = tracer.render(RawGaussians(raw2), make_batch(view, "cuda"))
    assert torch.equal(a["pred_rgb"], b["pred_rgb"]) and torch.equal(a["pred_dist"], b["pred_dist"])
    assert a["pred_opacity"].max() > 0.1


def test_dist_optimizer_step_chains_activation_adjoint_and_adam(raster):
    """dist.optimizer_step: summed [N,12] / [N,48] gradients -> raw-parameter gradients -> one SelectiveAdam launch masked by
    the step's visibility union, against the oracle's restatement of the same chain."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    n = 4099
    raw_t = _raw(n, 77)
    raw = {"positions": raw_t["positions"], "rotation": raw_t["rotation"], "scale": raw_t["scale"], "density": raw_t["density"],
           "features_albedo": raw_t["albedo"], "features_specular": raw_t["specular"]}
    before = {k: v.cpu().numpy().copy() for k, v in raw.items()}
    packed = raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"])
    ex = gdist.DeferredExchange(n, 48, 1, "cuda", world=1, track_visibility=True)
    gen = torch.Generator(device="cuda").manual_seed(3)
    ex.density_grad.copy_(torch.randn((n, 12), device="cuda", generator=gen))
    ex.radiance_grad.copy_(torch.randn((n, 48), device="cuda", generator=gen))
    ex.visibility.copy_((torch.rand((n, 1), device="cuda", generator=gen) < 0.7).float())
    state = {k: (torch.zeros_like(v), torch.zeros_like(v)) for k, v in raw.items()}
    lrs = {k: 10.0 ** (-2 - i) for i, k in enumerate(gdist.PARAMETER_GROUPS)}
    grads = gdist.optimizer_step(raster, ex, raw, packed, state, lrs)
    g = mo.activate_pack_backward(before["rotation"], packed.cpu().numpy(), ex.density_grad.cpu().numpy())
    ga, gs = mo.unpack_feature_gradient(ex.radiance_grad.cpu().numpy())
    want_g = dict(zip(gdist.PARAMETER_GROUPS, (g[0], g[1], g[2], g[3], ga, gs)))
    vis = ex.visibility.cpu().numpy()
    for k in gdist.PARAMETER_GROUPS:
        assert rel_l2(grads[k].cpu().numpy(), want_g[k]) <= TOL_ACT, k
        # Adam from the kernel's own gradient (the update itself is bit-exact, the activation adjoint is not)
        p, m, v = mo.selective_adam_update(before[k], grads[k].cpu().numpy(), np.zeros_like(before[k]), np.zeros_like(before[k]), vis,
                                           lrs[k], 0.9, 0.999, 1e-15)
        assert np.array_equal(raw[k].cpu().numpy(), p), k
        assert np.array_equal(state[k][0].cpu().numpy(), m) and np.array_equal(state[k][1].cpu().numpy(), v), k
        assert np.array_equal(raw[k].cpu().numpy()[vis[:, 0] == 0], before[k][vis[:, 0] == 0]), k


def test_optimizer_step_refuses_an_exchange_without_visibility(raster):
    """An exchange that did not track visibility holds an all-zero mask: SelectiveAdam would update nothing, silently."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    ex = gdist.DeferredExchange(16, 48, 1, "cuda", world=1)
    with pytest.raises(ValueError, match="track_visibility"):
        gdist.optimizer_step(raster, ex, {}, None, {}, {})


def test_peer_step_then_optimizer_step_changes_the_parameters(raster):
    """training_step_peer (the bench default on more than one GPU; world 1 here) records the visibility union when asked,
    so the optimiser step behind it really moves the visible particles."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    from tests.helpers import camera_of

    n, w, h = 5000, 96, 64
    raw_t = _raw(n, 21)
    raw_t["positions"] = (torch.rand((n, 3), device="cuda") * 3 - 1.5)
    raw_t["scale"] = torch.full((n, 3), -2.5, device="cuda")
    raw = {"positions": raw_t["positions"], "rotation": raw_t["rotation"], "scale": raw_t["scale"], "density": raw_t["density"],
           "features_albedo": raw_t["albedo"], "features_specular": raw_t["specular"]}
    before = {k: v.clone() for k, v in raw.items()}
    packed = raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"])
    sph = raster.pack_features(raw["features_albedo"], raw["features_specular"])
    view = syn.make_view(w, h)
    views = [{"rays_ori": view["rays_ori"].cuda(), "rays_dir": view["rays_dir"].cuda(), "camera": camera_of(view)}]
    up = [(torch.rand((h, w, 4), device="cuda"), torch.rand((h, w, 1), device="cuda") * 0.01)]
    try:
        ex = gdist.PeerExchange(raster, n, 48, 1, "cuda", 0, world=1, track_visibility=True)
    except RuntimeError as e:
        pytest.skip(f"CUDA IPC unavailable: {e}")
    try:
        gdist.training_step_peer(raster, ex, views, 3, packed, sph, up, rank=0)
        vis = ex.visibility.clone()
        assert 0 < vis.sum() < n
        state = {k: (torch.zeros_like(v), torch.zeros_like(v)) for k, v in raw.items()}
        lrs = {k: 1e-2 for k in gdist.PARAMETER_GROUPS}
        gdist.optimizer_step(raster, ex, raw, packed, state, lrs)
        torch.cuda.synchronize()
        moved = (raw["positions"] != before["positions"]).any(dim=1)
        assert moved.any() and not moved[vis[:, 0] == 0].any()
    finally:
        ex.close()


@pytest.mark.parametrize("max_deg", [0, 1, 2])
def test_feature_packing_for_every_sh_degree(max_deg):
    """gut_pack_features / gut_unpack_feature_gradient with particle_radiance_sph_degree 0..2 (1, 4, 9 coefficients):
    torch.cat((albedo, specular), dim=1) and its split, bit for bit."""
    native = importlib.import_module("3dgrut_b200.native")
    r = native.SplatRaster(cfg_from_conf(make_conf(particle_radiance_sph_degree=max_deg)))
    n, ns = 10_007, 3 * ((max_deg + 1) ** 2 - 1)
    g = torch.Generator(device="cuda").manual_seed(max_deg)
    alb = torch.randn((n, 3), device="cuda", generator=g); spc = torch.randn((n, ns), device="cuda", generator=g)
    packed = r.pack_features(alb, spc)
    assert torch.equal(packed, torch.cat((alb, spc), dim=1))
    a2, s2 = r.unpack_feature_gradient(packed)
    assert torch.equal(a2, alb) and torch.equal(s2, spc)


@pytest.mark.parametrize("kind", ["deferred", "peer", "peer-compact"])
def test_optimizer_epilogue_leaves_the_bits_of_the_unfused_chain(raster, kind):
    """dist.optimizer_epilogue (gut_density_adam + gut_sh_expand_adam: activation adjoints, SH expansion, SelectiveAdam and the
    densification statistics as the epilogue of the step's gradient kernels, no gradient tensors in between) against the
    unfused chain optimizer_step + densify_stats on the same three-view step, two steps in a row: parameters, both moments of all
    six groups and the statistics BIT for bit; invisible rows untouched."""
    gdist = importlib.import_module("3dgrut_b200.dist")
    from tests.helpers import camera_of

    n, w, h, nv, deg = 6000, 112, 80, 3, 3
    raw_t = _raw(n, 23)
    raw_t["positions"] = (torch.rand((n, 3), device="cuda") * 3 - 1.5)
    raw_t["scale"] = torch.full((n, 3), -2.5, device="cuda") + 0.3 * torch.rand((n, 3), device="cuda")
    names = {"positions": "positions", "rotation": "rotation", "scale": "scale", "density": "density", "features_albedo": "albedo",
             "features_specular": "specular"}
    views, ups = [], []
    g = torch.Generator().manual_seed(6)
    for k in range(nv):
        hv = syn.make_view(w, h, k=k, n_views=nv)
        views.append({"rays_ori": hv["rays_ori"].cuda(), "rays_dir": hv["rays_dir"].cuda(), "camera": camera_of(hv)})
        ups.append(((torch.rand((h, w, 4), generator=g) * 2 - 1).cuda(), ((torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01).cuda()))
    sensor = [0.3, -0.2, 4.0]
    lrs = {"positions": 1.6e-4, "rotation": 1e-3, "scale": 5e-3, "density": 5e-2, "features_albedo": 2.5e-3, "features_specular": 1.25e-4}

    raw = {k: raw_t[v].clone() for k, v in names.items()}
    state = {k: (torch.zeros_like(v), torch.zeros_like(v)) for k, v in raw.items()}
    accum, denom = torch.zeros((n, 1), device="cuda"), torch.zeros((n, 1), device="cuda")
    if kind.startswith("peer"):
        try:
            ex = gdist.PeerExchange(raster, n, 48, nv, "cuda", 0, world=1, track_visibility=True, compact=kind == "peer-compact")
        except RuntimeError as e:
            pytest.skip(f"CUDA IPC unavailable: {e}")
    else:
        ex = gdist.DeferredExchange(n, 48, nv, "cuda", world=1, track_visibility=True)
    try:
        ex.use(raster)        # the slot layout the handle writes and reads (dense, or the compact one)
        for _step in range(2):
            packed = raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"])
            sph = raster.pack_features(raw["features_albedo"], raw["features_specular"])
            ex.begin_step(0)
            for i, (v, (rg, dg)) in enumerate(zip(views, ups)):
                rgba, dist_, hits, vis = raster.trace(0, deg, packed, sph, v["rays_ori"], v["rays_dir"], None, v["camera"])
                c = raster.trace_bwd_colour(deg, packed, sph, v["rays_ori"], v["rays_dir"], v["camera"], rgba, rg, dist_, dg, ex.slot(i))
                ex.add_view(c, vis)
            # the step's gradients, once (two backward passes never agree to the bit: both arms read THESE buffers)
            _, none = ex.finish_step(raster, deg, packed, expand=False)
            assert none is None
            torch.cuda.synchronize()
            vis_union = ex.visibility.clone()
            assert 0 < float(vis_union.sum()) < n
            # arm B: the epilogue, on copies
            raw_b = {k: v.clone() for k, v in raw.items()}
            state_b = {k: (m.clone(), v.clone()) for k, (m, v) in state.items()}
            accum_b, denom_b = accum.clone(), denom.clone()
            gdist.optimizer_epilogue(raster, ex, raw_b, packed, state_b, lrs, deg, sensor_position=sensor, grad_norm_accum=accum_b,
                                     grad_norm_denom=denom_b)
            # arm A: expansion into [N,48], activation adjoints, feature split, SelectiveAdam, statistics (from the positions
            # BEFORE the update: post_backward runs before optimizer.step in the reference)
            raster.sh_gradient_expand_gather(deg, packed, ex.centres_all, ex.colour_ptrs(), ex.radiance_grad)
            pos_before = raw["positions"].clone()
            before = {k: v.clone() for k, v in raw.items()}
            grads = gdist.optimizer_step(raster, ex, raw, packed, state, lrs)
            raster.densify_stats(pos_before, grads["positions"], sensor, accum, denom)
            torch.cuda.synchronize()
            inv = vis_union[:, 0] == 0
            for k in names:
                assert torch.equal(raw[k], raw_b[k]), (k, _step)
                assert torch.equal(state[k][0], state_b[k][0]) and torch.equal(state[k][1], state_b[k][1]), (k, _step)
                assert not torch.equal(raw[k], before[k]), k                      # the step moved something
                assert torch.equal(raw_b[k][inv], before[k][inv]), k              # invisible rows untouched
            assert torch.equal(accum, accum_b) and torch.equal(denom, denom_b) and float(denom.sum()) > 0
    finally:
        raster.set_colour_layout(0)
        if kind.startswith("peer"):
            ex.close()


def test_two_handles_on_two_devices_in_one_process():
    """Kernel attributes (dynamic shared memory opt-in of the per-tile sort) are per device: a second handle on another GPU
    of the same process must render too."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    native = importlib.import_module("3dgrut_b200.native")
    from tests.helpers import camera_of

    scene = syn.make_scene(20000, 3, seed=3)
    view = syn.make_view(160, 96)
    outs = []
    for dev in ("cuda:0", "cuda:1"):
        r = native.SplatRaster(cfg_from_conf(make_conf()), device=dev)
        with torch.cuda.device(dev):
            out = r.trace(0, 3, syn.pack_density(scene).to(dev), scene["features"].to(dev), view["rays_ori"].to(dev),
                          view["rays_dir"].to(dev), None, camera_of(view))
            torch.cuda.synchronize()
        outs.append(out[0].cpu())
    assert torch.equal(outs[0], outs[1]) and outs[0].abs().max() > 0


def _loss(out, seed):
    gen = torch.Generator(device="cuda").manual_seed(seed)
    return (out["pred_rgb"] * torch.rand(out["pred_rgb"].shape, device="cuda", generator=gen)).sum() + \
        (out["pred_opacity"] * torch.rand(out["pred_opacity"].shape, device="cuda", generator=gen)).sum() + \
        0.01 * (out["pred_dist"] * torch.rand(out["pred_dist"].shape, device="cuda", generator=gen)).sum()


@pytest.mark.parametrize("fused,nonleaf,lanes", [(True, False, 2), (False, False, 2), (False, True, 2), (True, False, 1), (False, False, 3)])
def test_multi_view_step_equals_per_view_render(fused, nonleaf, lanes):
    """Tracer.begin_step / step.render / step.finish (deferred gradient conversions, one per step) against the reference-shaped
    loop Tracer.render + backward per view: same images, same summed gradients on the parameters -- raw parameters through
    the fused activation kernels, activated leaf tensors, and activated NON-leaf tensors (the model's own torch activations:
    gradients must continue through autograd to the raw parameters)."""
    pkg = importlib.import_module("3dgrut_b200")
    n, w, h, nv = 6000, 112, 80, 3
    scene = syn.make_scene(n, 3, seed=44)
    views = [syn.make_view(w, h, k=k, n_views=nv) for k in range(nv)]
    conf = make_conf()
    conf["render"]["fused_activations"] = fused
    tracer = pkg.Tracer(conf)

    def model():
        if fused or nonleaf:
            return RawGaussians(_raw_from_scene({k: v.cuda() for k, v in scene.items()}))
        return SyntheticGaussians(scene, "cuda", n_active_features=3, requires_grad=True)

    def params(g):
        return g.parameters() if hasattr(g, "parameters") else [g.positions, g._rotation, g._scale, g._density, g._features]

    ga = model()
    imgs_a = []
    for k, v in enumerate(views):
        out = tracer.render(ga, make_batch(v, "cuda"), train=True)
        _loss(out, k).backward()
        imgs_a.append(out["pred_rgb"].detach().clone())
    gb = model()
    step = tracer.begin_step(gb, views_per_rank=nv, lanes=lanes)     # views dealt to `lanes` handles on streams of their own
    assert len(step.lanes) == lanes
    imgs_b = []
    for k, v in enumerate(views):
        out = step.render(make_batch(v, "cuda"))
        if k % 2:
            _loss(out, k).backward()          # plain autograd call ...
        else:
            step.backward(_loss(out, k))      # ... and the one that keeps the caller's stream free of the view's backward
        imgs_b.append(out["pred_rgb"].detach().clone())
    vis = step.finish()
    torch.cuda.synchronize()
    for a, b in zip(imgs_a, imgs_b):
        assert torch.equal(a, b)
    for pa, pb in zip(params(ga), params(gb)):
        assert pb.grad is not None and pa.grad.shape == pb.grad.shape
        # two summation orders of the same fp32 terms.  The rotation gradient is the difference of two large halves (origin
        # fold, gut_project.cu) that are summed over the views separately in one order and per view in the other: 1e-4
        assert rel_l2(pb.grad.cpu().numpy(), pa.grad.cpu().numpy()) <= 1e-4
    assert vis is not None and 0 < float(vis.sum()) <= n
    # protocol errors: a second render before the backward, render after finish
    step2 = tracer.begin_step(model(), views_per_rank=nv)
    step2.render(make_batch(views[0], "cuda"))
    with pytest.raises(RuntimeError, match="backward"):
        step2.render(make_batch(views[1], "cuda"))
    step2.finish()
    with pytest.raises(RuntimeError):
        step2.render(make_batch(views[1], "cuda"))
