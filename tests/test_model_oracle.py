"""CPU checks of oracle/model_oracle.py.

The activations of the reference model ARE torch functions (threedgrut/utils/misc.py:48-53), so the oracle's forward and
backward are compared with torch.sigmoid / torch.exp / torch.nn.functional.normalize + torch.cat and their autograd:
that pins the activation oracle.  SelectiveAdam (a CUDA kernel, optimizers.cu:47-84) is checked on closed-form first
steps, against a float64 evaluation of the same formulas, and for the "invisible rows untouched" property."""
import numpy as np
import torch

from oracle import model_oracle as mo


def _raw_scene(n, seed):
    g = torch.Generator().manual_seed(seed)
    return {
        "positions": torch.randn(n, 3, generator=g),
        "rotation": torch.rand(n, 4, generator=g) + 0.05,
        "scale": torch.randn(n, 3, generator=g) * 0.7 - 3.0,
        "density": torch.randn(n, 1, generator=g) * 2.0,
        "albedo": torch.rand(n, 3, generator=g),
        "specular": torch.randn(n, 45, generator=g) * 0.05,
    }


def _torch_reference(raw):
    """The literal reference chain: model.py:74-93 then tracer.py:176-178."""
    rot = torch.nn.functional.normalize(raw["rotation"])
    scl = torch.exp(raw["scale"])
    dns = torch.sigmoid(raw["density"])
    packed = torch.cat([raw["positions"], dns, rot, scl, torch.zeros_like(dns)], dim=1)
    feats = torch.cat((raw["albedo"], raw["specular"]), dim=1)
    return packed, feats


def test_activate_pack_matches_torch_functions():
    raw = _raw_scene(4096, 1)
    packed_t, feats_t = _torch_reference(raw)
    packed_o = mo.activate_pack(*(raw[k].numpy() for k in ("positions", "rotation", "scale", "density")))
    feats_o = mo.pack_features(raw["albedo"].numpy(), raw["specular"].numpy())
    np.testing.assert_allclose(packed_o, packed_t.numpy(), rtol=2e-6, atol=1e-7)   # fp32, libm vs torch vectorised exp
    np.testing.assert_array_equal(feats_o, feats_t.numpy())
    assert (packed_o[:, 11] == 0).all()


def test_activate_pack_backward_matches_autograd():
    raw = {k: v.clone().requires_grad_(True) for k, v in _raw_scene(2048, 2).items()}
    packed_t, feats_t = _torch_reference(raw)
    g = torch.Generator().manual_seed(3)
    g12 = torch.randn(packed_t.shape, generator=g)
    g48 = torch.randn(feats_t.shape, generator=g)
    (packed_t * g12).sum().backward(retain_graph=True)
    (feats_t * g48).sum().backward()
    pos_g, rot_g, scl_g, dns_g = mo.activate_pack_backward(raw["rotation"].detach().numpy(), packed_t.detach().numpy(), g12.numpy())
    alb_g, spc_g = mo.unpack_feature_gradient(g48.numpy())
    for name, got in (("positions", pos_g), ("rotation", rot_g), ("scale", scl_g), ("density", dns_g), ("albedo", alb_g), ("specular", spc_g)):
        want = raw[name].grad.numpy()
        denom = max(np.linalg.norm(want), 1e-30)
        assert np.linalg.norm(got - want) / denom < 2e-6, name


def test_normalize_below_eps_follows_torch_clamp():
    rot = torch.tensor([[0.0, 0.0, 0.0, 0.0], [1e-20, 0.0, 0.0, 0.0]], requires_grad=True)
    y = torch.nn.functional.normalize(rot)
    gq = torch.tensor([[1.0, 2.0, 3.0, 4.0], [0.5, -1.0, 0.25, 2.0]])
    (y * gq).sum().backward()
    packed = np.zeros((2, 12), np.float32)
    g12 = np.zeros((2, 12), np.float32)
    g12[:, 4:8] = gq.numpy()
    _, rot_g, _, _ = mo.activate_pack_backward(rot.detach().numpy(), packed, g12, mo.ACT_NONE, mo.ACT_NONE)
    np.testing.assert_allclose(rot_g, rot.grad.numpy(), rtol=1e-6)


def test_selective_adam_first_step_closed_form():
    # from a zero state: m = (1-b1) g, v = (1-b2) g^2  ->  step = -lr (1-b1) g / (sqrt(1-b2) |g| + eps)
    g = np.array([[0.5, -2.0, 1e-3]], np.float32)
    p0 = np.array([[1.0, 2.0, 3.0]], np.float32)
    z = np.zeros_like(p0)
    lr, b1, b2, eps = 0.01, 0.9, 0.999, 1e-15
    p, m, v = mo.selective_adam_update(p0, g, z, z, None, lr, b1, b2, eps)
    # the kernel forms 1 - b in fp32 (`1.0f - b1`): 1 - 0.999f is 1.0000467e-3, not 1e-3
    c1, c2 = float(np.float32(1) - np.float32(b1)), float(np.float32(1) - np.float32(b2))
    want = p0.astype(np.float64) - lr * c1 * g / (np.sqrt(c2) * np.abs(g) + eps)
    np.testing.assert_allclose(p, want, rtol=1e-6)
    np.testing.assert_allclose(m, c1 * g, rtol=1e-6)
    np.testing.assert_allclose(v, c2 * g.astype(np.float64) ** 2, rtol=1e-6)


def test_selective_adam_many_steps_against_float64_and_visibility():
    rng = np.random.default_rng(5)
    n, w = 257, 7
    p = rng.normal(size=(n, w)).astype(np.float32)
    m = np.zeros_like(p)
    v = np.zeros_like(p)
    p64, m64, v64 = p.astype(np.float64), m.astype(np.float64), v.astype(np.float64)
    lr, b1, b2, eps = 2e-3, 0.9, 0.999, 1e-15
    frozen = np.zeros(n, bool)
    frozen[::5] = True      # never visible
    for step in range(20):
        g = rng.normal(size=(n, w)).astype(np.float32)
        vis = (rng.random(n) < 0.7) & ~frozen
        p, m, v = mo.selective_adam_update(p, g, m, v, vis.astype(np.float32)[:, None], lr, b1, b2, eps)
        mk = vis[:, None]
        c1, c2 = float(np.float32(1) - np.float32(b1)), float(np.float32(1) - np.float32(b2))
        m_n = float(np.float32(b1)) * m64 + c1 * g
        v_n = float(np.float32(b2)) * v64 + c2 * g.astype(np.float64) ** 2
        p_n = p64 - lr * m_n / (np.sqrt(v_n) + eps)
        p64, m64, v64 = np.where(mk, p_n, p64), np.where(mk, m_n, m64), np.where(mk, v_n, v64)
    np.testing.assert_allclose(p, p64, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(m, m64, rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(v, v64, rtol=1e-5, atol=1e-9)
    assert (m[frozen] == 0).all() and (v[frozen] == 0).all()


def test_update_gradient_buffer_matches_literal_torch():
    g = torch.Generator().manual_seed(9)
    n = 1000
    pos = torch.randn(n, 3, generator=g)
    grad = torch.randn(n, 3, generator=g)
    grad[::3] = 0.0
    sensor = torch.tensor([0.3, -4.0, 1.5])
    accum = torch.rand(n, 1, generator=g)
    denom = torch.randint(0, 5, (n, 1), generator=g).float()
    a_o, d_o = mo.update_gradient_buffer(pos.numpy(), grad.numpy(), sensor.numpy(), accum.numpy(), denom.numpy())
    # strategy/gs.py:104-115, literally
    mask = (grad != 0).max(dim=1)[0]
    dist = (pos[mask] - sensor).norm(dim=1, keepdim=True)
    accum[mask] += torch.norm(grad[mask] * dist, dim=-1, keepdim=True) / 2
    denom[mask] += 1
    np.testing.assert_allclose(a_o, accum.numpy(), rtol=1e-6)
    np.testing.assert_array_equal(d_o, denom.numpy())
