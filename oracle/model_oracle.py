"""CPU oracle (TEST INFRASTRUCTURE ONLY -- never imported by 3dgrut_b200/) for the model-side passes of SURVEY 8f:
activation + packing at the Tracer boundary, SelectiveAdam and the densification statistics.  numpy, float32.

Pinning: the activations ARE torch functions in the reference (threedgrut/utils/misc.py:48-53: torch.sigmoid, torch.exp,
torch.nn.functional.normalize), so tests/test_model_oracle.py checks this restatement -- forward and backward -- against
those torch functions and their autograd on CPU: pinned.  SelectiveAdam (threedgrut/optimizers/optimizers.cu:47-84) is
restated with the rounding of the reference's compiled kernel and pinned on that kernel: oracle/ref_build builds
optimizers.cu / optimizers.cpp unchanged (oracle/_ref/lib_optimizers_cc.so), tests/test_gpu_model.py compares bit for bit on
the GPU box, and tests/golden/reference_adam.npz holds vectors that kernel produced for the CPU suite.
"""
import numpy as np

F32 = np.float32
NORMALIZE_EPS = F32(1e-12)   # torch.nn.functional.normalize default eps

ACT_NONE, ACT_EXP, ACT_SIGMOID = 0, 1, 2


def _activate(kind, x):
    x = x.astype(F32)
    if kind == ACT_EXP:
        return np.exp(x).astype(F32)            # torch.exp (misc.py:50)
    if kind == ACT_SIGMOID:
        return (F32(1) / (F32(1) + np.exp(-x))).astype(F32)   # torch.sigmoid (misc.py:49)
    return x


def _activate_grad(kind, y, g):
    if kind == ACT_EXP:
        return (g * y).astype(F32)
    if kind == ACT_SIGMOID:
        return (g * (y * (F32(1) - y))).astype(F32)
    return g.astype(F32)


def activate_pack(positions, rotation_raw, scale_raw, density_raw, scale_act=ACT_EXP, density_act=ACT_SIGMOID):
    """model.py:77-93 (get_scale / get_rotation / get_density) followed by tracer.py:176-178
    (particle_density = cat([pos, density, rotation, scale, 0]))."""
    n = positions.shape[0]
    q = rotation_raw.astype(F32)
    norm = np.sqrt((q * q).sum(axis=1, dtype=F32)).astype(F32)
    norm = np.maximum(norm, NORMALIZE_EPS)[:, None]
    out = np.zeros((n, 12), dtype=F32)
    out[:, 0:3] = positions
    out[:, 3:4] = _activate(density_act, density_raw)
    out[:, 4:8] = q / norm
    out[:, 8:11] = _activate(scale_act, scale_raw)
    return out


def activate_pack_backward(rotation_raw, packed, grad12, scale_act=ACT_EXP, density_act=ACT_SIGMOID):
    """What autograd derives for the chain above: gradients of the raw (positions, rotation, scale, density)."""
    g = grad12.astype(F32)
    pos_g = g[:, 0:3].copy()
    dns_g = _activate_grad(density_act, packed[:, 3:4].astype(F32), g[:, 3:4])
    scl_g = _activate_grad(scale_act, packed[:, 8:11].astype(F32), g[:, 8:11])
    q = rotation_raw.astype(F32)
    gq = g[:, 4:8]
    nn = np.sqrt((q * q).sum(axis=1, dtype=F32)).astype(F32)[:, None]
    safe = np.where(nn > NORMALIZE_EPS, nn, F32(1))
    y = q / safe
    dt = (y * gq).sum(axis=1, dtype=F32)[:, None]
    rot_g = np.where(nn > NORMALIZE_EPS, (gq - y * dt) / safe, gq / NORMALIZE_EPS).astype(F32)
    return pos_g, rot_g, scl_g, dns_g


def pack_features(albedo, specular):
    """model.py:74-75: torch.cat((features_albedo, features_specular), dim=1)."""
    return np.concatenate([albedo, specular], axis=1).astype(F32)


def unpack_feature_gradient(grad48):
    return grad48[:, :3].copy(), grad48[:, 3:].copy()


def selective_adam_update(param, grad, exp_avg, exp_avg_sq, visibility, lr, b1, b2, eps):
    """selective_adam_update_kernel (optimizers.cu:47-84), one scalar per thread, all float32:
        if (!visibility[g_idx]) return;
        m = b1 * m + (1 - b1) * g;  v = b2 * v + (1 - b2) * g * g;  param += -lr * m / (sqrt(v) + eps)
    No bias correction.  Returns new (param, exp_avg, exp_avg_sq); rows with visibility == 0 are unchanged.
    The arithmetic runs in oracle/gut_oracle.c (gut_oracle_selective_adam): the reference's compiled kernel contracts
    the two moment updates into FMAs, and a single-rounding fmaf is not expressible in numpy."""
    import ctypes as C

    from oracle import gut_oracle as O

    p, g, m, v = (np.ascontiguousarray(a, dtype=F32).copy() for a in (param, grad, exp_avg, exp_avg_sq))
    n = p.shape[0]
    width = int(p.size // max(n, 1))
    vis_p = C.c_void_p(0)
    if visibility is not None:
        vis = np.ascontiguousarray((np.asarray(visibility).reshape(n) != 0).astype(np.uint8))   # visibility.bool(), __init__.py:126
        vis_p = vis.ctypes.data_as(C.c_void_p)
    fp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    O.lib().gut_oracle_selective_adam(C.c_uint64(n), C.c_uint32(width), fp(p), fp(g), fp(m), fp(v), vis_p,
                                      C.c_float(lr), C.c_float(b1), C.c_float(b2), C.c_float(eps))
    return p, m, v


def update_gradient_buffer(positions, positions_grad, sensor_position, accum, denom):
    """GSStrategy.update_gradient_buffer (strategy/gs.py:104-115):
        mask = (grad != 0).max(dim=1)
        dist = |positions[mask] - sensor_position|
        accum[mask] += |grad[mask] * dist| / 2 ;  denom[mask] += 1"""
    g = positions_grad.astype(F32)
    mask = (g != 0).any(axis=1)
    d = positions.astype(F32) - np.asarray(sensor_position, dtype=F32)[None, :]
    dist = np.sqrt((d * d).sum(axis=1, dtype=F32)).astype(F32)[:, None]
    a = g * dist
    nrm = (np.sqrt((a * a).sum(axis=1, dtype=F32)).astype(F32) / F32(2)).reshape(accum.shape)
    m = mask.reshape(accum.shape)
    return np.where(m, accum + nrm, accum).astype(F32), np.where(m, denom + F32(1), denom).astype(F32)
