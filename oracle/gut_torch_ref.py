"""Second, independent restatement of the 3DGUT path: vectorised PyTorch on the CPU, autograd for gradients.

TEST INFRASTRUCTURE ONLY (tests/ and tests/golden/make_golden.py import it; the product never does).
PARITY UNPINNED: the reference ships no vectors for this path; this file exists so that two restatements
written differently (scalar C with a hand-derived backward, oracle/gut_oracle.c; tensor code with autograd,
here) have to agree before either is trusted.

Written from the reference sources, paths relative to /root/reference/threedgut_tracer/:
  projection   include/3dgut/kernels/cuda/renderers/gutProjector.cuh:32-322
  cameras      include/3dgut/kernels/cuda/sensors/cameraProjections.cuh:22-185
  SH           include/3dgut/kernels/cuda/models/gaussianParticles.cuh:57-96
  keys/ranges  gutProjector.cuh:324-388, src/gutRenderer.cu:46-94
  rays         include/3dgut/kernels/cuda/common/rayPayload.cuh:75-129
  hit/blend    include/3dgut/kernels/slang/models/gaussianParticles.slang:96-254,
               include/3dgut/kernels/cuda/renderers/gutKBufferRenderer.cuh:153-292
Where the reference's backward is NOT the derivative of its forward the forward below is
written so that autograd reproduces the reference: the alpha clamp is a straight-through min(), the SH view
direction is detached, the ray is not differentiated, and the hit-distance Jacobian is diagonal-only
(gaussianParticles.cuh:553-567).
"""
import math

import numpy as np
import torch

TILE = 16
SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
SH_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
         1.445305721320277, -0.5900435899266435)


# ----------------------------------------------------------------------------------------------------------
# small algebra
# ----------------------------------------------------------------------------------------------------------
def quat_wxyz_to_matrix(q):
    """tcnn::to_mat3 (thirdparty/tiny-cuda-nn/include/tiny-cuda-nn/vec.h:1172-1183); q[...,4] = (w,x,y,z), NOT renormalised.
    Returns R[...,3,3] with R[..., :, i] = image of axis i."""
    w, x, y, z = q.unbind(-1)
    r00 = 1 - 2 * (y * y + z * z); r01 = 2 * (x * y - w * z); r02 = 2 * (x * z + w * y)
    r10 = 2 * (x * y + w * z); r11 = 1 - 2 * (x * x + z * z); r12 = 2 * (y * z - w * x)
    r20 = 2 * (x * z - w * y); r21 = 2 * (y * z + w * x); r22 = 1 - 2 * (x * x + y * y)
    return torch.stack([torch.stack([r00, r01, r02], -1), torch.stack([r10, r11, r12], -1), torch.stack([r20, r21, r22], -1)], -2)


def _slerp(q0, q1, t):
    """tcnn::slerp (vec.h:1133-1153) on (w,x,y,z) tensors [...,4]; t [...]."""
    c = (q0 * q1).sum(-1)
    q1 = torch.where((c < 0).unsqueeze(-1), -q1, q1)
    c = c.abs()
    lin = q0 * (1 - t).unsqueeze(-1) + q1 * t.unsqueeze(-1)
    ang = torch.acos(c.clamp(max=1.0))
    sd = torch.sin(ang).clamp_min(1e-30)
    sph = (torch.sin((1 - t) * ang) / sd).unsqueeze(-1) * q0 + (torch.sin(t * ang) / sd).unsqueeze(-1) * q1
    eps = torch.finfo(torch.float32).eps
    return torch.where((c > 1 - eps).unsqueeze(-1), lin, sph)


def _pose(p, dtype):
    p = torch.as_tensor(np.asarray(p, np.float64), dtype=dtype)
    return p[:3], torch.stack([p[6], p[3], p[4], p[5]])  # t, q(wxyz)


def frame_constants(cam, dtype=torch.float64):
    """Mid-shutter pose and its inverse (src/gutRenderer.cu:259-267, include/3dgut/sensors/sensors.h:44-73)."""
    t0, q0 = _pose(list(cam.start_pose), dtype)
    t1, q1 = _pose(list(cam.end_pose), dtype)
    half = torch.tensor(0.5, dtype=dtype)
    qm = _slerp(q0, q1, half)
    tm = t0 * 0.5 + t1 * 0.5
    R = quat_wxyz_to_matrix(qm)
    return {"R_w2c": R, "t_w2c": tm, "R_c2w": R.T, "t_c2w": -(R.T @ tm), "t0": t0, "q0": q0, "t1": t1, "q1": q1}


# ----------------------------------------------------------------------------------------------------------
# cameras
# ----------------------------------------------------------------------------------------------------------
def _within(cam_w, cam_h, tol, u):
    return (u[..., 0] > -tol * cam_w) & (u[..., 0] < cam_w + tol * cam_w) & (u[..., 1] > -tol * cam_h) & (u[..., 1] < cam_h + tol * cam_h)


def _project_point(cam, W, H, tol, p):
    """cameraProjections.cuh:57-144.  p [...,3] sensor space -> (u [...,2], valid [...])."""
    dtype = p.dtype
    f = torch.tensor(list(cam.focal_length), dtype=dtype)
    c = torch.tensor(list(cam.principal_point), dtype=dtype)
    k = [float(v) for v in cam.radial_coeffs]
    if cam.model == 0:  # OpenCV pinhole
        pt = [float(v) for v in cam.tangential_coeffs]
        s = [float(v) for v in cam.thin_prism_coeffs]
        z = p[..., 2]
        front = z > 0
        zs = torch.where(front, z, torch.ones_like(z))
        uv = p[..., :2] / zs.unsqueeze(-1)
        r2 = (uv * uv).sum(-1)
        icd = (1 + r2 * (k[0] + r2 * (k[1] + r2 * k[2]))) / (1 + r2 * (k[3] + r2 * (k[4] + r2 * k[5])))
        a1 = 2 * uv[..., 0] * uv[..., 1]
        a2 = r2 + 2 * uv[..., 0] ** 2
        a3 = r2 + 2 * uv[..., 1] ** 2
        dx = pt[0] * a1 + pt[1] * a2 + r2 * (s[0] + r2 * s[1])
        dy = pt[0] * a3 + pt[1] * a1 + r2 * (s[2] + r2 * s[3])
        ok_rad = (icd > 0.8) & (icd < 1.2)
        good = (icd.unsqueeze(-1) * uv + torch.stack([dx, dy], -1)) * f + c
        clampf = math.hypot(W, H) / torch.sqrt(r2.clamp_min(1e-300))
        bad = clampf.unsqueeze(-1) * uv + c
        u = torch.where(ok_rad.unsqueeze(-1), good, bad)
        u = torch.where(front.unsqueeze(-1), u, torch.zeros_like(u))
        return u, front & ok_rad & _within(W, H, tol, u)
    # OpenCV fisheye
    eps = torch.finfo(torch.float32).eps
    rho = torch.sqrt(p[..., 0] ** 2 + p[..., 1] ** 2).clamp_min(eps)
    theta_full = torch.atan2(rho, p[..., 2])
    theta = theta_full.clamp(max=float(cam.max_angle))
    t2 = theta * theta
    poly = k[0] + t2 * (k[1] + t2 * (k[2] + t2 * k[3]))
    delta = theta * (1 + t2 * poly) / rho
    u = f * p[..., :2] * delta.unsqueeze(-1) + c
    return u, (theta < float(cam.max_angle)) & _within(W, H, tol, u)


def _shutter_time(cam, W, H, u):
    s = cam.shutter
    if s == 0: return torch.floor(u[..., 1]) / (H - 1.0)
    if s == 1: return torch.floor(u[..., 0]) / (W - 1.0)
    if s == 2: return (H - torch.ceil(u[..., 1])) / (H - 1.0)
    if s == 3: return (W - torch.ceil(u[..., 0])) / (W - 1.0)
    return torch.full_like(u[..., 0], 0.5)


def project_with_shutter(cam, fc, W, H, n_iter, tol, x):
    """projectPointWithShutter (cameraProjections.cuh:146-185); x [...,3] world points."""
    def to_sensor(q, t, pts):
        return (quat_wxyz_to_matrix(q) @ pts.unsqueeze(-1)).squeeze(-1) + t
    u, ok = _project_point(cam, W, H, tol, to_sensor(fc["q0"], fc["t0"], x))
    if cam.shutter == 4:
        return u, ok
    u1, ok1 = _project_point(cam, W, H, tol, to_sensor(fc["q1"], fc["t1"], x))
    u = torch.where(ok.unsqueeze(-1), u, u1)
    both_bad = ~(ok | ok1)
    ok = ok | ok1
    for _ in range(n_iter):
        a = _shutter_time(cam, W, H, u)
        q = _slerp(fc["q0"].expand(*a.shape, 4), fc["q1"].expand(*a.shape, 4), a)
        t = fc["t0"] * (1 - a).unsqueeze(-1) + fc["t1"] * a.unsqueeze(-1)
        un, okn = _project_point(cam, W, H, tol, to_sensor(q, t, x))
        u = torch.where(both_bad.unsqueeze(-1), u, un)
        ok = torch.where(both_bad, torch.zeros_like(okn), okn)
    return u, ok


# ----------------------------------------------------------------------------------------------------------
# K1
# ----------------------------------------------------------------------------------------------------------
def sh_basis(deg, d):
    x, y, z = d.unbind(-1)
    b = [torch.full_like(x, SH_C0)]
    if deg > 0:
        b += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2 * zz - xx - yy), SH_C2[3] * xz, SH_C2[4] * (xx - yy)]
        if deg > 2:
            b += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
                  SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy), SH_C3[5] * z * (xx - yy),
                  SH_C3[6] * x * (xx - 3 * yy)]
    return torch.stack(b, -1)


def features_from_sh(cfg, fc, pos, sph, sph_degree):
    """radianceFromSpH + 0.5, unclamped, direction = normalize(pos - camera centre), no gradient through the direction."""
    n = pos.shape[0]
    ncoef = (cfg.sph_max_degree + 1) ** 2
    d = (pos.detach() - fc["t_c2w"])
    d = d / d.norm(dim=-1, keepdim=True)
    b = sh_basis(sph_degree, d)
    co = sph.reshape(n, ncoef, 3)[:, : b.shape[-1], :]
    return (b.unsqueeze(-1) * co).sum(1) + 0.5


def tile_min_power(tx, ty, conic, mean):
    """tileMinParticlePowerResponse (gutProjector.cuh:49-78); tx,ty arrays of tile coords; scalars conic (a,b,c), mean (x,y)."""
    a, b, c = conic
    tminx, tminy = TILE * tx, TILE * ty
    tmaxx, tmaxy = tminx + TILE, tminy + TILE
    lax, lay = (tminx > mean[0]), (tminy > mean[1])
    nrx, nry = lax | (mean[0] > tmaxx), lay | (mean[1] > tmaxy)
    px = np.where(lax, tminx, tmaxx); py = np.where(lay, tminy, tmaxy)
    dx = np.copysign(float(TILE), tminx - mean[0]); dy = np.copysign(float(TILE), tminy - mean[1])
    fx, fy = mean[0] - px, mean[1] - py
    with np.errstate(all="ignore"):
        ux = np.clip(np.nan_to_num((dx * a * fx + dx * b * fy) / (TILE * TILE * a)), 0, 1) * nry
        uy = np.clip(np.nan_to_num((dy * b * fx + dy * c * fy) / (TILE * TILE * c)), 0, 1) * nrx
    mx = mean[0] - (px + ux * dx); my = mean[1] - (py + uy * dy)
    val = 0.5 * (a * mx * mx + c * my * my) + b * mx * my
    return np.where(nrx | nry, val, 0.0)


def tile_lists_from_projection(cfg, W, H, proj_pos, conic_opacity, extent, valid=None):
    """Tile bbox + per-tile culling of K1/K2 (gutProjector.cuh:32-78, 282-289, 324-388) for every particle, fp64.
    Returns (tiles_count u32[N], list of u32 tile-id arrays in emission order = row major)."""
    n = proj_pos.shape[0]
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    mean_n = np.asarray(proj_pos, np.float64); ext_n = np.asarray(extent, np.float64); co = np.asarray(conic_opacity, np.float64)
    if valid is None:
        valid = ext_n[:, 0] > 1e-6
    tiles = np.zeros(n, np.uint32)
    tile_lists = [None] * n
    for i in np.nonzero(valid)[0]:
        pw = math.log(co[i, 3] / cfg.min_alpha) if co[i, 3] > 0 else -np.inf
        x0 = int(np.clip(np.floor((mean_n[i, 0] - 0.5 - ext_n[i, 0]) / TILE), 0, gx)); x1 = int(np.clip(np.ceil((mean_n[i, 0] - 0.5 + ext_n[i, 0]) / TILE), 0, gx))
        y0 = int(np.clip(np.floor((mean_n[i, 1] - 0.5 - ext_n[i, 1]) / TILE), 0, gy)); y1 = int(np.clip(np.ceil((mean_n[i, 1] - 0.5 + ext_n[i, 1]) / TILE), 0, gy))
        if x1 <= x0 or y1 <= y0:
            continue
        ty, tx = np.meshgrid(np.arange(y0, y1), np.arange(x0, x1), indexing="ij")
        keep = np.ones_like(tx, bool)
        if cfg.tile_based_culling:
            keep = tile_min_power(tx.astype(np.float64), ty.astype(np.float64), co[i, :3], mean_n[i]) < pw
        tiles[i] = int(keep.sum())
        tile_lists[i] = (ty[keep] * gx + tx[keep]).astype(np.uint32)  # row-major = emission order of K2
    return tiles, tile_lists


def project(cfg, cam, density12, sph, W, H, sph_degree, dtype=torch.float64):
    """K1 (gutProjector.cuh:118-322) vectorised over particles.  Returns numpy arrays named like oracle.project()."""
    d12 = torch.as_tensor(np.asarray(density12), dtype=dtype)
    sph_t = torch.as_tensor(np.asarray(sph), dtype=dtype)
    fc = frame_constants(cam, dtype)
    pos, dns, quat, scl = d12[:, 0:3], d12[:, 3], d12[:, 4:8], d12[:, 8:11]
    n = pos.shape[0]
    z = pos @ fc["R_w2c"][2] + fc["t_w2c"][2]
    pre = (dns >= cfg.min_alpha) & ~(z < 0.2)
    alpha, beta, kappa = cfg.ut_alpha, cfg.ut_beta, cfg.ut_kappa
    lam = alpha * alpha * (3 + kappa) - 3
    delta = math.sqrt(alpha * alpha * (3 + kappa))
    R = quat_wxyz_to_matrix(quat)
    offs = delta * scl.unsqueeze(1) * R  # [N,3(row),3(axis)] column i scaled by s_i
    pts = torch.cat([pos.unsqueeze(1), pos.unsqueeze(1) + offs.transpose(1, 2), pos.unsqueeze(1) - offs.transpose(1, 2)], 1)  # [N,7,3]
    u, ok = project_with_shutter(cam, fc, W, H, cfg.n_rolling_shutter_iterations, cfg.ut_in_image_margin_factor, pts)
    w_m = torch.full((7,), 1.0 / (2 * (3 + lam)), dtype=dtype); w_m[0] = lam / (3 + lam)
    w_c = w_m.clone(); w_c[0] = lam / (3 + lam) + (1 - alpha * alpha + beta)
    mean = (w_m.view(1, 7, 1) * u).sum(1)
    d = u - mean.unsqueeze(1)
    cxx = (w_c * d[..., 0] * d[..., 0]).sum(1); cxy = (w_c * d[..., 0] * d[..., 1]).sum(1); cyy = (w_c * d[..., 1] * d[..., 1]).sum(1)
    valid = pre & (ok.sum(1) > 0)
    # conic / extent (gutProjector.cuh:81-116)
    dxx, dyy = cxx + 0.3, cyy + 0.3
    ddet = dxx * dyy - cxy * cxy
    det = cxx * cyy - cxy * cxy
    safe = torch.where(ddet == 0, torch.ones_like(ddet), ddet)
    conic = torch.stack([dyy / safe, -cxy / safe, dxx / safe], -1)
    opac = dns * torch.sqrt(torch.clamp(det / safe, min=0.000025))
    valid = valid & (ddet != 0) & ~(opac < cfg.min_alpha)
    pw = torch.log(torch.where(valid, opac, torch.ones_like(opac)) / cfg.min_alpha)
    ef = torch.clamp(torch.sqrt(2 * pw.clamp_min(0)), max=3.33) if cfg.tight_opacity_bounding else torch.full_like(pw, 3.33)
    mid = 0.5 * (dxx + dyy)
    lmax = mid + torch.sqrt(torch.clamp(mid * mid - ddet, min=0.01))
    rad = ef * torch.sqrt(lmax)
    if cfg.rect_bounding:
        ext = torch.stack([torch.minimum(ef * torch.sqrt(dxx), rad), torch.minimum(ef * torch.sqrt(dyy), rad)], -1)
    else:
        ext = torch.stack([rad, rad], -1)
    valid = valid & (rad > 0)
    feats = features_from_sh(cfg, fc, pos, sph_t, sph_degree)

    mean_n, ext_n, conic_n, valid_n = mean.numpy(), ext.numpy(), conic.numpy(), valid.numpy()
    co_n = np.concatenate([conic_n, opac.numpy()[:, None]], 1)
    tiles, tile_lists = tile_lists_from_projection(cfg, W, H, mean_n, co_n, ext_n, valid_n)
    vis = tiles > 0
    zero = ~vis
    out = {
        "tiles_count": tiles, "proj_pos": mean_n.copy(), "conic_opacity": co_n,
        "extent": ext_n.copy(), "depth": z.numpy().copy(), "features": feats.detach().numpy().copy(), "tile_lists": tile_lists,
    }
    for k in ("proj_pos", "conic_opacity", "extent", "depth", "features"):
        out[k][zero] = 0
    return out


def bin_and_sort(W, H, tiles_count, tile_lists, depth32):
    """K2 + sort + K3 (gutProjector.cuh:324-388, gutRenderer.cu:46-94,356-365) with numpy: keys (tile<<32 | depth bits),
    stable sort, tile ranges.  depth32: float32 depths as stored by K1."""
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    T = gx * gy
    bits = np.asarray(depth32, np.float32).view(np.uint32).astype(np.uint64)
    ks, vs = [], []
    for i in np.nonzero(tiles_count)[0]:
        tl = tile_lists[i].astype(np.uint64)
        ks.append((tl << np.uint64(32)) | bits[i]); vs.append(np.full(tl.shape[0], i, np.uint32))
    keys = np.concatenate(ks) if ks else np.zeros(0, np.uint64)
    vals = np.concatenate(vs) if vs else np.zeros(0, np.uint32)
    order = np.argsort(keys, kind="stable")
    skeys, svals = keys[order], vals[order]
    ranges = np.zeros((T, 2), np.uint32)
    st = (skeys >> np.uint64(32)).astype(np.int64)
    for t in np.unique(st):
        idx = np.nonzero(st == t)[0]
        ranges[t] = (idx[0], idx[-1] + 1)
    return {"keys": keys, "values": vals, "sorted_keys": skeys, "sorted_values": svals, "tile_ranges": ranges, "M": int(keys.shape[0])}


# ----------------------------------------------------------------------------------------------------------
# K4 (+ K5/K6 through autograd)
# ----------------------------------------------------------------------------------------------------------
def _response(deg, g2):
    if deg == 8: return torch.exp(-0.000685871056241 * g2 ** 4)
    if deg == 5: return torch.exp(-0.0185185185185 * g2 * g2 * torch.sqrt(g2))
    if deg == 4: return torch.exp(-0.0555555555556 * g2 * g2)
    if deg == 3: return torch.exp(-0.166666666667 * g2 * torch.sqrt(g2))
    if deg == 1: return torch.exp(-1.5 * torch.sqrt(g2))
    if deg == 0: return torch.clamp(1 - 0.329630334487 * torch.sqrt(g2), min=0)
    return torch.exp(-0.5 * g2)


def composite(cfg, cam, pos, dns, quat, scl, feats, ray_ori, ray_dir, W, H, tile_ranges, sorted_values, dtype=torch.float64):
    """K4 for k_buffer_size == 0.  pos/dns/quat/scl/feats are torch tensors (may require grad); returns rgba [H,W,4],
    dist [H,W], hits [H,W] (torch).  One dense [pixels, list] evaluation per tile."""
    fc = frame_constants(cam, dtype)
    ro = torch.as_tensor(np.asarray(ray_ori), dtype=dtype).reshape(H, W, 3)
    rd = torch.as_tensor(np.asarray(ray_dir), dtype=dtype).reshape(H, W, 3)
    o_w = ro @ fc["R_c2w"].T + fc["t_c2w"]
    d_w = rd @ fc["R_c2w"].T
    # slab test against the +-1e6 box (include/3dgut/utils/bounding_box.h:79-128), tmin clamped at 0
    with torch.no_grad():
        inv = 1.0 / d_w
        ta, tb = (-1e6 - o_w) * inv, (1e6 - o_w) * inv
        tmin = torch.minimum(ta, tb).amax(-1).clamp_min(0)
        tmax = torch.maximum(ta, tb).amin(-1)
        valid_ray = tmax > tmin
    gx = (W + TILE - 1) // TILE
    rgba = torch.zeros((H, W, 4), dtype=dtype)
    dist = torch.full((H, W), 1e6, dtype=dtype)
    hits = torch.zeros((H, W), dtype=dtype)
    Rt = quat_wxyz_to_matrix(quat).transpose(-1, -2)  # rows of R^T
    rgba_parts, dist_parts = [], []
    for t in range(tile_ranges.shape[0]):
        b0, b1 = int(tile_ranges[t, 0]), int(tile_ranges[t, 1])
        ty, tx = divmod(t, gx)
        ys, xs = slice(ty * TILE, min((ty + 1) * TILE, H)), slice(tx * TILE, min((tx + 1) * TILE, W))
        o = o_w[ys, xs].reshape(-1, 3); d = d_w[ys, xs].reshape(-1, 3)
        vr = valid_ray[ys, xs].reshape(-1)
        npx = o.shape[0]
        if b1 <= b0:
            rgba_parts.append((ys, xs, torch.zeros((npx, 4), dtype=dtype) ))
            dist_parts.append((ys, xs, torch.where(vr, torch.zeros(npx, dtype=dtype), torch.full((npx,), 1e6, dtype=dtype)), torch.zeros(npx, dtype=dtype)))
            continue
        idx = torch.as_tensor(sorted_values[b0:b1].astype(np.int64))
        p, s, rt, dn, ft = pos[idx], scl[idx], Rt[idx], dns[idx], feats[idx]
        giscl = 1.0 / s
        gro = giscl.unsqueeze(0) * torch.einsum("lij,plj->pli", rt, o.unsqueeze(1) - p.unsqueeze(0))   # [P,L,3]
        grdu = giscl.unsqueeze(0) * torch.einsum("lij,pj->pli", rt, d)
        grd = grdu / grdu.norm(dim=-1, keepdim=True)
        cr = torch.linalg.cross(grd, gro, dim=-1)
        g2 = (cr * cr).sum(-1)
        resp = _response(cfg.kernel_degree, g2)
        a_raw = resp * dn.unsqueeze(0)
        alpha = a_raw + (torch.clamp(a_raw, max=cfg.max_alpha) - a_raw).detach()  # clamp not masked in the reference backward
        # hit distance |s * grd * dot(grd, -gro)|.  The reference's hand-written backward keeps only the DIAGONAL of
        # d grds / d grd and d grds / d gro (gaussianParticles.cuh:553-567: grdRayHitGrd, groRayHitGrd are element-wise),
        # so the cross terms of the dot product are detached here to make autograd reproduce it.
        prod = grd * gro
        dp = -(prod + (prod.sum(-1, keepdim=True) - prod).detach())
        grds = s.unsqueeze(0) * grd * dp
        tt = torch.sqrt((grds * grds).sum(-1))
        with torch.no_grad():
            acc = (resp > cfg.min_kernel_response) & (alpha > cfg.min_alpha) & (tt > tmin[ys, xs].reshape(-1, 1)) & (tt < tmax[ys, xs].reshape(-1, 1)) & vr.unsqueeze(1)
            one_minus = torch.where(acc, 1 - alpha, torch.ones_like(alpha))
            T_after = torch.cumprod(one_minus, dim=1)
            killed = T_after < cfg.min_transmittance
            # a hit is processed iff the ray was still alive before it: no earlier processed hit dropped T below the threshold
            dead_before = torch.cat([torch.zeros((npx, 1), dtype=torch.bool), killed[:, :-1]], 1)
            acc = acc & ~dead_before
        accf = acc.to(dtype)
        om = 1 - alpha * accf
        T_before = torch.cat([torch.ones((npx, 1), dtype=dtype), torch.cumprod(om, dim=1)[:, :-1]], 1)
        w = alpha * T_before * accf
        C = (w.unsqueeze(-1) * torch.clamp(ft, min=0).unsqueeze(0)).sum(1)
        # Distance: the reference's K = 0 backward takes d dist / d alpha_i = T_i * t_i only -- its residual-distance term is
        # identically 0 because of swapped arguments at the call site of processHitBwd (shRadiativeGaussianParticles.cuh:
        # 369-406 vs gaussianParticles.cuh:499-501; see oracle/gut_oracle.c, gut_oracle_render_bwd) -- i.e. the transmittance
        # in front of a hit carries no distance gradient back to the closer particles: detached here.
        D = (alpha * T_before.detach() * accf * tt).sum(1)
        T_final = torch.prod(om, dim=1)
        n_hits = (acc & (w > 0)).sum(1).to(dtype)
        px_rgba = torch.cat([C, (1 - T_final).unsqueeze(-1)], -1) * vr.unsqueeze(-1).to(dtype)
        rgba_parts.append((ys, xs, px_rgba))
        dist_parts.append((ys, xs, torch.where(vr, D, torch.full_like(D, 1e6)), n_hits))
    rows_rgba = torch.zeros((H, W, 4), dtype=dtype)
    # assemble without in-place writes on leaf-dependent tensors: build per-tile then index_put once
    rgba = rgba.clone(); dist = dist.clone()
    for (ys, xs, v), (_, _, dv, hv) in zip(rgba_parts, dist_parts):
        hh, ww = ys.stop - ys.start, xs.stop - xs.start
        rgba[ys, xs] = v.reshape(hh, ww, 4)
        dist[ys, xs] = dv.reshape(hh, ww)
        hits[ys, xs] = hv.reshape(hh, ww)
    del rows_rgba
    return rgba, dist, hits


def forward_backward(cfg, cam, density12, sph, ray_ori, ray_dir, W, H, sph_degree, tile_ranges, sorted_values, tiles_count,
                     rgba_grad=None, dist_grad=None, dtype=torch.float64):
    """K1's SH colour + K4, and (when upstream gradients are given) K5 + K6 by autograd.
    Binning (tile_ranges / sorted_values / tiles_count) is an input so that images are comparable pixel by pixel."""
    d12 = torch.as_tensor(np.asarray(density12), dtype=dtype)
    leaves = {"pos": d12[:, 0:3].clone(), "dns": d12[:, 3].clone(), "quat": d12[:, 4:8].clone(), "scl": d12[:, 8:11].clone(),
              "sph": torch.as_tensor(np.asarray(sph), dtype=dtype).clone()}
    want_grad = rgba_grad is not None
    for v in leaves.values():
        v.requires_grad_(want_grad)
    fc = frame_constants(cam, dtype)
    feats = features_from_sh(cfg, fc, leaves["pos"], leaves["sph"], sph_degree)
    feats = feats * torch.as_tensor(np.asarray(tiles_count) > 0, dtype=dtype).unsqueeze(-1)  # K1 writes zeros for culled particles
    rgba, dist, hits = composite(cfg, cam, leaves["pos"], leaves["dns"], leaves["quat"], leaves["scl"], feats, ray_ori, ray_dir, W, H,
                                 tile_ranges, sorted_values, dtype)
    out = {"rgba": rgba.detach().numpy(), "dist": dist.detach().numpy(), "hits": hits.numpy()}
    if want_grad:
        loss = (rgba * torch.as_tensor(np.asarray(rgba_grad), dtype=dtype)).sum() + (dist * torch.as_tensor(np.asarray(dist_grad), dtype=dtype)).sum()
        loss.backward()
        n = d12.shape[0]
        g12 = np.zeros((n, 12))
        g12[:, 0:3] = leaves["pos"].grad.numpy(); g12[:, 3] = leaves["dns"].grad.numpy()
        g12[:, 4:8] = leaves["quat"].grad.numpy(); g12[:, 8:11] = leaves["scl"].grad.numpy()
        out["density_grad"] = g12
        out["radiance_grad"] = leaves["sph"].grad.numpy()
    return out


# ---------------------------------------------------------------------------------------------------
# k_buffer_size > 0 (sorted compositing): gutKBufferRenderer.cuh:28-85 (HitParticleKBuffer), :217-292 (evalKBuffer)
# ---------------------------------------------------------------------------------------------------
def _kbuffer_order(K, alpha, tt, acc, min_t):
    """Order in which one ray processes its accepted list entries: a K-entry buffer sorted by hit distance; when full the
    closest entry is processed before the new one is inserted; the rest drains closest-first; processing stops when the
    transmittance falls below min_t.  alpha/tt/acc are 1-D numpy arrays over the tile's list."""
    buf = []          # (t, list position), ascending t; mirrors m_kbuffer[K - numHits ...]
    order = []
    T = 1.0
    for l in np.nonzero(acc)[0]:
        if len(buf) == K:
            t0, l0 = buf.pop(0)
            order.append(l0)
            T *= 1.0 - alpha[l0]
            if T < min_t:
                return order
        # insert: bubbles past entries with strictly smaller t (hitParticle.hitT > m_kbuffer[i].hitT)
        pos = len(buf)
        while pos > 0 and not (tt[l] > buf[pos - 1][0]):
            pos -= 1
        buf.insert(pos, (tt[l], l))
    for t0, l0 in buf:
        order.append(l0)
        T *= 1.0 - alpha[l0]
        if T < min_t:
            break
    return order


def forward_backward_kbuffer(cfg, cam, density12, sph, ray_ori, ray_dir, W, H, sph_degree, tile_ranges, sorted_values, tiles_count,
                             rgba_grad=None, dist_grad=None, dtype=torch.float64):
    """K4 / K5 / K6 for k_buffer_size > 0.  The forward is composited by torch in the k-buffer's order; the backward applies
    the reference's per-hit un-blend recurrences (shRadiativeParticles.slang:179-207, gaussianParticles.slang:394-451) to get
    dL/dalpha and dL/ddepth of every processed hit and lets autograd differentiate alpha and depth wrt the particle
    parameters -- which is what Slang's bwd_diff does there (true derivatives, clamp masked)."""
    K = int(cfg.k_buffer_size)
    d12 = torch.as_tensor(np.asarray(density12), dtype=dtype)
    pos, dns, quat, scl = d12[:, 0:3].clone(), d12[:, 3].clone(), d12[:, 4:8].clone(), d12[:, 8:11].clone()
    sphc = torch.as_tensor(np.asarray(sph), dtype=dtype).clone()
    want_grad = rgba_grad is not None
    for v in (pos, dns, quat, scl, sphc):
        v.requires_grad_(want_grad)
    fc = frame_constants(cam, dtype)
    feats = features_from_sh(cfg, fc, pos, sphc, sph_degree)
    feats = feats * torch.as_tensor(np.asarray(tiles_count) > 0, dtype=dtype).unsqueeze(-1)
    n = d12.shape[0]
    ro = torch.as_tensor(np.asarray(ray_ori), dtype=dtype).reshape(H, W, 3)
    rd = torch.as_tensor(np.asarray(ray_dir), dtype=dtype).reshape(H, W, 3)
    o_w = ro @ fc["R_c2w"].T + fc["t_c2w"]
    d_w = rd @ fc["R_c2w"].T
    inv = 1.0 / d_w
    ta, tb = (-1e6 - o_w) * inv, (1e6 - o_w) * inv
    tmin = torch.minimum(ta, tb).amax(-1).clamp_min(0)
    tmax = torch.maximum(ta, tb).amin(-1)
    valid_ray = tmax > tmin
    gx = (W + TILE - 1) // TILE
    Rt = quat_wxyz_to_matrix(quat).transpose(-1, -2)
    rgba = torch.zeros((H, W, 4), dtype=dtype)
    dist = torch.full((H, W), 1e6, dtype=dtype)
    hits = torch.zeros((H, W), dtype=dtype)
    g12 = np.zeros((n, 12)); fgrad_acc = torch.zeros((n, 3), dtype=dtype)
    leaves = [pos, dns, quat, scl]
    leaf_grads = [torch.zeros_like(v) for v in leaves]
    if want_grad:
        rg = torch.as_tensor(np.asarray(rgba_grad), dtype=dtype).reshape(H, W, 4)
        dg = torch.as_tensor(np.asarray(dist_grad), dtype=dtype).reshape(H, W)
    for t in range(tile_ranges.shape[0]):
        b0, b1 = int(tile_ranges[t, 0]), int(tile_ranges[t, 1])
        ty, tx = divmod(t, gx)
        ys, xs = slice(ty * TILE, min((ty + 1) * TILE, H)), slice(tx * TILE, min((tx + 1) * TILE, W))
        hh, ww = ys.stop - ys.start, xs.stop - xs.start
        vr = valid_ray[ys, xs].reshape(-1)
        npx = hh * ww
        if b1 <= b0:
            rgba[ys, xs] = 0.0
            dist[ys, xs] = torch.where(vr, torch.zeros(npx, dtype=dtype), torch.full((npx,), 1e6, dtype=dtype)).reshape(hh, ww)
            continue
        o = o_w[ys, xs].reshape(-1, 3); d = d_w[ys, xs].reshape(-1, 3)
        idx = torch.as_tensor(sorted_values[b0:b1].astype(np.int64))
        p, s, rt, dn = pos[idx], scl[idx], Rt[idx], dns[idx]
        giscl = 1.0 / s
        gro = giscl.unsqueeze(0) * torch.einsum("lij,plj->pli", rt, o.unsqueeze(1) - p.unsqueeze(0))
        grdu = giscl.unsqueeze(0) * torch.einsum("lij,pj->pli", rt, d)
        grd = grdu / grdu.norm(dim=-1, keepdim=True)
        cr = torch.linalg.cross(grd, gro, dim=-1)
        g2 = (cr * cr).sum(-1)
        resp = _response(cfg.kernel_degree, g2)
        alpha = torch.clamp(resp * dn.unsqueeze(0), max=cfg.max_alpha)
        grds = s.unsqueeze(0) * grd * (-(grd * gro).sum(-1, keepdim=True))
        tt = torch.sqrt((grds * grds).sum(-1))
        with torch.no_grad():
            acc = (resp > cfg.min_kernel_response) & (alpha > cfg.min_alpha) & (tt > tmin[ys, xs].reshape(-1, 1)) & \
                  (tt < tmax[ys, xs].reshape(-1, 1)) & vr.unsqueeze(1)
            a_np, t_np, acc_np = alpha.numpy(), tt.numpy(), acc.numpy()
            orders = [_kbuffer_order(K, a_np[q], t_np[q], acc_np[q], cfg.min_transmittance) for q in range(npx)]
            lmax = max(1, max(len(x) for x in orders))
            sel = torch.zeros((npx, lmax), dtype=torch.long)
            msk = torch.zeros((npx, lmax), dtype=torch.bool)
            for q, od in enumerate(orders):
                sel[q, :len(od)] = torch.as_tensor(od, dtype=torch.long)
                msk[q, :len(od)] = True
        mf = msk.to(dtype)
        a_sel = torch.gather(alpha, 1, sel) * mf
        t_sel = torch.gather(tt, 1, sel) * mf
        pidx = idx[sel]                                                    # [P, L] particle of each processed hit
        f_raw = feats[pidx]                                                # unclamped
        f_cl = torch.clamp(f_raw, min=0)
        T_before = torch.cat([torch.ones((npx, 1), dtype=dtype), torch.cumprod(1 - a_sel, dim=1)[:, :-1]], 1)
        w = a_sel * T_before
        C = (w.unsqueeze(-1) * f_cl).sum(1)
        D = (w * t_sel).sum(1)
        T_fin = torch.prod(1 - a_sel, dim=1)
        vrf = vr.to(dtype)
        rgba[ys, xs] = (torch.cat([C, (1 - T_fin).unsqueeze(-1)], -1) * vrf.unsqueeze(-1)).detach().reshape(hh, ww, 4)
        dist[ys, xs] = torch.where(vr, D, torch.full_like(D, 1e6)).detach().reshape(hh, ww)
        hits[ys, xs] = ((w > 0) & msk).sum(1).to(dtype).reshape(hh, ww)
        if not want_grad:
            continue
        with torch.no_grad():
            Tb = T_fin.clone(); Tg = -rg[ys, xs, 3].reshape(-1).clone()
            Db = D.clone(); Dg = dg[ys, xs].reshape(-1).clone()
            Fb = C.clone(); Fg = rg[ys, xs, :3].reshape(-1, 3).clone()
            d_alpha = torch.zeros_like(a_sel); d_depth = torch.zeros_like(a_sel)
            for kk in range(lmax):
                a = a_sel[:, kk]; on = (a > 0) & msk[:, kk]
                onf = on.to(dtype)
                wgt = 1.0 / (1.0 - a)
                fr = f_raw[:, kk]
                Fb_n = (Fb - fr * a.unsqueeze(-1)) * wgt.unsqueeze(-1)
                Fb = torch.where(on.unsqueeze(-1), Fb_n, Fb)
                ag = (Fg * (fr - Fb)).sum(-1) * onf
                fgrad_acc.index_add_(0, pidx[:, kk], Fg * (a * onf).unsqueeze(-1))
                Fg = Fg * (1 - a * onf).unsqueeze(-1)
                Tb = torch.where(on, Tb * wgt, Tb)
                Db = torch.where(on, (Db - t_sel[:, kk] * a) * wgt, Db)
                d_alpha[:, kk] = (ag + Dg * (t_sel[:, kk] - Db) - Tg * Tb) * onf
                d_depth[:, kk] = Dg * a * onf
                Dg = Dg * (1 - a * onf); Tg = Tg * (1 - a * onf)
        gr = torch.autograd.grad([a_sel, t_sel], leaves, grad_outputs=[d_alpha, d_depth], retain_graph=True, allow_unused=True)
        for acc_g, g_ in zip(leaf_grads, gr):
            if g_ is not None:
                acc_g += g_
    out = {"rgba": rgba.numpy(), "dist": dist.numpy(), "hits": hits.numpy()}
    if want_grad:
        g12[:, 0:3] = leaf_grads[0].numpy(); g12[:, 3] = leaf_grads[1].numpy()
        g12[:, 4:8] = leaf_grads[2].numpy(); g12[:, 8:11] = leaf_grads[3].numpy()
        # K6: mask by feature > 0, back through the SH evaluation (no position gradient through the view direction)
        fg = fgrad_acc * (feats.detach() > 0).to(dtype)
        sg, = torch.autograd.grad([feats], [sphc], grad_outputs=[fg])
        out["density_grad"] = g12
        out["radiance_grad"] = sg.numpy()
        out["feature_grad"] = fgrad_acc.numpy()
    return out
