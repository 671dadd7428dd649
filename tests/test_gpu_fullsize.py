"""GPU property tests at BASELINE.json's full sizes (the oracle cannot run there in seconds).

C2: 1M Gaussians, SH deg 3, one 1237x822 pinhole view.   C4: 1M Gaussians, OpenCV fisheye + rolling shutter, 1920x1080.
Size-independent properties of the domain:
  * binning: tile ranges partition [0, M); every tile's keys carry its tile id and ascending depth bits; the multiset of
    sorted particle indices equals the expansion of tiles_count (a checksum of counts per particle);
  * forward: alpha in [0, max], idempotent (bit-identical on a second call), forward-only and training forward agree;
  * backward: linear in the upstream gradient (g(a*x) = a*g(x)), accumulate mode adds, column 11 stays zero,
    particles that touch no tile get exactly zero gradient.
"""
import importlib

import numpy as np
import pytest
import torch

from tests.helpers import camera_of, cfg_from_conf, make_conf, rel_l2, syn

pytestmark = pytest.mark.gpu

CONFIGS = {
    "c2_pinhole_1237x822": dict(n=1_000_000, w=1237, h=822, model="pinhole", rolling=False, seed=1002),
    "c4_fisheye_rolling_1080p": dict(n=1_000_000, w=1920, h=1080, model="fisheye", rolling=True, seed=1004),
}


def _setup(c):
    native = importlib.import_module("3dgrut_b200.native")
    raster = native.SplatRaster(cfg_from_conf(make_conf()))
    scene = syn.make_scene(c["n"], 3, seed=c["seed"], kind="init")
    view = syn.make_view(c["w"], c["h"], model=c["model"], rolling=c["rolling"])
    d12 = syn.pack_density(scene).cuda()
    sph = scene["features"].cuda()
    return raster, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), camera_of(view)


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_binning_properties(name):
    c = CONFIGS[name]
    raster, d12, sph, ro, rd, cam = _setup(c)
    rgba, dist, hits, vis, st = raster.trace(0, 3, d12, sph, ro, rd, None, cam, stage_outputs=True)
    torch.cuda.synchronize()
    m = int(st["num_intersections"])
    tc = st["tiles_count"].long()
    assert int(tc.sum()) == m and m > c["n"]            # the scene covers the view: several tiles per particle
    ranges = st["tile_ranges"].view(torch.int32).long().reshape(-1, 2)
    t = ranges.shape[0]
    assert t == ((c["w"] + 15) // 16) * ((c["h"] + 15) // 16)
    # the ranges of the non-empty tiles (empty ones stay (0, 0) like the reference's, gutRenderer.cu:46-76) partition [0, M)
    assert bool((ranges[:, 1] >= ranges[:, 0]).all())
    ne = ranges[ranges[:, 1] > ranges[:, 0]]
    assert int(ne[0, 0]) == 0 and int(ne[-1, 1]) == m and bool((ne[1:, 0] == ne[:-1, 1]).all())
    keys = st["sorted_keys"].view(torch.int64)[:m]
    tile_of = keys >> 32
    depth_bits = keys & 0xFFFFFFFF
    owner = torch.repeat_interleave(torch.arange(t, device=keys.device), ranges[:, 1] - ranges[:, 0])
    assert bool((tile_of == owner).all())
    # (tile, depth) keys ascend over the whole list: sortedness of the radix sort output
    assert bool((keys[1:] >= keys[:-1]).all())
    assert bool((depth_bits >= 0).all())
    # checksum of counts: every particle appears exactly tiles_count times
    idx = st["sorted_particle_idx"].view(torch.int32).long()[:m]
    assert bool((torch.bincount(idx, minlength=c["n"]) == tc).all())
    # depth key of every entry is the particle's depth
    gd = st["global_depth"].view(torch.int32).long() & 0xFFFFFFFF
    assert bool((gd[idx] == depth_bits).all())
    # every particle that reaches a tile is flagged visible (the flag is set before tile culling, gutProjector.cuh:284-322)
    assert bool((vis.reshape(-1)[tc > 0] > 0).all())


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_forward_properties(name):
    c = CONFIGS[name]
    raster, d12, sph, ro, rd, cam = _setup(c)
    a = raster.trace(0, 3, d12, sph, ro, rd, None, cam)
    b = raster.trace(0, 3, d12, sph, ro, rd, None, cam)
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        assert torch.equal(x, y)                         # idempotent, bit for bit (no atomics on the image path)
    rgba, dist, hits, vis = a
    assert bool(torch.isfinite(rgba).all()) and bool(torch.isfinite(dist).all())
    assert float(rgba[..., 3].min()) >= 0.0 and float(rgba[..., 3].max()) <= 1.0
    assert float(rgba[..., :3].min()) >= 0.0
    assert float(hits.max()) > 0 and float(rgba[..., 3].mean()) > 0.05
    # a pixel without hits keeps alpha 0 and colour 0
    empty = hits.reshape(-1) == 0
    assert float(rgba.reshape(-1, 4)[empty].abs().sum()) == 0.0


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_backward_properties(name):
    c = CONFIGS[name]
    raster, d12, sph, ro, rd, cam = _setup(c)
    h, w = c["h"], c["w"]
    rgba, dist, hits, vis, st = raster.trace(0, 3, d12, sph, ro, rd, None, cam, stage_outputs=True)
    g = torch.Generator(device="cuda").manual_seed(c["seed"])
    rg = torch.rand((h, w, 4), device="cuda", generator=g) * 2 - 1
    dg = (torch.rand((h, w, 1), device="cuda", generator=g) * 2 - 1) * 0.01

    def bwd(scale, **kw):
        return raster.trace_bwd(0, 3, d12, sph, ro, rd, None, cam, 0, 1, None, None, rgba, rg * scale, dist, dg * scale, **kw)

    g12, gsph = bwd(1.0)
    g12b, gsphb = bwd(-0.5)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(g12).all()) and bool(torch.isfinite(gsph).all())
    assert float(g12[:, 11].abs().max()) == 0.0
    # linearity in the upstream gradient (atomics reorder sums: compare in relative L2)
    for lo, hi in ((0, 3), (3, 4), (4, 8), (8, 11)):
        assert rel_l2((-2.0 * g12b[:, lo:hi]).cpu().numpy(), g12[:, lo:hi].cpu().numpy()) <= 1e-4
    assert rel_l2((-2.0 * gsphb).cpu().numpy(), gsph.cpu().numpy()) <= 1e-4
    # particles that touch no tile receive exactly zero
    untouched = st["tiles_count"] == 0
    if bool(untouched.any()):
        assert float(g12[untouched].abs().max()) == 0.0 and float(gsph[untouched].abs().max()) == 0.0
    # accumulate mode adds into caller-owned buffers
    acc12, accsph = g12.clone(), gsph.clone()
    bwd(1.0, out_density_grad=acc12, out_radiance_grad=accsph, accumulate=True)
    torch.cuda.synchronize()
    assert rel_l2(acc12.cpu().numpy(), (2.0 * g12).cpu().numpy()) <= 1e-4
    assert rel_l2(accsph.cpu().numpy(), (2.0 * gsph).cpu().numpy()) <= 1e-4
