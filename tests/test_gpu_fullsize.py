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


def _lexicographic(st):
    m = int(st["num_intersections"])
    keys = st["sorted_keys"].view(torch.int64)[:m]
    idx = st["sorted_particle_idx"].view(torch.int32).long()[:m]
    up = keys[1:] > keys[:-1]
    tie = (keys[1:] == keys[:-1]) & (idx[1:] > idx[:-1])
    return bool((up | tie).all()), m, idx


SORT_CASES = {
    # bucket sizes per 16x16 tile of a 64x64 view: the three code paths of the per-tile sort
    "buckets_up_to_4096": dict(n=12_000, piled=False),
    "buckets_up_to_12288": dict(n=50_000, piled=False),
    "buckets_above_12288": dict(n=400_000, piled=False),
    # depths piled up on a few neighbouring float values plus outliers: the counting sort hands the bucket to the radix sort
    "piled_depths": dict(n=30_000, piled=True),
}


@pytest.mark.parametrize("name", sorted(SORT_CASES))
def test_sort_paths(name):
    c = SORT_CASES[name]
    native = importlib.import_module("3dgrut_b200.native")
    raster = native.SplatRaster(cfg_from_conf(make_conf()))
    n, w, h = c["n"], 64, 64
    scene = syn.make_scene(n, 3, seed=77, kind="init")
    view = syn.make_view(w, h)
    if c["piled"]:
        c2w = view["T_to_world"][0]
        right, down, fwd, eye = c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3]
        g = torch.Generator().manual_seed(5)
        z = 3.0 + (torch.arange(n) % 61).float() * 2.4e-7          # a few dozen neighbouring floats -> ties and near ties
        far = torch.rand(n, generator=g) < 0.02
        z[far] = 0.5 + 20.0 * torch.rand(int(far.sum()), generator=g)   # outliers stretch the key range
        a = (torch.rand(n, generator=g) - 0.5) * 0.9 * z
        b = (torch.rand(n, generator=g) - 0.5) * 0.9 * z
        scene["positions"] = (eye[None] + z[:, None] * fwd[None] + a[:, None] * right[None] + b[:, None] * down[None]).contiguous()
        scene["scale"] = torch.full((n, 3), 0.01)
    d12 = syn.pack_density(scene).cuda()
    sph = scene["features"].cuda()
    rgba, dist, hits, vis, st = raster.trace(0, 3, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, camera_of(view),
                                             stage_outputs=True)
    torch.cuda.synchronize()
    ok, m, idx = _lexicographic(st)
    longest = int(raster.frame_stats()["max_tile_list_length"])
    assert ok, f"{name}: (tile, depth, particle) order broken (M={m}, longest bucket {longest})"
    assert bool((torch.bincount(idx, minlength=n) == st["tiles_count"].long()).all())
    ranges = st["tile_ranges"].view(torch.int32).long().reshape(-1, 2)
    sizes = ranges[:, 1] - ranges[:, 0]
    if name == "buckets_up_to_4096":
        assert int(sizes.max()) <= 4096
    elif name == "buckets_up_to_12288":
        assert 4096 < int(sizes.max()) <= 12288
    elif name == "buckets_above_12288":
        assert int(sizes.max()) > 12288
    assert bool(torch.isfinite(rgba).all())


CULL_CASES = {
    "c2_isotropic": dict(n=1_000_000, w=1237, h=822, model="pinhole", kind="init", seed=1002),
    "anisotropic_dense": dict(n=300_000, w=640, h=480, model="pinhole", kind="dense", seed=31),
    "fisheye": dict(n=300_000, w=640, h=480, model="fisheye", kind="dense", seed=32),
}


@pytest.mark.parametrize("name", sorted(CULL_CASES))
def test_block_cull_changes_no_bit(name):
    """The forward drops, per 8x8 block, list entries that no ray of the block can accept (gut_render.cu, block_cull).
    The bound is conservative: image, distance, hit counts and the gradients (through the hit masks) are bit-identical /
    equal to walking every entry."""
    c = CULL_CASES[name]
    native = importlib.import_module("3dgrut_b200.native")
    raster = native.SplatRaster(cfg_from_conf(make_conf()))
    scene = syn.make_scene(c["n"], 3, seed=c["seed"], kind=c["kind"])
    view = syn.make_view(c["w"], c["h"], model=c["model"])
    d12 = syn.pack_density(scene).cuda()
    sph = scene["features"].cuda()
    ro, rd, cam = view["rays_ori"].cuda(), view["rays_dir"].cuda(), camera_of(view)
    g = torch.Generator(device="cuda").manual_seed(7)
    rg = torch.rand((c["h"], c["w"], 4), device="cuda", generator=g) * 2 - 1
    dg = (torch.rand((c["h"], c["w"], 1), device="cuda", generator=g) * 2 - 1) * 0.01

    def run():
        out = raster.trace(0, 3, d12, sph, ro, rd, None, cam)
        t, hh = raster.measure_work(ro, rd)
        g12, gsph = raster.trace_bwd(0, 3, d12, sph, ro, rd, None, cam, 0, 1, None, None, out[0], rg, out[1], dg)
        torch.cuda.synchronize()
        return [x.clone() for x in out], t, hh, g12.clone(), gsph.clone()

    culled, t_c, h_c, g12_c, gsph_c = run()
    raster.set_block_cull(False)
    full, t_f, h_f, g12_f, gsph_f = run()
    raster.set_block_cull(True)
    for a, b in zip(culled, full):
        assert torch.equal(a, b)
    assert h_c == h_f and t_c <= t_f
    assert rel_l2(g12_c.cpu().numpy(), g12_f.cpu().numpy()) <= 1e-5      # same hits; atomics reorder the sums
    assert rel_l2(gsph_c.cpu().numpy(), gsph_f.cpu().numpy()) <= 1e-5
    print(f"{name}: tests per ray {t_f} -> {t_c} ({100.0 * (1 - t_c / max(t_f, 1)):.1f} % dropped)")


# ---------------------------------------------------------------------------------------------------------------------
# Oracle checks at full size (VERDICT r01 "What's weak" 3).  The oracle cannot composite a whole full-size frame in seconds,
# but it can (i) bin and sort the whole frame given libgut_b200's projected records -- the north_star clause "bit-exact
# given identical projected inputs", here on lists of several thousand entries that reach every tier of the per-tile sort
# -- and (ii) composite a SAMPLE of tiles forward and backward (its tile_list argument, the machinery bench.py times).
# The sample holds the longest lists of the frame and random ones.  Upstream gradients are zero outside the sample, so
# the whole-frame CUDA backward is the sample's gradient.
# ---------------------------------------------------------------------------------------------------------------------
ORACLE_CONFIGS = dict(CONFIGS)
ORACLE_CONFIGS["c3_bench_3m_pinhole_1237x822"] = dict(n=3_000_000, w=1237, h=822, model="pinhole", rolling=False, seed=1003)
ORACLE_CONFIGS["c2_opencv_distorted_left_to_right"] = dict(n=1_000_000, w=1237, h=822, model="opencv_pinhole", rolling=True,
                                                           shutter="ROLLING_LEFT_TO_RIGHT", seed=1005)


@pytest.mark.parametrize("name", sorted(ORACLE_CONFIGS))
def test_oracle_on_a_tile_sample_at_full_size(name):
    from oracle import gut_oracle as O
    from tests.helpers import BORDERLINE_MARGIN, abi, compare_images

    c = ORACLE_CONFIGS[name]
    n, w, h, deg = c["n"], c["w"], c["h"], 3
    native = importlib.import_module("3dgrut_b200.native")
    cfg = cfg_from_conf(make_conf())
    raster = native.SplatRaster(cfg)
    scene = syn.make_scene(n, 3, seed=c["seed"], kind="init")
    view = syn.make_view(w, h, model=c["model"], rolling=c["rolling"], shutter=c.get("shutter"))
    cam = camera_of(view)
    d12 = syn.pack_density(scene).cuda(); sph = scene["features"].cuda()
    ro, rd = view["rays_ori"].cuda(), view["rays_dir"].cuda()
    O.set_threads(len(__import__("os").sched_getaffinity(0)))
    rgba, dist, hits, vis, st = raster.trace(0, deg, d12, sph, ro, rd, None, cam, stage_outputs=True)
    torch.cuda.synchronize()
    m = int(st["num_intersections"])
    assert m > n

    # (i) whole-frame binning given libgut_b200's projected records: bit-exact keys, particle order and ranges
    proj = {"tiles_count": st["tiles_count"].cpu().numpy().view(np.uint32), "proj_pos": st["proj_position"].cpu().numpy(),
            "conic_opacity": st["proj_conic_opacity"].cpu().numpy(), "extent": st["proj_extent"].cpu().numpy(),
            "depth": st["global_depth"].cpu().numpy()}
    bins = O.bin_and_sort(cfg, w, h, proj)
    assert bins["M"] == m
    assert np.array_equal(bins["sorted_keys"], st["sorted_keys"].cpu().numpy().view(np.uint64))
    assert np.array_equal(bins["sorted_values"], st["sorted_particle_idx"].cpu().numpy().view(np.uint32))
    assert np.array_equal(bins["tile_ranges"], st["tile_ranges"].cpu().numpy().view(np.uint32).reshape(-1, 2))
    lengths = (bins["tile_ranges"][:, 1].astype(np.int64) - bins["tile_ranges"][:, 0])
    # K1 itself against the oracle's projection of all particles
    op = O.project(cfg, cam, d12.cpu().numpy(), sph.cpu().numpy(), w, h, deg)
    assert float((op["tiles_count"] == proj["tiles_count"]).mean()) >= (0.995 if c["rolling"] else 0.9999)

    # (ii) a sample of tiles: the 12 longest lists, 20 random ones
    rng = np.random.default_rng(c["seed"])
    gx = (w + 15) // 16
    heavy = np.argsort(-lengths)[:12]
    rest = rng.choice(np.setdiff1d(np.arange(lengths.size), heavy), size=20, replace=False)
    tiles = np.sort(np.concatenate([heavy, rest])).astype(np.uint32)
    assert lengths[heavy[0]] > 1000     # full size: lists the small cases never see
    in_sample = np.zeros((h, w), bool)
    for t in tiles:
        tx, ty = int(t) % gx, int(t) // gx
        in_sample[ty * 16:min(h, ty * 16 + 16), tx * 16:min(w, tx * 16 + 16)] = True
    d12_np, sph_np = d12.cpu().numpy(), sph.cpu().numpy()
    ro_np, rd_np = view["rays_ori"].numpy(), view["rays_dir"].numpy()
    feat = st["features"].cpu().numpy()[:, :3].copy()
    orc = O.render(cfg, cam, d12_np, feat, ro_np, rd_np, w, h, bins["tile_ranges"], bins["sorted_values"], tile_list=tiles)
    border = orc["margin"] < BORDERLINE_MARGIN
    g_hits = hits.cpu().numpy().reshape(h, w)
    flipped = (g_hits != orc["hits"]) & ~border & in_sample
    # rays that took another accept / kill decision than the oracle with the oracle's margin above BORDERLINE_MARGIN: the small
    # cases allow 2e-4 of the rays at ~300 tests per ray; the rays of the sample's heavy tiles see 4000-6500 tests each: 1e-3
    assert int(flipped.sum()) <= max(1, int(1e-3 * in_sample.sum())), int(flipped.sum())
    skip = border | flipped | ~in_sample
    cmp = compare_images(rgba.cpu().numpy().reshape(h, w, 4), orc["rgba"], skip)
    assert cmp["max_abs_clean"] <= 1e-4, cmp
    clean = ~skip
    assert clean.sum() > 0.8 * in_sample.sum()
    derr = np.abs(dist.cpu().numpy().reshape(h, w) - orc["dist"])[clean].max() / max(1.0, float(np.abs(orc["dist"][clean]).max()))
    assert derr <= 1e-4, derr

    # backward: seeded upstream gradients on the clean rays of the sample, zero elsewhere
    rg = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32); dg = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    rg[~clean] = 0; dg[~clean] = 0
    g12, gsph = raster.trace_bwd(0, deg, d12, sph, ro, rd, None, cam, 0, 1, None, None, rgba, torch.from_numpy(rg).cuda(), dist,
                                 torch.from_numpy(dg).cuda().reshape(h, w, 1))
    torch.cuda.synchronize()
    cproj = dict(proj); cproj["features"] = feat
    fwd = {"proj": cproj, "bins": {"M": m, "tile_ranges": bins["tile_ranges"], "sorted_values": bins["sorted_values"]},
           "rgba": rgba.cpu().numpy().reshape(h, w, 4), "dist": dist.cpu().numpy().reshape(h, w)}
    o12, osph, _ = O.backward(cfg, cam, d12_np, sph_np, ro_np, rd_np, w, h, deg, fwd, rg, dg, tile_list=tiles)
    g12 = g12.cpu().numpy()
    assert np.abs(o12).max() > 0
    for lo, hi, what in ((0, 3, "position"), (3, 4, "density"), (4, 8, "rotation"), (8, 11, "scale")):
        e = rel_l2(g12[:, lo:hi], o12[:, lo:hi])
        assert e <= 1e-3, (name, what, e)
    assert rel_l2(gsph.cpu().numpy(), osph) <= 1e-3
    print(name, "M", m, "longest list", int(lengths.max()), "rgba", cmp["max_abs_clean"], "rotation rel L2",
          rel_l2(g12[:, 4:8], o12[:, 4:8]), "scale", rel_l2(g12[:, 8:11], o12[:, 8:11]))
