"""GPU: gut_project_noalloc (include/gut_b200.h) -- the projection stage with all of its scratch in a file-scope
`static __device__` array of libgut_b200, for callers that cannot allocate.  Called through the C-ABI (ctypes on the loaded
library) and compared with the existing entry point, gut_forward with stage outputs, on two calls in a row: the second call
(another view, fewer particles, another image size) finds the first call's contents in the shared array.

Bar: the same kernel on the same inputs -> per-particle records, visibility and tile ranges BIT-EXACT (the tolerance the
binning tests of tests/test_gpu_parity.py use for integer and index work; the float records are the same bits too).

The file is named to be collected last.  It keeps no state: it passes alone in a fresh
process and after the other tests.
"""
import ctypes as C
import importlib

import numpy as np
import pytest
import torch

from tests.helpers import abi, camera_of, cfg_from_conf, make_conf, syn

pytestmark = pytest.mark.gpu

MAX_PARTICLES = 65536   # GUT_NOALLOC_MAX_PARTICLES
MAX_TILES = 8192        # GUT_NOALLOC_MAX_TILES

K1_FIELDS = ("tiles_count", "proj_position", "proj_conic_opacity", "proj_extent", "global_depth", "features", "tile_ranges")


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _raster():
    native = importlib.import_module("3dgrut_b200.native")
    return native.SplatRaster(cfg_from_conf(make_conf()))


def _noalloc_outputs(n, w, h):
    """Caller-owned output buffers, filled with a pattern the call has to overwrite."""
    tiles = ((w + 15) // 16) * ((h + 15) // 16)
    dev = "cuda"
    st = {
        "tiles_count": torch.full((n,), -7, dtype=torch.int32, device=dev),
        "proj_position": torch.full((n, 2), -7.0, device=dev),
        "proj_conic_opacity": torch.full((n, 4), -7.0, device=dev),
        "proj_extent": torch.full((n, 2), -7.0, device=dev),
        "global_depth": torch.full((n,), -7.0, device=dev),
        "features": torch.full((n, 3), -7.0, device=dev),
        "tile_ranges": torch.full((tiles, 2), -7, dtype=torch.int32, device=dev),
    }
    so = abi.GutStageOutputs(
        d_tiles_count=st["tiles_count"].data_ptr(), d_proj_position=st["proj_position"].data_ptr(),
        d_proj_conic_opacity=st["proj_conic_opacity"].data_ptr(), d_proj_extent=st["proj_extent"].data_ptr(),
        d_global_depth=st["global_depth"].data_ptr(), d_features=st["features"].data_ptr(), d_sorted_keys=None,
        d_sorted_particle_idx=None, d_tile_ranges=st["tile_ranges"].data_ptr(), capacity_intersections=0, num_intersections=0)
    return so, st


def _case(n, w, h, deg, seed, kind, model, shutter):
    scene = syn.make_scene(n, 3, seed=seed, kind=kind)
    view = syn.make_view(w, h, model=model, shutter=shutter)
    return scene, view, camera_of(view), syn.pack_density(scene).cuda(), scene["features"].cuda(), deg


def _existing(raster, view, cam, d12, sph, deg):
    """The existing entry point: gut_forward with stage outputs."""
    out = raster.trace(0, deg, d12, sph, view["rays_ori"].cuda(), view["rays_dir"].cuda(), None, cam, stage_outputs=True)
    torch.cuda.synchronize()
    return out[3], out[4]


def _new(raster, cam, d12, sph, deg, w, h):
    """gut_project_noalloc through the C-ABI."""
    n = d12.shape[0]
    so, st = _noalloc_outputs(n, w, h)
    vis = torch.full((n, 1), -7.0, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = raster.lib.gut_project_noalloc(raster._h, stream, C.byref(cam), n, deg, _ptr(d12), _ptr(sph), w, h, _ptr(vis), C.byref(so))
    assert rc == 0, raster.lib.gut_last_error(raster._h).decode()
    torch.cuda.synchronize()
    return vis, st


def _assert_same(vis_new, st_new, vis_old, st_old, what):
    assert torch.equal(vis_new, vis_old), what
    for key in K1_FIELDS:
        a, b = st_new[key].cpu().numpy(), st_old[key].cpu().numpy()
        assert a.shape == b.shape, (what, key)
        # bit-exact, floats included (NaN-safe: compare the words)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (what, key, int((a.view(np.uint32) != b.view(np.uint32)).sum()))


def test_no_alloc_projection_equals_gut_forward_on_two_calls_in_a_row():
    raster = _raster()
    # first call: a dense view that fills many tile counters; second: fewer particles, smaller ragged image, fisheye under
    # a rolling shutter -- whatever the first call left in the shared array must not leak into it
    cases = [_case(20000, 320, 208, 3, 31, "dense", "pinhole", "GLOBAL"),
             _case(5000, 150, 70, 1, 32, "init", "fisheye", "ROLLING_TOP_TO_BOTTOM")]
    got = []
    for scene, view, cam, d12, sph, deg in cases:      # the two new calls back to back, nothing in between
        h, w = view["rays_ori"].shape[-3], view["rays_ori"].shape[-2]
        got.append(_new(raster, cam, d12, sph, deg, w, h))
    for i, (scene, view, cam, d12, sph, deg) in enumerate(cases):
        vis_old, st_old = _existing(raster, view, cam, d12, sph, deg)
        assert int(st_old["num_intersections"]) > 0
        _assert_same(got[i][0], got[i][1], vis_old, st_old, f"call {i}")
        r = got[i][1]["tile_ranges"].cpu().numpy().astype(np.int64)
        assert int((r[:, 1] - r[:, 0]).sum()) == int(st_old["num_intersections"])
    # and once more after the existing entry point ran: same answer as the first time
    scene, view, cam, d12, sph, deg = cases[0]
    vis3, st3 = _new(raster, cam, d12, sph, deg, 320, 208)
    _assert_same(vis3, st3, got[0][0], got[0][1], "repeat of call 0")


def test_no_alloc_projection_leaves_the_forward_context_alone():
    """A gut_backward after [gut_forward, gut_project_noalloc of another view] still belongs to the gut_forward."""
    raster = _raster()
    scene, view, cam, d12, sph, deg = _case(4000, 96, 64, 2, 33, "init", "pinhole", "GLOBAL")
    ro, rd = view["rays_ori"].cuda(), view["rays_dir"].cuda()
    rgba, dist, hits, vis = raster.trace(0, deg, d12, sph, ro, rd, None, cam)
    rng = np.random.default_rng(33)
    rg = torch.from_numpy(rng.uniform(-1, 1, (64, 96, 4)).astype(np.float32)).cuda()
    dg = torch.from_numpy(rng.uniform(-0.01, 0.01, (64, 96, 1)).astype(np.float32)).cuda()
    g12_a, gsph_a = raster.trace_bwd(0, deg, d12, sph, ro, rd, None, cam, 0, 1, None, None, rgba, rg, dist, dg)
    rgba, dist, hits, vis = raster.trace(0, deg, d12, sph, ro, rd, None, cam)
    _, _, cam2, d12_2, sph_2, deg2 = _case(3000, 128, 80, 1, 34, "dense", "pinhole", "GLOBAL")
    _new(raster, cam2, d12_2, sph_2, deg2, 128, 80)
    g12_b, gsph_b = raster.trace_bwd(0, deg, d12, sph, ro, rd, None, cam, 0, 1, None, None, rgba, rg, dist, dg)
    torch.cuda.synchronize()
    # float atomics land in another order between two runs: 1e-5 relative L2, as tests/test_gpu_parity.py has it
    for a, b in ((g12_a, g12_b), (gsph_a, gsph_b)):
        assert float((a - b).norm() / b.norm().clamp_min(1e-30)) <= 1e-5


def test_no_alloc_projection_rejects_sizes_beyond_the_fixed_scratch():
    raster = _raster()
    view = syn.make_view(64, 64)
    cam = camera_of(view)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    n = MAX_PARTICLES + 1
    d12 = torch.zeros((n, 12), device="cuda"); sph = torch.zeros((n, 48), device="cuda"); vis = torch.zeros((n, 1), device="cuda")
    rc = raster.lib.gut_project_noalloc(raster._h, stream, C.byref(cam), n, 0, _ptr(d12), _ptr(sph), 64, 64, _ptr(vis), None)
    assert rc == int(abi.GutStatus.BAD_INPUT)
    assert "GUT_NOALLOC_MAX_PARTICLES" in raster.lib.gut_last_error(raster._h).decode()
    # 2048 x 1040 = 128 x 65 = 8320 tiles > 8192
    rc = raster.lib.gut_project_noalloc(raster._h, stream, C.byref(cam), 16, 0, _ptr(d12), _ptr(sph), 2048, 1040, _ptr(vis), None)
    assert rc == int(abi.GutStatus.BAD_INPUT)
    assert "GUT_NOALLOC_MAX_TILES" in raster.lib.gut_last_error(raster._h).decode()
    # the limits themselves are accepted: 65536 particles on 1920 x 1080 (8160 tiles)
    n = MAX_PARTICLES
    scene = syn.make_scene(n, 3, seed=35)
    view = syn.make_view(1920, 1080)
    cam = camera_of(view)
    d12 = syn.pack_density(scene).cuda(); sph = scene["features"].cuda()
    vis_new, st_new = _new(raster, cam, d12, sph, 3, 1920, 1080)
    vis_old, st_old = _existing(raster, view, cam, d12, sph, 3)
    _assert_same(vis_new, st_new, vis_old, st_old, "at the limits")
    # sorted outputs are not produced: asking for them is an error, not a silent no-op
    so, st = _noalloc_outputs(16, 64, 64)
    keys = torch.zeros(16, dtype=torch.int64, device="cuda")
    so.d_sorted_keys = keys.data_ptr()
    rc = raster.lib.gut_project_noalloc(raster._h, stream, C.byref(cam), 16, 0, _ptr(d12), _ptr(sph), 64, 64, _ptr(vis), C.byref(so))
    assert rc == int(abi.GutStatus.BAD_INPUT)
