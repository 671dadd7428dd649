"""World-size-2 CPU test (gloo) of the view-sharded data-parallel plumbing (3dgrut_b200/dist.py):
views are split round-robin, each rank accumulates the gradients of its views in place, one sum-allreduce closes the
step -- the result must equal the single-process sum over all views.  The rasteriser itself is replaced by a
deterministic stand-in with the SplatRaster.trace / trace_bwd signature (no GPU here)."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

gdist = importlib.import_module("3dgrut_b200.dist")

N, NSPH, NVIEWS, H, W = 257, 48, 5, 6, 7


class FakeRaster:
    """Stand-in with the SplatRaster call signature: gradients are a fixed linear function of the view's upstream gradient."""

    def trace(self, frame, deg, density, radiance, ray_o, ray_d, ray_t, camera):
        rgba = ray_o.sum() * torch.ones((H, W, 4)); d = torch.ones((H, W, 1))
        vis = torch.zeros((density.shape[0], 1))
        vis[int(ray_o.flatten()[0])::NVIEWS + 2] = 1.0      # view k sees particles k, k + 7, k + 14, ...
        return rgba, d, torch.zeros((H, W, 1)), vis

    def trace_bwd(self, frame, deg, density, radiance, ray_o, ray_d, ray_t, camera, ts0, ts1, p0, p1, rgba, rgba_grad, d, d_grad,
                  out_density_grad=None, out_radiance_grad=None, accumulate=False):
        g1 = rgba_grad.sum() * density + d_grad.sum()
        g2 = rgba_grad.mean() * radiance
        if accumulate:
            out_density_grad += g1; out_radiance_grad += g2
        else:
            out_density_grad.copy_(g1); out_radiance_grad.copy_(g2)
        return out_density_grad, out_radiance_grad


    # deferred-SH variants (gut_backward_colour / gut_sh_gradient_expand): the stand-in's "SH gradient" of a view is
    # centre_x * colour gradient broadcast over the 48 columns, so that the expansion needs both gathered inputs
    def trace_bwd_colour(self, deg, density, radiance, ray_o, ray_d, camera, rgba, rgba_grad, d, d_grad, out_colour_grad,
                         out_density_grad=None):
        if out_density_grad is not None:    # one view per rank: the view's gradient is written directly, nothing pending
            out_density_grad.copy_(rgba_grad.sum() * density + d_grad.sum())
        else:
            self._pending = getattr(self, "_pending", 0) + (rgba_grad.sum() * density + d_grad.sum())
        out_colour_grad.copy_(rgba_grad.mean() * radiance[:, :3])
        return [float(ray_o.flatten()[0]) + 1.0, 0.0, 0.0]

    def density_gradient_finish(self, density, out, accumulate=False):
        out.copy_(getattr(self, "_pending", 0) + torch.zeros_like(out))
        self._pending = 0
        return out

    def sh_gradient_expand(self, deg, density, centres, colour_all, out, accumulate=False):
        val = (centres[:, 0].reshape(-1, 1, 1) * colour_all[:, :, :3]).sum(0)          # [N,3]
        out.copy_(val.repeat(1, NSPH // 3))
        return out


def _data(nviews=None):
    nviews = NVIEWS if nviews is None else nviews
    g = torch.Generator().manual_seed(0)
    density = torch.randn((N, 12), generator=g); radiance = torch.randn((N, NSPH), generator=g)
    views = [{"rays_ori": torch.full((1, H, W, 3), float(k)), "rays_dir": torch.ones((1, H, W, 3)), "camera": None} for k in range(nviews)]
    ups = [(torch.randn((H, W, 4), generator=g), torch.randn((H, W, 1), generator=g)) for _ in range(nviews)]
    return density, radiance, views, ups


def _step(rank, world):
    density, radiance, views, ups = _data()
    ids = gdist.shard_views(NVIEWS, rank, world)
    ex = gdist.GradientExchange(N, NSPH, "cpu", world=world)
    gdist.training_step(FakeRaster(), ex, [views[i] for i in ids], 3, density, radiance, [ups[i] for i in ids])
    return ex.density_grad.clone(), ex.radiance_grad.clone()


def _step_deferred(rank, world, nviews=None):
    nviews = NVIEWS if nviews is None else nviews
    density, radiance, views, ups = _data(nviews)
    ids = gdist.shard_views(nviews, rank, world)
    vpr = (nviews + world - 1) // world
    ex = gdist.DeferredExchange(N, NSPH, vpr, "cpu", world=world, track_visibility=True)
    ex.density_grad.fill_(123.0)     # stale content a step must not keep
    gdist.training_step_deferred(FakeRaster(), ex, [views[i] for i in ids], 3, density, radiance, [ups[i] for i in ids], rank=rank)
    # the visibility SelectiveAdam masks with is the union over ALL views of the step, on every rank
    want = torch.zeros((N, 1))
    for k in range(nviews):
        want[k::NVIEWS + 2] = 1.0
    assert torch.equal(ex.visibility, want), "visibility union"
    return ex.density_grad.clone(), ex.radiance_grad.clone()


def _worker_deferred(rank, world, port, q, nviews=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g1, g2 = _step_deferred(rank, world, nviews)
        q.put((rank, g1.numpy(), g2.numpy()))
    finally:
        dist.destroy_process_group()


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g1, g2 = _step(rank, world)
        q.put((rank, g1.numpy(), g2.numpy()))
    finally:
        dist.destroy_process_group()


def test_shard_views_partition():
    for world in (1, 2, 3, 4, 8):
        got = sorted(v for r in range(world) for v in gdist.shard_views(8, r, world))
        assert got == list(range(8))
        sizes = [len(gdist.shard_views(8, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(120)
def test_two_rank_step_equals_single_process():
    ref1, ref2 = _step(0, 1)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    res = [q.get(timeout=90) for _ in procs]
    for p in procs: p.join(30)
    assert all(p.exitcode == 0 for p in procs)
    for _, g1, g2 in res:   # every rank holds the full sum after the allreduce
        torch.testing.assert_close(torch.from_numpy(g1), ref1, rtol=1e-5, atol=1e-4)
        torch.testing.assert_close(torch.from_numpy(g2), ref2, rtol=1e-5, atol=1e-5)


@pytest.mark.timeout(120)
def test_two_rank_deferred_step_equals_single_process():
    """all-gather of per-view colour gradients + allreduce of [N,12] + local expansion == the single-process result
    (ranks hold 3 and 2 of the 5 views: the missing slot of rank 1 must contribute zeros)."""
    ref1, ref2 = _step_deferred(0, 1)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_deferred, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    res = [q.get(timeout=90) for _ in procs]
    for p in procs: p.join(30)
    assert all(p.exitcode == 0 for p in procs)
    for _, g1, g2 in res:
        torch.testing.assert_close(torch.from_numpy(g1), ref1, rtol=1e-5, atol=1e-4)
        torch.testing.assert_close(torch.from_numpy(g2), ref2, rtol=1e-5, atol=1e-5)


def test_rank_without_views_contributes_zero():
    # more ranks than views: the empty rank's buffers must be zeroed before the allreduce
    density, radiance, views, ups = _data()
    ex = gdist.GradientExchange(N, NSPH, "cpu", world=1)
    ex.density_grad.fill_(3.0); ex.radiance_grad.fill_(4.0)
    gdist.training_step(FakeRaster(), ex, [], 3, density, radiance, [])
    assert float(ex.density_grad.abs().max()) == 0.0 and float(ex.radiance_grad.abs().max()) == 0.0


@pytest.mark.timeout(120)
@pytest.mark.parametrize("nviews", [2, 1])
def test_two_rank_step_with_one_view_per_rank(nviews):
    """views_per_rank == 1: every rank's backward writes its [N,12] gradient directly (DeferredExchange.density_out) and no
    finish pass runs; with ONE view for two ranks the rank without a view contributes zeros, not its buffer's old content."""
    ref1, ref2 = _step_deferred(0, 1, nviews)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_deferred, args=(r, 2, port, q, nviews)) for r in range(2)]
    for p in procs: p.start()
    res = [q.get(timeout=90) for _ in procs]
    for p in procs: p.join(30)
    assert all(p.exitcode == 0 for p in procs)
    for _rank, g1, g2 in res:
        assert np.allclose(g1, ref1.numpy(), rtol=1e-5, atol=1e-5)
        assert np.allclose(g2, ref2.numpy(), rtol=1e-5, atol=1e-5)


def test_exchange_slot_helpers_on_cpu():
    """Slot bookkeeping shared by the exchanges: clear_slot / moved_rows on the dense layout, and use() handing the exchange's
    slot layout to handles that know about layouts (stand-ins without the call are left alone)."""
    ex = gdist.DeferredExchange(N, NSPH, 2, "cpu", world=1)
    ex.begin_step(0)
    assert ex.colour_layout == 0 and ex.moved_rows(0) == 0
    ex.slot(0)[3] = torch.tensor([0.0, 2.0, 0.0])
    ex.slot(0)[9] = torch.tensor([1.0, 0.0, -1.0])
    ex.slot(1)[0, 0] = 5.0
    assert ex.moved_rows(0) == 2 and ex.moved_rows(1) == 1
    ex.clear_slot(0)
    assert ex.moved_rows(0) == 0 and ex.moved_rows(1) == 1

    class Layouted:
        seen = None

        def set_colour_layout(self, layout):
            self.seen = layout

    a, b = Layouted(), Layouted()
    ex.use([a, FakeRaster(), b])
    ex.use(FakeRaster())
    assert a.seen == 0 and b.seen == 0
