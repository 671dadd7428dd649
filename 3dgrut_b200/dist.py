"""View-sharded data parallelism for the 3DGUT training step.

The reference is single-GPU, batch size 1 (threedgrut/trainer.py:143-149) and has no collective at all.
Here the particle set is replicated on every rank, the views of a step are split over the ranks, every rank
sums the gradients of its views in place (``accumulate`` mode of gut_backward, no re-zeroing between views)
and ONE fp32 sum-allreduce of the two per-Gaussian gradient buffers ([N,12] and [N,48]) over NCCL / NVLink
closes the step.  The [N,12] buffer is final as soon as the last view's compositing backward (K5) has run,
so its allreduce is issued on a side stream and overlaps the SH backward (K6) of that view; the [N,48]
allreduce follows K6.  There is no other exchange: images and hit counts stay rank local.
"""
import torch
import torch.distributed as dist


def shard_views(n_views, rank, world):
    """Views of this rank: round-robin, so every rank gets ceil/floor(n_views/world) of them."""
    return list(range(rank, n_views, world))


class GradientExchange:
    """Owns the rank-local gradient accumulators and the allreduce of a step."""

    def __init__(self, n_particles, n_sph_floats, device, world=None):
        self.world = world if world is not None else (dist.get_world_size() if dist.is_initialized() else 1)
        self.density_grad = torch.zeros((n_particles, 12), dtype=torch.float32, device=device)
        self.radiance_grad = torch.zeros((n_particles, n_sph_floats), dtype=torch.float32, device=device)
        self.comm_stream = torch.cuda.Stream(device=device) if (self.world > 1 and torch.device(device).type == "cuda") else None
        self._pending = []

    def bytes_per_rank(self):
        return (self.density_grad.numel() + self.radiance_grad.numel()) * 4

    def allreduce(self):
        """Sum both buffers over the ranks.  Enqueued behind the work already on the current stream; the current
        stream waits for the result before anything later runs (call ``finish`` to make that explicit)."""
        if self.world <= 1:
            return
        if self.comm_stream is None:  # CPU / gloo path used by the tests
            dist.all_reduce(self.density_grad, op=dist.ReduceOp.SUM)
            dist.all_reduce(self.radiance_grad, op=dist.ReduceOp.SUM)
            return
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        with torch.cuda.stream(self.comm_stream):
            self.comm_stream.wait_event(ev)
            dist.all_reduce(self.density_grad, op=dist.ReduceOp.SUM)
            dist.all_reduce(self.radiance_grad, op=dist.ReduceOp.SUM)
            done = torch.cuda.Event()
            done.record(self.comm_stream)
        cur.wait_event(done)


def training_step(raster, exchange, views, sph_degree, particle_density, particle_radiance, upstream):
    """fwd+bwd over this rank's views, gradients summed in ``exchange``, then the allreduce.

    views: list of dicts with device tensors rays_ori, rays_dir [1,H,W,3] and a GutCamera ``camera``.
    upstream: list of (rgba_grad [H,W,4], dist_grad [H,W,1]) device tensors, one per view.
    Returns the list of (rgba, dist) outputs of the local views.
    """
    outs = []
    for i, (v, (rg, dg)) in enumerate(zip(views, upstream)):
        rgba, dist_, hits, vis = raster.trace(0, sph_degree, particle_density, particle_radiance, v["rays_ori"], v["rays_dir"], None,
                                              v["camera"])
        raster.trace_bwd(0, sph_degree, particle_density, particle_radiance, v["rays_ori"], v["rays_dir"], None, v["camera"], 0, 1,
                         None, None, rgba, rg, dist_, dg, out_density_grad=exchange.density_grad,
                         out_radiance_grad=exchange.radiance_grad, accumulate=(i > 0))
        outs.append((rgba, dist_))
    if not views:  # a rank without views contributes zeros
        exchange.density_grad.zero_()
        exchange.radiance_grad.zero_()
    exchange.allreduce()
    return outs
