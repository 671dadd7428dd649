"""View-sharded data parallelism for the 3DGUT training step.

The reference is single-GPU, batch size 1 (threedgrut/trainer.py:143-149) and has no collective at all.
Here the particle set is replicated on every rank, the views of a step are split over the ranks, every rank
sums the gradients of its views in place (``accumulate`` mode of gut_backward, no re-zeroing between views)
and ONE fp32 sum-allreduce of the two per-Gaussian gradient buffers ([N,12] and [N,48]) over NCCL / NVLink
closes the step.  There is no other exchange: images and hit counts stay rank local.

``training_step_deferred`` is the cheaper formulation of the same step.  The SH gradient of one view is
basis(view direction) x colour gradient -- 3 numbers per particle and view -- so the ranks all-gather the per-view
colour gradients ([N,3] each, 12 B per particle and view instead of 192 B per particle), allreduce only the [N,12]
buffer, and every rank expands all views of the step into [N,48] once (gut_sh_gradient_expand).  The [N,12] part is
deferred too: the backward's per-particle accumulators are view independent, so they sum over the rank's views and are
converted once (gut_density_gradient_finish).  The result is the same summed gradient on every rank; on one GPU it
replaces V read-modify-writes of both gradient buffers by one write each.
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


class DeferredExchange:
    """Buffers of the deferred-SH step: [N,12] accumulator, per-view colour gradients of this rank and of all ranks."""

    def __init__(self, n_particles, n_sph_floats, views_per_rank, device, world=None, track_visibility=False, gather_buffer=True):
        self.world = world if world is not None else (dist.get_world_size() if dist.is_initialized() else 1)
        self.views_per_rank = views_per_rank
        self.track_visibility = track_visibility   # needed by SelectiveAdam only: one more (small) collective per step
        self.density_grad = torch.zeros((n_particles, 12), dtype=torch.float32, device=device)
        self.radiance_grad = torch.zeros((n_particles, n_sph_floats), dtype=torch.float32, device=device)
        self.colour_all = (torch.zeros((self.world * views_per_rank, n_particles, 3), dtype=torch.float32, device=device)
                           if gather_buffer else None)
        self.centres_all = torch.zeros((self.world * views_per_rank, 3), dtype=torch.float32, device=device)
        # union over all views of the step of the tracer's mog_visibility: what SelectiveAdam masks its update with
        # (trainer.py:747-749 passes the visibility of the step's single view; with V views it is their union)
        self.visibility = torch.zeros((n_particles, 1), dtype=torch.float32, device=device)

    def local_slot(self, rank, i):
        return rank * self.views_per_rank + i

    def bytes_per_rank(self):
        """Bytes this rank contributes to the collectives of a step."""
        return self.density_grad.numel() * 4 + self.views_per_rank * self.colour_all.shape[1] * 12

    # ---- step protocol shared with PeerExchange (used by training_step_* below and by Tracer.begin_step) --------------
    def begin_step(self, rank=0):
        """Start a step: returns nothing; ``slot(i)`` is where view i's colour gradient goes."""
        self._rank = rank
        self._centres = []
        if self.track_visibility:
            self.visibility.zero_()

    def slot(self, i):
        """[N,3] buffer that receives the masked colour gradient of this rank's view i (gut_backward_colour)."""
        return self.colour_all[self.local_slot(self._rank, i)]

    colour_layout = 0     # GUT_COLOUR_DENSE: every slot is an [N,3] tensor (PeerExchange may hold compact slots instead)

    def use(self, rasters):
        """Give the rasteriser handles that write or read this exchange's slots its slot layout (gut_set_colour_layout)."""
        for r in (rasters if isinstance(rasters, (list, tuple)) else [rasters]):
            if hasattr(r, "set_colour_layout"):
                r.set_colour_layout(self.colour_layout)

    def clear_slot(self, i):
        """Make this rank's view slot i of the current step contribute nothing (a view without a backward)."""
        self.slot(i).zero_()

    def moved_rows(self, i):
        """Particles with a colour gradient in this rank's view slot i (synchronises; bench / diagnostics)."""
        return int((self.slot(i) != 0).any(dim=1).sum())

    def density_out(self, i):
        """[N,12] buffer the backward of view i may write its gradient into directly (gut_backward_colour_density), or None.
        A rank that renders ONE view per step has nothing to sum over views: its backward converts the accumulators itself
        and the step needs neither the per-view fold pass nor gut_density_gradient_finish."""
        return self.density_grad if self.views_per_rank == 1 else None

    def add_view(self, centre, visibility=None):
        """Record a finished view: the world-space view centre its backward used and (optionally) its mog_visibility."""
        self._centres.append([float(c) for c in centre])
        if self.track_visibility and visibility is not None:
            torch.maximum(self.visibility, visibility.reshape(self.visibility.shape), out=self.visibility)

    def colour_ptrs(self):
        """One device pointer per view slot of the step (all ranks), what gut_sh_gradient_expand_gather / gut_sh_expand_adam take."""
        return [self.colour_all[v].data_ptr() for v in range(self.colour_all.shape[0])]

    def finish_step(self, raster, sph_degree, particle_density, radiance_out=None, expand=True):
        """Close the step: [N,12] conversion of the view-summed accumulators, exchange with the other ranks, [N,48] expansion
        over all views of the step.  Leaves the summed gradients in ``density_grad`` / ``radiance_grad`` on every rank
        (``radiance_out``: write the [N,48] gradient there instead of into the exchange's own buffer).
        ``expand=False`` stops before the expansion -- every rank's colour gradients and centres are in place and the [N,12]
        all-reduce has completed -- for ``optimizer_epilogue``, which expands inside the optimiser kernel."""
        radiance = self.radiance_grad if radiance_out is None else radiance_out
        lo = self.local_slot(self._rank, 0)
        mine = self.colour_all[lo:lo + self.views_per_rank]
        centres = list(self._centres)
        rasters = list(raster) if isinstance(raster, (list, tuple)) else [raster]
        raster = rasters[0]
        if self.views_per_rank > 1:
            for j, r in enumerate(rasters):      # every lane's view-summed accumulators; zeros when this rank had no view
                r.density_gradient_finish(particle_density, self.density_grad, accumulate=j > 0)
        elif not centres:
            self.density_grad.zero_()   # one view per rank and none for this rank in this step
        for i in range(len(centres), self.views_per_rank):
            mine[i].zero_()
            centres.append([0.0, 0.0, 0.0])
        my_centres = self.centres_all[lo:lo + self.views_per_rank]
        my_centres.copy_(torch.tensor(centres, dtype=torch.float32), non_blocking=True)
        pending = None
        if self.world > 1:
            _all_gather(self.colour_all, mine)
            _all_gather(self.centres_all, my_centres)
            # the [N,12] allreduce runs on the collective stream while this stream expands the SH gradient
            pending = [dist.all_reduce(self.density_grad, op=dist.ReduceOp.SUM, async_op=True)]
            if self.track_visibility:   # union of the ranks' visibilities, 4 B/particle
                pending.append(dist.all_reduce(self.visibility, op=dist.ReduceOp.MAX, async_op=True))
        if expand:
            raster.sh_gradient_expand(sph_degree, particle_density, self.centres_all, self.colour_all, radiance, accumulate=False)
        if pending is not None:
            for work in pending:
                work.wait()
        return self.density_grad, (radiance if expand else None)


def _lanes(raster):
    """``raster`` may be ONE SplatRaster or a list of them ("lanes"): the views of a step are then dealt round-robin to the
    lanes, each on a CUDA stream of its own, so that the short latency-bound kernels of one view (projection, scan, emit, sort,
    per-view fold) run under the compositing kernels of another and fill the ends of their grids.  Every lane sums its views'
    accumulators in its own handle; ``finish_step`` converts them one after the other into the same [N,12] buffer."""
    lanes = list(raster) if isinstance(raster, (list, tuple)) else [raster]
    if len(lanes) > 1:
        for r in lanes:
            if getattr(r, "_lane_stream", None) is None:
                r._lane_stream = torch.cuda.Stream(device=r.device)
    return lanes


def training_step_deferred(raster, ex, views, sph_degree, particle_density, particle_radiance, upstream, rank=0, finish=True):
    """Same result as ``training_step`` (summed [N,12] and [N,48] gradients on every rank) with the SH expansion deferred.
    Every rank must hold ``ex.views_per_rank`` views or fewer (missing ones contribute zeros).  Works with a
    ``DeferredExchange`` (NCCL all-gather of the colour gradients) and with a ``PeerExchange`` (peer-memory loads).
    ``raster``: one SplatRaster, or a list of them to overlap consecutive views on separate streams (``_lanes``).
    ``finish=False`` leaves ``ex.finish_step(raster, ...)`` to the caller (e.g. with ``expand=False`` before ``optimizer_epilogue``)."""
    outs = []
    ex.begin_step(rank)
    lanes = _lanes(raster)
    ex.use(lanes)
    if len(lanes) == 1 or len(views) < 2:
        r = lanes[0]
        for i, (v, (rg, dg)) in enumerate(zip(views, upstream)):
            rgba, dist_, hits, vis = r.trace(0, sph_degree, particle_density, particle_radiance, v["rays_ori"], v["rays_dir"], None,
                                             v["camera"])
            c = r.trace_bwd_colour(sph_degree, particle_density, particle_radiance, v["rays_ori"], v["rays_dir"], v["camera"],
                                   rgba, rg, dist_, dg, ex.slot(i), out_density_grad=ex.density_out(i))
            ex.add_view(c, vis)
            outs.append((rgba, dist_))
        if finish:
            ex.finish_step(r, sph_degree, particle_density)
        return outs
    cur = torch.cuda.current_stream()
    for r in lanes:
        r._lane_stream.wait_stream(cur)      # the step's inputs are ready
    seen = []
    for i, (v, (rg, dg)) in enumerate(zip(views, upstream)):
        r = lanes[i % len(lanes)]
        with torch.cuda.stream(r._lane_stream):
            rgba, dist_, hits, vis = r.trace(0, sph_degree, particle_density, particle_radiance, v["rays_ori"], v["rays_dir"], None,
                                             v["camera"])
            c = r.trace_bwd_colour(sph_degree, particle_density, particle_radiance, v["rays_ori"], v["rays_dir"], v["camera"],
                                   rgba, rg, dist_, dg, ex.slot(i))
            for t in (rgba, dist_, vis):
                t.record_stream(cur)
        seen.append((c, vis))
        outs.append((rgba, dist_))
    for r in lanes:
        cur.wait_stream(r._lane_stream)
    for c, vis in seen:
        ex.add_view(c, vis)
    if finish:
        ex.finish_step(lanes, sph_degree, particle_density)
    return outs


class _DeviceArray:
    """A raw device allocation as something torch.as_tensor can wrap without copying (__cuda_array_interface__)."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class PeerExchange(DeferredExchange):
    """DeferredExchange whose per-view colour gradients live in device buffers every rank of the node has mapped (CUDA IPC over
    NVLink / NVSwitch): the SH expansion reads the other ranks' slices directly (gut_sh_gradient_expand_gather) and the
    all-gather pass disappears.  Two buffer sets alternate between steps, so a rank may write step s+1 while a slower peer
    still reads step s; the all-gather of the view centres (a few bytes) is the step's cross-rank ordering point.

    One process per GPU, all on one node; ``raster`` provides the native handle.  Call ``close()`` before the process group
    is destroyed."""

    SETS = 2

    def __init__(self, raster, n_particles, n_sph_floats, views_per_rank, device, rank, world=None, track_visibility=False, compact=None):
        super().__init__(n_particles, n_sph_floats, views_per_rank, device, world=world, track_visibility=track_visibility,
                         gather_buffer=False)   # peer buffers instead
        self.raster, self.rank = raster, rank
        # compact slots (include/gut_b200.h, GUT_COLOUR_COMPACT): only the rows a view touched cross NVLink, at the price of a
        # second, dependent load in the expansion.  Off by default: measured equal to the dense layout at 1 and 2 GPUs (3M
        # particles, 8 views), not yet measured where the expansion is bound by NVLink ingress (8 GPUs).  Every rank must choose
        # the same.
        self.compact = bool(compact)
        self.colour_layout = 1 if self.compact else 0
        self.slot_floats = raster.colour_slot_floats(n_particles, self.colour_layout)       # 16-byte aligned slots
        self.slot_header_floats = self.slot_floats - (3 * n_particles + 3) // 4 * 4          # compact: row cursor + table
        floats = self.SETS * views_per_rank * self.slot_floats
        # Set-up either succeeds on every rank or raises on every rank: a rank that cannot export or map a buffer still
        # takes part in both exchanges below, so nobody is left waiting in a collective.
        self.my_ptr, handle, problem = None, None, None
        try:
            self.my_ptr, handle = raster.peer_buffer_create(floats * 4)
        except Exception as e:   # noqa: BLE001 -- reported to all ranks below
            problem = f"rank {rank}: {e}"
        handles = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(handles, handle)
        else:
            handles[0] = handle
        self.ptrs = [None] * self.world
        if problem is None and all(h is not None for h in handles):
            try:
                for r in range(self.world):
                    self.ptrs[r] = self.my_ptr if r == rank else raster.peer_buffer_open(handles[r])
            except Exception as e:   # noqa: BLE001
                problem = f"rank {rank}: {e}"
        elif problem is None:
            problem = f"rank {rank}: a peer could not export its buffer"
        problems = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(problems, problem)
        else:
            problems[0] = problem
        if any(problems):
            for r, q in enumerate(self.ptrs):
                if q is not None and r != rank:
                    raster.peer_buffer_close(q)
            if self.my_ptr is not None:
                raster.peer_buffer_destroy(self.my_ptr)
            self.ptrs = []
            raise RuntimeError("peer-memory exchange unavailable: " + "; ".join(q for q in problems if q))
        self.mine = torch.as_tensor(_DeviceArray(self.my_ptr, floats), device=device).view(self.SETS, views_per_rank, self.slot_floats)
        self.n_particles = n_particles
        self.my_centres = torch.zeros((views_per_rank, 3), dtype=torch.float32, device=device)
        self.step = 0
        self.profile, self._marks = None, []
        if self.world > 1:
            dist.barrier()       # every rank has mapped every buffer before anyone writes

    def slot_tensor(self, s, i):
        """This rank's slot i of buffer set s: an [N,3] view, or the whole slot (flat) in the compact layout."""
        if self.compact:
            return self.mine[s, i]
        return self.mine[s, i, :3 * self.n_particles].view(self.n_particles, 3)

    def clear_slot(self, i, s=None):
        s = self._set if s is None else s
        if self.compact:
            self.mine[s, i, :self.slot_header_floats].zero_()   # no row in use, every mask empty
        else:
            self.slot_tensor(s, i).zero_()

    def moved_rows(self, i):
        if self.compact:
            return int(self.mine[self._set, i, :1].view(torch.int32)[0])
        return super().moved_rows(i)

    def slot_ptr(self, r, s, i):
        return self.ptrs[r] + 4 * ((s * self.views_per_rank + i) * self.slot_floats)

    def bytes_per_rank(self):
        return self.density_grad.numel() * 4 + self.views_per_rank * self.n_particles * 12

    def begin_step(self, rank=None):
        super().begin_step(self.rank if rank is None else rank)
        self._set = self.step % self.SETS
        self.step += 1

    def slot(self, i):
        return self.slot_tensor(self._set, i)

    def _mark(self, name):
        """Optional phase timing of finish_step (``self.profile = {}`` switches it on): CUDA events on the compute stream."""
        if self.profile is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._marks.append((name, ev))

    def phase_times(self):
        """Mean milliseconds per phase of finish_step since the last call (needs ``profile``); synchronises."""
        torch.cuda.synchronize()
        self._fold_marks()
        out = {k: sum(v) / len(v) for k, v in (self.profile or {}).items() if v}
        if self.profile is not None:
            self.profile.clear()
        return out

    def _fold_marks(self):
        if self.profile is not None:
            for (_n0, e0), (n1, e1) in zip(self._marks[:-1], self._marks[1:]):
                self.profile.setdefault(n1, []).append(e0.elapsed_time(e1))
        self._marks = []

    def colour_ptrs(self):
        return [self.slot_ptr(r, self._set, i) for r in range(self.world) for i in range(self.views_per_rank)]

    def finish_step(self, raster, sph_degree, particle_density, radiance_out=None, expand=True):
        radiance = self.radiance_grad if radiance_out is None else radiance_out
        s = self._set
        if self.profile is not None and self._marks:     # fold the previous step's events into the sums
            torch.cuda.synchronize()
            self._fold_marks()
        self._mark("start")
        centres = list(self._centres)
        for i in range(len(centres), self.views_per_rank):
            self.clear_slot(i, s)
            centres.append([0.0, 0.0, 0.0])
        self.my_centres.copy_(torch.tensor(centres, dtype=torch.float32), non_blocking=True)
        gathered, reduced = None, None
        if self.world > 1:
            # ordering point of the step: this all-gather completes on a rank only after every rank's colour gradients are written
            gathered = dist.all_gather_into_tensor(self.centres_all.view(-1), self.my_centres.view(-1), async_op=True)
        else:
            self.centres_all.copy_(self.my_centres)
        rasters = list(raster) if isinstance(raster, (list, tuple)) else [raster]
        raster = rasters[0]
        self.use(raster)
        if self.views_per_rank > 1:
            for j, r in enumerate(rasters):      # every lane's view-summed accumulators; overlaps the ordering point
                r.density_gradient_finish(particle_density, self.density_grad, accumulate=j > 0)
        elif not self._centres:
            self.density_grad.zero_()   # one view per rank and none for this rank in this step
        self._mark("density_finish")
        if self.world > 1:
            reduced = [dist.all_reduce(self.density_grad, op=dist.ReduceOp.SUM, async_op=True)]
            if self.track_visibility:   # union of the ranks' visibilities, 4 B/particle
                reduced.append(dist.all_reduce(self.visibility, op=dist.ReduceOp.MAX, async_op=True))
            gathered.wait()
        self._mark("wait_for_all_ranks")
        if expand:
            raster.sh_gradient_expand_gather(sph_degree, particle_density, self.centres_all, self.colour_ptrs(), radiance, accumulate=False)
        self._mark("sh_expand_gather")
        if reduced is not None:
            for work in reduced:
                work.wait()
        self._mark("allreduce_remainder")
        return self.density_grad, (radiance if expand else None)

    def close(self):
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()       # nobody still reads a buffer that is about to be unmapped
        for r, p in enumerate(self.ptrs):
            if r != self.rank:
                self.raster.peer_buffer_close(p)
        self.mine = None
        self.raster.peer_buffer_destroy(self.my_ptr)
        self.ptrs = []


def training_step_peer(raster, ex, views, sph_degree, particle_density, particle_radiance, upstream, rank=0):
    """``training_step_deferred`` with the colour-gradient all-gather folded into the expansion kernel (peer-memory loads):
    the same loop over a ``PeerExchange``.  Same summed [N,12] and [N,48] gradients on every rank."""
    return training_step_deferred(raster, ex, views, sph_degree, particle_density, particle_radiance, upstream, rank=rank)


PARAMETER_GROUPS = ("positions", "rotation", "scale", "density", "features_albedo", "features_specular")


def optimizer_step(raster, ex, raw, particle_density, state, lrs, betas=(0.9, 0.999), eps=1e-15, scale_activation=1,
                   density_activation=2):
    """The step after the exchange, identical on every rank: the summed [N,12] / [N,48] gradients of
    ``ex`` are taken back to the model's RAW parameters (activation adjoints, gut_activate_pack_backward; feature split)
    and SelectiveAdam updates all six groups in ONE launch, masked by the step's visibility union.

    raw / state / lrs: dicts keyed by PARAMETER_GROUPS -- raw parameter tensors (updated in place), (exp_avg, exp_avg_sq)
    pairs, learning rates.  ``particle_density`` is the packed record the step rendered with (gut_activate_pack).
    Returns the raw-parameter gradients (dict).

    The exchange must have been built with ``track_visibility=True`` (its ``visibility`` buffer is what masks the update;
    an untracked buffer is all zeros and SelectiveAdam would silently update nothing): anything else raises."""
    if not getattr(ex, "track_visibility", False):
        raise ValueError("optimizer_step needs an exchange built with track_visibility=True: SelectiveAdam is masked by the "
                         "visibility union of the step's views, which this exchange did not record")
    pos_g, rot_g, scl_g, dns_g = raster.activate_pack_backward(raw["rotation"], particle_density, ex.density_grad, scale_activation,
                                                               density_activation)
    alb_g, spc_g = raster.unpack_feature_gradient(ex.radiance_grad)
    grads = dict(zip(PARAMETER_GROUPS, (pos_g, rot_g, scl_g, dns_g, alb_g, spc_g)))
    groups = [(raw[k], grads[k], state[k][0], state[k][1], lrs[k]) for k in PARAMETER_GROUPS]
    raster.selective_adam_step(groups, ex.visibility, betas[0], betas[1], eps)
    return grads


def optimizer_epilogue(raster, ex, raw, particle_density, state, lrs, sph_degree, betas=(0.9, 0.999), eps=1e-15, scale_activation=1,
                       density_activation=2, sensor_position=None, grad_norm_accum=None, grad_norm_denom=None):
    """``optimizer_step`` as the EPILOGUE of the step's gradient kernels, after ``ex.finish_step(..., expand=False)``:
    two launches, no gradient tensors in between.
      * gut_density_adam: the all-reduced [N,12] gradient -> activation adjoints -> SelectiveAdam on positions / rotation / scale /
        density, and GSStrategy.update_gradient_buffer on the way when ``sensor_position`` and the two statistics buffers are given;
      * gut_sh_expand_adam: the views' colour gradients (local or peer memory) are expanded per particle in registers and applied
        to features_albedo / features_specular -- the [N,48] gradient is never written.
    Same arguments and the same resulting bits as ``optimizer_step`` (+ ``densify_stats``)."""
    if not getattr(ex, "track_visibility", False):
        raise ValueError("optimizer_epilogue needs an exchange built with track_visibility=True (see optimizer_step)")
    n_slots = ex.world * ex.views_per_rank
    if n_slots > 16:
        raise ValueError(f"optimizer_epilogue expands at most 16 views in its one launch, the step holds {n_slots} slots: use optimizer_step")
    triple = lambda k: (raw[k], state[k][0], state[k][1])
    raster.density_adam(particle_density, ex.density_grad, triple("positions"), triple("rotation"), triple("scale"), triple("density"),
                        [lrs[k] for k in ("positions", "rotation", "scale", "density")], ex.visibility, betas[0], betas[1], eps,
                        scale_activation, density_activation, sensor_position, grad_norm_accum, grad_norm_denom)
    ex.use(raster)
    raster.sh_expand_adam(sph_degree, particle_density, ex.centres_all, ex.colour_ptrs(), triple("features_albedo"),
                          triple("features_specular"), (lrs["features_albedo"], lrs["features_specular"]), ex.visibility, betas[0],
                          betas[1], eps)


def _all_gather(out, mine):
    """All-gather of equal slices along dim 0 (``mine`` may be this rank's slice of ``out``: in-place)."""
    try:
        dist.all_gather_into_tensor(out.view(-1), mine.reshape(-1))
    except (RuntimeError, NotImplementedError):   # backend without the flat variant
        parts = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
        dist.all_gather(parts, mine.contiguous())
        out.copy_(torch.cat(parts, 0))
