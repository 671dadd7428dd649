"""``Tracer`` -- the renderer plugin the 3DGRUT model drives, backed by libgut_b200.

Drop-in for ``threedgut_tracer.Tracer`` (reference: threedgut_tracer/tracer.py:158-431):

    Tracer(conf)                                   # conf.render.* as in configs/render/3dgut.yaml
    Tracer.build_acc(gaussians, rebuild=True)      # no-op for 3DGUT
    Tracer.render(gaussians, gpu_batch, train=False, frame_id=0) -> dict
    Tracer.timings                                 # {"forward_render": ms, "backward_render": ms}

called from ``MixtureOfGaussians`` (threedgrut/model/model.py:184-185, 596, 609) and
``Trainer3DGRUT`` (threedgrut/trainer.py:786-789).  ``gaussians`` is duck-typed:
``num_gaussians``, ``n_active_features``, ``positions``, ``get_rotation()``,
``get_scale()``, ``get_density()``, ``get_features()``, ``background(...)``.
"""
import torch

from . import abi
from .camera import camera_from_batch
from .native import SplatRaster


def _get(conf, path, default=None):
    """Read ``a.b.c`` from an OmegaConf DictConfig, a plain dict, or an attribute namespace."""
    cur = conf
    for key in path.split("."):
        if cur is None:
            return default
        if isinstance(cur, dict):
            cur = cur.get(key, None)
        else:
            try:
                cur = getattr(cur, key)
            except (AttributeError, KeyError):
                try:
                    cur = cur[key]
                except Exception:
                    return default
    return default if cur is None else cur


def config_from_conf(conf) -> abi.GutConfig:
    """``conf.render.*`` -> GutConfig.  The reference turns the same keys into -D defines
    (threedgut_tracer/setup_3dgut.py:47-70); missing keys take the reference defaults."""
    d = abi.default_config()
    r = "render."
    d.kernel_degree = int(_get(conf, r + "particle_kernel_degree", d.kernel_degree))
    d.min_kernel_response = float(_get(conf, r + "particle_kernel_min_response", d.min_kernel_response))
    d.min_alpha = float(_get(conf, r + "particle_kernel_min_alpha", d.min_alpha))
    d.max_alpha = float(_get(conf, r + "particle_kernel_max_alpha", d.max_alpha))
    d.min_transmittance = float(_get(conf, r + "min_transmittance", d.min_transmittance))
    d.sph_max_degree = int(_get(conf, r + "particle_radiance_sph_degree", d.sph_max_degree))
    d.enable_hitcounts = int(bool(_get(conf, r + "enable_hitcounts", bool(d.enable_hitcounts))))
    d.enable_kernel_timings = int(bool(_get(conf, r + "enable_kernel_timings", bool(d.enable_kernel_timings))))
    s = r + "splat."
    d.rect_bounding = int(bool(_get(conf, s + "rect_bounding", True)))
    d.tight_opacity_bounding = int(bool(_get(conf, s + "tight_opacity_bounding", True)))
    d.tile_based_culling = int(bool(_get(conf, s + "tile_based_culling", True)))
    d.n_rolling_shutter_iterations = int(_get(conf, s + "n_rolling_shutter_iterations", d.n_rolling_shutter_iterations))
    d.ut_alpha = float(_get(conf, s + "ut_alpha", d.ut_alpha))
    d.ut_beta = float(_get(conf, s + "ut_beta", d.ut_beta))
    d.ut_kappa = float(_get(conf, s + "ut_kappa", d.ut_kappa))
    d.ut_in_image_margin_factor = float(_get(conf, s + "ut_in_image_margin_factor", d.ut_in_image_margin_factor))
    d.ut_require_all_sigma_points_valid = int(bool(_get(conf, s + "ut_require_all_sigma_points_valid", False)))
    d.k_buffer_size = int(_get(conf, s + "k_buffer_size", d.k_buffer_size))
    d.global_z_order = int(bool(_get(conf, s + "global_z_order", True)))
    return d


_KNOWN_ACTIVATIONS = {torch.exp: abi.Activation.EXP, torch.sigmoid: abi.Activation.SIGMOID}


def _raw_parameters(gaussians):
    """(scale_activation, density_activation) when ``gaussians`` exposes the reference model's raw parameters and its
    activations are ones libgut_b200 implements (MixtureOfGaussians: model.py:163-167, 480-485; table utils/misc.py:48-53);
    None otherwise, in which case render() calls get_rotation() / get_scale() / get_density() / get_features() as the
    reference does."""
    for name in ("rotation", "scale", "density", "features_albedo", "features_specular"):
        if not torch.is_tensor(getattr(gaussians, name, None)):
            return None
    if getattr(gaussians, "rotation_activation", None) is not torch.nn.functional.normalize:
        return None
    scale_act = _KNOWN_ACTIVATIONS.get(getattr(gaussians, "scale_activation", None))
    density_act = _KNOWN_ACTIVATIONS.get(getattr(gaussians, "density_activation", None))
    if scale_act is None or density_act is None:
        return None
    n = gaussians.positions.shape[0]
    if gaussians.features_albedo.shape != (n, 3) or gaussians.features_specular.shape != (n, 45):
        return None
    return int(scale_act), int(density_act)


class Tracer:
    class _Autograd(torch.autograd.Function):
        """reference: Tracer._Autograd (tracer.py:159-286)."""

        @staticmethod
        def forward(ctx, tracer_wrapper, frame_id, n_active_features, ray_ori, ray_dir, mog_pos, mog_rot, mog_scl, mog_dns, mog_sph,
                    sensor):
            # [pos, density, rot, scale, 0] -> [N,12]  (tracer.py:176-178's torch.cat, as one packing kernel)
            particle_density = tracer_wrapper.pack_particles(mog_pos, mog_dns, mog_rot, mog_scl)
            particle_radiance = mog_sph.contiguous()
            rgba, dist, hits, vis = tracer_wrapper.trace(frame_id, n_active_features, particle_density, particle_radiance,
                                                         ray_ori, ray_dir, None, sensor)
            ctx.save_for_backward(ray_ori, ray_dir, rgba, dist, particle_density, particle_radiance)
            ctx.frame_id = frame_id
            ctx.n_active_features = n_active_features
            ctx.sensor = sensor
            ctx.tracer_wrapper = tracer_wrapper
            ctx.mark_non_differentiable(hits, vis)
            return rgba, dist, hits, vis

        @staticmethod
        def backward(ctx, rgba_grad, dist_grad, _hits_grad_unused, _vis_grad_unused):
            ray_ori, ray_dir, rgba, dist, particle_density, particle_radiance = ctx.saved_tensors
            if rgba_grad is None:
                rgba_grad = torch.zeros_like(rgba)
            if dist_grad is None:
                dist_grad = torch.zeros_like(dist)
            density_grad, radiance_grad = ctx.tracer_wrapper.trace_bwd(
                ctx.frame_id, ctx.n_active_features, particle_density, particle_radiance, ray_ori, ray_dir, None, ctx.sensor,
                0, 1, None, None, rgba, rgba_grad.contiguous(), dist, dist_grad.contiguous())
            pos_g, dns_g, rot_g, scl_g = ctx.tracer_wrapper.unpack_particle_gradient(density_grad)
            return (None, None, None, None, None, pos_g, rot_g, scl_g, dns_g, radiance_grad, None)

    class _AutogradRaw(torch.autograd.Function):
        """The same call with the model's RAW parameters: the activations of model.py:77-93, the feature
        cat of model.py:74-75 and the [N,12] packing of tracer.py:176-178 are one kernel each way instead of ~10 torch
        kernels and their autograd mirror images.  Gradients come back for the raw nn.Parameters."""

        @staticmethod
        def forward(ctx, tracer_wrapper, frame_id, n_active_features, ray_ori, ray_dir, mog_pos, rot_raw, scl_raw, dns_raw, albedo, specular,
                    sensor, scale_act, density_act):
            particle_density = tracer_wrapper.activate_pack(mog_pos, rot_raw, scl_raw, dns_raw, scale_act, density_act)
            particle_radiance = tracer_wrapper.pack_features(albedo, specular)
            rgba, dist, hits, vis = tracer_wrapper.trace(frame_id, n_active_features, particle_density, particle_radiance,
                                                         ray_ori, ray_dir, None, sensor)
            ctx.save_for_backward(ray_ori, ray_dir, rgba, dist, particle_density, particle_radiance, rot_raw)
            ctx.frame_id = frame_id
            ctx.n_active_features = n_active_features
            ctx.sensor = sensor
            ctx.tracer_wrapper = tracer_wrapper
            ctx.activations = (scale_act, density_act)
            ctx.mark_non_differentiable(hits, vis)
            return rgba, dist, hits, vis

        @staticmethod
        def backward(ctx, rgba_grad, dist_grad, _hits_grad_unused, _vis_grad_unused):
            ray_ori, ray_dir, rgba, dist, particle_density, particle_radiance, rot_raw = ctx.saved_tensors
            if rgba_grad is None:
                rgba_grad = torch.zeros_like(rgba)
            if dist_grad is None:
                dist_grad = torch.zeros_like(dist)
            density_grad, radiance_grad = ctx.tracer_wrapper.trace_bwd(
                ctx.frame_id, ctx.n_active_features, particle_density, particle_radiance, ray_ori, ray_dir, None, ctx.sensor,
                0, 1, None, None, rgba, rgba_grad.contiguous(), dist, dist_grad.contiguous())
            pos_g, rot_g, scl_g, dns_g = ctx.tracer_wrapper.activate_pack_backward(rot_raw, particle_density, density_grad, *ctx.activations)
            alb_g, spc_g = ctx.tracer_wrapper.unpack_feature_gradient(radiance_grad)
            return (None, None, None, None, None, pos_g, rot_g, scl_g, dns_g, alb_g, spc_g, None, None, None)

    class _SplitRGBA(torch.autograd.Function):
        """[H,W,4] -> pred_rgb [1,H,W,3], pred_opacity [1,H,W,1] (tracer.py:332-333 slices + contiguous()).  As plain torch ops the
        adjoint is two zero-filled [H,W,4] images, two strided copies and an add per view; here it is one concatenation."""

        @staticmethod
        def forward(ctx, rgba):
            return rgba[..., :3].unsqueeze(0).contiguous(), rgba[..., 3:].unsqueeze(0).contiguous()

        @staticmethod
        def backward(ctx, rgb_grad, opacity_grad):
            if rgb_grad is None and opacity_grad is None:
                return None
            if rgb_grad is None:
                rgb_grad = opacity_grad.new_zeros(opacity_grad.shape[:-1] + (3,))
            if opacity_grad is None:
                opacity_grad = rgb_grad.new_zeros(rgb_grad.shape[:-1] + (1,))
            return torch.cat([rgb_grad[0], opacity_grad[0]], dim=-1)

    class _AutogradView(torch.autograd.Function):
        """One view of a multi-view step (Tracer.begin_step): the forward renders from the step's packed parameters, the
        backward is gut_backward_colour -- it leaves the view's colour gradient in the step's slot and keeps summing the
        density-part accumulators inside the handle; the parameter gradients are formed once, by TracerStep.finish()."""

        @staticmethod
        def forward(ctx, step, view_index, anchor, ray_ori, ray_dir, sensor):
            w = step.lane(view_index)
            rgba, dist, hits, vis = w.trace(view_index, step.n_active_features, step.particle_density, step.particle_radiance,
                                            ray_ori, ray_dir, None, sensor)
            ctx.save_for_backward(ray_ori, ray_dir, rgba, dist)
            ctx.step, ctx.view_index, ctx.sensor, ctx.vis = step, view_index, sensor, vis
            ctx.mark_non_differentiable(hits, vis)
            return rgba, dist, hits, vis

        @staticmethod
        def backward(ctx, rgba_grad, dist_grad, _hits_grad_unused, _vis_grad_unused):
            ray_ori, ray_dir, rgba, dist = ctx.saved_tensors
            step = ctx.step
            step._view_backward(ctx.view_index, ctx.sensor, ray_ori, ray_dir, rgba, torch.zeros_like(rgba) if rgba_grad is None else rgba_grad,
                                dist, torch.zeros_like(dist) if dist_grad is None else dist_grad, ctx.vis)
            return (None, None, torch.zeros(1, device=rgba.device), None, None, None)

    def __init__(self, conf):
        self.device = "cuda"
        self.conf = conf
        # render.fused_activations (not a reference key; default on): take the model's raw parameters when it exposes them
        self.fused_activations = bool(_get(conf, "render.fused_activations", True))
        self._normals, self._normals_key = None, None
        torch.zeros(1, device=self.device)  # force CUDA context creation, as the reference does (tracer.py:292)
        self.tracer_wrapper = SplatRaster(config_from_conf(conf))
        self._lane_wrappers = [self.tracer_wrapper]

    @property
    def timings(self):
        return self.tracer_wrapper.collect_times()

    def _constant_normals(self, like):
        """The reference returns normalize(ones_like(pred_rgb)) -- the constant 1/sqrt(3) -- every frame (tracer.py:347), five
        full-image torch kernels.  The constant is built once per image shape and handed out as a read-only expanded view."""
        key = (tuple(like.shape), like.device)
        if self._normals_key != key:
            self._normals = torch.nn.functional.normalize(torch.ones((1, 1, 1, like.shape[3]), dtype=like.dtype, device=like.device), dim=3)
            self._normals_key = key
        return self._normals.expand(like.shape)

    def build_acc(self, gaussians, rebuild=True):
        pass  # 3DGUT has no acceleration structure (tracer.py:301-302)

    def _step_lanes(self, n):
        """The step's rasteriser handles ("lanes", 3dgrut_b200.dist._lanes): the plugin's own plus n - 1 more, created on first use."""
        gdist = __import__("importlib").import_module(__package__ + ".dist")
        while len(self._lane_wrappers) < n:
            self._lane_wrappers.append(SplatRaster(config_from_conf(self.conf)))
        return gdist._lanes(self._lane_wrappers[:n])

    def begin_step(self, gaussians, exchange=None, views_per_rank=None, rank=0, lanes=None):
        """Multi-view training step (this build's extension; the reference renders one view per step, trainer.py:143-149):

            step = tracer.begin_step(gaussians, exchange)      # parameters activated + packed ONCE for all views
            for batch in batches:
                out = step.render(batch)                        # same dict as Tracer.render
                step.backward(loss_fn(out))                     # per view, BEFORE the next step.render (= loss.backward() that
                                                                # does not hold the caller's stream back, see TracerStep.backward)
            step.finish()                                       # gradients of all views summed into the parameters' .grad

        The views' backward passes leave 12 B per particle and view (dL/dcolour) plus view-independent accumulators inside the
        rasteriser; finish() turns them into [N,12] / [N,48] once (gut_density_gradient_finish, gut_sh_gradient_expand), runs the
        data-parallel exchange when ``exchange`` spans several ranks (3dgrut_b200.dist.DeferredExchange / PeerExchange), applies
        the activation adjoints and accumulates into ``.grad`` like autograd would.  Without ``exchange`` a single-rank one is
        created for ``views_per_rank`` views (default 8) and reused across steps.
        ``lanes`` (default ``conf.render.step_lanes``, 2): consecutive views go to that many rasteriser handles on CUDA streams of
        their own, so the projection / binning kernels of a view run under the backward of the previous one; 1 = one stream."""
        if lanes is None:
            lanes = int(_get(self.conf, "render.step_lanes", 2))
        return TracerStep(self, gaussians, exchange, views_per_rank, rank, max(1, lanes))

    def render(self, gaussians, gpu_batch, train=False, frame_id=0):
        rays_o = gpu_batch.rays_ori
        rays_d = gpu_batch.rays_dir
        sensor = camera_from_batch(gpu_batch)
        num_gaussians = gaussians.num_gaussians
        # under torch.no_grad() (validation / rendering) no backward can follow: skip the forward's hit-mask recording
        self.tracer_wrapper.set_training(torch.is_grad_enabled())
        raw = _raw_parameters(gaussians) if self.fused_activations else None
        with torch.cuda.nvtx.range(f"model.forward({num_gaussians} gaussians)"):
            if raw is not None:
                scale_act, density_act = raw
                pred_rgba, pred_dist, hits_count, mog_visibility = Tracer._AutogradRaw.apply(
                    self.tracer_wrapper, frame_id, gaussians.n_active_features, rays_o.contiguous(), rays_d.contiguous(),
                    gaussians.positions.contiguous(), gaussians.rotation.contiguous(), gaussians.scale.contiguous(),
                    gaussians.density.contiguous(), gaussians.features_albedo.contiguous(), gaussians.features_specular.contiguous(),
                    sensor, scale_act, density_act)
            else:
                pred_rgba, pred_dist, hits_count, mog_visibility = Tracer._Autograd.apply(
                    self.tracer_wrapper, frame_id, gaussians.n_active_features, rays_o.contiguous(), rays_d.contiguous(),
                    gaussians.positions.contiguous(), gaussians.get_rotation().contiguous(), gaussians.get_scale().contiguous(),
                    gaussians.get_density().contiguous(), gaussians.get_features().contiguous(), sensor)
            pred_rgb, pred_opacity = Tracer._SplitRGBA.apply(pred_rgba)
            pred_dist = pred_dist.unsqueeze(0).contiguous()
            hits_count = hits_count.unsqueeze(0).contiguous()
            pred_rgb, pred_opacity = gaussians.background(gpu_batch.T_to_world.contiguous(), rays_d, pred_rgb, pred_opacity, train)
            timings = self.tracer_wrapper.collect_times()
        return {
            "pred_rgb": pred_rgb,
            "pred_opacity": pred_opacity,
            "pred_dist": pred_dist,
            "pred_normals": self._constant_normals(pred_rgb),
            "hits_count": hits_count,
            "frame_time_ms": timings["forward_render"] if "forward_render" in timings else 0.0,
            "mog_visibility": mog_visibility,
        }


class TracerStep:
    """One multi-view step in flight (see Tracer.begin_step)."""

    def __init__(self, tracer, gaussians, exchange, views_per_rank, rank, lanes=1):
        from . import dist as gdist

        self.tracer, self.gaussians, self.rank = tracer, gaussians, rank
        w = tracer.tracer_wrapper
        w.set_training(True)
        self._n_lanes = lanes
        self.n_active_features = gaussians.n_active_features
        self.raw = _raw_parameters(gaussians) if tracer.fused_activations else None
        if self.raw is not None:
            self.leaves = (gaussians.positions, gaussians.rotation, gaussians.scale, gaussians.density, gaussians.features_albedo,
                           gaussians.features_specular)
        else:
            # activated tensors, possibly non-leaf (the model's torch activations, evaluated WITH autograd recording): finish()
            # hands them their gradients and autograd carries on to the raw parameters
            self.leaves = (gaussians.positions, gaussians.get_rotation(), gaussians.get_scale(), gaussians.get_density(),
                           gaussians.get_features())
        with torch.no_grad():
            if self.raw is not None:
                self.particle_density = w.activate_pack(gaussians.positions, gaussians.rotation, gaussians.scale, gaussians.density, *self.raw)
                self.particle_radiance = w.pack_features(gaussians.features_albedo, gaussians.features_specular)
            else:
                pos, rot, scl, dns, feat = self.leaves
                self.particle_density = w.pack_particles(pos.contiguous(), dns.contiguous(), rot.contiguous(), scl.contiguous())
                self.particle_radiance = feat.detach().contiguous()
        n, n_sph = self.particle_density.shape[0], self.particle_radiance.shape[1]
        if exchange is None:
            vpr = views_per_rank or 8
            cached = getattr(tracer, "_step_exchange", None)
            if cached is None or cached.density_grad.shape[0] != n or cached.views_per_rank < vpr or cached.radiance_grad.shape[1] != n_sph:
                cached = gdist.DeferredExchange(n, n_sph, vpr, self.particle_density.device, world=1, track_visibility=True)
                tracer._step_exchange = cached
            exchange = cached
        self.exchange = exchange
        # one view per rank has nothing to overlap (and its backward writes [N,12] directly)
        self.lanes = tracer._step_lanes(self._n_lanes if exchange.views_per_rank > 1 else 1)
        for r in self.lanes[1:]:
            r.set_training(True)
        self.anchor = torch.zeros(1, device=self.particle_density.device, requires_grad=True)
        self._seen = []           # (view centre, visibility) of the views whose backward has run, in order
        self._anchors = {}        # per lane: the leaf that makes autograd call the view's backward
        self.n_views = 0          # forwards issued
        self.n_backward = 0       # backwards done
        self._open = None         # view whose forward context the rasteriser currently holds
        self.finished = False
        for r in self.lanes:
            r.reset_gradients()   # nothing pending from an abandoned step
        if hasattr(exchange, "use"):
            exchange.use(self.lanes)   # slot layout of the exchange (dense [N,3] or compact, 3dgrut_b200.dist.PeerExchange)
        exchange.begin_step(rank)

    def lane(self, view_index):
        return self.lanes[view_index % len(self.lanes)]

    def render(self, gpu_batch, train=True):
        if self.finished:
            raise RuntimeError("TracerStep.render after finish()")
        if self.n_views >= self.exchange.views_per_rank:
            raise RuntimeError(f"the step's exchange holds {self.exchange.views_per_rank} views per rank")
        if self._open is not None:
            raise RuntimeError("TracerStep.render: the previous view's backward has not run; the rasteriser holds ONE forward "
                               "context (as the reference, gutRenderer.cu:413-417), so call backward() per view before the next render")
        t = self.tracer
        sensor = camera_from_batch(gpu_batch)
        i = self.n_views
        self.n_views += 1
        self._open = i
        ray_ori, ray_dir = gpu_batch.rays_ori.contiguous(), gpu_batch.rays_dir.contiguous()
        if len(self.lanes) > 1:
            # the view runs on its lane's stream: it starts once the caller's stream has produced its inputs and the caller's
            # stream continues once the view's forward is done.  Autograd runs the view's backward on the lane's stream too, so
            # the NEXT view's forward (other lane) is not queued behind it.
            cur, side = torch.cuda.current_stream(), self.lane(i)._lane_stream
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                lane_id = i % len(self.lanes)
                if lane_id not in self._anchors:   # a leaf of the lane's own stream: its gradient accumulation runs there too
                    self._anchors[lane_id] = torch.zeros(1, device=self.particle_density.device, requires_grad=True)
                rgba, dist, hits, vis = Tracer._AutogradView.apply(self, i, self._anchors[lane_id], ray_ori, ray_dir, sensor)
            cur.wait_stream(side)
            for t_ in (rgba, dist, hits, vis):
                t_.record_stream(cur)
        else:
            rgba, dist, hits, vis = Tracer._AutogradView.apply(self, i, self.anchor, ray_ori, ray_dir, sensor)
        pred_rgb, pred_opacity = Tracer._SplitRGBA.apply(rgba)
        pred_rgb, pred_opacity = self.gaussians.background(gpu_batch.T_to_world.contiguous(), gpu_batch.rays_dir, pred_rgb, pred_opacity, train)
        timings = t.tracer_wrapper.collect_times()
        return {"pred_rgb": pred_rgb, "pred_opacity": pred_opacity, "pred_dist": dist.unsqueeze(0).contiguous(),
                "pred_normals": t._constant_normals(pred_rgb), "hits_count": hits.unsqueeze(0).contiguous(),
                "frame_time_ms": timings["forward_render"] if "forward_render" in timings else 0.0, "mog_visibility": vis}

    def backward(self, loss, **kwargs):
        """``loss.backward()`` for the view rendered last, with the view's lane as the ambient stream.  torch makes the stream
        that is current when ``backward()`` is called wait for every stream its graph ran on; called on the caller's stream
        that wait would put the next view's forward behind this view's backward kernels and undo the lanes' overlap.  With one
        lane it is plain ``loss.backward()``.
        Returns a CUDA event recorded behind the view's backward kernels: the view's input tensors (rays, upstream gradients
        the caller built) may be overwritten once it has fired -- the caller's stream alone no longer tells."""
        done = torch.cuda.Event()
        if len(self.lanes) > 1 and self.n_views > 0:
            side = self.lane(self.n_views - 1)._lane_stream
            with torch.cuda.stream(side):
                loss.backward(**kwargs)
                done.record(side)
        else:
            loss.backward(**kwargs)
            done.record(torch.cuda.current_stream())
        return done

    def _view_backward(self, i, sensor, ray_ori, ray_dir, rgba, rgba_grad, dist, dist_grad, vis):
        if self._open != i:
            raise RuntimeError(f"backward of view {i} after another view was rendered: its forward context is gone")
        w = self.lane(i)
        if len(self.lanes) > 1:
            # the kernels below run on the lane's stream; tensors that were allocated on another one (the upstream gradients
            # autograd built on the caller's stream, the rays) must not return to that stream's pool while they are read
            side = w._lane_stream
            rgba_grad, dist_grad = rgba_grad.contiguous(), dist_grad.contiguous()
            for t_ in (ray_ori, ray_dir, rgba, dist, rgba_grad, dist_grad):
                t_.record_stream(side)
        c = w.trace_bwd_colour(self.n_active_features, self.particle_density, self.particle_radiance, ray_ori, ray_dir, sensor,
                               rgba, rgba_grad.contiguous(), dist, dist_grad.contiguous(), self.exchange.slot(i),
                               out_density_grad=self.exchange.density_out(i))
        self._seen.append((c, vis))      # handed to the exchange by finish(): the lanes' streams must not race on its buffers
        self.n_backward += 1
        self._open = None

    def finish(self):
        """Sum the views' gradients into the parameters' ``.grad`` (all ranks of the exchange must call it).  Returns the
        step's visibility union [N,1] (what SelectiveAdam masks with) when the exchange tracks it."""
        if self.finished:
            raise RuntimeError("TracerStep.finish called twice")
        if self._open is not None:     # a view rendered without a backward contributes nothing: drop its slot
            self.exchange.clear_slot(self._open)
            direct = self.exchange.density_out(self._open)
            if direct is not None:     # one view per rank: no backward wrote the step's [N,12] gradient
                direct.zero_()
            self._seen.append(([0.0, 0.0, 0.0], None))
            self._open = None
        self.finished = True
        w = self.tracer.tracer_wrapper
        if len(self.lanes) > 1:
            cur = torch.cuda.current_stream()
            for r in self.lanes:
                cur.wait_stream(r._lane_stream)      # every view's backward has been queued on its lane: wait for them here
        for c, vis in self._seen:
            self.exchange.add_view(c, vis)
        # activated leaves take the [N,48] gradient as it is: the expansion writes into a fresh tensor that becomes the
        # parameter's .grad (the exchange's own buffer is reused by the next step, a copy out of it was 576 MB per step)
        fresh = None
        if self.raw is None and self.leaves[4].requires_grad:
            fresh = torch.empty_like(self.particle_radiance)
        g12, g48 = self.exchange.finish_step(self.lanes, self.n_active_features, self.particle_density, radiance_out=fresh)
        if self.raw is not None:
            pos, rot, scl, dns, alb, spc = self.leaves
            pos_g, rot_g, scl_g, dns_g = w.activate_pack_backward(rot, self.particle_density, g12, *self.raw)
            alb_g, spc_g = w.unpack_feature_gradient(g48)
            pairs = zip(self.leaves, (pos_g, rot_g, scl_g, dns_g, alb_g, spc_g))
        else:
            pos_g, dns_g, rot_g, scl_g = w.unpack_particle_gradient(g12)
            pairs = zip(self.leaves, (pos_g, rot_g, scl_g, dns_g, g48 if fresh is not None else g48.clone()))
        tensors, grads = [], []
        for t, g in pairs:
            if not t.requires_grad:
                continue
            g = g.reshape(t.shape)
            if t.is_leaf:      # what AccumulateGrad does, without its defensive copy of a gradient it cannot prove unshared
                if t.grad is None:
                    t.grad = g
                else:
                    t.grad.add_(g)
            else:
                tensors.append(t)
                grads.append(g)
        if tensors:
            torch.autograd.backward(tensors, grads)
        return self.exchange.visibility if getattr(self.exchange, "track_visibility", False) else None
