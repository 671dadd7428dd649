"""ctypes binding of libgut_b200.so and the ``SplatRaster`` object the Tracer drives.

Mirrors the reference's pybind11 class ``lib3dgut_cc.SplatRaster``
(threedgut_tracer/bindings.cpp:79-86, threedgut_tracer/src/splatRaster.cpp): same
method names (``trace``, ``trace_bwd``, ``collect_times``), same argument meaning,
torch tensors in and out -- but underneath it only passes raw device pointers and
the current CUDA stream through the C ABI of ``include/gut_b200.h``.

There is no CPU fallback: if the library is missing or was not built for the
device, loading raises.
"""
import ctypes as C
import os
import subprocess

import torch

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# GUT_B200_LIBRARY: developer hook to A/B a differently tuned build of the same library (scripts/ab_build.sh)
LIB_PATH = os.environ.get("GUT_B200_LIBRARY") or os.path.join(_HERE, "libgut_b200.so")
_lib = None


class GutError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libgut_b200: {abi.GutStatus(code).name}: {message}")
        self.code = code


def build_library(force=False):
    """Compile the CUDA library in-tree for sm_100a (``make -C csrc``)."""
    src_dir = os.path.join(_HERE, "csrc")
    args = ["make", "-C", src_dir, "-s"]
    if force:
        subprocess.check_call(["make", "-C", src_dir, "-s", "clean"])
    subprocess.check_call(args)
    return LIB_PATH


def load_library():
    """dlopen the C-ABI library and declare its prototypes.  Raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C 3dgrut_b200/csrc`).  There is no CPU fallback for the 3DGUT rasteriser."
        )
    lib = C.CDLL(LIB_PATH)
    vp, u32, i32, fp = C.c_void_p, C.c_uint32, C.c_int32, C.c_void_p
    lib.gut_abi_version.restype = C.c_int
    lib.gut_default_config.argtypes = [C.POINTER(abi.GutConfig)]
    lib.gut_default_config.restype = None
    lib.gut_create.argtypes = [C.POINTER(abi.GutConfig), C.c_int, C.POINTER(vp)]
    lib.gut_create.restype = C.c_int
    lib.gut_destroy.argtypes = [vp]
    lib.gut_destroy.restype = None
    lib.gut_forward.argtypes = [vp, vp, C.POINTER(abi.GutCamera), u32, i32, fp, fp, u32, u32, fp, fp, fp, fp, fp, fp,
                                C.POINTER(abi.GutStageOutputs)]
    lib.gut_forward.restype = C.c_int
    lib.gut_backward.argtypes = [vp, vp, C.POINTER(abi.GutCamera), u32, i32, fp, fp, u32, u32, fp, fp, fp, fp, fp, fp, fp, fp, i32]
    lib.gut_backward.restype = C.c_int
    lib.gut_backward_colour.argtypes = [vp, vp, C.POINTER(abi.GutCamera), u32, i32, fp, fp, u32, u32, fp, fp, fp, fp, fp, fp, fp,
                                        C.POINTER(C.c_float)]
    lib.gut_backward_colour.restype = C.c_int
    lib.gut_backward_colour_density.argtypes = [vp, vp, C.POINTER(abi.GutCamera), u32, i32, fp, fp, u32, u32, fp, fp, fp, fp, fp, fp, fp, fp,
                                                C.POINTER(C.c_float), i32]
    lib.gut_backward_colour_density.restype = C.c_int
    lib.gut_set_training.argtypes = [vp, i32]
    lib.gut_set_training.restype = C.c_int
    lib.gut_set_colour_layout.argtypes = [vp, i32]
    lib.gut_set_colour_layout.restype = C.c_int
    lib.gut_colour_slot_floats.argtypes = [u32, i32]
    lib.gut_colour_slot_floats.restype = C.c_uint64
    lib.gut_set_block_cull.argtypes = [vp, i32]
    lib.gut_set_block_cull.restype = C.c_int
    lib.gut_reset_gradients.argtypes = [vp, vp]
    lib.gut_reset_gradients.restype = C.c_int
    lib.gut_density_gradient_finish.argtypes = [vp, vp, u32, fp, fp, i32]
    lib.gut_density_gradient_finish.restype = C.c_int
    lib.gut_sh_gradient_expand.argtypes = [vp, vp, u32, i32, fp, u32, fp, fp, fp, i32]
    lib.gut_sh_gradient_expand.restype = C.c_int
    lib.gut_sh_gradient_expand_gather.argtypes = [vp, vp, u32, i32, fp, u32, fp, C.POINTER(C.c_void_p), fp, i32]
    lib.gut_sh_gradient_expand_gather.restype = C.c_int
    lib.gut_peer_buffer_create.argtypes = [vp, C.c_size_t, C.POINTER(vp), C.c_char_p]
    lib.gut_peer_buffer_create.restype = C.c_int
    lib.gut_peer_buffer_destroy.argtypes = [vp, vp]
    lib.gut_peer_buffer_destroy.restype = C.c_int
    lib.gut_peer_buffer_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    lib.gut_peer_buffer_open.restype = C.c_int
    lib.gut_peer_buffer_close.argtypes = [vp, vp]
    lib.gut_peer_buffer_close.restype = C.c_int
    lib.gut_pack_particles.argtypes = [vp, vp, u32, fp, fp, fp, fp, fp]
    lib.gut_pack_particles.restype = C.c_int
    lib.gut_unpack_particle_gradient.argtypes = [vp, vp, u32, fp, fp, fp, fp, fp]
    lib.gut_unpack_particle_gradient.restype = C.c_int
    lib.gut_activate_pack.argtypes = [vp, vp, u32, fp, fp, fp, fp, i32, i32, fp]
    lib.gut_activate_pack.restype = C.c_int
    lib.gut_activate_pack_backward.argtypes = [vp, vp, u32, fp, fp, fp, i32, i32, fp, fp, fp, fp]
    lib.gut_activate_pack_backward.restype = C.c_int
    lib.gut_pack_features.argtypes = [vp, vp, u32, fp, fp, fp]
    lib.gut_pack_features.restype = C.c_int
    lib.gut_unpack_feature_gradient.argtypes = [vp, vp, u32, fp, fp, fp]
    lib.gut_unpack_feature_gradient.restype = C.c_int
    lib.gut_selective_adam_step.argtypes = [vp, vp, u32, C.POINTER(abi.GutAdamGroup), u32, fp, C.c_float, C.c_float, C.c_float, i32]
    lib.gut_selective_adam_step.restype = C.c_int
    at = C.POINTER(abi.GutAdamTensors)
    lib.gut_density_adam.argtypes = [vp, vp, u32, fp, fp, at, at, at, at, i32, i32, fp, C.c_float, C.c_float, C.c_float,
                                     C.POINTER(C.c_float), fp, fp]
    lib.gut_density_adam.restype = C.c_int
    lib.gut_sh_expand_adam.argtypes = [vp, vp, u32, i32, fp, u32, fp, C.POINTER(C.c_void_p), at, at, fp, C.c_float, C.c_float, C.c_float]
    lib.gut_sh_expand_adam.restype = C.c_int
    lib.gut_densify_stats.argtypes = [vp, vp, u32, fp, fp, C.POINTER(C.c_float), fp, fp]
    lib.gut_densify_stats.restype = C.c_int
    lib.gut_forward_host.argtypes = [vp, vp, C.POINTER(abi.GutCamera), u32, i32, fp, fp, u32, u32, fp, fp, fp, fp, fp, fp]
    lib.gut_forward_host.restype = C.c_int
    lib.gut_backward_host.argtypes = [vp, vp, C.POINTER(abi.GutCamera), u32, i32, fp, fp, u32, u32, fp, fp, fp, fp, i32]
    lib.gut_backward_host.restype = C.c_int
    lib.gut_timings.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.gut_timings.restype = C.c_int
    lib.gut_num_stages.restype = C.c_int
    lib.gut_stage_name.argtypes = [C.c_int]
    lib.gut_stage_name.restype = C.c_char_p
    lib.gut_stage_timings.argtypes = [vp, C.POINTER(C.c_float)]
    lib.gut_stage_timings.restype = C.c_int
    lib.gut_frame_stats.argtypes = [vp, C.POINTER(abi.GutFrameStats)]
    lib.gut_frame_stats.restype = C.c_int
    lib.gut_measure_work.argtypes = [vp, vp, fp, fp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.gut_measure_work.restype = C.c_int
    lib.gut_project_noalloc.argtypes = [vp, vp, C.POINTER(abi.GutCamera), u32, i32, fp, fp, u32, u32, fp, C.POINTER(abi.GutStageOutputs)]
    lib.gut_project_noalloc.restype = C.c_int
    lib.gut_last_error.argtypes = [vp]
    lib.gut_last_error.restype = C.c_char_p
    for name in ("gut_sizeof_config", "gut_sizeof_camera", "gut_sizeof_stage_outputs", "gut_sizeof_frame_stats",
                 "gut_sizeof_adam_group", "gut_sizeof_adam_tensors"):
        getattr(lib, name).restype = C.c_size_t
    if lib.gut_abi_version() != abi.GUT_ABI_VERSION:
        raise ImportError(f"libgut_b200 ABI {lib.gut_abi_version()} != python mirror {abi.GUT_ABI_VERSION}")
    if lib.gut_sizeof_config() != C.sizeof(abi.GutConfig) or lib.gut_sizeof_camera() != C.sizeof(abi.GutCamera):
        raise ImportError("libgut_b200 struct layout differs from 3dgrut_b200/abi.py")
    if lib.gut_sizeof_adam_group() != C.sizeof(abi.GutAdamGroup):
        raise ImportError("libgut_b200 GutAdamGroup layout differs from 3dgrut_b200/abi.py")
    if lib.gut_sizeof_adam_tensors() != C.sizeof(abi.GutAdamTensors):
        raise ImportError("libgut_b200 GutAdamTensors layout differs from 3dgrut_b200/abi.py")
    _lib = lib
    return lib


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _check_cuda_f32(name, t, device):
    if not t.is_cuda or t.dtype != torch.float32:
        raise GutError(abi.GutStatus.BAD_INPUT, f"{name} must be a float32 CUDA tensor (got {t.dtype} on {t.device})")
    if t.device != device:
        raise GutError(abi.GutStatus.BAD_INPUT, f"{name} is on {t.device}, expected {device}")
    return t.contiguous()


class SplatRaster:
    """One rasteriser instance = one native handle + its grow-only device workspace.

    reference: class SplatRaster (include/3dgut/splatRaster.h:27-90).
    """

    def __init__(self, config: abi.GutConfig, device=None):
        self.lib = load_library()
        if not torch.cuda.is_available():
            raise RuntimeError("3dgrut_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError(f"3dgrut_b200 runs on CUDA devices only (got {self.device})")
        if self.device.index is None:   # "cuda" without an index: the current device
            self.device = torch.device("cuda", torch.cuda.current_device())
        major, _minor = torch.cuda.get_device_capability(self.device)
        if major != 10:
            raise RuntimeError(f"libgut_b200 is built for sm_100a only; device capability is {major}.{_minor}")
        self.config = config
        h = C.c_void_p()
        rc = self.lib.gut_create(C.byref(config), self.device.index, C.byref(h))
        if rc != 0:
            raise GutError(rc, "gut_create failed (unsupported configuration?)")
        self._h = h
        self.colour_layout = abi.GUT_COLOUR_DENSE
        self.n_coeffs = (config.sph_max_degree + 1) ** 2

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                self.lib.gut_destroy(h)
            except Exception:
                pass
            self._h = None

    def _raise(self, rc):
        raise GutError(rc, self.lib.gut_last_error(self._h).decode())

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ---- reference-shaped entry points ------------------------------------------------------------------
    def trace(self, frame_number, num_active_features, particle_density, particle_radiance, ray_origin, ray_direction,
              ray_timestamp, sensor: abi.GutCamera, start_timestamp=0, end_timestamp=1, start_pose=None, end_pose=None,
              stage_outputs=False):
        """reference: SplatRaster::trace (splatRaster.cpp:173-244).

        ``sensor`` already carries the two poses (set by camera.camera_from_batch); ``start_pose``/``end_pose``
        override them when given, for callers that follow the reference argument list literally.
        ``ray_timestamp`` is accepted and ignored, like the reference's kernels ignore it.
        Returns (rayRadianceDensity [H,W,4], rayHitDistance [H,W,1], rayHitCount [H,W,1], particleVisibility [N,1])
        and, when ``stage_outputs`` is set, a dict of the intermediate stage tensors as a fifth element.
        """
        del frame_number, ray_timestamp
        dev = self.device
        if start_pose is not None:
            abi.set_poses(sensor, [float(v) for v in start_pose], [float(v) for v in (end_pose if end_pose is not None else start_pose)],
                          (start_timestamp, end_timestamp))
        dens = _check_cuda_f32("particle_density", particle_density, dev)
        rad = _check_cuda_f32("particle_radiance", particle_radiance, dev)
        ro = _check_cuda_f32("ray_origin", ray_origin, dev)
        rd = _check_cuda_f32("ray_direction", ray_direction, dev)
        if ro.dim() != 4 or ro.shape[0] != 1 or ro.shape[3] != 3 or rd.shape != ro.shape:
            raise GutError(abi.GutStatus.BAD_INPUT, f"rays must be [1,H,W,3], got {tuple(ro.shape)} / {tuple(rd.shape)}")
        n = dens.shape[0]
        if dens.dim() != 2 or dens.shape[1] != 12:
            raise GutError(abi.GutStatus.BAD_INPUT, f"particle_density must be [N,12], got {tuple(dens.shape)}")
        if rad.shape[0] != n or rad.numel() != n * self.n_coeffs * 3:
            raise GutError(abi.GutStatus.BAD_INPUT, f"particle_radiance must be [N,{self.n_coeffs * 3}], got {tuple(rad.shape)}")
        height, width = ro.shape[1], ro.shape[2]
        # the library initialises all four outputs itself (no torch.zeros fill kernels)
        rgba = torch.empty((height, width, 4), dtype=torch.float32, device=dev)
        dist = torch.empty((height, width, 1), dtype=torch.float32, device=dev)
        hits = torch.empty((height, width, 1), dtype=torch.float32, device=dev)
        vis = torch.empty((n, 1), dtype=torch.float32, device=dev)
        so = None
        stage = None
        if stage_outputs:
            so, stage = self._alloc_stage_outputs(n, width, height)
        rc = self.lib.gut_forward(self._h, self._stream(), C.byref(sensor), n, int(num_active_features), _ptr(dens), _ptr(rad),
                                  width, height, _ptr(ro), _ptr(rd), _ptr(rgba), _ptr(dist), _ptr(hits), _ptr(vis),
                                  C.byref(so) if so is not None else None)
        if rc != 0 and stage_outputs and "stage outputs hold" in self.lib.gut_last_error(self._h).decode():
            # grow the debug buffers to the frame's M and run again
            so, stage = self._alloc_stage_outputs(n, width, height, capacity=int(so.num_intersections))
            rc = self.lib.gut_forward(self._h, self._stream(), C.byref(sensor), n, int(num_active_features), _ptr(dens), _ptr(rad),
                                      width, height, _ptr(ro), _ptr(rd), _ptr(rgba), _ptr(dist), _ptr(hits), _ptr(vis), C.byref(so))
        if rc != 0:
            self._raise(rc)
        if stage_outputs:
            m = int(so.num_intersections)
            stage["num_intersections"] = m
            stage["sorted_keys"] = stage["sorted_keys"][:m]
            stage["sorted_particle_idx"] = stage["sorted_particle_idx"][:m]
            return rgba, dist, hits, vis, stage
        return rgba, dist, hits, vis

    def _alloc_stage_outputs(self, n, width, height, capacity=None):
        dev = self.device
        tiles = ((width + 15) // 16) * ((height + 15) // 16)
        cap = capacity if capacity is not None else max(1, 16 * n)
        st = {
            "tiles_count": torch.zeros(n, dtype=torch.int32, device=dev),
            "proj_position": torch.zeros((n, 2), dtype=torch.float32, device=dev),
            "proj_conic_opacity": torch.zeros((n, 4), dtype=torch.float32, device=dev),
            "proj_extent": torch.zeros((n, 2), dtype=torch.float32, device=dev),
            "global_depth": torch.zeros(n, dtype=torch.float32, device=dev),
            "features": torch.zeros((n, 3), dtype=torch.float32, device=dev),
            "sorted_keys": torch.zeros(cap, dtype=torch.int64, device=dev),
            "sorted_particle_idx": torch.zeros(cap, dtype=torch.int32, device=dev),
            "tile_ranges": torch.zeros((tiles, 2), dtype=torch.int32, device=dev),
        }
        so = abi.GutStageOutputs(
            d_tiles_count=st["tiles_count"].data_ptr(), d_proj_position=st["proj_position"].data_ptr(),
            d_proj_conic_opacity=st["proj_conic_opacity"].data_ptr(), d_proj_extent=st["proj_extent"].data_ptr(),
            d_global_depth=st["global_depth"].data_ptr(), d_features=st["features"].data_ptr(),
            d_sorted_keys=st["sorted_keys"].data_ptr(), d_sorted_particle_idx=st["sorted_particle_idx"].data_ptr(),
            d_tile_ranges=st["tile_ranges"].data_ptr(), capacity_intersections=cap, num_intersections=0,
        )
        return so, st

    def trace_bwd(self, frame_number, num_active_features, particle_density, particle_radiance, ray_origin, ray_direction,
                  ray_timestamp, sensor: abi.GutCamera, start_timestamp, end_timestamp, start_pose, end_pose,
                  ray_radiance_density, ray_radiance_density_grad, ray_hit_distance, ray_hit_distance_grad,
                  out_density_grad=None, out_radiance_grad=None, accumulate=False):
        """reference: SplatRaster::traceBwd (splatRaster.cpp:246-332).

        ``out_*`` + ``accumulate`` are this build's extension: gradients are added into caller-owned buffers
        (sum over the views of a rank before the data-parallel allreduce).
        """
        del frame_number, ray_timestamp, start_timestamp, end_timestamp, start_pose, end_pose
        dev = self.device
        dens = _check_cuda_f32("particle_density", particle_density, dev)
        rad = _check_cuda_f32("particle_radiance", particle_radiance, dev)
        ro = _check_cuda_f32("ray_origin", ray_origin, dev)
        rd = _check_cuda_f32("ray_direction", ray_direction, dev)
        rgba = _check_cuda_f32("ray_radiance_density", ray_radiance_density, dev)
        rgba_g = _check_cuda_f32("ray_radiance_density_grad", ray_radiance_density_grad, dev)
        dist = _check_cuda_f32("ray_hit_distance", ray_hit_distance, dev)
        dist_g = _check_cuda_f32("ray_hit_distance_grad", ray_hit_distance_grad, dev)
        n = dens.shape[0]
        height, width = ro.shape[1], ro.shape[2]
        if rgba_g.numel() != height * width * 4 or dist_g.numel() != height * width:
            raise GutError(abi.GutStatus.BAD_INPUT, "gradient shapes do not match the view")
        if out_density_grad is None:
            out_density_grad = torch.empty((n, 12), dtype=torch.float32, device=dev)
            out_radiance_grad = torch.empty((n, self.n_coeffs * 3), dtype=torch.float32, device=dev)
            accumulate = False
        rc = self.lib.gut_backward(self._h, self._stream(), C.byref(sensor), n, int(num_active_features), _ptr(dens), _ptr(rad),
                                   width, height, _ptr(ro), _ptr(rd), _ptr(rgba), _ptr(rgba_g), _ptr(dist), _ptr(dist_g),
                                   _ptr(out_density_grad), _ptr(out_radiance_grad), 1 if accumulate else 0)
        if rc != 0:
            self._raise(rc)
        return out_density_grad, out_radiance_grad

    def trace_bwd_colour(self, num_active_features, particle_density, particle_radiance, ray_origin, ray_direction, sensor: abi.GutCamera,
                         ray_radiance_density, ray_radiance_density_grad, ray_hit_distance, ray_hit_distance_grad, out_colour_grad,
                         out_density_grad=None, accumulate=False):
        """gut_backward_colour: traceBwd of one view of a deferred step.  Writes the view's masked colour gradient into
        ``out_colour_grad`` [N,3], keeps summing the density-part accumulators inside the handle (``density_gradient_finish``
        converts them once per step) and returns the world-space view centre (3 floats) ``sh_gradient_expand`` needs.
        With ``out_density_grad`` [N,12] (gut_backward_colour_density: the handle's only view of the step) the view's [N,12]
        gradient is written at once and nothing is left pending."""
        dev = self.device
        dens = _check_cuda_f32("particle_density", particle_density, dev)
        rad = _check_cuda_f32("particle_radiance", particle_radiance, dev)
        ro = _check_cuda_f32("ray_origin", ray_origin, dev)
        rd = _check_cuda_f32("ray_direction", ray_direction, dev)
        rgba = _check_cuda_f32("ray_radiance_density", ray_radiance_density, dev)
        rgba_g = _check_cuda_f32("ray_radiance_density_grad", ray_radiance_density_grad, dev)
        dist = _check_cuda_f32("ray_hit_distance", ray_hit_distance, dev)
        dist_g = _check_cuda_f32("ray_hit_distance_grad", ray_hit_distance_grad, dev)
        n = dens.shape[0]
        height, width = ro.shape[1], ro.shape[2]
        if rgba_g.numel() != height * width * 4 or dist_g.numel() != height * width:
            raise GutError(abi.GutStatus.BAD_INPUT, "gradient shapes do not match the view")
        if self.colour_layout == abi.GUT_COLOUR_COMPACT:
            if out_colour_grad.numel() != self.colour_slot_floats(n) or not out_colour_grad.is_contiguous() or out_colour_grad.data_ptr() % 16:
                raise GutError(abi.GutStatus.BAD_INPUT, "out_colour_grad must be a 16-byte aligned contiguous slot of "
                                                        "colour_slot_floats(N) floats (compact colour layout)")
        elif out_colour_grad.shape != (n, 3) or not out_colour_grad.is_contiguous():
            raise GutError(abi.GutStatus.BAD_INPUT, "out_colour_grad must be a contiguous [N,3]")
        centre = (C.c_float * 3)()
        if out_density_grad is not None:
            if out_density_grad.shape != (n, 12) or not out_density_grad.is_contiguous():
                raise GutError(abi.GutStatus.BAD_INPUT, "out_density_grad must be a contiguous [N,12]")
            rc = self.lib.gut_backward_colour_density(self._h, self._stream(), C.byref(sensor), n, int(num_active_features), _ptr(dens),
                                                      _ptr(rad), width, height, _ptr(ro), _ptr(rd), _ptr(rgba), _ptr(rgba_g), _ptr(dist),
                                                      _ptr(dist_g), _ptr(out_density_grad), _ptr(out_colour_grad), centre,
                                                      1 if accumulate else 0)
        else:
            rc = self.lib.gut_backward_colour(self._h, self._stream(), C.byref(sensor), n, int(num_active_features), _ptr(dens), _ptr(rad),
                                              width, height, _ptr(ro), _ptr(rd), _ptr(rgba), _ptr(rgba_g), _ptr(dist), _ptr(dist_g),
                                              _ptr(out_colour_grad), centre)
        if rc != 0:
            self._raise(rc)
        return [centre[0], centre[1], centre[2]]

    def density_gradient_finish(self, particle_density, out_density_grad, accumulate=False):
        """gut_density_gradient_finish: out_density_grad [N,12] (+)= the gradient of all views since the last finish."""
        dens = _check_cuda_f32("particle_density", particle_density, self.device)
        if out_density_grad.shape != (dens.shape[0], 12) or not out_density_grad.is_contiguous():
            raise GutError(abi.GutStatus.BAD_INPUT, "out_density_grad must be a contiguous [N,12]")
        rc = self.lib.gut_density_gradient_finish(self._h, self._stream(), dens.shape[0], _ptr(dens), _ptr(out_density_grad),
                                                  1 if accumulate else 0)
        if rc != 0:
            self._raise(rc)
        return out_density_grad

    def sh_gradient_expand(self, num_active_features, particle_density, view_centres, colour_grads, out_radiance_grad, accumulate=False):
        """gut_sh_gradient_expand: out_radiance_grad [N,48] (+)= sum over views of basis(direction) x colour gradient.
        ``colour_grads`` is a contiguous [V,N,3] tensor, ``view_centres`` a [V,3] float32 CUDA tensor."""
        dev = self.device
        dens = _check_cuda_f32("particle_density", particle_density, dev)
        cg = _check_cuda_f32("colour_grads", colour_grads, dev)
        vc = _check_cuda_f32("view_centres", view_centres, dev)
        n = dens.shape[0]
        nv = vc.shape[0]
        if vc.shape != (nv, 3) or cg.shape != (nv, n, 3) or out_radiance_grad.shape != (n, self.n_coeffs * 3) or not out_radiance_grad.is_contiguous():
            raise GutError(abi.GutStatus.BAD_INPUT, "colour_grads must be [V,N,3] and out_radiance_grad a contiguous [N,48]")
        rc = self.lib.gut_sh_gradient_expand(self._h, self._stream(), n, int(num_active_features), _ptr(dens), nv, _ptr(vc), _ptr(cg),
                                             _ptr(out_radiance_grad), 1 if accumulate else 0)
        if rc != 0:
            self._raise(rc)
        return out_radiance_grad

    def sh_gradient_expand_gather(self, num_active_features, particle_density, view_centres, colour_ptrs, out_radiance_grad, accumulate=False):
        """gut_sh_gradient_expand_gather: like sh_gradient_expand, with one device pointer per view (int addresses; local slices or
        buffers of peer GPUs opened with peer_buffer_open).  ``view_centres`` is a [V,3] float32 CUDA tensor."""
        dev = self.device
        dens = _check_cuda_f32("particle_density", particle_density, dev)
        vc = _check_cuda_f32("view_centres", view_centres, dev)
        n, nv = dens.shape[0], len(colour_ptrs)
        if vc.shape != (nv, 3) or out_radiance_grad.shape != (n, self.n_coeffs * 3) or not out_radiance_grad.is_contiguous():
            raise GutError(abi.GutStatus.BAD_INPUT, "view_centres must be [V,3] with one pointer per view, out_radiance_grad a contiguous [N,48]")
        table = (C.c_void_p * max(nv, 1))(*[int(x) for x in colour_ptrs])
        rc = self.lib.gut_sh_gradient_expand_gather(self._h, self._stream(), n, int(num_active_features), _ptr(dens), nv, _ptr(vc), table,
                                                    _ptr(out_radiance_grad), 1 if accumulate else 0)
        if rc != 0:
            self._raise(rc)
        return out_radiance_grad

    def peer_buffer_create(self, nbytes):
        """gut_peer_buffer_create -> (device address, 64-byte CUDA IPC handle) of a zeroed device allocation peers can map."""
        ptr = C.c_void_p()
        handle = C.create_string_buffer(64)
        rc = self.lib.gut_peer_buffer_create(self._h, int(nbytes), C.byref(ptr), handle)
        if rc != 0:
            self._raise(rc)
        return int(ptr.value), handle.raw

    def peer_buffer_open(self, handle):
        ptr = C.c_void_p()
        rc = self.lib.gut_peer_buffer_open(self._h, bytes(handle), C.byref(ptr))
        if rc != 0:
            self._raise(rc)
        return int(ptr.value)

    def peer_buffer_close(self, ptr):
        self.lib.gut_peer_buffer_close(self._h, C.c_void_p(int(ptr)))

    def peer_buffer_destroy(self, ptr):
        self.lib.gut_peer_buffer_destroy(self._h, C.c_void_p(int(ptr)))

    def pack_particles(self, positions, density, rotation, scale):
        """gut_pack_particles: [N,3], [N,1], [N,4], [N,3] -> the [N,12] particle_density record (tracer.py:176-178's torch.cat)."""
        dev = self.device
        pos = _check_cuda_f32("positions", positions, dev)
        dns = _check_cuda_f32("density", density, dev)
        rot = _check_cuda_f32("rotation", rotation, dev)
        scl = _check_cuda_f32("scale", scale, dev)
        n = pos.shape[0]
        if pos.shape != (n, 3) or dns.shape != (n, 1) or rot.shape != (n, 4) or scl.shape != (n, 3):
            raise GutError(abi.GutStatus.BAD_INPUT, "expected positions [N,3], density [N,1], rotation [N,4], scale [N,3]")
        out = torch.empty((n, 12), dtype=torch.float32, device=dev)
        rc = self.lib.gut_pack_particles(self._h, self._stream(), n, _ptr(pos), _ptr(dns), _ptr(rot), _ptr(scl), _ptr(out))
        if rc != 0:
            self._raise(rc)
        return out

    def unpack_particle_gradient(self, density_grad):
        """gut_unpack_particle_gradient: [N,12] gradient -> (positions [N,3], density [N,1], rotation [N,4], scale [N,3]) gradients."""
        dev = self.device
        g = _check_cuda_f32("density_grad", density_grad, dev)
        n = g.shape[0]
        if g.shape != (n, 12):
            raise GutError(abi.GutStatus.BAD_INPUT, "density_grad must be [N,12]")
        pos = torch.empty((n, 3), dtype=torch.float32, device=dev)
        dns = torch.empty((n, 1), dtype=torch.float32, device=dev)
        rot = torch.empty((n, 4), dtype=torch.float32, device=dev)
        scl = torch.empty((n, 3), dtype=torch.float32, device=dev)
        rc = self.lib.gut_unpack_particle_gradient(self._h, self._stream(), n, _ptr(g), _ptr(pos), _ptr(dns), _ptr(rot), _ptr(scl))
        if rc != 0:
            self._raise(rc)
        return pos, dns, rot, scl

    # ---- model-side passes ---------------------------------------------------------
    def activate_pack(self, positions, rotation_raw, scale_raw, density_raw, scale_activation=abi.Activation.EXP,
                      density_activation=abi.Activation.SIGMOID):
        """gut_activate_pack: raw parameters -> [N,12] record with normalize / scale / density activations applied
        (model.py:77-93 + tracer.py:176-178 in one pass)."""
        dev = self.device
        pos = _check_cuda_f32("positions", positions, dev)
        rot = _check_cuda_f32("rotation", rotation_raw, dev)
        scl = _check_cuda_f32("scale", scale_raw, dev)
        dns = _check_cuda_f32("density", density_raw, dev)
        n = pos.shape[0]
        if pos.shape != (n, 3) or dns.shape != (n, 1) or rot.shape != (n, 4) or scl.shape != (n, 3):
            raise GutError(abi.GutStatus.BAD_INPUT, "expected positions [N,3], rotation [N,4], scale [N,3], density [N,1]")
        out = torch.empty((n, 12), dtype=torch.float32, device=dev)
        rc = self.lib.gut_activate_pack(self._h, self._stream(), n, _ptr(pos), _ptr(rot), _ptr(scl), _ptr(dns), int(scale_activation),
                                        int(density_activation), _ptr(out))
        if rc != 0:
            self._raise(rc)
        return out

    def activate_pack_backward(self, rotation_raw, particle_density, density_grad, scale_activation=abi.Activation.EXP,
                               density_activation=abi.Activation.SIGMOID):
        """gut_activate_pack_backward: [N,12] gradient -> gradients of the RAW (positions, rotation, scale, density)."""
        dev = self.device
        rot = _check_cuda_f32("rotation", rotation_raw, dev)
        pk = _check_cuda_f32("particle_density", particle_density, dev)
        g = _check_cuda_f32("density_grad", density_grad, dev)
        n = rot.shape[0]
        if rot.shape != (n, 4) or pk.shape != (n, 12) or g.shape != (n, 12):
            raise GutError(abi.GutStatus.BAD_INPUT, "expected rotation [N,4], particle_density [N,12], density_grad [N,12]")
        pos_g = torch.empty((n, 3), dtype=torch.float32, device=dev)
        rot_g = torch.empty((n, 4), dtype=torch.float32, device=dev)
        scl_g = torch.empty((n, 3), dtype=torch.float32, device=dev)
        dns_g = torch.empty((n, 1), dtype=torch.float32, device=dev)
        rc = self.lib.gut_activate_pack_backward(self._h, self._stream(), n, _ptr(rot), _ptr(pk), _ptr(g), int(scale_activation),
                                                 int(density_activation), _ptr(pos_g), _ptr(rot_g), _ptr(scl_g), _ptr(dns_g))
        if rc != 0:
            self._raise(rc)
        return pos_g, rot_g, scl_g, dns_g

    def pack_features(self, features_albedo, features_specular):
        """gut_pack_features: cat((albedo [N,3], specular [N,45]), dim=1) -> [N,48] (model.py:74-75)."""
        dev = self.device
        alb = _check_cuda_f32("features_albedo", features_albedo, dev)
        spc = _check_cuda_f32("features_specular", features_specular, dev)
        n = alb.shape[0]
        ns = 3 * (self.n_coeffs - 1)
        if alb.shape != (n, 3) or spc.shape != (n, ns):
            raise GutError(abi.GutStatus.BAD_INPUT, f"expected features_albedo [N,3] and features_specular [N,{ns}]")
        out = torch.empty((n, 3 * self.n_coeffs), dtype=torch.float32, device=dev)
        rc = self.lib.gut_pack_features(self._h, self._stream(), n, _ptr(alb), _ptr(spc), _ptr(out))
        if rc != 0:
            self._raise(rc)
        return out

    def unpack_feature_gradient(self, radiance_grad):
        """gut_unpack_feature_gradient: [N,48] gradient -> (albedo [N,3], specular [N,45]) gradients."""
        dev = self.device
        g = _check_cuda_f32("radiance_grad", radiance_grad, dev)
        n = g.shape[0]
        if g.shape != (n, 3 * self.n_coeffs):
            raise GutError(abi.GutStatus.BAD_INPUT, f"radiance_grad must be [N,{3 * self.n_coeffs}]")
        alb = torch.empty((n, 3), dtype=torch.float32, device=dev)
        spc = torch.empty((n, 3 * (self.n_coeffs - 1)), dtype=torch.float32, device=dev)
        rc = self.lib.gut_unpack_feature_gradient(self._h, self._stream(), n, _ptr(g), _ptr(alb), _ptr(spc))
        if rc != 0:
            self._raise(rc)
        return alb, spc

    def selective_adam_step(self, groups, visibility, beta1, beta2, eps, zero_grad=False):
        """gut_selective_adam_step.  ``groups`` is a list of (param, grad, exp_avg, exp_avg_sq, lr); every tensor a contiguous
        float32 CUDA [N, ...] tensor, updated in place.  ``visibility`` is the tracer's mog_visibility ([N,1] float) or None."""
        if len(groups) == 0:
            return
        if len(groups) > abi.GUT_ADAM_MAX_GROUPS:
            raise GutError(abi.GutStatus.BAD_INPUT, f"at most {abi.GUT_ADAM_MAX_GROUPS} parameter groups per call")
        dev = self.device
        n = groups[0][0].shape[0]
        arr = (abi.GutAdamGroup * len(groups))()
        for i, (p, g, m, v, lr) in enumerate(groups):
            for name, t in (("param", p), ("grad", g), ("exp_avg", m), ("exp_avg_sq", v)):
                if not t.is_cuda or t.dtype != torch.float32 or t.device != dev or not t.is_contiguous():
                    raise GutError(abi.GutStatus.BAD_INPUT, f"group {i}: {name} must be a contiguous float32 tensor on {dev}")
                if t.shape != p.shape:
                    raise GutError(abi.GutStatus.BAD_INPUT, f"group {i}: {name} has shape {tuple(t.shape)}, param {tuple(p.shape)}")
            if p.shape[0] != n:
                raise GutError(abi.GutStatus.BAD_INPUT, f"group {i} has {p.shape[0]} rows, group 0 has {n}")
            arr[i].param, arr[i].grad, arr[i].exp_avg, arr[i].exp_avg_sq = p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr()
            arr[i].width = p.numel() // max(n, 1)
            arr[i].lr = float(lr)
        vis = None
        if visibility is not None:
            vis = visibility.reshape(-1)
            if vis.dtype != torch.float32:
                vis = vis.to(torch.float32)
            vis = _check_cuda_f32("visibility", vis, dev)
            if vis.shape[0] != n:
                raise GutError(abi.GutStatus.BAD_INPUT, f"visibility has {vis.shape[0]} rows, parameters {n}")
        rc = self.lib.gut_selective_adam_step(self._h, self._stream(), n, arr, len(groups), _ptr(vis), float(beta1), float(beta2), float(eps),
                                              1 if zero_grad else 0)
        if rc != 0:
            self._raise(rc)

    def _adam_tensors(self, name, triple, lr, n):
        p, m, v = triple
        for what, t in (("param", p), ("exp_avg", m), ("exp_avg_sq", v)):
            if not t.is_cuda or t.dtype != torch.float32 or t.device != self.device or not t.is_contiguous() or t.shape != p.shape:
                raise GutError(abi.GutStatus.BAD_INPUT, f"{name}: {what} must be a contiguous float32 tensor on {self.device} shaped like the parameter")
        if p.shape[0] != n:
            raise GutError(abi.GutStatus.BAD_INPUT, f"{name} has {p.shape[0]} rows, expected {n}")
        return abi.GutAdamTensors(param=p.data_ptr(), exp_avg=m.data_ptr(), exp_avg_sq=v.data_ptr(), lr=float(lr))

    def _vis(self, visibility, n):
        if visibility is None:
            return None
        vis = visibility.reshape(-1)
        if vis.dtype != torch.float32:
            vis = vis.to(torch.float32)
        vis = _check_cuda_f32("visibility", vis, self.device)
        if vis.shape[0] != n:
            raise GutError(abi.GutStatus.BAD_INPUT, f"visibility has {vis.shape[0]} rows, parameters {n}")
        return vis

    def density_adam(self, particle_density, density_grad, positions, rotation, scale, density, lrs, visibility, beta1, beta2, eps,
                     scale_activation=abi.Activation.EXP, density_activation=abi.Activation.SIGMOID, sensor_position=None,
                     grad_norm_accum=None, grad_norm_denom=None):
        """gut_density_adam: the summed [N,12] gradient -> activation adjoints -> SelectiveAdam on the four raw parameters
        (each a (param, exp_avg, exp_avg_sq) triple, ``lrs`` their four learning rates) in one pass; with ``sensor_position`` and
        the two statistics buffers also GSStrategy.update_gradient_buffer."""
        dens = _check_cuda_f32("particle_density", particle_density, self.device)
        g = _check_cuda_f32("density_grad", density_grad, self.device)
        n = dens.shape[0]
        if dens.shape != (n, 12) or g.shape != (n, 12):
            raise GutError(abi.GutStatus.BAD_INPUT, "particle_density and density_grad must be [N,12]")
        ts = [self._adam_tensors(k, t, lr, n) for k, t, lr in zip(("positions", "rotation", "scale", "density"),
                                                                   (positions, rotation, scale, density), lrs)]
        vis = self._vis(visibility, n)
        sensor = None
        if sensor_position is not None:
            sensor = (C.c_float * 3)(*[float(x) for x in sensor_position])
            for t in (grad_norm_accum, grad_norm_denom):
                if t is None or t.numel() != n or not t.is_contiguous() or t.dtype != torch.float32:
                    raise GutError(abi.GutStatus.BAD_INPUT, "grad_norm_accum / grad_norm_denom must be contiguous float32 [N,1]")
        rc = self.lib.gut_density_adam(self._h, self._stream(), n, _ptr(dens), _ptr(g), C.byref(ts[0]), C.byref(ts[1]), C.byref(ts[2]),
                                       C.byref(ts[3]), int(scale_activation), int(density_activation), _ptr(vis), float(beta1), float(beta2),
                                       float(eps), sensor, _ptr(grad_norm_accum) if sensor is not None else None,
                                       _ptr(grad_norm_denom) if sensor is not None else None)
        if rc != 0:
            self._raise(rc)

    def sh_expand_adam(self, num_active_features, particle_density, view_centres, colour_ptrs, features_albedo, features_specular, lrs,
                       visibility, beta1, beta2, eps):
        """gut_sh_expand_adam: expansion of the step's per-view colour gradients (one device pointer per view, as
        sh_gradient_expand_gather) with SelectiveAdam on features_albedo / features_specular as its epilogue."""
        dens = _check_cuda_f32("particle_density", particle_density, self.device)
        cen = _check_cuda_f32("view_centres", view_centres, self.device)
        n, nv = dens.shape[0], len(colour_ptrs)
        if cen.numel() < 3 * nv:
            raise GutError(abi.GutStatus.BAD_INPUT, "view_centres holds fewer rows than there are colour pointers")
        alb = self._adam_tensors("features_albedo", features_albedo, lrs[0], n)
        spc = self._adam_tensors("features_specular", features_specular, lrs[1], n)
        if features_albedo[0].numel() != 3 * n or features_specular[0].numel() != 45 * n:
            raise GutError(abi.GutStatus.BAD_INPUT, "features_albedo must be [N,3] and features_specular [N,45]")
        arr = (C.c_void_p * max(nv, 1))(*[int(p) for p in colour_ptrs])
        rc = self.lib.gut_sh_expand_adam(self._h, self._stream(), n, int(num_active_features), _ptr(dens), nv, _ptr(cen), arr,
                                         C.byref(alb), C.byref(spc), _ptr(self._vis(visibility, n)), float(beta1), float(beta2), float(eps))
        if rc != 0:
            self._raise(rc)

    def densify_stats(self, positions, positions_grad, sensor_position, grad_norm_accum, grad_norm_denom):
        """gut_densify_stats: GSStrategy.update_gradient_buffer (strategy/gs.py:104-115), in place on accum / denom [N,1]."""
        dev = self.device
        pos = _check_cuda_f32("positions", positions, dev)
        g = _check_cuda_f32("positions_grad", positions_grad, dev)
        n = pos.shape[0]
        for name, t in (("grad_norm_accum", grad_norm_accum), ("grad_norm_denom", grad_norm_denom)):
            if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != n:
                raise GutError(abi.GutStatus.BAD_INPUT, f"{name} must be a contiguous float32 CUDA tensor with N elements")
        sp = (C.c_float * 3)(*[float(x) for x in sensor_position])
        rc = self.lib.gut_densify_stats(self._h, self._stream(), n, _ptr(pos), _ptr(g), sp, _ptr(grad_norm_accum), _ptr(grad_norm_denom))
        if rc != 0:
            self._raise(rc)

    def set_training(self, training: bool):
        """gut_set_training: training=False skips the hit-mask recording of the forward (inference); no backward may follow."""
        self.lib.gut_set_training(self._h, 1 if training else 0)

    def set_colour_layout(self, layout: int):
        """gut_set_colour_layout: GUT_COLOUR_DENSE ([N,3] per view) or GUT_COLOUR_COMPACT (touched rows behind a table) for the
        colour-gradient slots this handle writes (trace_bwd_colour) and reads (sh_gradient_expand_gather, sh_expand_adam)."""
        if int(layout) == self.colour_layout:
            return
        rc = self.lib.gut_set_colour_layout(self._h, int(layout))
        if rc != 0:
            self._raise(rc)
        self.colour_layout = int(layout)

    def colour_slot_floats(self, n_particles, layout=None):
        """gut_colour_slot_floats: floats one view's colour-gradient slot takes for n_particles in the given (default: current) layout."""
        return int(self.lib.gut_colour_slot_floats(int(n_particles), self.colour_layout if layout is None else int(layout)))

    def set_block_cull(self, enabled: bool):
        """gut_set_block_cull: test control, see include/gut_b200.h (outputs are bit-identical either way)."""
        self.lib.gut_set_block_cull(self._h, 1 if enabled else 0)

    def reset_gradients(self):
        """gut_reset_gradients: forget the pending sums of an abandoned multi-view step."""
        rc = self.lib.gut_reset_gradients(self._h, self._stream())
        if rc != 0:
            self._raise(rc)

    def collect_times(self):
        """reference: SplatRaster::collectTimes (splatRaster.cpp:334-364) -> {tag: mean ms}."""
        f, b = C.c_float(-1.0), C.c_float(-1.0)
        self.lib.gut_timings(self._h, C.byref(f), C.byref(b))
        out = {}
        if f.value >= 0:
            out["forward_render"] = f.value
        if b.value >= 0:
            out["backward_render"] = b.value
        return out

    def stage_times(self):
        n = self.lib.gut_num_stages()
        arr = (C.c_float * n)()
        self.lib.gut_stage_timings(self._h, arr)
        return {self.lib.gut_stage_name(i).decode(): arr[i] for i in range(n) if arr[i] >= 0}

    def frame_stats(self):
        st = abi.GutFrameStats()
        self.lib.gut_frame_stats(self._h, C.byref(st))
        return {"num_intersections": st.num_intersections, "num_tiles": st.num_tiles,
                "max_tile_list_length": st.max_tile_list_length, "kernel_launches": st.kernel_launches}

    def measure_work(self, ray_origin, ray_direction):
        """(tests, hits) of the last forward's compositing walk -- roofline accounting only."""
        t, hh = C.c_uint64(0), C.c_uint64(0)
        rc = self.lib.gut_measure_work(self._h, self._stream(), _ptr(ray_origin.contiguous()), _ptr(ray_direction.contiguous()),
                                       C.byref(t), C.byref(hh))
        if rc != 0:
            self._raise(rc)
        return t.value, hh.value

    # ---- host-buffer entry points (end-to-end path) --------------------------------------------------------
    def trace_host(self, num_active_features, particle_density, particle_radiance, h_ray_origin, h_ray_direction, sensor,
                   h_rgba, h_dist, h_hits, visibility):
        n = particle_density.shape[0]
        height, width = h_ray_origin.shape[1], h_ray_origin.shape[2]
        rc = self.lib.gut_forward_host(self._h, self._stream(), C.byref(sensor), n, int(num_active_features), _ptr(particle_density),
                                       _ptr(particle_radiance), width, height, _ptr(h_ray_origin), _ptr(h_ray_direction),
                                       _ptr(h_rgba), _ptr(h_dist), _ptr(h_hits), _ptr(visibility))
        if rc != 0:
            self._raise(rc)

    def trace_bwd_host(self, num_active_features, particle_density, particle_radiance, width, height, sensor, h_rgba_grad,
                       h_dist_grad, out_density_grad, out_radiance_grad, accumulate=False):
        n = particle_density.shape[0]
        rc = self.lib.gut_backward_host(self._h, self._stream(), C.byref(sensor), n, int(num_active_features), _ptr(particle_density),
                                        _ptr(particle_radiance), width, height, _ptr(h_rgba_grad), _ptr(h_dist_grad),
                                        _ptr(out_density_grad), _ptr(out_radiance_grad), 1 if accumulate else 0)
        if rc != 0:
            self._raise(rc)
