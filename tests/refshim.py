"""Loader for the "reference + shim" builds under oracle/_ref (made by oracle/ref_build/Makefile in the container that has
/root/reference; the .so files travel to the GPU box, the sources do not).

TEST INFRASTRUCTURE.  Two artefacts:
  * ``lib3dgut_cc_k<K>``  the reference's own pybind11 module (bindings.cpp unchanged): ``SplatRaster(conf).trace / trace_bwd``
  * ``libgut_ref_stages_k<K>.so``  probe.cu: GUTRenderer::renderForward / renderBackward behind a C ABI + read-out of the
    forward context (tiles_count ... sorted keys, ranges)
Both run the reference's kernels; the only non-reference code inside is oracle/ref_build/threedgutSlang.cuh.
"""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def available(k=0):
    return os.path.exists(os.path.join(REF_DIR, f"lib3dgut_cc_k{k}.so"))


def probe_available(k=0):
    return os.path.exists(os.path.join(REF_DIR, f"libgut_ref_stages_k{k}.so"))


def module(k=0):
    """The reference's extension module for k_buffer_size = k (a compile-time constant there, setup_3dgut.py:57)."""
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    return importlib.import_module(f"lib3dgut_cc_k{k}")


_PROBE = {}


def probe(k=0):
    if k not in _PROBE:
        lib = C.CDLL(os.path.join(REF_DIR, f"libgut_ref_stages_k{k}.so"))
        lib.gutref_forward.restype = C.c_int
        lib.gutref_backward.restype = C.c_int
        lib.gutref_read.restype = C.c_longlong
        lib.gutref_read.argtypes = [C.c_char_p, C.c_void_p, C.c_ulonglong]
        assert lib.gutref_k_buffer_size() == k
        _PROBE[k] = lib
    return _PROBE[k]


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def sensor_of(mod, cam):
    """abi.GutCamera -> the module's CameraModelParameters through the reference's own factory functions
    (bindings.cpp:33-77), as tracer.py:386-427 does."""
    shutter = [mod.ShutterType.ROLLING_TOP_TO_BOTTOM, mod.ShutterType.ROLLING_LEFT_TO_RIGHT, mod.ShutterType.ROLLING_BOTTOM_TO_TOP,
               mod.ShutterType.ROLLING_RIGHT_TO_LEFT, mod.ShutterType.GLOBAL][cam.shutter]
    if cam.model == 0:
        return mod.fromOpenCVPinholeCameraModelParameters(
            resolution=[0, 0], shutter_type=shutter, principal_point=list(cam.principal_point), focal_length=list(cam.focal_length),
            radial_coeffs=list(cam.radial_coeffs), tangential_coeffs=list(cam.tangential_coeffs),
            thin_prism_coeffs=list(cam.thin_prism_coeffs))
    return mod.fromOpenCVFisheyeCameraModelParameters(
        resolution=[0, 0], shutter_type=shutter, principal_point=list(cam.principal_point), focal_length=list(cam.focal_length),
        radial_coeffs=list(cam.radial_coeffs)[:4], max_angle=float(cam.max_angle))


class ReferenceRaster:
    """The reference's SplatRaster called the way threedgut_tracer/tracer.py:159-286 calls it."""

    def __init__(self, k=0, enable_kernel_timings=False):
        self.mod = module(k)
        self.raster = self.mod.SplatRaster({"render": {"enable_kernel_timings": bool(enable_kernel_timings)}})

    def _args(self, sph_degree, d12, sph, ray_ori, ray_dir, cam):
        h, w = ray_ori.shape[-3], ray_ori.shape[-2]
        ray_time = torch.zeros((1, h, w, 1), dtype=torch.long, device=d12.device)   # tracer.py:181-186 (dead input)
        sp = torch.tensor(list(cam.start_pose), dtype=torch.float32)
        ep = torch.tensor(list(cam.end_pose), dtype=torch.float32)
        return (0, int(sph_degree), d12.contiguous(), sph.contiguous(), ray_ori.reshape(1, h, w, 3).contiguous(),
                ray_dir.reshape(1, h, w, 3).contiguous(), ray_time, sensor_of(self.mod, cam), int(cam.start_timestamp_us),
                int(cam.end_timestamp_us), sp, ep)

    def trace(self, sph_degree, d12, sph, ray_ori, ray_dir, cam):
        return self.raster.trace(*self._args(sph_degree, d12, sph, ray_ori, ray_dir, cam))

    def trace_bwd(self, sph_degree, d12, sph, ray_ori, ray_dir, cam, rgba, rgba_grad, dist, dist_grad):
        return self.raster.trace_bwd(*self._args(sph_degree, d12, sph, ray_ori, ray_dir, cam), rgba.contiguous(),
                                     rgba_grad.contiguous(), dist.contiguous(), dist_grad.contiguous())

    def collect_times(self):
        return dict(self.raster.collect_times())


def probe_forward(cam, sph_degree, d12, sph, ray_ori, ray_dir, w, h, k=0):
    """Reference forward through the probe; returns (rgba, dist, hits, visibility, stages-dict of numpy arrays)."""
    lib = probe(k)
    dev = d12.device
    n = d12.shape[0]
    rgba = torch.zeros((h, w, 4), device=dev)
    dist = torch.full((h, w, 1), 1e6, device=dev)
    hits = torch.zeros((h, w, 1), device=dev)
    vis = torch.zeros((n, 1), device=dev)
    m = C.c_uint32(0)
    torch.cuda.synchronize()
    rc = lib.gutref_forward(C.byref(cam), C.c_uint32(n), C.c_int(sph_degree), _ptr(d12), _ptr(sph), C.c_uint32(w), C.c_uint32(h),
                            _ptr(ray_ori), _ptr(ray_dir), _ptr(rgba), _ptr(dist), _ptr(hits), _ptr(vis), C.byref(m))
    assert rc == 0, rc
    M = int(m.value)
    t = ((w + 15) // 16) * ((h + 15) // 16)
    shapes = {"tiles_count": ((n,), np.uint32), "tiles_offset": ((n,), np.uint32), "proj_pos": ((n, 2), np.float32),
              "conic_opacity": ((n, 4), np.float32), "extent": ((n, 2), np.float32), "depth": ((n,), np.float32),
              "features": ((n, 3), np.float32)}
    if M:
        shapes.update({"unsorted_keys": ((M,), np.uint64), "unsorted_idx": ((M,), np.uint32), "sorted_keys": ((M,), np.uint64),
                       "sorted_idx": ((M,), np.uint32), "ranges": ((t, 2), np.uint32)})
    st = {"M": M}
    for name, (shape, dt) in shapes.items():
        a = np.zeros(shape, dt)
        got = lib.gutref_read(name.encode(), a.ctypes.data_as(C.c_void_p), C.c_ulonglong(a.nbytes))
        assert got == a.nbytes, (name, got, a.nbytes)
        st[name] = a
    return rgba, dist, hits, vis, st


def probe_backward(cam, sph_degree, d12, sph, ray_ori, ray_dir, w, h, rgba, rgba_grad, dist, dist_grad, k=0):
    lib = probe(k)
    n = d12.shape[0]
    g12 = torch.zeros((n, 12), device=d12.device)
    gsph = torch.zeros_like(sph)
    torch.cuda.synchronize()
    rc = lib.gutref_backward(C.byref(cam), C.c_uint32(n), C.c_int(sph_degree), _ptr(d12), _ptr(sph), C.c_uint32(w), C.c_uint32(h),
                             _ptr(ray_ori), _ptr(ray_dir), _ptr(rgba), _ptr(rgba_grad), _ptr(dist), _ptr(dist_grad), _ptr(g12),
                             _ptr(gsph))
    assert rc == 0, rc
    return g12, gsph


def optimizers_available():
    return os.path.exists(os.path.join(REF_DIR, "lib_optimizers_cc.so"))


def optimizers_module():
    """threedgrut/optimizers/optimizers.{cu,cpp} built unchanged (plain CUDA + torch, no Slang): ``selective_adam_update``."""
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    return importlib.import_module("lib_optimizers_cc")
