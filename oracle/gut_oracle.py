"""ctypes front-end of the C oracle (oracle/gut_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  Never by the product.
PARITY UNPINNED (see gut_oracle.c header): no reference golden vectors exist.

All arrays are numpy, float32 / uint32 / uint64, C-contiguous.
"""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np

abi = importlib.import_module("3dgrut_b200.abi")

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "libgut_oracle.so")
    src = os.path.join(_HERE, "gut_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.gut_oracle_logf.restype = C.c_float
        _LIB.gut_oracle_logf.argtypes = [C.c_float]
        _LIB.gut_oracle_scan.restype = C.c_uint32
        _LIB.gut_oracle_higher_msb.restype = C.c_uint32
        _LIB.gut_oracle_higher_msb.argtypes = [C.c_uint32]
        _LIB.gut_oracle_sort.restype = C.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def set_threads(n):
    """OpenMP threads of the oracle's loops (overrides OMP_NUM_THREADS, which torchrun sets to 1)."""
    lib().gut_oracle_set_threads(C.c_int(int(n)))


def threads_used():
    """Threads a parallel region of the oracle really gets (counted inside the library)."""
    lib().gut_oracle_threads_used.restype = C.c_int
    return int(lib().gut_oracle_threads_used())


def higher_msb(n):
    return int(lib().gut_oracle_higher_msb(int(n)))


def logf(x):
    return float(lib().gut_oracle_logf(float(x)))


def frame_consts(cam, W, H):
    out = np.zeros(41, np.float32)
    lib().gut_oracle_frame_consts(C.byref(cam), C.c_uint32(W), C.c_uint32(H), _p(out))
    return {"pose_mid": out[:7], "pose_mid_inv": out[7:14], "view": out[14:26], "c2w": out[26:38], "cam_centre": out[38:41]}


def project(cfg, cam, density12, sph, W, H, sph_degree):
    """K1. Returns dict of per-particle stage arrays."""
    density12 = _f32(density12); sph = _f32(sph)
    N = density12.shape[0]
    o = {
        "tiles_count": np.zeros(N, np.uint32),
        "proj_pos": np.zeros((N, 2), np.float32),
        "conic_opacity": np.zeros((N, 4), np.float32),
        "extent": np.zeros((N, 2), np.float32),
        "depth": np.zeros(N, np.float32),
        "features": np.zeros((N, 3), np.float32),
        "visibility": np.zeros(N, np.float32),
    }
    lib().gut_oracle_project(C.byref(cfg), C.byref(cam), C.c_uint32(N), C.c_int(sph_degree), _p(density12), _p(sph),
                             C.c_uint32(W), C.c_uint32(H), _p(o["tiles_count"]), _p(o["proj_pos"]),
                             _p(o["conic_opacity"]), _p(o["extent"]), _p(o["depth"]), _p(o["features"]), _p(o["visibility"]))
    return o


def bin_and_sort(cfg, W, H, proj):
    """L1 + K2 + L2 + K3 from K1's outputs. Returns keys/values (unsorted, sorted) and tile ranges."""
    N = proj["tiles_count"].shape[0]
    offs = np.zeros(N, np.uint32)
    M = int(lib().gut_oracle_scan(C.c_uint32(N), _p(proj["tiles_count"]), _p(offs)))
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    keys = np.zeros(M, np.uint64); vals = np.zeros(M, np.uint32)
    skeys = np.zeros(M, np.uint64); svals = np.zeros(M, np.uint32)
    ranges = np.zeros((T, 2), np.uint32)
    if M:
        lib().gut_oracle_expand(C.byref(cfg), C.c_uint32(N), C.c_uint32(W), C.c_uint32(H), _p(offs),
                                _p(proj["proj_pos"]), _p(proj["conic_opacity"]), _p(proj["extent"]), _p(proj["depth"]),
                                _p(keys), _p(vals))
        end_bit = 32 + higher_msb(T)
        rc = lib().gut_oracle_sort(C.c_uint64(M), C.c_uint32(end_bit), _p(keys), _p(vals), _p(skeys), _p(svals))
        assert rc == 0
        lib().gut_oracle_tile_ranges(C.c_uint64(M), _p(skeys), C.c_uint32(T), _p(ranges))
    return {"tiles_offset": offs, "M": M, "keys": keys, "values": vals, "sorted_keys": skeys,
            "sorted_values": svals, "tile_ranges": ranges}


def _tiles(tile_list):
    if tile_list is None:
        return C.c_void_p(0), C.c_uint32(0), None
    tl = np.ascontiguousarray(tile_list, dtype=np.uint32)
    return _p(tl), C.c_uint32(tl.shape[0]), tl


def render(cfg, cam, density12, features, ray_ori, ray_dir, W, H, tile_ranges, sorted_values, tile_list=None):
    density12 = _f32(density12); features = _f32(features)
    ray_ori = _f32(ray_ori).reshape(-1, 3); ray_dir = _f32(ray_dir).reshape(-1, 3)
    rgba = np.zeros((H, W, 4), np.float32)
    dist = np.full((H, W), 1e6, np.float32)
    hits = np.zeros((H, W), np.float32)
    tests = C.c_uint64(0)
    margin = np.full((H, W), np.inf, np.float32)
    tl_p, tl_n, _keep = _tiles(tile_list)
    lib().gut_oracle_render(C.byref(cfg), C.byref(cam), C.c_uint32(density12.shape[0]), C.c_uint32(W), C.c_uint32(H),
                            _p(density12), _p(features), _p(ray_ori), _p(ray_dir), _p(tile_ranges), _p(sorted_values),
                            _p(rgba), _p(dist), _p(hits), C.byref(tests), tl_p, tl_n, _p(margin))
    return {"rgba": rgba, "dist": dist, "hits": hits, "tests": tests.value, "margin": margin}


def forward(cfg, cam, density12, sph, ray_ori, ray_dir, W, H, sph_degree, tile_list=None):
    """Whole forward (K1..K4) as GUTRenderer::renderForward sequences it (src/gutRenderer.cu:241-398)."""
    proj = project(cfg, cam, density12, sph, W, H, sph_degree)
    bins = bin_and_sort(cfg, W, H, proj)
    if bins["M"] == 0:  # early return, gutRenderer.cu:323-325
        out = {"rgba": np.zeros((H, W, 4), np.float32), "dist": np.full((H, W), 1e6, np.float32),
               "hits": np.zeros((H, W), np.float32), "tests": 0, "margin": np.full((H, W), np.inf, np.float32)}
    else:
        out = render(cfg, cam, density12, proj["features"], ray_ori, ray_dir, W, H, bins["tile_ranges"], bins["sorted_values"],
                     tile_list=tile_list)
    out.update(proj=proj, bins=bins, visibility=proj["visibility"])
    return out


def backward(cfg, cam, density12, sph, ray_ori, ray_dir, W, H, sph_degree, fwd, rgba_grad, dist_grad, tile_list=None):
    """K5 + K6 (src/gutRenderer.cu:400-497). k_buffer_size == 0: evalBackwardNoKBuffer (gutKBufferRenderer.cuh:294-386);
    k_buffer_size > 0: evalKBuffer<Backward> (:217-292).  Returns ([N,12], [N,ncoef*3], [N,3]) float32."""
    density12 = _f32(density12); sph = _f32(sph)
    N = density12.shape[0]
    ncoef = (cfg.sph_max_degree + 1) ** 2
    dg = np.zeros((N, 12), np.float64); fg = np.zeros((N, 3), np.float64)
    sg = np.zeros((N, ncoef * 3), np.float32)
    if fwd["bins"]["M"] == 0:
        return dg.astype(np.float32), sg, fg.astype(np.float32)
    ray_ori = _f32(ray_ori).reshape(-1, 3); ray_dir = _f32(ray_dir).reshape(-1, 3)
    rgba_grad = _f32(rgba_grad); dist_grad = _f32(dist_grad)
    proj, bins = fwd["proj"], fwd["bins"]
    tl_p, tl_n, _keep = _tiles(tile_list)
    fn = lib().gut_oracle_render_bwd if cfg.k_buffer_size == 0 else lib().gut_oracle_render_bwd_kbuffer
    fn(C.byref(cfg), C.byref(cam), C.c_uint32(N), C.c_uint32(W), C.c_uint32(H), _p(density12),
                                _p(proj["features"]), _p(ray_ori), _p(ray_dir), _p(bins["tile_ranges"]),
                                _p(bins["sorted_values"]), _p(fwd["rgba"]), _p(rgba_grad), _p(fwd["dist"]), _p(dist_grad),
                                _p(dg), _p(fg), tl_p, tl_n)
    fg32 = fg.astype(np.float32)
    lib().gut_oracle_project_bwd(C.byref(cfg), C.byref(cam), C.c_uint32(N), C.c_int(sph_degree), C.c_uint32(W), C.c_uint32(H),
                                 _p(density12), _p(proj["tiles_count"]), _p(proj["features"]), _p(fg32), _p(sg))
    return dg.astype(np.float32), sg, fg32
