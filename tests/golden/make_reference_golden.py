#!/usr/bin/env python
"""Generates tests/golden/reference_*.npz: outputs of the REFERENCE's own CUDA tracer ("reference + shim": the unchanged
sources of /root/reference/threedgut_tracer built by oracle/ref_build/Makefile against the hand-written
threedgutSlang.cuh) on the seeded golden cases.  Needs a GPU and oracle/_ref:

    gpurun -- 'python tests/golden/make_reference_golden.py'      # writes gpurun_out/reference_<case>.npz
    cp gpurun_out/reference_*.npz tests/golden/

These are the vectors that pin the CPU oracle without a GPU (tests/test_reference_golden.py): every array below was
produced by the reference's kernels (K1 projectOnTiles, CUB scan, K2 expandTileProjections, CUB radix sort,
K3 computeSortedTileRangeIndices, K4 render, K5 renderBackward, K6 projectBackward).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from tests import refshim  # noqa: E402
from tests.golden.make_golden import CASES, case_inputs  # noqa: E402
from tests.helpers import camera_of, syn  # noqa: E402

EXTRA = {"c1_small": (2500, 64, 64, 0, "init", 1001, "pinhole", False)}


def run_reference(name, k=0):
    scene, view, (n, w, h, deg), rgba_grad, dist_grad = case_inputs(name)
    cam = camera_of(view)
    d12 = syn.pack_density(scene).cuda(); sph = scene["features"].cuda()
    ro = view["rays_ori"].cuda().contiguous(); rd = view["rays_dir"].cuda().contiguous()
    rgba, dist, hits, vis, st = refshim.probe_forward(cam, deg, d12, sph, ro, rd, w, h, k=k)
    g12, gsph = refshim.probe_backward(cam, deg, d12, sph, ro, rd, w, h, rgba, torch.from_numpy(rgba_grad).cuda(), dist,
                                       torch.from_numpy(dist_grad).cuda().reshape(h, w, 1), k=k)
    torch.cuda.synchronize()
    out = {k2: v for k2, v in st.items() if k2 != "M"}
    out.update(rgba=rgba.cpu().numpy(), dist=dist.cpu().numpy().reshape(h, w), hits=hits.cpu().numpy().reshape(h, w).astype(np.uint16),
               visibility=(vis.cpu().numpy().reshape(-1).view(np.int32) != 0), density_grad=g12.cpu().numpy(),
               radiance_grad=gsph.cpu().numpy(), k_buffer_size=np.int32(k))
    return out


def run_reference_adam():
    """Three SelectiveAdam steps of the reference's compiled kernel (oracle/_ref/lib_optimizers_cc.so) on three groups."""
    ref = refshim.optimizers_module()
    n = 3000
    gen = torch.Generator(device="cuda").manual_seed(11)
    out = {}
    b1, b2, eps = 0.9, 0.999, 1e-15
    out["hyper"] = np.array([b1, b2, eps], np.float64)
    for gi, (w, lr) in enumerate(((3, 1.6e-4), (1, 5e-2), (45, 1.25e-4))):
        p = torch.randn((n, w), device="cuda", generator=gen)
        m = torch.randn((n, w), device="cuda", generator=gen) * 0.01
        v = torch.rand((n, w), device="cuda", generator=gen) * 1e-4
        out[f"g{gi}_lr"] = np.float64(lr)
        out[f"g{gi}_param0"], out[f"g{gi}_m0"], out[f"g{gi}_v0"] = p.cpu().numpy(), m.cpu().numpy(), v.cpu().numpy()
        for step in range(3):
            vis = torch.rand((n, 1), device="cuda", generator=gen) < 0.6
            g = torch.randn((n, w), device="cuda", generator=gen) * 10.0 ** (-step)
            ref.selective_adam_update(p, g, m, v, vis.squeeze(1), lr, b1, b2, eps)
            out[f"g{gi}_s{step}_vis"], out[f"g{gi}_s{step}_grad"] = vis.cpu().numpy(), g.cpu().numpy()
            out[f"g{gi}_s{step}_param"], out[f"g{gi}_s{step}_m"], out[f"g{gi}_s{step}_v"] = p.cpu().numpy(), m.cpu().numpy(), v.cpu().numpy()
    return out


if __name__ == "__main__":
    CASES.update(EXTRA)
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    if refshim.optimizers_available():
        path = os.path.join(out_dir, "reference_adam.npz")
        np.savez_compressed(path, **run_reference_adam())
        print("adam", os.path.getsize(path) // 1024, "KiB")
    for name in CASES:
        for k in (0, 16):
            if k and name not in ("pinhole_dense_sh2",):
                continue
            out = run_reference(name, k)
            suffix = "" if k == 0 else f"_k{k}"
            path = os.path.join(out_dir, f"reference_{name}{suffix}.npz")
            np.savez_compressed(path, **out)
            print(name, k, "M =", out["sorted_keys"].shape[0], os.path.getsize(path) // 1024, "KiB")
