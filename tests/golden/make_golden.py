#!/usr/bin/env python
"""Generates the frozen fixtures tests/golden/*.npz from the CPU oracle (oracle/gut_oracle.c) on seeded synthetic inputs.

The reference ships no golden vectors for this path and cannot be built or imported here, so these
are regression fixtures of the oracle (PARITY UNPINNED), cross-validated by the independent torch transcription in
tests/test_oracle.py.  Re-run only when the oracle is deliberately changed:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import gut_oracle as O  # noqa: E402
from tests.helpers import BORDERLINE_MARGIN, abi, camera_of, syn  # noqa: E402

# name -> (N, W, H, active SH degree, scene kind, seed, camera model, rolling shutter)
CASES = {
    "pinhole_init_sh3": (1200, 80, 48, 3, "init", 101, "pinhole", False),
    "pinhole_dense_sh2": (900, 64, 64, 2, "dense", 102, "pinhole", False),
    "fisheye_rolling_sh1": (1200, 96, 54, 1, "init", 103, "fisheye", True),
}


def case_inputs(name):
    n, w, h, deg, kind, seed, model, rolling = CASES[name]
    scene = syn.make_scene(n, 3, seed=seed, kind=kind)
    view = syn.make_view(w, h, model=model, rolling=rolling)
    rng = np.random.default_rng(seed)
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    return scene, view, (n, w, h, deg), rgba_grad, dist_grad


def run_oracle(name):
    scene, view, (n, w, h, deg), rgba_grad, dist_grad = case_inputs(name)
    cfg = abi.default_config()
    cam = camera_of(view)
    d12 = syn.pack_density(scene).numpy(); sph = scene["features"].numpy()
    ro, rd = view["rays_ori"].numpy(), view["rays_dir"].numpy()
    fwd = O.forward(cfg, cam, d12, sph, ro, rd, w, h, deg)
    g12, gsph, _ = O.backward(cfg, cam, d12, sph, ro, rd, w, h, deg, fwd, rgba_grad, dist_grad)
    return {
        "tiles_count": fwd["proj"]["tiles_count"], "proj_pos": fwd["proj"]["proj_pos"], "conic_opacity": fwd["proj"]["conic_opacity"],
        "extent": fwd["proj"]["extent"], "depth": fwd["proj"]["depth"], "features": fwd["proj"]["features"],
        "sorted_keys": fwd["bins"]["sorted_keys"], "sorted_values": fwd["bins"]["sorted_values"], "tile_ranges": fwd["bins"]["tile_ranges"],
        "rgba": fwd["rgba"], "dist": fwd["dist"], "hits": fwd["hits"].astype(np.uint16),
        "borderline": fwd["margin"] < BORDERLINE_MARGIN, "density_grad": g12, "radiance_grad": gsph.astype(np.float32),
    }


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    for name in CASES:
        out = run_oracle(name)
        path = os.path.join(here, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, "M =", out["sorted_keys"].shape[0], "max hits =", int(out["hits"].max()), os.path.getsize(path) // 1024, "KiB")
