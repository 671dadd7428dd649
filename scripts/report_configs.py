#!/usr/bin/env python
"""Single-GPU figures for the BASELINE.json configurations that are not the bench line (C2, C4, C5; C3 is bench.py itself).
Device-resident inputs, CUDA events, median of `--iters` iterations after 5 warm-ups.  Writes a markdown table.

    python scripts/report_configs.py --out gpurun_out/configs.md
"""
import argparse
import importlib
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.helpers import abi, camera_of, syn  # noqa: E402


def timed(fn, iters):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return statistics.median(ms)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/configs.md")
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    native = importlib.import_module("3dgrut_b200.native")
    dev = torch.device("cuda", 0)
    rows = []

    def fwd_bwd(name, n, w, h, model, rolling, seed):
        scene = syn.make_scene(n, 3, seed=seed)
        view = syn.make_view(w, h, model=model, rolling=rolling)
        d12, sph = syn.pack_density(scene).to(dev), scene["features"].to(dev)
        ro, rd, cam = view["rays_ori"].to(dev), view["rays_dir"].to(dev), camera_of(view)
        raster = native.SplatRaster(abi.default_config(enable_kernel_timings=1), device=dev)
        g = torch.Generator(device=dev).manual_seed(seed)
        rg = torch.rand((h, w, 4), device=dev, generator=g) * 2 - 1
        dg = (torch.rand((h, w, 1), device=dev, generator=g) * 2 - 1) * 0.01

        def step():
            o = raster.trace(0, 3, d12, sph, ro, rd, None, cam)
            raster.trace_bwd(0, 3, d12, sph, ro, rd, None, cam, 0, 1, None, None, o[0], rg, o[1], dg)
        ms = timed(step, args.iters)
        st, fs = raster.stage_times(), raster.frame_stats()
        rows.append(f"| {name} | {n} | {w}x{h} | fwd+bwd | {ms:.3f} | {w * h / ms / 1e3:.1f} Mpix/s | M = {fs['num_intersections']}, max list "
                    f"{fs['max_tile_list_length']}; " + ", ".join(f"{k.replace('render', 'r')} {v:.3f}" for k, v in st.items()) + " |")

    fwd_bwd("C2: pinhole", 1_000_000, 1237, 822, "pinhole", False, 1002)
    fwd_bwd("C4: OpenCV fisheye + rolling shutter", 1_000_000, 1920, 1080, "fisheye", True, 1004)

    view = syn.make_view(1920, 1080)
    ro, rd, cam = view["rays_ori"].to(dev), view["rays_dir"].to(dev), camera_of(view)
    for n in (500_000, 1_000_000, 2_000_000, 3_000_000, 6_000_000):
        scene = syn.make_scene(n, 3, seed=1005)
        d12, sph = syn.pack_density(scene).to(dev), scene["features"].to(dev)
        raster = native.SplatRaster(abi.default_config(), device=dev)
        raster.set_training(False)
        ms = timed(lambda: raster.trace(0, 3, d12, sph, ro, rd, None, cam), args.iters)
        fs = raster.frame_stats()
        rows.append(f"| C5: forward-only sweep | {n} | 1920x1080 | fwd | {ms:.3f} | {1e3 / ms:.0f} frames/s ({1920 * 1080 / ms / 1e3:.0f} Mpix/s) | "
                    f"M = {fs['num_intersections']}, max list {fs['max_tile_list_length']} |")
        del raster, d12, sph

    with open(args.out, "w") as fh:
        fh.write("| config | particles | image | pass | ms (median) | throughput | work, stage ms |\n|---|---|---|---|---|---|---|\n" + "\n".join(rows) + "\n")
    print("\n".join(rows))


if __name__ == "__main__":
    main()
