#!/usr/bin/env python
"""bench.py -- 3DGUT fwd+bwd throughput (Mpix/s) of the B200-native rasteriser.

Workload (BASELINE.json metric "3DGUT fwd+bwd Mpix/s at 3M Gaussians, 1/2/4/8 B200", configs[2]):
one training step = forward + backward of a batch of 8 synthetic 1237x822 pinhole views over a replicated
set of 3M random-init Gaussians (SH degree 3), the views sharded over the N ranks, per-Gaussian gradients
summed with one NCCL allreduce.  Total work per step is fixed, so scaling is "strong".

    python bench.py --gpus 1 --steps K --warmup W            # this implementation
    python bench.py --impl reference ...                     # CPU oracle port on the host cores (the reference has no CPU path)

Prints ONE JSON line on rank 0 (see the task contract for the keys).
"""
import argparse
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "3dgut_fwd_bwd_mpix_per_s"
UNIT = "Mpix/s"
W_IMG, H_IMG = 1237, 822
N_VIEWS = 8
SPH_DEGREE = 3
# algorithmic work per unit (restated in DESIGN.md)
FLOP_FWD_TEST, FLOP_FWD_HIT = 71.0, 32.0
FLOP_BWD_TEST, FLOP_BWD_HIT = 100.0, 250.0


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--particles", type=int, default=3_000_000, help="override only for smoke runs; the metric is quoted at 3M")
    p.add_argument("--views", type=int, default=N_VIEWS)
    p.add_argument("--width", type=int, default=W_IMG)
    p.add_argument("--height", type=int, default=H_IMG)
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--colour-slots", choices=("auto", "dense", "compact"), default="auto",
                   help="layout of the peer exchange's per-view colour-gradient slots (auto = the exchange's default: dense)")
    p.add_argument("--lanes", type=int, default=2, help="handles / CUDA streams a rank deals its views to (1 = one stream)")
    p.add_argument("--exchange", choices=("auto", "peer", "deferred", "allreduce"), default="auto",
                   help="gradient exchange of the step: deferred = allreduce [N,12] + all-gather per-view colour gradients, SH expansion "
                        "once per step (3dgrut_b200/dist.py training_step_deferred); peer = the same with the all-gather folded into the "
                        "expansion kernel (loads from the other ranks' buffers over NVLink, training_step_peer); allreduce = per-view "
                        "[N,48] accumulation + one allreduce of [N,12]+[N,48]; auto = peer on more than one GPU, deferred on one")
    p.add_argument("--no-aux", action="store_true", help="skip the auxiliary measurements (sorted K=16, model-side passes, reference + shim)")
    p.add_argument("--no-sweep", action="store_true", help="skip the C5 forward-only 1080p sweep over 0.5M-6M particles")
    p.add_argument("--cpu-sample-tiles", type=int, default=24)
    return p.parse_args()


def workload_config(args, world):
    exchange = ("peer" if world > 1 else "deferred") if args.exchange == "auto" else args.exchange
    return {
        "workload": f"3DGUT training step: {args.particles} random-init Gaussians (S-init), SH deg {SPH_DEGREE}, batch of {args.views} "
                    f"pinhole views {args.width}x{args.height}, k_buffer_size=0, views sharded over {world} GPU(s), "
                    + ("NCCL allreduce of the [N,12] fp32 gradients + all-gather of the per-view colour gradients [N,3], SH gradient "
                       "[N,48] expanded once per step on every rank" if exchange == "deferred" else
                       "NCCL allreduce of the [N,12] fp32 gradients; per-view colour gradients [N,3] read from the other ranks' buffers over "
                       "NVLink inside the kernel that expands the SH gradient [N,48] (peer memory, no all-gather pass)"
                       if exchange == "peer" else "NCCL allreduce of [N,12]+[N,48] fp32 gradients"),
        "particles": args.particles, "views_per_step": args.views, "width": args.width, "height": args.height,
        "sph_degree": SPH_DEGREE, "parallelism": f"view-sharded dp{world}, particles replicated",
        "l2_policy": "inputs larger than L2: every view streams the particle parameters and SH (3M x 240 B = 720 MB), the [N,16] gradient "
                     "accumulators (192 MB) and the [N,12]+[N,48] gradients (720 MB) against a 126 MB L2",
    }


def ncu_traffic(stage):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's kernel, from the committed `ncu --set full` capture
    (profiles/traffic.json, written by scripts/ncu_summarise.py); None when no capture is committed."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return t["stages"][stage]["dram_bytes_per_launch"]
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.sm_max = None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.sm_max = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            }
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.05)
        except Exception as e:  # NVML unavailable: report that instead of failing the bench
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# ------------------------------------------------------------------------------------------------------
# CPU oracle arm (cpu_baseline of our line, and the whole --impl reference line)
# ------------------------------------------------------------------------------------------------------
def cpu_oracle_sample(args, n_tiles_sample, repeats=1):
    """fwd+bwd of ONE view of the workload with the C oracle on all host cores, compositing restricted to a
    sample of tiles (projection, binning and SH backward run over the whole frame).  Returns Mpix/s extrapolated
    as sample_pixels / (t_composite_sample + t_whole_frame_stages * sample_pixels / frame_pixels)."""
    from oracle import gut_oracle as O
    from tests.helpers import abi, camera_of, syn

    import ctypes as C

    # torchrun exports OMP_NUM_THREADS=1 to its workers: set the thread count explicitly and report what the library got
    want = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    O.set_threads(want)
    threads = O.threads_used()
    cfg = abi.default_config()
    w, h, n = args.width, args.height, args.particles
    scene = syn.make_scene(n, 3, seed=1003)
    view = syn.make_view(w, h, k=0, n_views=args.views)
    cam = camera_of(view)
    d12 = syn.pack_density(scene).numpy()
    sph = scene["features"].numpy()
    ro, rd = view["rays_ori"].numpy(), view["rays_dir"].numpy()
    gx, gy = (w + 15) // 16, (h + 15) // 16
    rng = np.random.default_rng(0)
    tiles = np.sort(rng.choice(gx * gy, size=min(n_tiles_sample, gx * gy), replace=False)).astype(np.uint32)
    tx, ty = tiles % gx, tiles // gx
    sample_pixels = int(((np.minimum((tx + 1) * 16, w) - tx * 16) * (np.minimum((ty + 1) * 16, h) - ty * 16)).sum())
    rgba_grad = rng.uniform(-1, 1, (h, w, 4)).astype(np.float32)
    dist_grad = rng.uniform(-0.01, 0.01, (h, w)).astype(np.float32)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        proj = O.project(cfg, cam, d12, sph, w, h, SPH_DEGREE)
        bins = O.bin_and_sort(cfg, w, h, proj)
        t1 = time.perf_counter()
        out = O.render(cfg, cam, d12, proj["features"], ro, rd, w, h, bins["tile_ranges"], bins["sorted_values"], tile_list=tiles)
        out.update(proj=proj, bins=bins)
        t2 = time.perf_counter()
        O.backward(cfg, cam, d12, sph, ro, rd, w, h, SPH_DEGREE, out, rgba_grad, dist_grad, tile_list=tiles)
        t3 = time.perf_counter()
        whole, comp = (t1 - t0), (t3 - t1)
        frac = sample_pixels / float(w * h)
        est = sample_pixels / (comp + whole * frac) / 1e6
        rec = {"mpix_s": est, "t_whole_frame_stages_s": whole, "t_composite_sample_s": comp, "tests": out["tests"], "M": bins["M"]}
        if best is None or rec["mpix_s"] > best["mpix_s"]:
            best = rec
    best["cores"] = threads
    best["host_cpus"] = os.cpu_count()
    best["sample"] = (f"one {w}x{h} view of the workload: projection+binning+SH-backward over all {n} particles, compositing fwd+bwd on "
                      f"{len(tiles)} random tiles ({sample_pixels} px) with OpenMP on {threads} host threads (counted inside the library); "
                      f"extrapolated per pixel: sample_px / (t_composite_sample + t_whole_frame_stages * sample_px / frame_px)")
    best["sample_pixels"] = sample_pixels
    return best


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    vals = []
    last = None
    for _ in range(max(1, args.warmup // 3)):
        cpu_oracle_sample(args, max(4, args.cpu_sample_tiles // 4))
    t_begin = time.perf_counter()
    for _ in range(args.steps):
        last = cpu_oracle_sample(args, args.cpu_sample_tiles)
        vals.append(last["mpix_s"])
    total = time.perf_counter() - t_begin
    v = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / max(1, args.steps), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": last["cores"], "kind": "port", "sample": last["sample"],
                         "t_whole_frame_stages_s": last["t_whole_frame_stages_s"], "t_composite_sample_s": last["t_composite_sample_s"],
                         "sample_pixels": last["sample_pixels"], "host_cpus": last["host_cpus"]},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference ships no CPU implementation of this path (SURVEY.md 8c); this arm times the C oracle port of it "
                "(pinned on the reference's own kernels: tests/test_reference_golden.py, tests/test_gpu_reference.py)",
    }
    if not args.no_aux:
        try:
            line["reference_shim_gpu"] = reference_shim_gpu(args)
        except Exception as e:   # no GPU in this process / oracle/_ref not built: the line stands without it
            line["reference_shim_gpu"] = {"unavailable": repr(e)[:300]}
    print(json.dumps(line), flush=True)


def reference_shim_gpu(args):
    """Context figure for the reference arm only (never loaded by the GPU arm of this file): the reference's OWN CUDA tracer --
    its unchanged sources built against the hand-written threedgutSlang.cuh of oracle/ref_build ("reference + shim",
    oracle/_ref/lib3dgut_cc_k0.so, through the reference's pybind API) -- forward + backward of ONE view of the workload on
    cuda:0, timed with CUDA events after warm-up."""
    import torch

    from tests import refshim
    from tests.helpers import camera_of, syn

    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device")
    if not refshim.available(0):
        raise RuntimeError("oracle/_ref/lib3dgut_cc_k0.so not built")
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    w, h, n = args.width, args.height, args.particles
    scene = syn.make_scene(n, 3, seed=1003)
    view = syn.make_view(w, h, k=0, n_views=args.views)
    cam = camera_of(view)
    d12 = syn.pack_density(scene).to(dev); sph = scene["features"].to(dev)
    ro, rd = view["rays_ori"].to(dev), view["rays_dir"].to(dev)
    g = torch.Generator().manual_seed(2000)
    rg = (torch.rand((h, w, 4), generator=g) * 2 - 1).to(dev)
    dg = ((torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01).to(dev)
    ref = refshim.ReferenceRaster(0, enable_kernel_timings=True)

    def one():
        rgba, dist, hits, vis = ref.trace(SPH_DEGREE, d12, sph, ro, rd, cam)
        ref.trace_bwd(SPH_DEGREE, d12, sph, ro, rd, cam, rgba, rg, dist, dg)

    for _ in range(2):
        one()
    torch.cuda.synchronize()
    ref.collect_times()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        one()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    t = ref.collect_times()
    return {"label": "reference + shim (the reference's CUDA sources unchanged, Slang-generated header hand-written; K=0)",
            "ms_per_view_fwd_bwd": ms, "mpix_per_s": w * h / (ms * 1e-3) / 1e6, "forward_render_ms": t.get("forward_render"),
            "backward_render_ms": t.get("backward_render"), "particles": n, "width": w, "height": h,
            "includes": "torch.zeros of the outputs and of the [N,12]+[N,48] gradients per call, as SplatRaster::trace / traceBwd do"}


# ------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    from tests.helpers import SyntheticGaussians, abi, camera_of, make_batch, make_conf, syn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    native = importlib.import_module("3dgrut_b200.native")
    gdist = importlib.import_module("3dgrut_b200.dist")
    pkg = importlib.import_module("3dgrut_b200")

    n, w, h = args.particles, args.width, args.height
    P = w * h
    scene = syn.make_scene(n, 3, seed=1003)
    d12 = syn.pack_density(scene).to(dev)
    sph = scene["features"].to(dev)
    my_view_ids = gdist.shard_views(args.views, rank, world)
    host_views = {k: syn.make_view(w, h, k=k, n_views=args.views) for k in my_view_ids}
    views, upstream, host_up = [], [], []
    for k in my_view_ids:
        hv = host_views[k]
        views.append({"rays_ori": hv["rays_ori"].to(dev), "rays_dir": hv["rays_dir"].to(dev), "camera": camera_of(hv)})
        g = torch.Generator().manual_seed(2000 + k)
        rg = torch.rand((h, w, 4), generator=g) * 2 - 1
        dg = (torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01
        host_up.append((rg.pin_memory(), dg.pin_memory()))
        upstream.append((rg.to(dev), dg.to(dev)))

    cfg = abi.default_config(enable_kernel_timings=1)
    raster = native.SplatRaster(cfg, device=dev)
    # a second handle ("lane", 3dgrut_b200/dist.py _lanes): consecutive views of a rank's share run on two streams, so that the
    # latency-bound kernels of one view sit under the compositing kernels of the other.  Nothing to overlap with one view per rank.
    n_my_views = len(my_view_ids)
    lanes = [raster] + ([native.SplatRaster(cfg, device=dev)] if (args.lanes > 1 and n_my_views > 1) else [])
    if args.exchange == "auto":
        args.exchange = "peer" if world > 1 else "deferred"
    # colour-gradient slots of the peer exchange: dense [N,3], or compact (touched rows only: less NVLink traffic, one more dependent load)
    compact_slots = {"auto": None, "dense": False, "compact": True}[args.colour_slots]
    if args.exchange == "peer":
        try:
            exchange = gdist.PeerExchange(raster, n, sph.shape[1], (args.views + world - 1) // world, dev, rank, world=world,
                                          compact=compact_slots)
            ok = 1
        except RuntimeError as e:   # CUDA IPC unavailable on this box: the constructor raises on every rank together
            print(f"[bench] rank {rank}: peer buffers unavailable ({e!r})", file=sys.stderr, flush=True)
            exchange, ok = None, 0
        if not ok:   # PeerExchange fails on every rank or on none
            args.exchange = "deferred"
    if args.exchange == "deferred":
        exchange = gdist.DeferredExchange(n, sph.shape[1], (args.views + world - 1) // world, dev, world=world)
    elif args.exchange == "allreduce":
        exchange = gdist.GradientExchange(n, sph.shape[1], dev, world=world)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def step(use=None):
        use = lanes if use is None else use
        if args.exchange == "peer":
            return gdist.training_step_peer(use, exchange, views, SPH_DEGREE, d12, sph, upstream, rank=rank)
        if args.exchange == "deferred":
            return gdist.training_step_deferred(use, exchange, views, SPH_DEGREE, d12, sph, upstream, rank=rank)
        return gdist.training_step(raster, exchange, views, SPH_DEGREE, d12, sph, upstream)

    for _ in range(args.warmup):
        step()
    barrier()
    raster.stage_times()  # drop warm-up timings
    launches0 = sum(r.frame_stats()["kernel_launches"] for r in lanes)
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    launches = sum(r.frame_stats()["kernel_launches"] for r in lanes) - launches0
    stage_ms = raster.stage_times()
    ms_one_lane = None
    if len(lanes) > 1:
        # With two lanes kernels of different views overlap and a kernel's CUDA events measure the contention too: the per-stage
        # times (roofline, stage_ms) come from the same steps run on ONE lane right here, and that step time is reported beside
        # the timed one.
        for _ in range(2):
            step([raster])
        barrier()
        raster.stage_times()
        o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        o0.record()
        for _ in range(args.steps):
            step([raster])
        o1.record()
        barrier()
        to = torch.tensor([o0.elapsed_time(o1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(to, op=dist.ReduceOp.MAX)
        ms_one_lane = float(to.item()) / args.steps
        stage_ms = raster.stage_times()
    # what a step spends outside its views' kernels (gradient conversion, exchange, waiting for the slowest rank): the step time
    # minus this rank's per-view stage times; and, untimed, three more steps with CUDA events between the phases of the exchange
    step_tail_ms = (ms_one_lane if ms_one_lane is not None else ms_total / args.steps) - len(views) * sum(stage_ms.values())
    exchange_phases = None
    if hasattr(exchange, "phase_times"):
        exchange.profile = {}
        for _ in range(3):
            step()
        exchange_phases = exchange.phase_times()
        exchange.profile = None
        raster.stage_times()
    if world > 1:
        tails = [None] * world
        dist.all_gather_object(tails, {"rank": rank, "views": len(views), "views_ms": len(views) * sum(stage_ms.values()),
                                       "tail_ms": step_tail_ms, "exchange_phases_ms": exchange_phases})
    else:
        tails = [{"rank": 0, "views": len(views), "views_ms": len(views) * sum(stage_ms.values()), "tail_ms": step_tail_ms,
                  "exchange_phases_ms": exchange_phases}]
    stats = raster.frame_stats()
    value = args.views * P * args.steps / (ms_total * 1e-3) / 1e6

    # ---- roofline of the dominant kernel (rank 0's views) ----------------------------------------------
    tests = hits = tests_exec = 0
    m_sum = 0
    for v in views:
        raster.trace(0, SPH_DEGREE, d12, sph, v["rays_ori"], v["rays_dir"], None, v["camera"])
        # algorithmic work = what the reference's walk does: every list entry of the tile for every live ray.  The forward's
        # conservative block cull (gut_render_common.cuh, stage_culled) skips part of it; the executed count is reported too.
        raster.set_block_cull(False)
        t_, h_ = raster.measure_work(v["rays_ori"], v["rays_dir"])
        raster.set_block_cull(True)
        te_, _ = raster.measure_work(v["rays_ori"], v["rays_dir"])
        tests += t_; hits += h_; tests_exec += te_
        m_sum += raster.frame_stats()["num_intersections"]
    nv = max(1, len(views))
    tests_v, hits_v, m_v = tests / nv, hits / nv, m_sum / nv
    clk = clocks["sm_mhz"] or 1965.0
    fp32_peak = 148 * 128 * 2 * clk * 1e6 / 1e12
    hbm_peak = 6555.2
    try:
        hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        hbm_peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    vis_frac = 1.0
    stages = []

    def add_stage(tag, bound, work, unit_scale, peak, unit, what):
        ms = stage_ms.get(tag)
        if ms is None or ms <= 0:
            return
        ach = work / (ms * 1e-3) / unit_scale
        stages.append({"stage": tag, "bound": bound, "ms_per_launch": ms, "achieved": ach, "peak": peak, "unit": unit,
                       "frac": ach / peak, "algorithmic_work": what})

    add_stage("render::project", "hbm", n * (48 + 72 + 192 * vis_frac), 1e9, hbm_peak, "GB/s",
              "48 B parameters + 192 B SH in, 72 B projected record out per particle")
    add_stage("render::expand", "hbm", n * 16 + m_v * 8, 1e9, hbm_peak, "GB/s", "16 B culled-box record per particle + 8 B per intersection")
    add_stage("render::sort", "hbm", m_v * (8 + 4), 1e9, hbm_peak, "GB/s",
              "8 B bucket read + 4 B sorted-list write per intersection (the counting pass runs in shared memory)")
    add_stage("render::render", "fp32", tests_v * FLOP_FWD_TEST + hits_v * FLOP_FWD_HIT, 1e12, fp32_peak, "TFLOP/s",
              "71 FLOP/test + 32 FLOP/hit")
    add_stage("render-backward::render", "fp32", tests_v * FLOP_BWD_TEST + hits_v * FLOP_BWD_HIT, 1e12, fp32_peak, "TFLOP/s",
              "100 FLOP/test + 250 FLOP/hit")
    if args.exchange in ("deferred", "peer"):
        # rows the view's rays pushed on: those with a colour gradient in the view's slot of the exchange (left by the timed steps)
        moved = 0.0
        for i_ in range(len(views)):
            moved += float(exchange.moved_rows(i_))
        moved_v = moved / nv
        if exchange.views_per_rank == 1:
            add_stage("render-backward::project", "hbm", n * (48 + 4 + 16 + 64 + 48 + 12) + moved_v * 64, 1e9, hbm_peak, "GB/s",
                      "one view per rank (gut_project_backward_kernel<density, colour>): 48 B parameters + 4 B tile count + 16 B colour + 64 B "
                      "accumulators in, 48 B [N,12] gradient + 12 B colour gradient out per particle; 64 B of accumulators reset per particle "
                      "the view's rays pushed on")
        else:
            add_stage("render-backward::project", "hbm", n * (16 + 48 + 12) + moved_v * (80 + 80), 1e9, hbm_peak, "GB/s",
                      "per view (gut_view_fold_kernel): 16 B colour + three 16 B accumulator vectors in, 12 B colour gradient out per particle; "
                      "per particle the view's rays pushed on: + 48 B parameters, 16 B accumulators, 16 B A sum in and 80 B of accumulators / "
                      "A sum out (the [N,12] / [N,48] conversions run once per step, outside the per-view stages)")
    else:
        add_stage("render-backward::project", "hbm", n * (48 + 16 + 64 + 48 + 192), 1e9, hbm_peak, "GB/s",
                  "48 B params + 16 B colour + 64 B accumulators in, 48 + 192 B gradients out per particle (accumulation re-reads not counted)")
    # The unit that actually binds each kernel, with ncu's own percent-of-peak for it (profiles/ncu_full_r02d.md, one view of this
    # workload).  Static annotations of a committed capture, not measured in this run: for K1, K2 and the sort
    # the HBM roofline above is the contract's yardstick, not their limiter.
    limiters = {
        "render::project": ("instruction issue (IEEE --fmad=false arithmetic for bit-exact tile counts)", "smsp__issue_active", 78.7),
        "render::expand": ("L2 request rate: one returning atomic + one scattered 8 B store per intersection", "lts__throughput", 63.5),
        "render::sort": ("shared memory: random-bin atomics and scatter of the per-tile counting sort", "l1tex__throughput", 67.8),
        "render::render": ("instruction issue", "smsp__issue_active", 75.5),
        "render-backward::render": ("FMA pipe (packed FP32 occupies it two cycles per instruction)", "sm__pipe_fma_cycles_active", 66.9),
        "render-backward::project": ("DRAM (per-view fold kernel; ncu r02c)", "gpu__dram_throughput", 71.1),
    }
    for st_ in stages:
        lim = limiters.get(st_["stage"])
        if lim:
            st_["limiter"] = {"unit": lim[0], "ncu_metric": lim[1], "pct_of_peak": lim[2], "source": "profiles/ncu_full_r02d.md"}
    dominant = max(stages, key=lambda s: s["ms_per_launch"]) if stages else None
    roofline = None
    if dominant:
        roofline = {"kernel": dominant["stage"], "bound": dominant["bound"], "achieved": dominant["achieved"], "peak": dominant["peak"],
                    "unit": dominant["unit"], "frac": dominant["frac"], "traffic": ncu_traffic(dominant["stage"]),
                    "ms_per_launch": dominant["ms_per_launch"],
                    "peak_source": (f"derived: 148 SM x 128 lanes x 2 x {clk:.0f} MHz sampled under load (FP32 is not in MEASURED_PEAKS.json)"
                                    if dominant["bound"] == "fp32" else peak_src),
                    "note": "compositing is FP32-issue bound, not HBM or tensor bound (no stage is a dense contraction); "
                            "HBM-bound stages are listed in roofline_stages"}

    # ---- end to end through the public Tracer API with host inputs --------------------------------------
    e2e = None
    if not args.no_e2e:
        tracer = pkg.Tracer(make_conf())
        gauss = SyntheticGaussians(scene, dev, n_active_features=SPH_DEGREE, requires_grad=True)
        pinned = []
        for k in my_view_ids:
            hv = host_views[k]
            pinned.append({kk: (vv.pin_memory() if hasattr(vv, "pin_memory") else vv) for kk, vv in hv.items()})
        out_host = torch.empty((h, w, 4), dtype=torch.float32).pin_memory()
        h2d = d2h = 0
        # Inputs of view k+1 are copied host->device on a copy stream while view k computes (two device buffer sets);
        # the rendered RGBA goes back as ONE contiguous device->host copy per view.  All copies are inside the timed region.
        copy_stream = torch.cuda.Stream(device=dev)     # host -> device
        back_stream = torch.cuda.Stream(device=dev)     # device -> host: its own stream, the two directions have their own engines
        nbuf = 3   # a set is free again only when its view's BACKWARD has run (two views later with two lanes)
        dev_in = [{"rays_ori": torch.empty((1, h, w, 3), device=dev), "rays_dir": torch.empty((1, h, w, 3), device=dev),
                   "T_to_world": torch.empty((1, 4, 4), device=dev), "rg": torch.empty((h, w, 4), device=dev),
                   "dg": torch.empty((h, w, 1), device=dev)} for _ in range(nbuf)]
        rgba_dev = [torch.empty((h, w, 4), device=dev) for _ in range(nbuf)]
        free_ev = [None] * nbuf   # recorded on the compute stream when a buffer set may be overwritten
        back_ev = [None] * nbuf   # recorded on the read-back stream when an image buffer has been copied out
        prefetched = []
        loss_slots = [{"host": torch.zeros(1, dtype=torch.float32).pin_memory(), "event": None} for _ in range(2)]
        loss_turn = [0]

        seq = [0]   # views staged so far, over all steps: the buffer sets alternate along this sequence, so that the inputs of
                    # the NEXT view -- of this step or, for its last view, of the next step -- travel while the current one renders

        def stage(vi):
            nonlocal h2d
            bi = seq[0] % nbuf
            seq[0] += 1
            pv, (rg_h, dg_h) = pinned[vi], host_up[vi]
            with torch.cuda.stream(copy_stream):
                if free_ev[bi] is not None:
                    copy_stream.wait_event(free_ev[bi])
                for kk in ("rays_ori", "rays_dir"):
                    dev_in[bi][kk].copy_(pv[kk], non_blocking=True)
                    h2d += pv[kk].numel() * 4
                dev_in[bi]["rg"].copy_(rg_h, non_blocking=True)
                dev_in[bi]["dg"].copy_(dg_h, non_blocking=True)
                h2d += (rg_h.numel() + dg_h.numel()) * 4
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            return ev, bi

        # the step's gradient exchange is the one the device-resident loop above used (peer memory / NCCL), now driven by
        # Tracer.begin_step ... finish(); a plain per-tensor allreduce has no step protocol, so that mode takes the deferred one
        e2e_exchange = exchange if args.exchange in ("peer", "deferred") else gdist.DeferredExchange(
            n, sph.shape[1], (args.views + world - 1) // world, dev, world=world)

        def e2e_step():
            nonlocal h2d, d2h
            h2d = d2h = 0
            for p in (gauss.positions, gauss._rotation, gauss._scale, gauss._density, gauss._features):
                p.grad = None
            checksum = None
            tstep = tracer.begin_step(gauss, exchange=e2e_exchange, rank=rank)
            cur = torch.cuda.current_stream()
            # the inputs of this step's first view were put in flight at the end of the previous step (a data loader's prefetch)
            ready = prefetched.pop() if prefetched else (stage(0) if pinned else None)
            for vi in range(len(pinned)):
                ready_ev, bi = ready
                cur.wait_event(ready_ev)
                # next view in flight now: the following view of this step, or the first view of the next step (a data loader's
                # prefetch); stage() waits for its buffer set's free event
                if vi + 1 < len(pinned):
                    ready = stage(vi + 1)
                else:
                    prefetched.append(stage(0))
                # the pose stays a host tensor: the rasteriser consumes it on the host (kernel parameters), and a device copy
                # would be read straight back with a stream-wide synchronisation (the reference does that, tracer.py:376)
                dv = dict(pinned[vi])
                for kk in ("rays_ori", "rays_dir"):
                    dv[kk] = dev_in[bi][kk]
                rg, dg = dev_in[bi]["rg"], dev_in[bi]["dg"]
                o = tstep.render(make_batch(dv, dev, pose_device="cpu"))
                # (batch axis kept: indexing it away costs a zero-filled image and a copy per output in the adjoint)
                loss = (o["pred_rgb"] * rg[None, ..., :3]).sum() + (o["pred_opacity"] * rg[None, ..., 3:]).sum() + (o["pred_dist"] * dg[None]).sum()
                # loss.backward() with the view's lane as ambient stream (TracerStep.backward); the event fires behind the view's
                # backward kernels: only then may this buffer set (rays, upstream gradients) be overwritten
                free_ev[bi] = tstep.backward(loss)
                if back_ev[bi] is not None:
                    cur.wait_event(back_ev[bi])       # the read-back that last used this image buffer has ended
                torch.cat([o["pred_rgb"][0].detach(), o["pred_opacity"][0].detach()], dim=-1, out=rgba_dev[bi])
                done = torch.cuda.Event()
                done.record(cur)
                with torch.cuda.stream(back_stream):
                    back_stream.wait_event(done)
                    out_host.copy_(rgba_dev[bi], non_blocking=True)
                    back_ev[bi] = torch.cuda.Event()
                    back_ev[bi].record(back_stream)
                d2h += out_host.numel() * 4
                checksum = loss.detach() if checksum is None else checksum + loss.detach()
            # gradients of all views (and of all ranks) -> .grad of the parameters: one [N,12] conversion, the exchange, one [N,48]
            # expansion per step
            tstep.finish()
            # the read-back of the last view's image may still be in flight (it is ordered before the next step's read-backs on the
            # copy stream); the timed region ends with cur.wait_stream(copy_stream) below
            if checksum is not None:
                # device->host read of the step's scalar result, the way a training loop logs it: copied to pinned memory behind
                # the step, consumed one step later (the previous step's value is read here), so the host keeps queueing work
                prev, slot = loss_slots[(loss_turn[0] + 1) % 2], loss_slots[loss_turn[0] % 2]
                loss_turn[0] += 1
                slot["host"].copy_(checksum.reshape(1), non_blocking=True)
                slot["event"] = torch.cuda.Event()
                slot["event"].record(cur)
                if prev["event"] is not None:
                    prev["event"].synchronize()
                    float(prev["host"][0])
                d2h += 4

        for _ in range(min(args.warmup, 3)):
            e2e_step()
        if os.environ.get("GUT_BENCH_E2E_TRACE"):   # developer hook: kernel table of one end-to-end step (rank 0 writes it)
            from torch.profiler import ProfilerActivity, profile
            if rank == 0:
                with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
                    e2e_step()
                    torch.cuda.synchronize()
                with open(os.environ["GUT_BENCH_E2E_TRACE"], "w") as fh:
                    fh.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=60, max_name_column_width=90))
            else:
                e2e_step()   # the step holds collectives: every rank runs it
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        k_e2e = args.steps
        for _ in range(k_e2e):
            e2e_step()
        torch.cuda.current_stream().wait_stream(back_stream)   # every read-back of the timed steps has landed (the host->device
        torch.cuda.current_stream().wait_stream(copy_stream)   # copies too: one view of a step after the timed ones is in flight)
        f1.record()
        barrier()
        for slot in loss_slots:                                 # the last steps' results (their copies ended before f1)
            float(slot["host"][0])
        ms = f0.elapsed_time(f1)
        tt = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e = {"value": args.views * P * k_e2e / (float(tt.item()) * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h),
               "path": "Tracer.begin_step / step.render + autograd backward per view / step.finish (the multi-view entry of the plugin: "
                       "gradient conversions and the data-parallel exchange once per step); rays and upstream gradients copied from pinned host memory "
                       "(copy stream, one view ahead -- across step boundaries too, like a data loader's prefetch; read-backs on a second stream), the 4x4 pose handed over as a host tensor (it becomes kernel parameters), rendered RGBA of every view and the scalar loss read back every step (the loss through pinned memory, consumed one step later); particles stay "
                       "device resident (they are the model)"}

    # ---- auxiliary: forward-only frame time at 1920x1080 (second half of BASELINE.json's metric; reference convention:
    # CUDA events around the forward only, splatRaster.cpp:208-241 / threedgrut/render.py:241-251) -------------------
    fwd_1080p = None
    if rank == 0 and not args.no_aux:
        hv = syn.make_view(1920, 1080, k=0, n_views=args.views)
        ro, rd, cam = hv["rays_ori"].to(dev), hv["rays_dir"].to(dev), camera_of(hv)
        raster.set_training(False)   # inference: no backward follows (Tracer.render does this under torch.no_grad())
        for _ in range(3):
            raster.trace(0, SPH_DEGREE, d12, sph, ro, rd, None, cam)
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        a0.record()
        for _ in range(reps):
            raster.trace(0, SPH_DEGREE, d12, sph, ro, rd, None, cam)
        a1.record()
        torch.cuda.synchronize()
        ms = a0.elapsed_time(a1) / reps
        fwd_1080p = {"fps": 1e3 / ms, "ms_per_frame": ms, "particles": n, "intersections_M": raster.frame_stats()["num_intersections"]}
        raster.set_training(True)

    # ---- C5 (BASELINE.json configs[4], metric "fwd-only FPS @1080p"): forward-only render sweep 0.5M-6M Gaussians, 1080p views
    # dealt round-robin to the ranks, no collective (render.py eval throughput).  Reference timing convention: CUDA events around
    # the forward only (splatRaster.cpp:208-241, threedgrut/render.py:231-259), frames per second over the whole job = views /
    # (max over ranks of the time a rank needs for its views). ---------------------------------------------------------------
    fwd_sweep = None
    if not args.no_sweep:
        fwd_sweep = {"width": 1920, "height": 1080, "views": args.views, "sph_degree": SPH_DEGREE, "n_gpus": world,
                     "convention": "CUDA events around gut_forward only, inference mode (no hit-mask recording), inputs device resident",
                     "by_particles": {}}
        sweep_views = []
        for k in my_view_ids:
            hv = syn.make_view(1920, 1080, k=k, n_views=args.views)
            sweep_views.append((hv["rays_ori"].to(dev), hv["rays_dir"].to(dev), camera_of(hv)))
        raster.set_training(False)
        for n_s in (500_000, 1_000_000, 2_000_000, 3_000_000, 6_000_000):
            if n_s == n:
                d12_s, sph_s = d12, sph
            else:
                sc = syn.make_scene(n_s, 3, seed=1003)
                d12_s, sph_s = syn.pack_density(sc).to(dev), sc["features"].to(dev)
                del sc
            for ro_, rd_, cam_ in sweep_views:      # warm-up: grows the workspace to this size
                raster.trace(0, SPH_DEGREE, d12_s, sph_s, ro_, rd_, None, cam_)
            barrier()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 3
            a0.record()
            m_tot = 0
            for _ in range(reps):
                for ro_, rd_, cam_ in sweep_views:
                    raster.trace(0, SPH_DEGREE, d12_s, sph_s, ro_, rd_, None, cam_)
                    m_tot += raster.frame_stats()["num_intersections"]
            a1.record()
            barrier()
            ms_rank = a0.elapsed_time(a1) / reps
            tt = torch.tensor([ms_rank], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms_job = float(tt.item())
            fwd_sweep["by_particles"][str(n_s)] = {
                "fps": args.views / (ms_job * 1e-3) if ms_job > 0 else None, "ms_per_frame_on_a_gpu": ms_rank / max(1, len(sweep_views)),
                "intersections_M_per_frame": m_tot / max(1, reps * len(sweep_views)) / 1e6}
            if n_s != n:
                del d12_s, sph_s
        raster.set_training(True)
        del sweep_views

    # ---- auxiliary: the WHOLE training step -- raw parameters -> activation + packing -> the views' forward and
    # backward -> gradient exchange -> SelectiveAdam on all six groups + densification statistics -- with the optimiser as separate
    # kernels behind the exchange (dist.optimizer_step) and as the epilogue of the gradient kernels (dist.optimizer_epilogue).
    # Every rank takes part (the step holds collectives). ----------------------------------------------------------------
    train_step = None
    if not args.no_aux and (args.views + world - 1) // world * world <= 16:
        sc = {k: v.to(dev) for k, v in scene.items()}   # the raw parameters a MixtureOfGaussians would hold for this scene
        rawd = {"positions": sc["positions"].clone(), "rotation": (sc["rotation"] * 1.7).contiguous(), "scale": torch.log(sc["scale"]),
                "density": torch.log(sc["density"] / (1 - sc["density"])), "features_albedo": sc["features"][:, :3].contiguous(),
                "features_specular": sc["features"][:, 3:].contiguous()}
        state = {k: (torch.zeros_like(v), torch.zeros_like(v)) for k, v in rawd.items()}
        # learning rates 0: the update runs in full (moments change, parameters are read and written) but the scene stays the one
        # the gradient-only figure was measured on, so the three step times are comparable
        lrs = {k_: 0.0 for k_ in gdist.PARAMETER_GROUPS}
        accum, denom = torch.zeros((n, 1), device=dev), torch.zeros((n, 1), device=dev)
        vpr = (args.views + world - 1) // world
        ex2 = None
        try:
            if args.exchange == "peer":
                ex2 = gdist.PeerExchange(raster, n, sph.shape[1], vpr, dev, rank, world=world, track_visibility=True, compact=compact_slots)
            else:
                ex2 = gdist.DeferredExchange(n, sph.shape[1], vpr, dev, world=world, track_visibility=True)
        except RuntimeError as e:
            print(f"[bench] rank {rank}: train_step measurement skipped ({e!r})", file=sys.stderr, flush=True)
        if ex2 is not None:
            sensor = [0.0, 0.0, 4.0]

            def whole_step(fused):
                packed = raster.activate_pack(rawd["positions"], rawd["rotation"], rawd["scale"], rawd["density"])
                sphp = raster.pack_features(rawd["features_albedo"], rawd["features_specular"])
                gdist.training_step_deferred(lanes, ex2, views, SPH_DEGREE, packed, sphp, upstream, rank=rank, finish=not fused)
                if fused:
                    ex2.finish_step(lanes, SPH_DEGREE, packed, expand=False)
                    gdist.optimizer_epilogue(raster, ex2, rawd, packed, state, lrs, SPH_DEGREE, sensor_position=sensor,
                                             grad_norm_accum=accum, grad_norm_denom=denom)
                else:
                    gr = gdist.optimizer_step(raster, ex2, rawd, packed, state, lrs)
                    raster.densify_stats(rawd["positions"], gr["positions"], sensor, accum, denom)

            def time_steps(fused, reps=5):
                for _ in range(2):
                    whole_step(fused)
                barrier()
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
                for _ in range(reps):
                    whole_step(fused)
                t1.record()
                barrier()
                tt_ = torch.tensor([t0.elapsed_time(t1) / reps], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(tt_, op=dist.ReduceOp.MAX)
                return float(tt_.item())

            ms_unfused, ms_fused = time_steps(False), time_steps(True)
            train_step = {"what": "raw parameters -> activation + packing -> 8 views forward + backward -> exchange -> SelectiveAdam (6 groups, "
                                  "visibility-masked, learning rates 0 so that the scene stays the measured one) + densification statistics; device-resident inputs",
                          "ms_per_step_optimizer_as_separate_kernels": ms_unfused, "ms_per_step_optimizer_as_epilogue": ms_fused,
                          "mpix_per_s_optimizer_as_epilogue": args.views * P / (ms_fused * 1e-3) / 1e6,
                          "ms_per_step_gradients_only": ms_total / args.steps}
            if hasattr(ex2, "close"):
                ex2.close()
            raster.stage_times()

    # ---- auxiliary: the "sorted" configuration (k_buffer_size = 16, configs/paper/3dgut/base_sorted.yaml; SURVEY 8d's second
    # config): forward + backward of ONE view of the same workload, device resident ---------------------------------------
    sorted_k16 = None
    if rank == 0 and not args.no_aux and views:
        rk = native.SplatRaster(abi.default_config(enable_kernel_timings=1, k_buffer_size=16), device=dev)
        v0 = views[0]
        g_rgba = torch.rand((h, w, 4), device=dev) * 2 - 1
        g_dist = (torch.rand((h, w, 1), device=dev) * 2 - 1) * 0.01

        def step_k():
            o = rk.trace(0, SPH_DEGREE, d12, sph, v0["rays_ori"], v0["rays_dir"], None, v0["camera"])
            rk.trace_bwd(0, SPH_DEGREE, d12, sph, v0["rays_ori"], v0["rays_dir"], None, v0["camera"], 0, 1, None, None,
                         o[0], g_rgba, o[1], g_dist)
        step_k()
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 3
        a0.record()
        for _ in range(reps):
            step_k()
        a1.record()
        torch.cuda.synchronize()
        ms = a0.elapsed_time(a1) / reps
        sorted_k16 = {"mpix_per_s": P / (ms * 1e-3) / 1e6, "ms_per_view": ms, "stage_ms": rk.stage_times()}

    # ---- auxiliary: the model-side passes either side of the path at the workload's N: each is one
    # HBM-bound pass; algorithmic bytes per particle / measured time against the measured copy bandwidth -------------------
    model_side = None
    if rank == 0 and not args.no_aux:
        gen = torch.Generator(device=dev).manual_seed(5)
        raw = {"positions": d12[:, 0:3].contiguous(), "rotation": (d12[:, 4:8] * 1.7).contiguous(), "scale": torch.log(d12[:, 8:11]).contiguous(),
               "density": torch.logit(d12[:, 3:4]).contiguous(), "albedo": sph[:, :3].contiguous(), "specular": sph[:, 3:].contiguous()}
        g12 = torch.randn((n, 12), device=dev, generator=gen)
        g48 = torch.randn((n, 48), device=dev, generator=gen)
        packed = raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"])
        groups = []
        for k_, lr_ in (("positions", 1.6e-4), ("rotation", 1e-3), ("scale", 5e-3), ("density", 5e-2), ("albedo", 2.5e-3), ("specular", 1.25e-4)):
            p_ = raw[k_].clone()
            groups.append((p_, torch.randn(p_.shape, device=dev, generator=gen) * 1e-3, torch.zeros_like(p_), torch.zeros_like(p_), lr_))
        vis_all = torch.ones((n, 1), device=dev)
        stat_a, stat_d = torch.zeros((n, 1), device=dev), torch.zeros((n, 1), device=dev)

        def torch_chain():   # what the reference runs per forward+backward: model.py:74-93, tracer.py:176-178 and autograd of it
            leaves = {k_: v_.detach().requires_grad_(True) for k_, v_ in raw.items()}
            dn = torch.sigmoid(leaves["density"])
            pk = torch.cat([leaves["positions"], dn, torch.nn.functional.normalize(leaves["rotation"]), torch.exp(leaves["scale"]),
                            torch.zeros_like(dn)], dim=1)
            ft = torch.cat((leaves["albedo"], leaves["specular"]), dim=1)
            torch.autograd.backward([pk, ft], [g12, g48])

        def timed(fn, reps=5):
            fn()
            torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            for _ in range(reps):
                fn()
            a1.record()
            torch.cuda.synchronize()
            return a0.elapsed_time(a1) / reps

        passes = {
            "activate_pack": (lambda: raster.activate_pack(raw["positions"], raw["rotation"], raw["scale"], raw["density"]), 44 + 48),
            "activate_pack_backward": (lambda: raster.activate_pack_backward(raw["rotation"], packed, g12), 48 + 32 + 16 + 44),
            "pack_features": (lambda: raster.pack_features(raw["albedo"], raw["specular"]), 192 + 192),
            "unpack_feature_gradient": (lambda: raster.unpack_feature_gradient(g48), 192 + 192),
            "selective_adam_step(6 groups, all visible)": (lambda: raster.selective_adam_step(groups, vis_all, 0.9, 0.999, 1e-15), 59 * 28),
            "densify_stats": (lambda: raster.densify_stats(groups[0][0], groups[0][1], (0.0, 0.0, 4.0), stat_a, stat_d), 24 + 16),
        }
        model_side = {"particles": n, "hbm_peak_gbs": hbm_peak, "passes": {}}
        for name_, (fn_, bytes_) in passes.items():
            ms_ = timed(fn_)
            model_side["passes"][name_] = {"ms": ms_, "bytes_per_particle": bytes_, "gbs": n * bytes_ / (ms_ * 1e-3) / 1e9,
                                           "frac": n * bytes_ / (ms_ * 1e-3) / 1e9 / hbm_peak}
        fused_ms = sum(model_side["passes"][k_]["ms"] for k_ in ("activate_pack", "activate_pack_backward", "pack_features", "unpack_feature_gradient"))
        model_side["activations_fwd_bwd_ms"] = {"fused_kernels": fused_ms, "torch_chain_of_the_reference": timed(torch_chain)}
        del groups, g12, g48, packed, raw

    cpu_baseline = None
    if rank == 0 and not args.no_cpu_baseline:
        try:
            c = cpu_oracle_sample(args, args.cpu_sample_tiles if world == 1 else max(8, args.cpu_sample_tiles // 2))
            cpu_baseline = {"value": c["mpix_s"], "unit": UNIT, "cores": c["cores"], "kind": "port", "sample": c["sample"],
                            "t_whole_frame_stages_s": c["t_whole_frame_stages_s"], "t_composite_sample_s": c["t_composite_sample_s"],
                            "sample_pixels": c["sample_pixels"], "host_cpus": c["host_cpus"]}
        except Exception as e:
            cpu_baseline = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {e!r}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args, world), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "lanes": len(lanes), "colour_slots": "compact" if getattr(exchange, "compact", False) else "dense", "ms_per_step_one_lane": ms_one_lane, "step_breakdown": tails,
            "roofline": roofline, "roofline_stages": stages, "cpu_baseline": cpu_baseline,
            "work_per_view": {"intersections_M": m_v, "tiles": stats["num_tiles"], "mean_tile_list": m_v / max(1, stats["num_tiles"]),
                              "max_tile_list": stats["max_tile_list_length"], "tests_per_pixel": tests_v / P, "hits_per_pixel": hits_v / P,
                              "forward_tests_executed_per_pixel": tests_exec / nv / P},
            "stage_ms": stage_ms, "fwd_only_1080p": fwd_1080p, "fwd_sweep": fwd_sweep, "sorted_k16": sorted_k16, "model_side": model_side, "train_step": train_step,
        }
        print(json.dumps(line), flush=True)
    if args.exchange == "peer":
        exchange.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
