"""Experiment: the views of one step on TWO handles / TWO streams (the small latency-bound kernels of one view under the
compositing kernels of the other) against one stream.  Device-resident inputs, CUDA events.  Prints ms per 8-view step."""
import importlib, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.helpers import abi, camera_of, syn

native = importlib.import_module("3dgrut_b200.native")
dev = torch.device("cuda", 0)
n, w, h, nv = 3_000_000, 1237, 822, 8
scene = syn.make_scene(n, 3, seed=1003)
d12, sph = syn.pack_density(scene).to(dev), scene["features"].to(dev)
views, ups = [], []
for k in range(nv):
    hv = syn.make_view(w, h, k=k, n_views=nv)
    views.append((hv["rays_ori"].to(dev), hv["rays_dir"].to(dev), camera_of(hv)))
    g = torch.Generator().manual_seed(2000 + k)
    ups.append(((torch.rand((h, w, 4), generator=g) * 2 - 1).to(dev), ((torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01).to(dev)))
colour = torch.zeros((nv, n, 3), device=dev)
g12 = torch.zeros((n, 12), device=dev)

def run(n_streams, prio=False):
    rasters = [native.SplatRaster(abi.default_config(), device=dev) for _ in range(n_streams)]
    if prio:
        lo, hi = -1, 0
        streams = [torch.cuda.Stream(device=dev, priority=-1) for _ in range(n_streams)]
    else:
        streams = [torch.cuda.Stream(device=dev) for _ in range(n_streams)]
    def step():
        for i, (ro, rd, cam) in enumerate(views):
            r, s = rasters[i % n_streams], streams[i % n_streams]
            with torch.cuda.stream(s):
                rgba, dist, hits, vis = r.trace(0, 3, d12, sph, ro, rd, None, cam)
                r.trace_bwd_colour(3, d12, sph, ro, rd, cam, rgba, ups[i][0], dist, ups[i][1], colour[i])
        for j, (r, s) in enumerate(zip(rasters, streams)):
            with torch.cuda.stream(s):
                if j:
                    s.wait_stream(streams[0])
                r.density_gradient_finish(d12, g12, accumulate=j > 0)
        torch.cuda.current_stream().wait_stream(streams[-1])
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        step()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3

for ns in (1, 2, 3, 1, 2):
    print(ns, "stream(s):", round(run(ns), 3), "ms per step")
