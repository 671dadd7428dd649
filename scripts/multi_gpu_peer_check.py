"""Equivalence of the peer-memory exchange and the NCCL exchange of a data-parallel step, on 2+ GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/multi_gpu_peer_check.py

Every rank renders its share of 5 views of a small scene forward + backward; training_step_peer (colour gradients read from the
other ranks' buffers inside the SH expansion) must leave the same [N,12] / [N,48] gradients as training_step_deferred (NCCL
all-gather) on three consecutive steps (both buffer sets), up to the rounding noise of two separate backward passes (1e-5)."""
import importlib
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.helpers import abi, camera_of, syn  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    native = importlib.import_module("3dgrut_b200.native")
    gdist = importlib.import_module("3dgrut_b200.dist")
    n, w, h, nviews = 20000, 160, 96, 5
    scene = syn.make_scene(n, 3, seed=11, kind="dense")
    d12, sph = syn.pack_density(scene).to(dev), scene["features"].to(dev)
    raster = native.SplatRaster(abi.default_config(), device=dev)
    ids = gdist.shard_views(nviews, rank, world)
    views, ups = [], []
    g = torch.Generator().manual_seed(5)
    all_up = [(torch.rand((h, w, 4), generator=g) * 2 - 1, (torch.rand((h, w, 1), generator=g) * 2 - 1) * 0.01) for _ in range(nviews)]
    for k in ids:
        hv = syn.make_view(w, h, k=k, n_views=nviews)
        views.append({"rays_ori": hv["rays_ori"].to(dev), "rays_dir": hv["rays_dir"].to(dev), "camera": camera_of(hv)})
        ups.append((all_up[k][0].to(dev), all_up[k][1].to(dev)))
    vpr = (nviews + world - 1) // world
    ref = gdist.DeferredExchange(n, 48, vpr, dev, world=world)
    peer = gdist.PeerExchange(raster, n, 48, vpr, dev, rank, world=world)
    ok = True
    for step in range(3):
        gdist.training_step_deferred(raster, ref, views, 3, d12, sph, ups, rank=rank)
        gdist.training_step_peer(raster, peer, views, 3, d12, sph, ups, rank=rank)
        torch.cuda.synchronize()
        # two backward passes never agree to the bit (their REDs land in a different order): relative L2 of the difference
        e12 = float((ref.density_grad - peer.density_grad).norm() / ref.density_grad.norm())
        e48 = float((ref.radiance_grad - peer.radiance_grad).norm() / ref.radiance_grad.norm())
        nz = int((peer.radiance_grad != 0).any(dim=1).sum())
        print(f"rank {rank} step {step}: rel L2 [N,12] {e12:.2e} [N,48] {e48:.2e}; particles with SH gradient {nz}", flush=True)
        ok = ok and e12 < 1e-5 and e48 < 1e-5 and nz > 0
    peer.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
