"""All-reduce of the step's [N,12] fp32 gradient (144 MB at 3M particles) under the current NCCL_* environment: mean ms over
20 calls (CUDA events, max over ranks) and the NCCL kernel NCCL picked (torch profiler).  torchrun, one process per GPU."""
import os, sys
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
x = torch.randn((3_000_000, 12), device=dev)
for _ in range(5):
    dist.all_reduce(x)
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20):
    dist.all_reduce(x)
b.record(); torch.cuda.synchronize()
t = torch.tensor([a.elapsed_time(b) / 20], device=dev, dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
name = ""
if rank == 0:
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        dist.all_reduce(x); torch.cuda.synchronize()
    name = [e.key for e in prof.key_averages() if "nccl" in e.key.lower()]
else:
    dist.all_reduce(x); torch.cuda.synchronize()
if rank == 0:
    env = {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}
    print(f"world {world} env {env}: {float(t.item()):.3f} ms per 144 MB all-reduce ({144e6 * 2 * (world - 1) / world / (float(t.item()) * 1e-3) / 1e9:.0f} GB/s bus); kernel {name}", flush=True)
dist.destroy_process_group()
