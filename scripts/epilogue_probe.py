"""Times the two kernels of the fused optimiser epilogue and the unfused chain's kernels at N = 3M, 8 views (device resident)."""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.helpers import abi, syn
native = importlib.import_module("3dgrut_b200.native")
dev = torch.device("cuda", 0)
n, nv = 3_000_000, 8
r = native.SplatRaster(abi.default_config(), device=dev)
scene = syn.make_scene(n, 3, seed=1)
d12 = syn.pack_density(scene).to(dev)
g = torch.Generator(device=dev).manual_seed(0)
colour = torch.randn((nv, n, 3), device=dev, generator=g); colour[:, ::3] = 0
centres = torch.randn((nv, 3), device=dev, generator=g) * 4
ptrs = [colour[v].data_ptr() for v in range(nv)]
mk = lambda *s: (torch.randn(s, device=dev, generator=g), torch.zeros(s, device=dev), torch.zeros(s, device=dev))
alb, spc, pos, rot, scl, dns = mk(n, 3), mk(n, 45), mk(n, 3), mk(n, 4), mk(n, 3), mk(n, 1)
vis = (torch.rand((n, 1), device=dev, generator=g) > 0.2).float()
g12 = torch.randn((n, 12), device=dev, generator=g); g48 = torch.empty((n, 48), device=dev)
acc, den = torch.zeros((n, 1), device=dev), torch.zeros((n, 1), device=dev)
def timed(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
print("sh_expand_adam      ", round(timed(lambda: r.sh_expand_adam(3, d12, centres, ptrs, alb, spc, (0.0, 0.0), vis, 0.9, 0.999, 1e-15)), 4))
print("density_adam (+stats)", round(timed(lambda: r.density_adam(d12, g12, pos, rot, scl, dns, [0.0] * 4, vis, 0.9, 0.999, 1e-15, 1, 2, [0.0, 0.0, 4.0], acc, den)), 4))
print("sh_expand_gather    ", round(timed(lambda: r.sh_gradient_expand_gather(3, d12, centres, ptrs, g48)), 4))
print("unpack_feature_grad ", round(timed(lambda: r.unpack_feature_gradient(g48)), 4))
ag, sg = r.unpack_feature_gradient(g48)
print("adam albedo+specular", round(timed(lambda: r.selective_adam_step([(alb[0], ag, alb[1], alb[2], 0.0), (spc[0], sg, spc[1], spc[2], 0.0)], vis, 0.9, 0.999, 1e-15)), 4))
