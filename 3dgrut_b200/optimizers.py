"""``SelectiveAdam`` and the densification statistics, backed by libgut_b200.

Drop-in for ``threedgrut.optimizers.SelectiveAdam`` (reference: threedgrut/optimizers/__init__.py:45-133, kernel
threedgrut/optimizers/optimizers.cu:47-84): same constructor, same ``step(visibility)``, same state keys
(``step``, ``exp_avg``, ``exp_avg_sq``) so checkpoints load either way.  The reference launches one kernel per
parameter group; here every group of the model goes through ONE launch of ``gut_selective_adam_step``, and
``step(visibility, zero_grad=True)`` clears the gradients in the same pass (trainer.py:751 does it separately).

``update_gradient_buffer`` is ``GSStrategy.update_gradient_buffer`` (threedgrut/strategy/gs.py:104-115) as one kernel.

There is no CPU path: without the CUDA library these raise.
"""
import torch

from . import abi
from .native import SplatRaster

_shared = {}


def _raster(device) -> SplatRaster:
    """One native handle per device for the model-side kernels (they use no rasteriser workspace)."""
    key = torch.device(device).index
    if key not in _shared:
        _shared[key] = SplatRaster(abi.default_config(), device=device)
    return _shared[key]


class SelectiveAdam(torch.optim.Adam):
    def __init__(self, params, lr=0.001, betas=(0.9, 0.999), eps=1e-08):
        super().__init__(params=params, lr=lr, eps=eps, betas=betas)

    @torch.no_grad()
    def step(self, visibility, zero_grad=False):
        calls = {}   # (beta1, beta2, eps, rows, device) -> groups sharing one launch
        for group in self.param_groups:
            assert len(group["params"]) == 1, "More than one tensor in group is not supported"
            param = group["params"][0]
            if param.grad is None:
                continue
            state = self.state[param]
            if len(state) == 0:
                state["step"] = torch.tensor(0.0, dtype=torch.float32)
                state["exp_avg"] = torch.zeros_like(param, memory_format=torch.preserve_format)
                state["exp_avg_sq"] = torch.zeros_like(param, memory_format=torch.preserve_format)
            if not (param.is_contiguous() and param.grad.is_contiguous()):
                raise ValueError("SelectiveAdam updates parameters in place: parameters and gradients must be contiguous")
            beta1, beta2 = group["betas"]
            key = (beta1, beta2, group["eps"], param.shape[0], param.device)
            calls.setdefault(key, []).append((param.data, param.grad, state["exp_avg"], state["exp_avg_sq"], group["lr"]))
        for (beta1, beta2, eps, _rows, device), groups in calls.items():
            raster = _raster(device)
            for i in range(0, len(groups), abi.GUT_ADAM_MAX_GROUPS):
                raster.selective_adam_step(groups[i:i + abi.GUT_ADAM_MAX_GROUPS], visibility, beta1, beta2, eps, zero_grad=zero_grad)


@torch.no_grad()
def update_gradient_buffer(positions, densify_grad_norm_accum, densify_grad_norm_denom, sensor_position):
    """In place: accum[mask] += |positions.grad * |positions - sensor_position|| / 2, denom[mask] += 1, mask = grad != 0."""
    if torch.is_tensor(sensor_position):
        sensor_position = sensor_position.detach().reshape(3).cpu().tolist()
    _raster(positions.device).densify_stats(positions.detach(), positions.grad, sensor_position, densify_grad_norm_accum,
                                            densify_grad_norm_denom)
