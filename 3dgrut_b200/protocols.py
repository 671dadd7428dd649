"""Input contract of ``Tracer.render`` -- the per-view ``Batch``.

Field-compatible with the reference's ``threedgrut.datasets.protocols.Batch``
(threedgrut/datasets/protocols.py:23-58), so a reference ``Batch`` instance can
be passed to this Tracer unchanged (the Tracer only reads attributes).
``T_to_world_end`` is this build's optional extension carrying the
shutter-end pose for rolling-shutter cameras (the reference always passes the
same pose twice, threedgut_tracer/tracer.py:148-150).
"""
from dataclasses import dataclass
from typing import Optional

import torch


@dataclass
class Batch:
    rays_ori: torch.Tensor  # [1, H, W, 3] ray origins, sensor space
    rays_dir: torch.Tensor  # [1, H, W, 3] ray directions, sensor space
    T_to_world: torch.Tensor  # [1, 4, 4] sensor -> world
    rgb_gt: Optional[torch.Tensor] = None
    mask: Optional[torch.Tensor] = None
    intrinsics: Optional[list] = None  # [fx, fy, cx, cy]
    intrinsics_OpenCVPinholeCameraModelParameters: Optional[dict] = None
    intrinsics_OpenCVFisheyeCameraModelParameters: Optional[dict] = None
    T_to_world_end: Optional[torch.Tensor] = None

    def __post_init__(self):
        b = self.T_to_world.shape[0]
        if self.rays_ori.shape[0] != b or self.rays_dir.shape[0] != b:
            raise AssertionError("rays_ori / rays_dir must have the batch size of T_to_world")
        if self.rgb_gt is not None and (self.rgb_gt.ndim != 4 or self.rgb_gt.shape[0] != b):
            raise AssertionError("rgb_gt must be [B, H, W, 3]")
        if self.mask is not None and (self.mask.ndim != 4 or self.mask.shape[0] != b):
            raise AssertionError("mask must be [B, H, W, 1]")
        if self.intrinsics:
            if not isinstance(self.intrinsics, list) or len(self.intrinsics) != 4:
                raise AssertionError("intrinsics must be a list [fx, fy, cx, cy]")

    def keys(self):
        return [k for k, v in vars(self).items() if v is not None]
