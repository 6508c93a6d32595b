"""Ray construction on the GPU; same names and argument meaning as the reference's
``nerfs/ray_sampling.py`` (get_ray_directions :111-136, get_rays :50-108,
clamp_rays_near_far :139-176).  CUDA tensors only: the reference's dataset workers build
their rays on the CPU with the reference's own functions."""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from .. import ops
from .scene_box import SceneBox


def pack_rays(rays_o: Tensor, rays_d: Tensor, near: Tensor, far: Tensor) -> Tensor:
    return torch.cat([rays_o, rays_d, near, far], dim=-1)


def unpack_rays(rays: Tensor):
    assert rays.shape[-1] == 8, "packed rays must be (..., 8)"
    flat = rays.view(-1, 8).contiguous()
    return flat[:, :3], flat[:, 3:6], flat[:, 6:7], flat[:, 7:8]


def get_ray_directions(H: int, W: int, fx: float, fy: float, cx: float, cy: float, center_pixels: bool, device) -> Tensor:
    """Unit camera-frame directions (H,W,3), RUB camera looking down -z."""
    return ops.ray_directions(int(H), int(W), float(fx), float(fy), float(cx), float(cy), bool(center_pixels), torch.device(device))


def get_rays(directions: Tensor, c2w: Tensor, scene_box: Optional[SceneBox] = None, near: Optional[float] = None,
             far: Optional[float] = None, *, aabb_max_bound: float = 1e10, aabb_invalid_value: float = 1e10) -> Tensor:
    """(H,W,3) or (N,3) camera directions -> (H,W,8) or (N,8) rays [o, d, near, far]."""
    if not (directions.shape[-1] == 3 and directions.ndim in (2, 3)):
        raise ValueError(f"directions must be (H, W, 3) or (N, 3), got {tuple(directions.shape)}")
    if scene_box is None and (near is None or far is None):
        raise ValueError("Provide near/far when scene_box is None")
    aabb = None if scene_box is None else scene_box.aabb
    return ops.get_rays(directions, c2w, aabb, near, far, aabb_max_bound, aabb_invalid_value)


@torch.no_grad()
def clamp_rays_near_far(rays: Tensor, near_far_override: Optional[Tuple[Optional[float], Optional[float]]], *, eps: float = 1e-6,
                        invalid_value: float = float("inf")) -> Tuple[Tensor, Tensor]:
    """Returns (rays_clamped, valid_mask); ``near_far_override=None`` leaves the rays untouched."""
    return ops.clamp_near_far(rays, near_far_override, eps, invalid_value)
