"""Axis-aligned scene box; interface of the reference's ``nerfs/scene_box.py`` (SceneBox, :10-205)
restricted to what the hot path and its callers use.  The slab test runs in the CUDA library."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence, Tuple, Union

import torch
from torch import Tensor

from .. import ops


@dataclass
class SceneBox:
    aabb: Tensor  # (2,3) [min, max], world DRB

    @property
    def min(self) -> Tensor:
        return self.aabb[0]

    @property
    def max(self) -> Tensor:
        return self.aabb[1]

    @property
    def center(self) -> Tensor:
        return (self.aabb[0] + self.aabb[1]) * 0.5

    @property
    def extent(self) -> Tensor:
        return self.aabb[1] - self.aabb[0]

    def to(self, dev) -> "SceneBox":
        return SceneBox(aabb=self.aabb.to(torch.device(dev) if isinstance(dev, str) else dev))

    def __repr__(self) -> str:
        lo = ", ".join(f"{v:.3f}" for v in self.min.cpu().tolist())
        hi = ", ".join(f"{v:.3f}" for v in self.max.cpu().tolist())
        return f"SceneBox (min=[{lo}], max=[{hi}], diag={self.get_diagonal_length().item():.3f})"

    def ray_aabb_intersect(self, origins: Tensor, directions: Tensor, eps: float = 1e-8, max_bound: float = 1e10,
                           invalid_value: float = 1e10) -> Tuple[Tensor, Tensor]:
        """Slab test (reference scene_box.py:45-107) on CUDA rays; eps is fixed at 1e-8 in the kernel."""
        if eps != 1e-8:
            raise ValueError("the sm_100a slab kernel is compiled for eps=1e-8 (the only value the reference passes)")
        return ops.ray_aabb(origins.reshape(-1, 3), directions.reshape(-1, 3), self.aabb, max_bound, invalid_value)

    def within(self, pts: Tensor, inclusive: bool = False) -> Tensor:
        lo, hi = self.aabb[0].to(pts.device), self.aabb[1].to(pts.device)
        if inclusive:
            return (pts >= lo).all(dim=-1) & (pts <= hi).all(dim=-1)
        return (pts > lo).all(dim=-1) & (pts < hi).all(dim=-1)

    def get_diagonal_length(self) -> Tensor:
        return torch.linalg.norm(self.aabb[1] - self.aabb[0])

    @staticmethod
    def from_bound(aabb: Tensor) -> "SceneBox":
        assert isinstance(aabb, torch.Tensor) and aabb.shape == (2, 3)
        return SceneBox(aabb=aabb)

    def expand(self, pad: Union[float, Tensor, Sequence[float]]) -> "SceneBox":
        p = torch.as_tensor(pad, dtype=self.aabb.dtype, device=self.aabb.device)
        if p.ndim == 0:
            p = p.repeat(3)
        return SceneBox(aabb=torch.stack([self.aabb[0] - p, self.aabb[1] + p], dim=0))

    # ---- box utilities the dataset / clustering tools call (scene_box.py:119-147, 208-217) ----
    def get_centered_and_scaled_scene_box(self, scale_factor: Union[float, Tensor] = 1.0) -> "SceneBox":
        s = torch.as_tensor(scale_factor, device=self.aabb.device, dtype=self.aabb.dtype)
        return SceneBox(aabb=(self.aabb - self.center) * s)

    @staticmethod
    def get_normalized_positions(positions: Tensor, aabb: Tensor) -> Tensor:
        return (positions - aabb[0]) / (aabb[1] - aabb[0])

    @staticmethod
    def from_camera_poses(poses: Tensor, scale_factor: float = 1.0) -> "SceneBox":
        centres = poses[..., :3, -1]
        return SceneBox(aabb=torch.stack([centres.min(dim=-2).values, centres.max(dim=-2).values], dim=0) * scale_factor)

    def union(self, other: "SceneBox") -> "SceneBox":
        return SceneBox(aabb=torch.stack([torch.minimum(self.min, other.min), torch.maximum(self.max, other.max)], dim=0))

    @staticmethod
    def reduce_union(aabbs: Tensor) -> "SceneBox":
        return SceneBox(aabb=torch.stack([aabbs[:, 0, :].min(dim=0).values, aabbs[:, 1, :].max(dim=0).values], dim=0))
