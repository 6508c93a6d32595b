"""CPU oracle of the Voronoi mask builder (TEST INFRASTRUCTURE: imported by tests/ only).

``voronoi_orig`` restates scripts/create_clusters.py:559-634 (``compute_voronoi_orig``, the reference's CPU path: torch.cdist
distances, membership = min over the ray's samples of D / (min_c D + 1e-8) <= margin); ``voronoi_boxes`` restates the box /
count streams of ``compute_voronoi_opt`` (:386-556, squared-distance rule d2 <= m^2 min d2; margin 1: argmin).
**Parity: PINNED** for the masks - tests/golden/voronoi.npz was written by tests/golden/make_golden_voronoi.py from the
unmodified ``compute_voronoi_orig``; the box streams exist only in the reference's CUDA branch, which cannot run here."""
from __future__ import annotations

import torch

Tensor = torch.Tensor


def sample_positions(rays: Tensor, S: int) -> Tensor:
    z = torch.linspace(0.0, 1.0, S, dtype=torch.float32)
    t = torch.lerp(rays[:, 6:7], rays[:, 7:8], z[None, :])
    return rays[:, None, 0:3] + rays[:, None, 3:6] * t[..., None]  # (N,S,3)


def voronoi_orig(rays: Tensor, S: int, centroids: Tensor, cluster_2d: bool, margin: float, eps: float = 1e-8) -> Tensor:
    start = 1 if cluster_2d else 0
    xyz = sample_positions(rays.float(), S)
    D = torch.cdist(xyz.reshape(-1, 3)[:, start:], centroids[:, start:].float())
    ratio = D / (D.min(dim=1, keepdim=True).values + eps)
    return ratio.view(rays.shape[0], S, -1).amin(dim=1) <= margin


def voronoi_boxes(rays: Tensor, S: int, centroids: Tensor, cluster_2d: bool, margin: float):
    """member mask per sample (N,S,C) by the squared rule, and the per-expert (mins, maxs, counts) of the member positions"""
    start = 1 if cluster_2d else 0
    xyz = sample_positions(rays.float(), S)
    d2 = ((xyz[..., None, start:] - centroids[None, None, :, start:].float()) ** 2).sum(-1)
    if margin == 1.0:
        ok = torch.nn.functional.one_hot(d2.argmin(dim=-1), centroids.shape[0]).bool()
    else:
        ok = d2 <= (margin * margin) * d2.min(dim=-1, keepdim=True).values
    C = centroids.shape[0]
    mins, maxs, cnts = torch.full((C, 3), float("inf")), torch.full((C, 3), float("-inf")), torch.zeros(C, dtype=torch.int64)
    for c in range(C):
        x = xyz[ok[..., c]]
        if x.numel():
            mins[c], maxs[c], cnts[c] = x.min(0).values, x.max(0).values, x.shape[0]
    return ok, mins, maxs, cnts
