"""Ray -> micro-cell routing of the episode sampler on the GPU.

The reference's ``TaskDataset`` (data/task_dataset.py) routes every training ray of a region to one cell of an
``(nx, ny, nz)`` grid over the region box before the first step, on the CPU: ``_region_segment`` (:155-171),
``_dda_maxoverlap`` (:299-351) or ``_alpha_local_max_overlap`` (:353-418), then ``_route_and_bin`` (:544-627) sorts the rays by
cell and drops weak overlaps.  `EpisodeRouter` keeps the reference's names and decisions (bit-exact cell ids) with the per-ray
work in one kernel (csrc/episode.cu); the tensors may stay on the GPU or be copied back as the dataset's CPU storage."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from .. import ops
from .._lib import check, host_floats, i32, lib


def build_cell_bounds(aabb: Tensor, cells: Sequence[int], device) -> Tuple[Tensor, Tensor]:
    """TaskDataset._build_cell_bounds (:173-196): per-cell boxes (C,2,3) and sizes (C,3), cell id = (ix*ny + iy)*nz + iz."""
    nx, ny, nz = cells
    lo, hi = aabb[0].to(device), aabb[1].to(device)
    size = (hi - lo).clamp(min=1e-9)
    xs, ys, zs = (torch.linspace(0, 1, steps=n + 1, device=device) for n in (nx, ny, nz))
    X0, Y0, Z0 = torch.meshgrid(xs[:-1], ys[:-1], zs[:-1], indexing="ij")
    X1, Y1, Z1 = torch.meshgrid(xs[1:], ys[1:], zs[1:], indexing="ij")
    lo_n = torch.stack([X0, Y0, Z0], dim=-1).reshape(-1, 3)
    hi_n = torch.stack([X1, Y1, Z1], dim=-1).reshape(-1, 3)
    bounds = torch.stack([lo + size * lo_n, lo + size * hi_n], dim=1)
    return bounds, (bounds[:, 1] - bounds[:, 0]).abs()


class EpisodeRouter:
    def __init__(self, aabb: Tensor, cells: Sequence[int] = (1, 6, 6), routing_policy: str = "alpha", assignment_checkpoint: float = 0.7,
                 device="cuda"):
        if routing_policy not in ("alpha", "dda"):
            raise ValueError(f"routing_policy must be 'alpha' or 'dda', got {routing_policy!r}")
        self.device = torch.device(device)
        self.aabb = torch.as_tensor(aabb, dtype=torch.float32).reshape(2, 3).to(self.device)
        self.cells = tuple(int(c) for c in cells)
        self.routing_policy, self.alpha = routing_policy, float(assignment_checkpoint)
        # the constants are computed with the reference's own torch expressions (on the CPU, as the reference does), so that
        # the kernel compares against bit-identical boxes and tolerances
        cb, cs = build_cell_bounds(self.aabb.cpu(), self.cells, "cpu")
        self.cell_bounds, self.cell_sizes = cb.to(self.device), cs.to(self.device)
        size = (cb[:, 1] - cb[:, 0]).norm(dim=1)
        self._tol_cell = torch.maximum(1e-6 * size, torch.tensor(1e-9)).to(self.device).contiguous()
        cell_diag = cs.pow(2).sum(dim=1).sqrt().median().item()
        self._tol_abs = max(1e-6 * cell_diag, 1e-9)
        self._aabb_host = host_floats(self.aabb.cpu().reshape(-1).tolist())
        self._cells_host = (i32 * 3)(*self.cells)

    @torch.no_grad()
    def route(self, rays: Tensor, max_steps: int = 64):
        """valid (N,) bool, cid (N,) int64 (-1 where invalid), len_best (N,), keep (N,) bool for (N, >=6) rays."""
        rays = ops._f32(rays, "rays")
        N, ncol = rays.shape
        valid = torch.empty(N, dtype=torch.uint8, device=rays.device)
        keep = torch.empty(N, dtype=torch.uint8, device=rays.device)
        cid = torch.empty(N, dtype=torch.int32, device=rays.device)
        blen = torch.empty(N, device=rays.device)
        with ops._on(rays.device), ops._timed("episode_route"):
            check(lib().acn_episode_route(ops._ptr(rays), int(ncol), N, self._aabb_host, self._cells_host, ops._ptr(self.cell_bounds.reshape(-1, 6)),
                                          ops._ptr(self._tol_cell), 0 if self.routing_policy == "dda" else 1, self.alpha, float(self._tol_abs),
                                          int(max_steps), ops._ptr(valid), ops._ptr(cid), ops._ptr(blen), ops._ptr(keep), ops._stream()),
                  "acn_episode_route")
        ops._count(1)
        return valid.bool(), cid.long(), blen, keep.bool()

    @torch.no_grad()
    def route_and_bin(self, rays: Tensor) -> List[Tensor]:
        """TaskDataset._route_and_bin: one int64 index tensor per cell (global ray ids, original order inside a cell, weak
        overlaps dropped)."""
        nx, ny, nz = self.cells
        C_ = nx * ny * nz
        valid, cid, _, keep = self.route(rays)
        sel = (valid & keep).nonzero(as_tuple=False).reshape(-1)
        c = cid[sel]
        order = torch.argsort(c, stable=True)
        sel, c = sel[order], c[order]
        counts = torch.bincount(c, minlength=C_).tolist()
        return list(torch.split(sel, counts))
