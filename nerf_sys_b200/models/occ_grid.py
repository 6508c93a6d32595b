"""Multi-level occupancy grid with the interface and buffers of ``nerfacc.OccGridEstimator`` (nerfacc 0.5.3, the
third-party dependency of the reference's occupancy path: models/inr/meta_ngp.py:132-136, 345-352, 373-380, 430-442).

nerfacc is not vendored by the reference and is absent here, so this is a restatement of its documented behaviour, not a
copy: the buffers keep nerfacc's names and shapes (``resolution``, ``aabbs``, ``occs``, ``binaries``; checkpoints of
``--use_occ`` runs carry them as ``submodules.{k}.occ_grid.*``), the grid bookkeeping (EMA update, cell sampling,
visibility pre-marking) is index arithmetic in torch, and the per-ray work - grid traversal, visibility filtering - runs
in csrc/occ.cu.  **Parity: unpinned** (no nerfacc to compare with); oracle/acnerf_oracle.py restates the same traversal
on the CPU and the GPU tests hold the kernels to it.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple, Union

import torch
import torch.nn as nn
from torch import Tensor

from .. import ops


def enlarge_aabb(aabb: Tensor, factor: float) -> Tensor:
    center = (aabb[:3] + aabb[3:]) / 2
    half = (aabb[3:] - aabb[:3]) / 2
    return torch.cat([center - half * factor, center + half * factor])


class OccGridEstimator(nn.Module):
    DIM = 3

    def __init__(self, roi_aabb: Union[List[float], Tensor], resolution: Union[int, List[int], Tensor] = 128, levels: int = 1, **kwargs) -> None:
        super().__init__()
        if isinstance(resolution, int):
            resolution = [resolution] * self.DIM
        if isinstance(resolution, (list, tuple)):
            resolution = torch.tensor(resolution, dtype=torch.int32)
        if isinstance(roi_aabb, (list, tuple)):
            roi_aabb = torch.tensor(roi_aabb, dtype=torch.float32)
        assert resolution.shape[0] == self.DIM and roi_aabb.shape[0] == self.DIM * 2
        if int(resolution.min()) != int(resolution.max()):
            raise NotImplementedError("the sm_100a traversal kernel takes cubic grids (the reference builds resolution^3)")
        if not (1 <= levels <= ops.ACN_OCC_MAX_LEVELS):
            raise ValueError(f"levels must be in [1, {ops.ACN_OCC_MAX_LEVELS}]")
        aabbs = torch.stack([enlarge_aabb(roi_aabb.float().cpu(), 2 ** i) for i in range(levels)], dim=0)
        self.cells_per_lvl = int(resolution.prod().item())
        self.levels = int(levels)
        self.register_buffer("resolution", resolution)
        self.register_buffer("aabbs", aabbs)
        self.register_buffer("occs", torch.zeros(self.levels * self.cells_per_lvl))
        self.register_buffer("binaries", torch.zeros([levels] + resolution.tolist(), dtype=torch.bool))
        r = [torch.arange(int(n)) for n in resolution.tolist()]
        grid_coords = torch.stack(torch.meshgrid(r, indexing="ij"), dim=-1).long().reshape(self.cells_per_lvl, self.DIM)
        self.register_buffer("grid_coords", grid_coords, persistent=False)
        self.register_buffer("grid_indices", torch.arange(self.cells_per_lvl), persistent=False)
        self._aabbs_host = None

    # the level boxes travel in the kernel-parameter block
    def aabbs_host(self) -> List[float]:
        key = (self.aabbs.data_ptr(), self.aabbs._version)
        if self._aabbs_host is None or self._aabbs_host[0] != key:
            self._aabbs_host = (key, self.aabbs.detach().float().cpu().reshape(-1).tolist())
        return self._aabbs_host[1]

    # ---- sampling --------------------------------------------------------------------------------------------
    @torch.no_grad()
    def sampling(self, rays_o: Tensor, rays_d: Tensor, sigma_fn: Optional[Callable] = None, alpha_fn: Optional[Callable] = None,
                 near_plane: float = 0.0, far_plane: float = 1e10, t_min: Optional[Tensor] = None, t_max: Optional[Tensor] = None,
                 render_step_size: float = 1e-3, early_stop_eps: float = 1e-4, alpha_thre: float = 0.0, stratified: bool = False,
                 cone_angle: float = 0.0, max_samples_per_ray: int = 4096) -> Tuple[Tensor, Tensor, Tensor]:
        """(ray_indices (M,) int64, t_starts (M,), t_ends (M,)), packed ray by ray, front to back.  Samples tile the runs of
        occupied cells with length clamp(t * cone_angle, render_step_size, inf); with ``sigma_fn`` (training) samples behind
        a transmittance below ``early_stop_eps`` or with alpha below min(alpha_thre, mean occupancy) are dropped."""
        if alpha_fn is not None:
            raise NotImplementedError("alpha_fn: the reference passes sigma_fn (meta_ngp.py:414-428)")
        N = rays_o.shape[0]
        near = t_min.float() if t_min is not None else torch.full((N,), float(near_plane), device=rays_o.device)
        far = t_max.float() if t_max is not None else torch.full((N,), float(far_plane), device=rays_o.device)
        rays = torch.cat([rays_o.float(), rays_d.float(), near[:, None], far[:, None]], dim=-1).contiguous()
        jitter = torch.rand(N, device=rays.device) if stratified else None
        seg = ops.occ_march(rays, jitter, self.binaries, self.aabbs_host(), float(render_step_size), float(cone_angle), int(max_samples_per_ray))
        if (alpha_thre > 0.0 or early_stop_eps > 0.0) and sigma_fn is not None and seg.M > 0:
            alpha_thre = min(float(alpha_thre), float(self.occs.mean().item()))
            sigmas = sigma_fn(seg.t0, seg.t1, seg.ray_idx.long()).reshape(-1)
            keep = ops.occ_visibility(seg, sigmas, float(early_stop_eps), alpha_thre)
            return seg.ray_idx.long()[keep], seg.t0[keep], seg.t1[keep]
        return seg.ray_idx.long(), seg.t0, seg.t1

    # ---- grid maintenance ------------------------------------------------------------------------------------
    @torch.no_grad()
    def update_every_n_steps(self, step: int, occ_eval_fn: Callable, occ_thre: float = 1e-2, ema_decay: float = 0.95,
                             warmup_steps: int = 256, n: int = 16) -> None:
        if not self.training:
            raise RuntimeError("You should only call this function only during training. Please call _update() directly if you want "
                               "to update the field during inference.")
        if step % n == 0 and self.training:
            self._update(step=step, occ_eval_fn=occ_eval_fn, occ_thre=occ_thre, ema_decay=ema_decay, warmup_steps=warmup_steps)

    @torch.no_grad()
    def _get_all_cells(self) -> List[Tensor]:
        out = []
        for lvl in range(self.levels):
            cell_ids = lvl * self.cells_per_lvl + self.grid_indices
            out.append(self.grid_indices[self.occs[cell_ids] >= 0.0])
        return out

    @torch.no_grad()
    def _sample_uniform_and_occupied_cells(self, n: int) -> List[Tensor]:
        out = []
        for lvl in range(self.levels):
            uniform = torch.randint(self.cells_per_lvl, (n,), device=self.occs.device)
            uniform = uniform[self.occs[lvl * self.cells_per_lvl + uniform] >= 0.0]
            occupied = torch.nonzero(self.binaries[lvl].flatten())[:, 0]
            if n < len(occupied):
                occupied = occupied[torch.randint(len(occupied), (n,), device=self.occs.device)]
            out.append(torch.cat([uniform, occupied], dim=0))
        return out

    @torch.no_grad()
    def _update(self, step: int, occ_eval_fn: Callable, occ_thre: float = 0.01, ema_decay: float = 0.95, warmup_steps: int = 256) -> None:
        """EMA of the evaluated occupancy at one jittered point per visited cell; all cells during warm-up, then a quarter of
        the grid uniformly plus the occupied cells.  binaries = occs > min(mean of the valid cells, occ_thre)."""
        lvl_indices = self._get_all_cells() if step < warmup_steps else self._sample_uniform_and_occupied_cells(self.cells_per_lvl // 4)
        res = self.resolution.to(torch.float32)
        for lvl, indices in enumerate(lvl_indices):
            coords = self.grid_coords[indices]
            x = (coords + torch.rand_like(coords, dtype=torch.float32)) / res
            x = self.aabbs[lvl, :3] + x * (self.aabbs[lvl, 3:] - self.aabbs[lvl, :3])
            occ = occ_eval_fn(x).reshape(-1)
            cell_ids = lvl * self.cells_per_lvl + indices
            self.occs[cell_ids] = torch.maximum(self.occs[cell_ids] * ema_decay, occ)
        thre = torch.clamp(self.occs[self.occs >= 0].mean(), max=occ_thre)
        self.binaries = (self.occs > thre).view(self.binaries.shape)

    @torch.no_grad()
    def mark_invisible_cells(self, K: Tensor, c2w: Tensor, width: int, height: int, near_plane: float = 0.0, chunk: int = 32 ** 3) -> None:
        """Cells seen by no camera (pinhole K, OpenCV c2w) - or closer than near_plane to one - get occupancy -1 and are
        never sampled or updated again."""
        assert K.dim() == 3 and K.shape[1:] == (3, 3)
        assert c2w.dim() == 3 and (c2w.shape[1:] == (3, 4) or c2w.shape[1:] == (4, 4))
        assert K.shape[0] == c2w.shape[0] or K.shape[0] == 1
        n_cams = c2w.shape[0]
        w2c_R = c2w[:, :3, :3].transpose(2, 1)
        w2c_T = -w2c_R @ c2w[:, :3, 3:]
        res1 = (self.resolution - 1).to(torch.float32)
        for lvl, indices in enumerate(self._get_all_cells()):
            coords = self.grid_coords[indices]
            for i in range(0, len(indices), chunk):
                x = coords[i:i + chunk] / res1
                ids = indices[i:i + chunk]
                xyz_w = (self.aabbs[lvl, :3] + x * (self.aabbs[lvl, 3:] - self.aabbs[lvl, :3])).T
                uvd = K @ (w2c_R @ xyz_w + w2c_T)
                uv = uvd[:, :2] / uvd[:, 2:]
                in_image = (uvd[:, 2] >= 0) & (uv[:, 0] >= 0) & (uv[:, 0] < width) & (uv[:, 1] >= 0) & (uv[:, 1] < height)
                seen = ((uvd[:, 2] >= near_plane) & in_image).sum(0) / n_cams
                too_near = ((uvd[:, 2] < near_plane) & in_image).any(0)
                valid = (seen > 0) & (~too_near)
                self.occs[lvl * self.cells_per_lvl + ids] = torch.where(valid, 0.0, -1.0)
