"""CPU oracle of the episode sampler's ray -> micro-cell routing (TEST INFRASTRUCTURE: imported by tests/ only).

Restates, in vectorised torch on the CPU, data/task_dataset.py of the reference: ``_aabb_intersect`` (:130-153),
``_region_segment`` (:155-171), ``_build_cell_bounds`` (:173-196), ``_map_points_to_block_ids`` (:198-213),
``_overlap_len_with_cell`` (:216-232), ``_dda_transform / _dda_init / _dda_maxoverlap`` (:242-351),
``_alpha_local_max_overlap`` (:353-418) and the decisions of ``_route_and_bin`` (:544-627).
**Parity: PINNED** - tests/golden/episode.npz was written by tests/golden/make_golden_episode.py from the UNMODIFIED reference
(TaskDataset methods called on seeded rays); tests/test_oracle_golden.py holds this file to it exactly."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch

Tensor = torch.Tensor


def aabb_intersect(o: Tensor, d: Tensor, aabb: Tensor, eps: float = 1e-12):
    lo, hi = (aabb[0], aabb[1]) if aabb.dim() == 2 else (aabb[:, 0, :], aabb[:, 1, :])
    parallel = d.abs() < eps
    inv_d = 1.0 / d
    t0, t1 = (lo - o) * inv_d, (hi - o) * inv_d
    tmin, tmax = torch.minimum(t0, t1), torch.maximum(t0, t1)
    miss = (parallel & ~((o >= lo) & (o <= hi))).any(dim=1)
    t_entry, t_exit = tmin.max(dim=1).values, tmax.min(dim=1).values
    return (t_exit >= t_entry) & ~miss, t_entry, t_exit


def region_segment(rays: Tensor, aabb: Tensor):
    hit, te, tx = aabb_intersect(rays[:, 0:3], rays[:, 3:6], aabb)
    t0, t1 = torch.maximum(te, torch.zeros_like(te)), tx.clone()
    if rays.shape[1] >= 8:
        t0, t1 = torch.maximum(t0, rays[:, 6]), torch.minimum(t1, rays[:, 7])
    seg = t1 - t0
    return hit & (seg > 0), t0, t1, seg


def build_cell_bounds(aabb: Tensor, cells: Sequence[int]):
    nx, ny, nz = cells
    lo, hi = aabb[0], aabb[1]
    size = (hi - lo).clamp(min=1e-9)
    xs, ys, zs = (torch.linspace(0, 1, steps=n + 1) for n in (nx, ny, nz))
    A = torch.stack(torch.meshgrid(xs[:-1], ys[:-1], zs[:-1], indexing="ij"), dim=-1).reshape(-1, 3)
    B = torch.stack(torch.meshgrid(xs[1:], ys[1:], zs[1:], indexing="ij"), dim=-1).reshape(-1, 3)
    cb = torch.stack([lo + size * A, lo + size * B], dim=1)
    return cb, (cb[:, 1] - cb[:, 0]).abs()


def overlap_len(rays: Tensor, cell_aabb: Tensor) -> Tensor:
    hit, te, tx = aabb_intersect(rays[:, :3], rays[:, 3:6], cell_aabb)
    t0, t1 = torch.maximum(te, torch.zeros_like(te)), tx.clone()
    if rays.shape[1] >= 8:
        t0, t1 = torch.maximum(t0, rays[:, 6]), torch.minimum(t1, rays[:, 7])
    return torch.where(hit, (t1 - t0).clamp_min(0.0), torch.zeros_like(te))


def dda_maxoverlap(rays: Tensor, aabb: Tensor, cells: Sequence[int], t0: Tensor, t1: Tensor, max_steps: int = 64):
    o, d = rays[:, :3], rays[:, 3:6]
    cell = torch.clamp((aabb[1] - aabb[0]) / torch.tensor(cells, dtype=o.dtype), min=1e-12)
    g_o, g_d = (o - aabb[0]) / cell, d / cell
    p = g_o + g_d * (t0 + 1e-6).unsqueeze(-1)
    idx = torch.floor(p).to(torch.int64)
    step = torch.sign(g_d).to(torch.int64).clamp(min=-1, max=1)
    nb = torch.where(step > 0, torch.floor(p) + 1.0, torch.ceil(p) - 1.0)
    inv = 1.0 / g_d
    tMax = ((nb - p) * inv).nan_to_num(nan=1e30, posinf=1e30, neginf=1e30)
    tDelta = (step.to(p.dtype) * inv).nan_to_num(nan=1e30, posinf=1e30, neginf=1e30)
    n = torch.tensor(cells, dtype=torch.int64)
    nyz = cells[1] * cells[2]
    idx = torch.minimum(idx.clamp(min=0), n - 1)
    t = t0.clone()
    best_len = torch.zeros_like(t0)
    best_cid = idx[:, 0] * nyz + idx[:, 1] * cells[2] + idx[:, 2]
    for _ in range(max_steps):
        m = tMax.min(dim=1).values
        t_next = torch.minimum(m, t1)
        dt = (t_next - t).clamp_min(0.0)
        cid = idx[:, 0] * nyz + idx[:, 1] * cells[2] + idx[:, 2]
        improve = dt > best_len
        best_len, best_cid = torch.where(improve, dt, best_len), torch.where(improve, cid, best_cid)
        if bool((t_next >= t1).all()):
            break
        ax = (tMax[:, 0] <= tMax[:, 1]) & (tMax[:, 0] <= tMax[:, 2])
        ay = (~(tMax[:, 0] <= tMax[:, 1])) & (tMax[:, 1] <= tMax[:, 2])
        adv = torch.stack([ax, ay, ~(ax | ay)], dim=1)
        idx = torch.where(adv, torch.minimum((idx + step).clamp(min=0), n - 1), idx)
        tMax = torch.where(adv, tMax + tDelta, tMax)
        t = t_next
    return best_cid, best_len


def alpha_local_max_overlap(rays: Tensor, aabb: Tensor, cells: Sequence[int], alpha: float, t0: Tensor, t1: Tensor, seg: Tensor,
                            cell_bounds: Tensor, cell_sizes: Tensor):
    o, d = rays[:, :3], rays[:, 3:6]
    ta = t0 + alpha * (t1 - t0)
    ta = ta + 1e-6 * (t1 - t0)
    p = o + d * ta.unsqueeze(-1)
    n = torch.tensor(cells, dtype=torch.int64)
    rel = ((p - aabb[0]) / (aabb[1] - aabb[0]).clamp(min=1e-9)).clamp(0.0, 1.0 - 1e-7)
    ic = torch.minimum(torch.floor(rel * torch.tensor(cells)).to(torch.int64).clamp(min=0), n - 1)
    nyz = cells[1] * cells[2]
    primary = ic[:, 0] * nyz + ic[:, 1] * cells[2] + ic[:, 2]
    off = torch.tensor([[-1, 0, 0], [1, 0, 0], [0, -1, 0], [0, 1, 0], [0, 0, -1], [0, 0, 1], [0, 0, 0]])
    cand = torch.minimum((ic[:, None, :] + off[None]).clamp(min=0), n - 1)
    cand_cid = cand[..., 0] * nyz + cand[..., 1] * cells[2] + cand[..., 2]
    ov = torch.stack([overlap_len(rays, cell_bounds.index_select(0, cand_cid[:, k])) for k in range(7)], dim=1)
    best_k = ov.argmax(dim=1)
    ar = torch.arange(rays.shape[0])
    cid_best, len_best = cand_cid[ar, best_k], ov[ar, best_k]
    tol_abs = max(1e-6 * cell_sizes.pow(2).sum(dim=1).sqrt().median().item(), 1e-9)
    ok = len_best >= torch.maximum(torch.tensor(tol_abs), 1e-6 * seg)
    return torch.where(ok, cid_best, primary), len_best


def route(rays: Tensor, aabb: Tensor, cells: Sequence[int], policy: str, alpha: float = 0.7):
    """-> valid (N,) bool, cid (N,) int64 (-1 invalid), len_best (N,), keep (N,) bool: the decisions of _route_and_bin."""
    N = rays.shape[0]
    cb, cs = build_cell_bounds(aabb, cells)
    valid, t0, t1, seg = region_segment(rays, aabb)
    cid = torch.full((N,), -1, dtype=torch.int64)
    blen = torch.zeros(N)
    keep = torch.zeros(N, dtype=torch.bool)
    iv = valid.nonzero().reshape(-1)
    if iv.numel():
        rv = rays[iv]
        if policy == "dda":
            c, l = dda_maxoverlap(rv, aabb, cells, t0[iv], t1[iv])
        else:
            c, l = alpha_local_max_overlap(rv, aabb, cells, alpha, t0[iv], t1[iv], seg[iv], cb, cs)
        tol = torch.maximum(1e-6 * (cb[:, 1] - cb[:, 0]).norm(dim=1), torch.tensor(1e-9))
        cid[iv], blen[iv] = c, l
        keep[iv] = overlap_len(rv, cb.index_select(0, c)) >= tol.index_select(0, c)
    return valid, cid, blen, keep


def bins_of(valid: Tensor, cid: Tensor, keep: Tensor, n_cells: int) -> List[Tensor]:
    sel = (valid & keep).nonzero().reshape(-1)
    order = torch.argsort(cid[sel], stable=True)
    sel = sel[order]
    return list(torch.split(sel, torch.bincount(cid[sel], minlength=n_cells).tolist()))
