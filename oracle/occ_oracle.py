"""CPU oracle of the occupancy-guided renderer's per-ray work (TEST INFRASTRUCTURE: imported by tests/ only).

**Parity: nerfacc's arithmetic UNPINNED, the reference's own assembly PINNED.**  The reference delegates the traversal, the
visibility filter and the packed integration to nerfacc 0.5.3 (requirements.txt:8; call sites models/inr/meta_ngp.py:430-442,
nerfs/ray_rendering.py:477-487, 543-554), which is neither vendored under /root/reference nor installed here, so there is
nothing to run those against.  The code AROUND those calls is the reference's own (prefilter, union of the experts' segments,
midpoints, routing, blend, background): the functions at the end of this file restate it and are checked against vectors made
by the unmodified reference (tests/golden/make_golden_occ.py -> occ_assembly.npz, tests/test_oracle_golden.py).  Restated here:

* ``march`` - ``OccGridEstimator.sampling`` up to its visibility filter, i.e. nerfacc's ``traverse_grids``: nested level boxes
  (level l = the region of interest scaled by 2^l), the finest box containing the current position is traversed cell by cell,
  runs of occupied cells are tiled with samples of length ``max(t * cone_angle, step)`` (a sample is emitted while its
  midpoint lies before the end of the cell; a run continues seamlessly into the next occupied cell; a new run never starts
  before the end of the previous sample, so the samples of a ray do not overlap);
* ``visibility`` - ``render_visibility_from_density``;
* ``union`` - the reference's own ``_merge_segments_union`` (nerfs/ray_rendering.py:196-256), boundaries by ``torch.unique``;
* ``packed_render`` - ``render_weight_from_density`` + ``accumulate_along_rays`` in differentiable torch.

``march`` is written in scalar numpy float32 with exactly the operations, and the order of operations, of
nerf_sys_b200/csrc/occ.cu (compiled -fmad=false): cell ids, sample counts and ray ids are compared bit for bit.
"""
from __future__ import annotations

import math
from typing import Callable, List, Sequence, Tuple

import numpy as np
import torch

f32 = np.float32
INF = f32(np.inf)


def enlarge_aabb(aabb: torch.Tensor, factor: float) -> torch.Tensor:
    c, h = (aabb[:3] + aabb[3:]) / 2, (aabb[3:] - aabb[:3]) / 2
    return torch.cat([c - h * factor, c + h * factor])


def _box_span(o, d, bb):
    t0, t1 = -INF, INF
    for a in range(3):
        if d[a] == 0:
            if o[a] < bb[a] or o[a] > bb[3 + a]:
                return False, t0, t1
            continue
        inv = f32(1.0) / d[a]
        ta, tb = (bb[a] - o[a]) * inv, (bb[3 + a] - o[a]) * inv
        t0 = max(t0, min(ta, tb))
        t1 = min(t1, max(ta, tb))
    return bool(t1 > t0), t0, t1


def march(rays: torch.Tensor, binaries: torch.Tensor, aabbs: torch.Tensor, step: float, cone: float, near_jitter=None, max_samples: int = 4096):
    """-> ray_idx (M,) int64, t0 (M,), t1 (M,), cells (M,) int64 (level * R^3 + cell), counts (N,)"""
    R3 = np.ascontiguousarray(binaries.cpu().numpy().astype(np.uint8))
    levels, R = R3.shape[0], R3.shape[1]
    bbs = aabbs.cpu().numpy().astype(np.float32)
    ry = rays.cpu().numpy().astype(np.float32)
    step, cone = f32(step), f32(cone)
    out_r, out_a, out_b, out_c, counts = [], [], [], [], []
    with np.errstate(all="ignore"):
        for r in range(ry.shape[0]):
            o, d = ry[r, 0:3], ry[r, 3:6]
            near, far = ry[r, 6], ry[r, 7]
            if near_jitter is not None:
                near = f32(near + f32(near_jitter[r]) * step)
            n = 0
            if not (far > near):
                counts.append(0)
                continue
            lo, hi, hit = [], [], []
            for l in range(levels):
                h, a, b = _box_span(o, d, bbs[l])
                a, b = max(a, near), min(b, far)
                lo.append(a)
                hi.append(b)
                hit.append(h and b > a)
            continuous, t_last, cur = False, near, near
            while n < max_samples:
                lvl = -1
                for l in range(levels):
                    if hit[l] and lo[l] <= cur and cur < hi[l]:
                        lvl = l
                        break
                if lvl < 0:
                    end = INF
                    for l in range(levels):
                        if hit[l] and lo[l] > cur:
                            end = min(end, lo[l])
                    if not (end < INF):
                        break
                    cur, continuous = end, False
                    continue
                end = hi[lvl]
                for l in range(lvl):
                    if hit[l] and lo[l] > cur:
                        end = min(end, lo[l])
                bb = bbs[lvl]
                Rf = f32(R)
                cell, stepi, tmax, tdelta = [0, 0, 0], [0, 0, 0], [INF] * 3, [INF] * 3
                te = f32(cur + f32(f32(end - cur) * f32(1e-6)))
                for a in range(3):
                    ext = f32(bb[3 + a] - bb[a])
                    gpos = f32(f32(f32(f32(o[a] + f32(d[a] * te)) - bb[a]) / ext) * Rf)
                    c = int(np.floor(gpos))
                    c = 0 if c < 0 else (R - 1 if c > R - 1 else c)
                    cell[a] = c
                    cs = f32(ext / Rf)
                    if d[a] > 0:
                        stepi[a] = 1
                        tmax[a] = f32(f32(f32(bb[a] + f32(f32(c + 1) * cs)) - o[a]) / d[a])
                        tdelta[a] = f32(cs / d[a])
                    elif d[a] < 0:
                        stepi[a] = -1
                        tmax[a] = f32(f32(f32(bb[a] + f32(f32(c) * cs)) - o[a]) / d[a])
                        tdelta[a] = f32(f32(-cs) / d[a])
                ta = cur
                while ta < end and n < max_samples:
                    tcross = min(tmax[0], min(tmax[1], tmax[2]))
                    tb = min(tcross, end)
                    cid = (cell[0] * R + cell[1]) * R + cell[2]
                    if R3[lvl, cell[0], cell[1], cell[2]]:
                        while n < max_samples:
                            if continuous:
                                dt = max(f32(t_last * cone), step)
                                if f32(t_last + f32(dt * f32(0.5))) >= tb:
                                    break
                                t0, t1 = t_last, f32(t_last + dt)
                            else:
                                ts = max(ta, t_last)  # never before the end of the previous sample
                                dt = max(f32(ts * cone), step)
                                if f32(ts + f32(dt * f32(0.5))) >= tb:
                                    break
                                t0, t1 = ts, f32(ts + dt)
                            out_r.append(r)
                            out_a.append(t0)
                            out_b.append(t1)
                            out_c.append(lvl * R * R * R + cid)
                            n += 1
                            t_last, continuous = t1, True
                    else:
                        continuous = False
                    if not (tcross < end):
                        break
                    ax = 0 if (tmax[0] <= tmax[1] and tmax[0] <= tmax[2]) else (1 if tmax[1] <= tmax[2] else 2)
                    cell[ax] += stepi[ax]
                    tmax[ax] = f32(tmax[ax] + tdelta[ax])
                    ta = tcross
                    if cell[ax] < 0 or cell[ax] >= R:
                        break
                cur = end
            counts.append(n)
    return (torch.tensor(out_r, dtype=torch.int64), torch.tensor(np.array(out_a, dtype=np.float32)), torch.tensor(np.array(out_b, dtype=np.float32)),
            torch.tensor(out_c, dtype=torch.int64), torch.tensor(counts, dtype=torch.int64))


def visibility(ray_idx: torch.Tensor, t0: torch.Tensor, t1: torch.Tensor, sigma: torch.Tensor, n_rays: int, early_stop_eps: float, alpha_thre: float):
    """nerfacc.render_visibility_from_density: T_i >= early_stop_eps and alpha_i >= alpha_thre."""
    alpha = 1.0 - torch.exp(-sigma * (t1 - t0))
    vis = torch.zeros_like(alpha, dtype=torch.bool)
    for r in range(n_rays):
        idx = (ray_idx == r).nonzero().squeeze(1)
        T = torch.tensor(1.0)
        for i in idx.tolist():
            vis[i] = bool(T >= early_stop_eps) and bool(alpha[i] >= alpha_thre)
            T = T * (1.0 - alpha[i])
    return vis


def union(ray_lists: Sequence[torch.Tensor], t0_lists: Sequence[torch.Tensor], t1_lists: Sequence[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """nerfs/ray_rendering.py:196-256, with an exact (ray, t0) lexicographic sort in place of the reference's ri * 1e9 + t0 key."""
    ri, t0, t1 = torch.cat(list(ray_lists)), torch.cat(list(t0_lists)), torch.cat(list(t1_lists))
    out_r: List[torch.Tensor] = []
    out_a: List[torch.Tensor] = []
    out_b: List[torch.Tensor] = []
    for r in torch.unique(ri).tolist():
        m = ri == r
        bnd = torch.unique(torch.cat([t0[m], t1[m]]), sorted=True)
        if bnd.numel() < 2:
            continue
        out_r.append(torch.full((bnd.numel() - 1,), r, dtype=torch.int64))
        out_a.append(bnd[:-1])
        out_b.append(bnd[1:])
    if not out_a:
        return torch.zeros(0, dtype=torch.int64), torch.zeros(0), torch.zeros(0)
    return torch.cat(out_r), torch.cat(out_a), torch.cat(out_b)


def packed_render(ray_idx: torch.Tensor, t0: torch.Tensor, t1: torch.Tensor, sigma: torch.Tensor, rgb: torch.Tensor, n_rays: int):
    """render_weight_from_density (alpha = 1 - exp(-sigma dt), w = alpha * exclusive cumprod(1 - alpha)) and the segmented sums
    of accumulate_along_rays; differentiable w.r.t. sigma and rgb.  -> rgb_map (N,3), depth (N,), weights (M,), acc (N,)"""
    alpha = 1.0 - torch.exp(-sigma * (t1 - t0))
    w = torch.zeros_like(alpha)
    for r in torch.unique(ray_idx).tolist():
        idx = (ray_idx == r).nonzero().squeeze(1)
        a = alpha[idx]
        T = torch.cumprod(torch.cat([torch.ones(1), 1.0 - a]), 0)[:-1]
        w = w.index_add(0, idx, a * T)
    t_mid = 0.5 * (t0 + t1)
    rgb_map = torch.zeros(n_rays, 3).index_add(0, ray_idx, w[:, None] * rgb)
    depth = torch.zeros(n_rays).index_add(0, ray_idx, w * t_mid)
    acc = torch.zeros(n_rays).index_add(0, ray_idx, w)
    return rgb_map, depth, w, acc


# --------------------------------------------------------------------------------------------------------------------------
# The reference's OWN part of the occupancy renderer (nerfs/ray_rendering.py), pinned: tests/golden/occ_assembly.npz holds
# the outputs of the unmodified _intersect_rays_aabb / _merge_segments_union / render_rays_occ / render_expert_occ, run with
# nerfacc's three packed primitives supplied by the functions below (tests/golden/make_golden_occ.py).  What stays unpinned
# is nerfacc's arithmetic itself (march / visibility / the primitives).
# --------------------------------------------------------------------------------------------------------------------------
def pack_info(ray_idx: torch.Tensor, n_rays: int) -> torch.Tensor:
    """nerfacc.pack_info: (n_rays, 2) [start, count] of each ray's run in a ray-sorted sample list."""
    counts = torch.bincount(ray_idx, minlength=n_rays)
    starts = torch.cumsum(counts, 0) - counts
    return torch.stack([starts, counts], 1)


def weights_from_density(t0: torch.Tensor, t1: torch.Tensor, sigma: torch.Tensor, ray_idx: torch.Tensor):
    """nerfacc.render_weight_from_density -> (weights, transmittance, alphas)."""
    alpha = 1.0 - torch.exp(-sigma * (t1 - t0))
    T = torch.ones_like(alpha)
    for r in torch.unique(ray_idx).tolist():
        idx = (ray_idx == r).nonzero().squeeze(1)
        T = T.index_copy(0, idx, torch.cumprod(torch.cat([torch.ones(1), 1.0 - alpha[idx]]), 0)[:-1])
    return alpha * T, T, alpha


def accumulate(w: torch.Tensor, values, ray_idx: torch.Tensor, n_rays: int) -> torch.Tensor:
    """nerfacc.accumulate_along_rays: (n_rays, D) segmented sums of w * values (values None: of w, D = 1)."""
    src = w[:, None] if values is None else w[:, None] * values
    return torch.zeros(n_rays, src.shape[1], dtype=src.dtype).index_add(0, ray_idx, src)


def toy_field(xd: torch.Tensor, k: int) -> torch.Tensor:
    """A closed-form expert for the assembly goldens: (m,6) [x, d] -> (m,4) [rgb, sigma]."""
    x, d = xd[:, :3], xd[:, 3:6]
    c = 0.3 * torch.tensor([math.sin(k + 1.0), math.cos(2.0 * k + 1.0), math.sin(3.0 * k + 2.0)])
    sigma = 6.0 * torch.exp(-4.0 * ((x - c) ** 2).sum(-1)) + 0.05 * (k + 1)
    rgb = 0.5 + 0.5 * torch.sin(3.0 * x + (k + 1.0) * d)
    return torch.cat([rgb, sigma[:, None]], 1)


def toy_background(d: torch.Tensor) -> torch.Tensor:
    return 0.5 + 0.5 * d


def intersect_rays_aabb(rays: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor) -> torch.Tensor:
    """nerfs/ray_rendering.py:172-193: slab test against [near, far]; |d| <= 1e-9 uses 1/1e-9."""
    o, d = rays[:, :3], rays[:, 3:6]
    invd = torch.where(d.abs() > 1e-9, 1.0 / d, torch.full_like(d, 1.0 / 1e-9))
    a, b = (lo[None] - o) * invd, (hi[None] - o) * invd
    t_in = torch.maximum(torch.minimum(a, b).amax(-1), rays[:, 6])
    t_out = torch.minimum(torch.maximum(a, b).amin(-1), rays[:, 7])
    return t_out > t_in


def assemble_expert(rays: torch.Tensor, ray_idx: torch.Tensor, t0: torch.Tensor, t1: torch.Tensor, field: Callable, bg_rgb: torch.Tensor):
    """nerfs/ray_rendering.py:494-559 (render_expert_occ) from the marching result on: midpoints, one expert call, one packed
    integration, background.  -> rgb (N,3), depth (N,), weights (M,1), acc (N,)"""
    N = rays.shape[0]
    if t0.numel() == 0:
        return bg_rgb, torch.zeros(N), torch.zeros(1, 1), torch.zeros(N)
    t_mid = 0.5 * (t0 + t1)
    y = field(torch.cat([rays[ray_idx, :3] + rays[ray_idx, 3:6] * t_mid[:, None], rays[ray_idx, 3:6]], 1))
    w, _, _ = weights_from_density(t0, t1, y[:, 3], ray_idx)
    acc = accumulate(w, None, ray_idx, N)[:, 0]
    rgb = accumulate(w, y[:, :3], ray_idx, N) + (1.0 - acc)[:, None] * bg_rgb
    return rgb, accumulate(w, t_mid[:, None], ray_idx, N)[:, 0], w[:, None], acc


def assemble_container(rays: torch.Tensor, lists: Sequence[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]], route: Callable,
                       fields: Sequence[Callable], bg_rgb: torch.Tensor):
    """nerfs/ray_rendering.py:400-491 (render_rays_occ, full container) from the per-expert marching results on: union of the
    segments per ray, routing weights at the midpoints (route(x) -> (M,K)), experts evaluated where their weight > 1e-8, sigma
    and sigma-weighted rgb blended BEFORE one packed integration, background.  lists: per expert (global ray ids, t0, t1), the
    experts without a sample left out.  -> rgb (N,3), depth (N,), weights (M,1), acc (N,), (merged ray ids, t0, t1)"""
    N, K = rays.shape[0], len(fields)
    if len(lists) == 0:
        return bg_rgb, torch.zeros(N), torch.zeros(1, 1), torch.zeros(N), None
    ri, t0, t1 = union([l[0] for l in lists], [l[1] for l in lists], [l[2] for l in lists])
    t_mid = 0.5 * (t0 + t1)
    x, d = rays[ri, :3] + rays[ri, 3:6] * t_mid[:, None], rays[ri, 3:6]
    W = route(x)
    SIG, RGB = torch.zeros(len(ri), K), torch.zeros(len(ri), K, 3)
    for k in range(K):
        idx = (W[:, k] > 1e-8).nonzero().squeeze(1)
        if idx.numel():
            y = fields[k](torch.cat([x[idx], d[idx]], 1))
            SIG[idx, k], RGB[idx, k] = y[:, 3], y[:, :3]
    s_num = (W * SIG).sum(1, keepdim=True).clamp_min(1e-12)
    rgb_mix = (W[..., None] * SIG[..., None] * RGB).sum(1) / s_num
    w, _, _ = weights_from_density(t0, t1, s_num[:, 0], ri)
    acc = accumulate(w, None, ri, N)[:, 0]
    rgb = accumulate(w, rgb_mix, ri, N) + (1.0 - acc)[:, None] * bg_rgb
    return rgb, accumulate(w, t_mid[:, None], ri, N)[:, 0], w[:, None], acc, (ri, t0, t1)
