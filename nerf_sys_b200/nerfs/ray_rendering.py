"""The renderer, with the reference's entry points (``nerfs/ray_rendering.py``: render_rays :564-574,
render_rays_stratified :290-345, stratified_t_vals :262-287, volume_render :114-165,
render_image :577-627).

What changes is the data flow, not the arithmetic: sample positions are never materialised
(the (N*S,6) staging tensor of the reference, :317-319, is 3.2 GB at 1080p x 64 samples);
the kernels read the packed rays and the (N,S) sample distances directly."""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
from torch import Tensor

from .. import ops
from .ray_sampling import clamp_rays_near_far, get_ray_directions, get_rays  # noqa: F401


def get_bg_default_color(rgb_sigma, N: int, bg_color: str = "white") -> Optional[Tensor]:
    device = None if rgb_sigma is None else rgb_sigma.device
    dtype = None if rgb_sigma is None else rgb_sigma.dtype
    if bg_color == "none":
        return None
    if bg_color == "white":
        return torch.ones(N, 3, device=device, dtype=dtype)
    if bg_color == "black":
        return torch.zeros(N, 3, device=device, dtype=dtype)
    if bg_color == "random":
        return torch.rand(N, 3, device=device, dtype=dtype)
    if bg_color == "last_sample":
        if rgb_sigma is None or rgb_sigma.dim() != 3 or rgb_sigma.size(-1) < 3:
            raise ValueError("bg_color='last_sample' requires rgb_sigma of shape (N,S,4) or (N,S,>=3).")
        return rgb_sigma[:, -1, :3]
    raise ValueError(f"Unknown background policy: {bg_color}")


def _get_bg_rgb(model, dirs: Tensor, params, rgb_sigma, N: int, bg_color_default: str) -> Optional[Tensor]:
    if getattr(model, "use_bg_nerf", False):
        return model.background_color(dirs)
    return get_bg_default_color(rgb_sigma, N, bg_color_default)


def volume_render(rgb_sigma: Tensor, t_vals: Tensor, bg_rgb: Optional[Tensor] = None, *, raw_rgb: bool = False,
                  raw_sigma: bool = False, sigma_scale: float = 1.0, **kwargs) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """(N,S,4),(N,S),(N,3)|None -> rgb (N,3), depth (N,), weights (N,S), acc (N,).  No early termination."""
    if raw_rgb or raw_sigma:
        from ..models.trunc_exp import trunc_exp
        rgb = torch.sigmoid(rgb_sigma[..., :3]) if raw_rgb else rgb_sigma[..., :3]
        sig = trunc_exp(rgb_sigma[..., 3:4]) if raw_sigma else rgb_sigma[..., 3:4]
        rgb_sigma = torch.cat([rgb, sig], dim=-1)
    return ops.composite(rgb_sigma, t_vals, bg_rgb, sigma_scale)


@torch.no_grad()
def stratified_t_vals(near: Tensor, far: Tensor, ray_samples: int, randomized: bool = True, u: Optional[Tensor] = None) -> Tensor:
    """(N,),(N,) -> t (N,S).  ``u`` optionally injects the jitter (the reference draws torch.rand_like)."""
    rays = torch.zeros(near.shape[0], 8, device=near.device, dtype=torch.float32)
    rays[:, 6], rays[:, 7] = near, far
    if randomized and u is None:
        u = torch.rand(near.shape[0], ray_samples, device=near.device, dtype=torch.float32)
    return ops.stratified_t(rays, int(ray_samples), u if randomized else None)


@torch.no_grad()
def _evaluate_front_to_back(model, rays: Tensor, t_vals: Tensor, params, route_checks, eps: float, group: int, sigma_scale: float) -> Tensor:
    """(S*N, 4) [rgb, sigma] in sample-major order, evaluated group by group; terminated rays are skipped (see the caller)."""
    N, S = t_vals.shape
    out = torch.empty(S * N, 4, device=rays.device, dtype=torch.float32)
    delta = torch.empty_like(t_vals)
    delta[:, :-1] = (t_vals[:, 1:] - t_vals[:, :-1]).clamp_min(1e-4)  # ray_rendering.py:141-146
    delta[:, -1] = delta[:, -2]
    tau = torch.zeros(N, device=rays.device)  # optical depth so far
    alive = None
    for s0 in range(0, S, group):
        s1 = min(S, s0 + group)
        t_g = t_vals[:, s0:s1]
        if alive is not None:
            t_g = torch.where(alive[:, None], t_g, torch.full_like(t_g, float("nan")))
        pts_g = ops.PointSet(rays=rays, t=t_g.contiguous())
        pts_g.route_checks = route_checks
        y = model.evaluate(pts_g.sample_major(), params=params)
        out[s0 * N:s1 * N] = y
        if s1 < S:
            sig = y.view(s1 - s0, N, 4)[..., 3].clamp_min(0) * sigma_scale
            tau += (sig * delta[:, s0:s1].t()).sum(0)
            alive = tau < -math.log(eps)  # T = exp(-tau) >= eps
    return out


def render_rays_stratified(model, rays: Tensor, ray_samples: int, params=None, active_module: Optional[int] = None,
                           bg_color_default: str = "white", chunk: int = 1_000_000, sigma_scale=1.0, u: Optional[Tensor] = None,
                           coherent: bool = False, route_checks: Optional[list] = None, early_stop_eps: float = 0.0,
                           early_stop_group: int = 16, **kwargs):
    """Stratified renderer: rgb (N,3), depth (N,), weights (N,S), acc (N,).

    ``chunk`` is accepted for signature compatibility; nothing is staged per point, so there is
    nothing to chunk (callers bound memory by the number of rays they pass).  ``u`` (N,S) injects the
    stratified jitter for parity tests; by default it is drawn with torch.rand when model.training."""
    rays = ops._f32(rays, "rays")
    N, S = rays.shape[0], int(ray_samples)
    with torch.no_grad():
        if model.training and u is None:
            u = torch.rand(N, S, device=rays.device, dtype=torch.float32)
        t_vals = ops.stratified_t(rays, S, u if model.training else None)
    pts = ops.PointSet(rays=rays, t=t_vals)
    pts.route_checks = route_checks  # render_image verifies the routing capacity once per frame instead of once per block
    # as in the reference (:323-326): with active_module the expert is called directly and `params`
    # holds EXPERT-level keys (meta_core.extract_module_params); otherwise container-level keys
    model_eff = model.submodules[active_module] if active_module is not None else model
    # A block of image rows (render_image), no autograd: number the points sample-major, so that the lanes of the routing /
    # gather / MLP kernels are the same sample of adjacent pixels and share grid cells (measured, 1080p x 8 experts: frame
    # 53.9 -> 48 ms); the compositor reads the (S,N,4) result in place.  Every kernel is order-agnostic (a single expert's
    # identity slot list simply becomes sample-major too); sharded experts keep ray-major rows.
    sample_major = coherent and not torch.is_grad_enabled() and getattr(model, "_ep", None) is None
    # Transmittance early termination (additive: the reference declares `early_stop_eps` in its viewer presets,
    # viewer/viewer.py:46-57, and never passes it on; 0 = off = the reference's arithmetic).  Front to back in groups of
    # `early_stop_group` samples: a ray whose transmittance has fallen below eps gets NaN sample distances for the rest of
    # its samples, the routing kernel gives a non-finite position no expert (meta_container.py:119-126), so neither the
    # gather nor the MLPs see those samples and their [rgb, sigma] rows stay zero.  The image differs from the full
    # evaluation by at most eps per channel (the dropped weights sum to less than T < eps).
    if (early_stop_eps > 0.0 and sample_major and active_module is None and len(getattr(model, "submodules", ())) > 1
            and S > early_stop_group):
        rgb_sigma = _evaluate_front_to_back(model, rays, t_vals, params, route_checks, float(early_stop_eps), int(early_stop_group),
                                            float(sigma_scale))
        pts = pts.sample_major()
    else:
        if sample_major:
            pts = pts.sample_major()
        rgb_sigma = model_eff.evaluate(pts, params=params)
    bg_view = rgb_sigma.view(S, N, 4).transpose(0, 1) if sample_major else rgb_sigma.view(N, S, 4)  # only for the bg policy's shape / last sample
    bg_rgb = _get_bg_rgb(model, rays[:, 3:6], params, bg_view, N=N, bg_color_default=bg_color_default)
    if sample_major:
        return ops.composite_sample_major(rgb_sigma, pts.tT, bg_rgb, float(sigma_scale))
    return ops.composite(rgb_sigma.view(N, S, 4), t_vals, bg_rgb, float(sigma_scale))


# ============================== occupancy rendering (ray_rendering.py:171-256, 349-558) ===============================
@torch.no_grad()
def _intersect_rays_aabb(rays: Tensor, scene_box) -> Tensor:
    """(N,) bool: the ray's [near, far] overlaps the box (slab test with the reference's 1e-9 direction guard, :171-193)."""
    o, d = rays[:, :3], rays[:, 3:6]
    near, far = rays[:, 6:7], rays[:, 7:8]
    eps = 1e-9
    invd = torch.where(torch.abs(d) > eps, 1.0 / d, torch.full_like(d, 1.0 / eps))
    t0 = (scene_box.min[None, :].to(rays.device) - o) * invd
    t1 = (scene_box.max[None, :].to(rays.device) - o) * invd
    tmin = torch.minimum(t0, t1).amax(dim=-1, keepdim=True)
    tmax = torch.maximum(t0, t1).amin(dim=-1, keepdim=True)
    return (torch.minimum(tmax, far) > torch.maximum(tmin, near)).squeeze(-1)


@torch.no_grad()
def _merge_segments_union(ray_indices_list, t0_list, t1_list, n_rays: Optional[int] = None):
    """Per-ray union of the experts' segment boundaries (:196-256) - one kernel (csrc/occ.cu k_seg_union) instead of the
    reference's Python loop over rays.  ``n_rays`` defaults to max ray id + 1."""
    if len(t0_list) == 0:
        z = torch.zeros(0)
        return z.long(), z.float(), z.float()
    N = int(n_rays) if n_rays is not None else int(max(int(r.max().item()) for r in ray_indices_list if r.numel()) + 1)
    lists = [ops.pack_segments(ri, a, b, N) for ri, a, b in zip(ray_indices_list, t0_list, t1_list)]
    m = ops.seg_union(lists, N)
    return m.ray_idx.long(), m.t0, m.t1


def _packed_integrate(model, rays: Tensor, params, seg: "ops.PackedSegments", sigma: Tensor, rgb: Tensor, bg_color_default: str):
    """nerfacc.render_weight_from_density + accumulate_along_rays + background (:477-491, :543-558)."""
    N = rays.shape[0]
    rgb_map, depth, w, acc = ops.packed_render(sigma, rgb, seg)
    bg_rgb = _get_bg_rgb(model, rays[:, 3:6], params, rgb_map, N, bg_color_default)
    rgb_map = rgb_map + (1.0 - acc)[..., None] * bg_rgb
    return rgb_map, depth, w[:, None], acc


def _empty_occ_result(model, rays: Tensor, params, bg_color_default: str):
    N = rays.shape[0]
    acc = rays.new_zeros(N)
    bg_rgb = _get_bg_rgb(model, rays[:, 3:6], params, None, N, bg_color_default)
    return bg_rgb, acc.clone(), torch.zeros(1, 1, device=rays.device, dtype=rays.dtype), acc


def render_expert_occ(model, rays: Tensor, *, params=None, bg_color_default: str = "white", chunk: int = 1_000_000,
                      render_step_size=None, alpha_thre=None, cone_angle=None):
    """One expert through its own occupancy grid (:494-558): march, evaluate the expert at the segment midpoints, integrate."""
    rays = ops._f32(rays, "rays")
    N = rays.shape[0]
    ri, t0, t1 = model.occupancy_marching(rays, params=params, render_step_size=render_step_size, alpha_thre=alpha_thre, cone_angle=cone_angle)
    if t0.numel() == 0:
        return _empty_occ_result(model, rays, params, bg_color_default)
    seg = ops.pack_segments(ri, t0, t1, N)
    t_mid = 0.5 * (t0 + t1)
    x6 = torch.cat([rays[ri, :3] + rays[ri, 3:6] * t_mid[:, None], rays[ri, 3:6]], dim=-1)
    out = model(x6, params=params)
    return _packed_integrate(model, rays, params, seg, out[:, 3], out[:, :3], bg_color_default)


def render_rays_occ(model, rays: Tensor, *, params=None, bg_color_default: str = "white", chunk: int = 1_000_000, render_step_size=None,
                    alpha_thre=None, cone_angle=None, active_module: Optional[int] = None, **kwargs):
    """Occupancy-guided soft mixture-of-experts renderer (:349-491): per-expert AABB prefilter and grid marching, per-ray
    union of the experts' segments, experts evaluated at the midpoints where their routing weight is positive,
    sigma = sum_k w_k sigma_k and rgb = sum_k w_k sigma_k rgb_k / sigma blended BEFORE one integration.
    Returns rgb (N,3), depth (N,), weights (M,1) packed, acc (N,)."""
    rays = ops._f32(rays, "rays")
    N = rays.shape[0]
    if active_module is not None:
        return render_expert_occ(model.submodules[active_module], rays, params=params, bg_color_default=bg_color_default, chunk=chunk,
                                 render_step_size=render_step_size, alpha_thre=alpha_thre, cone_angle=cone_angle)
    lists = []
    for k, expert in enumerate(model.submodules):
        hit = _intersect_rays_aabb(rays, expert.scene_box)
        if not bool(hit.any()):
            continue
        hit_idx = hit.nonzero(as_tuple=False).squeeze(1)
        ri_k, t0_k, t1_k = expert.occupancy_marching(rays[hit_idx], params=(model.get_subdict(params, f"submodules.{k}") if params is not None else None),
                                                     render_step_size=render_step_size, alpha_thre=alpha_thre, cone_angle=cone_angle)
        if t0_k.numel() == 0:
            continue
        lists.append(ops.pack_segments(hit_idx[ri_k], t0_k, t1_k, N))
    if not lists:
        return _empty_occ_result(model, rays, params, bg_color_default)
    seg = ops.seg_union(lists, N)
    if seg.M == 0:
        return _empty_occ_result(model, rays, params, bg_color_default)
    ri = seg.ray_idx.long()
    t_mid = 0.5 * (seg.t0 + seg.t1)
    x_mid = rays[ri, :3] + rays[ri, 3:6] * t_mid[:, None]
    d_mid = rays[ri, 3:6]
    K = len(model.submodules)
    with torch.no_grad():
        W, hard = model._routing(x_mid)
        if W is None:
            W = torch.nn.functional.one_hot(hard.view(-1), K).to(x_mid.dtype)
    # sum_k W sigma_k (clamped at 1e-12) and sum_k W sigma_k rgb_k, accumulated expert by expert (the reference fills dense (M,K)
    # and (M,K,3) tensors first; the sums are the same terms in the same order k = 0..K-1)
    s_num = x_mid.new_zeros(seg.M)
    c_num = x_mid.new_zeros(seg.M, 3)
    for k, expert in enumerate(model.submodules):
        idx = torch.nonzero(W[:, k] > 1e-8, as_tuple=False).squeeze(1)
        if idx.numel() == 0:
            continue
        out = expert(torch.cat([x_mid[idx], d_mid[idx]], dim=-1), params=(model.get_subdict(params, f"submodules.{k}") if params is not None else None))
        ws = W[idx, k] * out[:, 3]
        s_num = s_num.index_add(0, idx, ws)
        c_num = c_num.index_add(0, idx, ws[:, None] * out[:, :3])
    sigma_mix = s_num.clamp_min(1e-12)
    rgb_mix = c_num / sigma_mix[:, None]
    return _packed_integrate(model, rays, params, seg, sigma_mix, rgb_mix, bg_color_default)


def render_rays_hierarchical(model, rays: Tensor, ray_samples: int, proposal_samples: int, params=None, active_module: Optional[int] = None,
                             bg_color_default: str = "white", sigma_scale=1.0, **kwargs):
    """Coarse-to-fine rendering with proposal resampling - ADDITIVE: the reference has no proposal sampler (SURVEY headline 4);
    north_star names one.  A no-grad coarse pass over `ray_samples` stratified samples gives the weights; `proposal_samples`
    new distances are drawn from them by inverse-CDF (ops.sample_pdf: stratified u in training, bin centres in eval) and merged
    with the coarse ones; the fine pass over the merged distances is the differentiable one.  Returns the fine pass's
    (rgb, depth, weights (N, S+S2), acc)."""
    rays = ops._f32(rays, "rays")
    N, S, S2 = rays.shape[0], int(ray_samples), int(proposal_samples)
    with torch.no_grad():
        u_c = torch.rand(N, S, device=rays.device) if model.training else None  # the coarse pass and the bins share one jitter
        _, _, w, _ = render_rays_stratified(model, rays, S, params=params, active_module=active_module, bg_color_default=bg_color_default,
                                            sigma_scale=sigma_scale, u=u_c)
        t = ops.stratified_t(rays, S, u_c)
        base = (torch.arange(S2, device=rays.device, dtype=torch.float32) + 0.5) / S2
        u = base[None, :].expand(N, S2)
        if model.training:
            u = (torch.arange(S2, device=rays.device, dtype=torch.float32)[None, :] + torch.rand(N, S2, device=rays.device)) / S2
        _, t_all = ops.sample_pdf(t, w.contiguous(), u.contiguous())
    model_eff = model.submodules[active_module] if active_module is not None else model
    pts = ops.PointSet(rays=rays, t=t_all)
    rgb_sigma = model_eff.evaluate(pts, params=params)
    bg_rgb = _get_bg_rgb(model, rays[:, 3:6], params, rgb_sigma.view(N, S + S2, 4), N=N, bg_color_default=bg_color_default)
    return ops.composite(rgb_sigma.view(N, S + S2, 4), t_all, bg_rgb, float(sigma_scale))


def render_rays(model, rays, *args, **kwargs):
    """Dispatch of ray_rendering.py:563-574: the occupancy renderer once the model's grids are ready, else stratified."""
    if getattr(model, "use_occ", False) and getattr(model, "occ_ready", False):
        kwargs.pop("ray_samples", None)
        for k in ("u", "coherent", "route_checks", "sigma_scale"):
            kwargs.pop(k, None)
        return render_rays_occ(model, rays, *args, **kwargs)
    return render_rays_stratified(model, rays, *args, **kwargs)


@torch.no_grad()
def render_image(model, *, H: int, W: int, fx: float, fy: float, cx: float, cy: float, c2w: Tensor, scene_box, params=None,
                 active_module: Optional[int] = None, ray_samples: int = 64, chunk_points: int = 1 << 16,
                 bg_color_default: str = "white", center_pixels: bool = True, use_amp: bool = False, rows_per_chunk: Optional[int] = None,
                 row_range: Optional[Tuple[int, int]] = None, early_stop_eps: float = 0.0):
    """Full frame: rays are generated on the device (fused a1+a2) and rendered in row blocks.
    ``row_range=(r0, r1)`` renders only those image rows (the ray-sharded multi-GPU render,
    nerf_sys_b200.parallel.render_image_sharded) and returns rgb ((r1-r0)*W, 3), depth, acc for them."""
    device = next(model.parameters()).device
    aabb = None if scene_box is None else scene_box.aabb
    if rows_per_chunk is None:
        rows_per_chunk = max(1, min(H, (1 << 24) // max(1, W * ray_samples)))  # ~16M sample points (~2.5 GB of staging) per block
    R0, R1 = (0, H) if row_range is None else (int(row_range[0]), int(row_range[1]))
    if not (0 <= R0 <= R1 <= H):
        raise ValueError(f"row_range {row_range} outside [0, {H}]")
    n_rows = R1 - R0
    rgb = torch.empty(n_rows * W, 3, device=device)
    depth = torch.empty(n_rows * W, device=device)
    acc = torch.empty(n_rows * W, device=device)
    # First pass: no host read inside the frame (the routing of every block keeps its speculative capacity and the whole
    # frame is verified with one read at the end); in the rare overflow (a very wide boundary margin) the frame is redone
    # with the per-block read that sizes the lists exactly.
    for checks in ([], None):
        for r0 in range(R0, R1, rows_per_chunk):
            rows = min(rows_per_chunk, R1 - r0)
            rays = ops.raygen(H, W, fx, fy, cx, cy, center_pixels, c2w, aabb, device, row0=r0, rows=rows)
            c, d, _, a = render_rays(model, rays, ray_samples=ray_samples, params=params, active_module=active_module,
                                     bg_color_default=bg_color_default, chunk=chunk_points, coherent=True, route_checks=checks,
                                     early_stop_eps=early_stop_eps)
            sl = slice((r0 - R0) * W, (r0 - R0 + rows) * W)
            rgb[sl], depth[sl], acc[sl] = c, d, a
        if checks is None or ops.route_capacity_ok(checks):
            break
    if row_range is not None:
        return rgb.clamp_(0, 1), depth, acc
    return rgb.view(H, W, 3).clamp_(0, 1), depth.view(-1), acc.view(-1)
