"""The renderer, with the reference's entry points (``nerfs/ray_rendering.py``: render_rays :564-574,
render_rays_stratified :290-345, stratified_t_vals :262-287, volume_render :114-165,
render_image :577-627).

What changes is the data flow, not the arithmetic: sample positions are never materialised
(the (N*S,6) staging tensor of the reference, :317-319, is 3.2 GB at 1080p x 64 samples);
the kernels read the packed rays and the (N,S) sample distances directly."""
from __future__ import annotations

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


def render_rays_stratified(model, rays: Tensor, ray_samples: int, params=None, active_module: Optional[int] = None,
                           bg_color_default: str = "white", chunk: int = 1_000_000, sigma_scale=1.0, u: Optional[Tensor] = None,
                           coherent: bool = False, route_checks: Optional[list] = None, **kwargs):
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
    if sample_major:
        pts = pts.sample_major()
    rgb_sigma = model_eff.evaluate(pts, params=params)
    bg_view = rgb_sigma.view(S, N, 4).transpose(0, 1) if sample_major else rgb_sigma.view(N, S, 4)  # only for the bg policy's shape / last sample
    bg_rgb = _get_bg_rgb(model, rays[:, 3:6], params, bg_view, N=N, bg_color_default=bg_color_default)
    if sample_major:
        return ops.composite_sample_major(rgb_sigma, pts.tT, bg_rgb, float(sigma_scale))
    return ops.composite(rgb_sigma.view(N, S, 4), t_vals, bg_rgb, float(sigma_scale))


def render_rays(model, rays, *args, **kwargs):
    if getattr(model, "use_occ", False) and getattr(model, "occ_ready", False):
        raise NotImplementedError("occupancy renderer (SURVEY 8(f)-1) is not built; the stratified renderer is the production path")
    return render_rays_stratified(model, rays, *args, **kwargs)


@torch.no_grad()
def render_image(model, *, H: int, W: int, fx: float, fy: float, cx: float, cy: float, c2w: Tensor, scene_box, params=None,
                 active_module: Optional[int] = None, ray_samples: int = 64, chunk_points: int = 1 << 16,
                 bg_color_default: str = "white", center_pixels: bool = True, use_amp: bool = False, rows_per_chunk: Optional[int] = None,
                 row_range: Optional[Tuple[int, int]] = None):
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
                                     bg_color_default=bg_color_default, chunk=chunk_points, coherent=True, route_checks=checks)
            sl = slice((r0 - R0) * W, (r0 - R0 + rows) * W)
            rgb[sl], depth[sl], acc[sl] = c, d, a
        if checks is None or ops.route_capacity_ok(checks):
            break
    if row_range is not None:
        return rgb.clamp_(0, 1), depth, acc
    return rgb.view(H, W, 3).clamp_(0, 1), depth.view(-1), acc.view(-1)
