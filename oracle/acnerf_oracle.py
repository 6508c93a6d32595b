"""CPU oracle for the Adaptive City NeRF per-ray hot path.

TEST INFRASTRUCTURE ONLY.  This file is a from-scratch restatement, in plain
fp32 PyTorch-on-CPU tensor arithmetic, of the algorithm the reference
(psklavos1/NeRF-Sys, ``adaptive_nerf/``) runs on its pure-PyTorch ("torch"
backend) path.  It is imported only by ``tests/``, ``__graft_entry__.smoke()``
and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.  Nothing
under ``nerf_sys_b200/`` may import it: the product path is CUDA-only.

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the real
reference from ``/root/reference`` (in the build container, where it exists),
runs it on seeded inputs and commits the input/output vectors as
``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every function
below against those vectors (integers exactly, fp32 to <=1e-6), and, when
``/root/reference`` is present, against the live reference as well.

All models are *functional*: parameters travel in a flat dict that uses the
reference's own ``state_dict`` keys, so the same dict
loads into the CUDA modules.  Gradients for backward parity come from torch
autograd over these functions.

Citations are ``file:line`` under ``/root/reference/adaptive_nerf/``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

HASH_PRIME_Y = 2654435761  # models/encodings.py:273
HASH_PRIME_Z = 805459861
EXP_CLAMP_F32 = 88.722839111  # models/trunc_exp.py:30-35


# --------------------------------------------------------------------------
# a1-a3  ray generation, AABB slab test, near/far clamp
# --------------------------------------------------------------------------
def ray_directions(H: int, W: int, fx: float, fy: float, cx: float, cy: float,
                   center_pixels: bool = True) -> Tensor:
    """Unit camera-frame pinhole directions (H,W,3), RUB camera looking down -z.

    nerfs/ray_sampling.py:111-136.
    """
    rows = torch.arange(H, dtype=torch.float32).view(H, 1).expand(H, W)
    cols = torch.arange(W, dtype=torch.float32).view(1, W).expand(H, W)
    if center_pixels:
        cols = cols + 0.5
        rows = rows + 0.5
    v = torch.stack([(cols - cx) / fx, -(rows - cy) / fy, -torch.ones(H, W)], dim=-1)
    return v / v.norm(dim=-1, keepdim=True).clamp_min(1e-12)


def aabb_near_far(o: Tensor, d: Tensor, aabb: Tensor, eps: float = 1e-8,
                  max_bound: float = 1e10, invalid_value: float = 1e10) -> Tuple[Tensor, Tensor]:
    """Slab test; nerfs/scene_box.py:45-107."""
    tiny = torch.full_like(d, eps)
    d_safe = torch.where(d.abs() < eps, torch.where(d >= 0, tiny, -tiny), d)
    inv = 1.0 / d_safe
    ta = (aabb[0][None, :] - o) * inv
    tb = (aabb[1][None, :] - o) * inv
    t_in = torch.minimum(ta, tb).amax(dim=-1).clamp(min=0.0, max=max_bound)
    t_out = torch.maximum(ta, tb).amin(dim=-1).clamp(min=0.0, max=max_bound)
    miss = t_out <= t_in
    bad = torch.tensor(invalid_value, dtype=o.dtype)
    return torch.where(miss, bad, t_in), torch.where(miss, bad, t_out)


def make_rays(dirs_cam: Tensor, c2w: Tensor, aabb: Optional[Tensor] = None,
              near: Optional[float] = None, far: Optional[float] = None) -> Tensor:
    """Camera dirs -> packed world rays (...,8)=[o,d,near,far].

    nerfs/ray_sampling.py:10-24 (rotation, no re-normalisation), :50-108.
    """
    shape = dirs_cam.shape
    flat = dirs_cam.reshape(-1, 3)
    d = flat @ c2w[:3, :3].T
    o = c2w[:3, 3].expand_as(d)
    if aabb is not None:
        n, f = aabb_near_far(o, d, aabb)
    else:
        n = torch.full((d.shape[0],), float(near))
        f = torch.full((d.shape[0],), float(far))
    rays = torch.cat([o, d, n[:, None], f[:, None]], dim=-1)
    return rays.reshape(*shape[:-1], 8)


def clamp_near_far(rays: Tensor, override: Optional[Tuple[Optional[float], Optional[float]]],
                   eps: float = 1e-6, invalid_value: float = float("inf")) -> Tuple[Tensor, Tensor]:
    """nerfs/ray_sampling.py:139-176."""
    near, far = rays[:, 6], rays[:, 7]
    if override is None:
        ok = torch.isfinite(near) & torch.isfinite(far) & (far > near + eps)
        return rays, ok
    lo, hi = override
    if lo is not None:
        near = torch.maximum(near, torch.tensor(float(lo)))
    if hi is not None:
        far = torch.minimum(far, torch.tensor(float(hi)))
    ok = torch.isfinite(near) & torch.isfinite(far) & (far > near + eps)
    fill = torch.full_like(near, float(invalid_value))
    out = rays.clone()
    out[:, 6] = torch.where(ok, near, fill)
    out[:, 7] = torch.where(ok, far, fill)
    return out, ok


# --------------------------------------------------------------------------
# a4-a5  stratified sampling and sample positions
# --------------------------------------------------------------------------
def stratified_t(near: Tensor, far: Tensor, S: int, u: Optional[Tensor] = None) -> Tensor:
    """t (N,S).  ``u`` (N,S) in [0,1) is the injected jitter (the reference draws
    it with torch.rand_like, ray_rendering.py:286); None = eval mode.

    nerfs/ray_rendering.py:262-287.
    """
    tau = torch.linspace(0.0, 1.0, S)[None, :]
    t = near[:, None] * (1.0 - tau) + far[:, None] * tau
    if u is not None:
        mid = 0.5 * (t[:, :-1] + t[:, 1:])
        lo = torch.cat([t[:, :1], mid], dim=1)
        hi = torch.cat([mid, t[:, -1:]], dim=1)
        t = lo + (hi - lo) * u
    return t


def sample_points(rays: Tensor, t: Tensor) -> Tensor:
    """(N,8),(N,S) -> (N*S,6) [xyz, dir]; nerfs/ray_rendering.py:317-319."""
    o, d = rays[:, :3], rays[:, 3:6]
    p = o[:, None, :] + d[:, None, :] * t[:, :, None]
    return torch.cat([p, d[:, None, :].expand_as(p)], dim=-1).reshape(-1, 6)


# --------------------------------------------------------------------------
# a6  routing
# --------------------------------------------------------------------------
def routing(pts: Tensor, centroids: Tensor, margin: float, cluster_2d: bool,
            direct: bool = False) -> Tuple[Optional[Tensor], Optional[Tensor]]:
    """Soft weights (P,K) when margin>1, else hard argmin (P,).

    models/inr/meta_container.py:97-134.  ``direct=True`` evaluates the
    Euclidean distance as sqrt(sum((x-c)^2)) instead of torch.cdist's
    matmul expansion; the CUDA kernel's contract is the direct form and the
    integer decisions (argmin, membership) of both are compared in the tests.
    """
    axes = [1, 2] if cluster_2d else [0, 1, 2]
    x = pts[:, axes].float()
    c = centroids[:, axes].float()
    if direct:
        diff = x[:, None, :] - c[None, :, :]
        dist = (diff * diff).sum(-1).sqrt()
    else:
        dist = torch.cdist(x, c)
    if margin > 1.0:
        dist = dist.clamp_min(1e-6)
        inv = 1.0 / dist
        nearest = dist.min(dim=1, keepdim=True).values
        inv = inv * (dist <= margin * nearest)
        return inv / inv.sum(dim=1, keepdim=True).clamp_min(1e-6), None
    return None, dist.argmin(dim=1)


# --------------------------------------------------------------------------
# a8-a9  unit-cube mapping and the multi-resolution hash grid
# --------------------------------------------------------------------------
def level_resolutions(levels: int = 16, min_res: int = 16, max_res: int = 4096) -> Tensor:
    """int32 (L,); models/encodings.py:203-215."""
    g = 1.0 if levels <= 1 else float(math.exp((math.log(max_res) - math.log(min_res)) / (levels - 1)))
    lv = torch.arange(levels, dtype=torch.float32)
    return torch.floor(min_res * (g ** lv)).to(torch.int32)


def world_to_unit(x: Tensor, box_min: Tensor, extent: Tensor) -> Tensor:
    """models/inr/meta_ngp.py:155-158."""
    eps = torch.tensor(1e-6, dtype=torch.float32)
    return ((x - box_min) / extent).clamp(eps, 1.0 - eps)


def hash_corner_index(ix: Tensor, iy: Tensor, iz: Tensor, log2_T: int) -> Tensor:
    """models/encodings.py:308-316 (int64 arithmetic, wrap-around included)."""
    h = ix.to(torch.int64) ^ (iy.to(torch.int64) * HASH_PRIME_Y) ^ (iz.to(torch.int64) * HASH_PRIME_Z)
    return h % (1 << log2_T)


def hash_indices(x01: Tensor, res: Tensor, log2_T: int) -> Tuple[Tensor, Tensor]:
    """Row indices into the (L*T,F) table for the 8 cell corners, plus the
    fractional position.  Returns idx (P,L,8) int64 (corner order
    c = 4*bx + 2*by + bz, bit set = "+1" on that axis; level offset included)
    and frac (P,L,3).  models/encodings.py:331-358.
    """
    L = res.numel()
    s = x01[:, None, :] * res.to(torch.float32).view(1, L, 1)
    base = torch.floor(s)
    frac = s - base
    i0 = base.to(torch.int64)
    i1 = i0 + 1
    off = (torch.arange(L, dtype=torch.int64) * (1 << log2_T)).view(1, L)
    corners = []
    for bx in (0, 1):
        for by in (0, 1):
            for bz in (0, 1):
                cx = (i1 if bx else i0)[..., 0]
                cy = (i1 if by else i0)[..., 1]
                cz = (i1 if bz else i0)[..., 2]
                corners.append(hash_corner_index(cx, cy, cz, log2_T) + off)
    return torch.stack(corners, dim=-1), frac


def hashgrid_encode(x01: Tensor, table: Tensor, res: Tensor, log2_T: int,
                    interpolation: str = "Linear") -> Tensor:
    """(P,3) in [0,1] -> (P, L*F); models/encodings.py:331-381."""
    P, L, Fdim = x01.shape[0], res.numel(), table.shape[1]
    if interpolation == "Nearest":
        s = x01[:, None, :] * res.to(torch.float32).view(1, L, 1)
        r = torch.round(s).to(torch.int64)
        off = (torch.arange(L, dtype=torch.int64) * (1 << log2_T)).view(1, L)
        return table[hash_corner_index(r[..., 0], r[..., 1], r[..., 2], log2_T) + off].reshape(P, L * Fdim)
    idx, w = hash_indices(x01, res, log2_T)
    f = table[idx]  # (P,L,8,F)
    wx, wy, wz = w[..., 0:1], w[..., 1:2], w[..., 2:3]
    if interpolation == "Smoothstep":
        wx = wx * wx * (3 - 2 * wx)
        wy = wy * wy * (3 - 2 * wy)
        wz = wz * wz * (3 - 2 * wz)
    # corner c = 4*bx+2*by+bz; interpolate x, then y, then z (encodings.py:373-379)
    c00 = f[:, :, 0] * (1 - wx) + f[:, :, 4] * wx
    c01 = f[:, :, 1] * (1 - wx) + f[:, :, 5] * wx
    c10 = f[:, :, 2] * (1 - wx) + f[:, :, 6] * wx
    c11 = f[:, :, 3] * (1 - wx) + f[:, :, 7] * wx
    c0 = c00 * (1 - wy) + c10 * wy
    c1 = c01 * (1 - wy) + c11 * wy
    return (c0 * (1 - wz) + c1 * wz).reshape(P, L * Fdim)


# --------------------------------------------------------------------------
# a10-a11  direction / frequency encodings
# --------------------------------------------------------------------------
def sh16(d: Tensor) -> Tensor:
    """Real SH up to l=3 of d/max(|d|,1e-9); models/encodings.py:27-81, 133-151."""
    d = d / d.norm(dim=-1, keepdim=True).clamp_min(1e-9)
    x, y, z = d[..., 0], d[..., 1], d[..., 2]
    xx, yy, zz = x * x, y * y, z * z
    out = d.new_zeros(*d.shape[:-1], 16)
    out[..., 0] = 0.28209479177387814
    out[..., 1] = 0.4886025119029199 * y
    out[..., 2] = 0.4886025119029199 * z
    out[..., 3] = 0.4886025119029199 * x
    out[..., 4] = 1.0925484305920792 * x * y
    out[..., 5] = 1.0925484305920792 * y * z
    out[..., 6] = 0.9461746957575601 * zz - 0.31539156525251999
    out[..., 7] = 1.0925484305920792 * x * z
    out[..., 8] = 0.5462742152960396 * (xx - yy)
    out[..., 9] = 0.5900435899266435 * y * (3 * xx - yy)
    out[..., 10] = 2.890611442640554 * x * y * z
    out[..., 11] = 0.4570457994644658 * y * (5 * zz - 1)
    out[..., 12] = 0.3731763325901154 * z * (5 * zz - 3)
    out[..., 13] = 0.4570457994644658 * x * (5 * zz - 1)
    out[..., 14] = 1.445305721320277 * z * (xx - yy)
    out[..., 15] = 0.5900435899266435 * x * (xx - 3 * yy)
    return out


def freq_encode(x: Tensor, n_freq: int, include_input: bool = True) -> Tensor:
    """[x, per input dim: cos(2^k x) k<n, sin(2^k x) k<n]; models/encodings.py:437-444."""
    bands = 2.0 ** torch.arange(n_freq, dtype=torch.float32)
    a = x[..., None] * bands
    pe = torch.cat([torch.cos(a), torch.sin(a)], dim=-1).reshape(*x.shape[:-1], -1)
    return torch.cat([x, pe], dim=-1) if include_input else pe


# --------------------------------------------------------------------------
# a12-a14  hash-grid expert (MetaNGP)
# --------------------------------------------------------------------------
class _TruncExp(torch.autograd.Function):
    """exp(clamp(x)) whose backward is g*exp(clamp(x)) (not gated by the clamp);
    models/trunc_exp.py:43-57."""

    @staticmethod
    def forward(ctx, x):
        xc = x.clamp(-EXP_CLAMP_F32, EXP_CLAMP_F32)
        ctx.save_for_backward(xc)
        return torch.exp(xc)

    @staticmethod
    def backward(ctx, g):
        (xc,) = ctx.saved_tensors
        return g * torch.exp(xc)


def trunc_exp(x: Tensor) -> Tensor:
    return _TruncExp.apply(x)


def _lin(x: Tensor, p: Dict[str, Tensor], key: str) -> Tensor:
    return x.matmul(p[key + ".weight"].t()) + p[key + ".bias"]


def _mlp_depth(p: Dict[str, Tensor], stem: str) -> int:
    n = 0
    while f"{stem}.{n}.linear.weight" in p:
        n += 1
    return n


def ngp_density(x: Tensor, p: Dict[str, Tensor], box_min: Tensor, hash_cfg: dict) -> Tuple[Tensor, Tensor]:
    """sigma (P,1), geo (P,15); models/inr/meta_ngp.py:192-224."""
    res = level_resolutions(hash_cfg["levels"], hash_cfg["min_res"], hash_cfg["max_res"])
    x01 = world_to_unit(x, box_min, p["aabb_extent"])
    h = hashgrid_encode(x01, p["xyz_encoder.hash_table"], res, hash_cfg["log2_hashmap_size"],
                        hash_cfg.get("interpolation", "Linear"))
    for i in range(_mlp_depth(p, "sigma_trunk")):
        h = torch.relu(_lin(h, p, f"sigma_trunk.{i}.linear"))
    return trunc_exp(_lin(h, p, "sigma_head")), _lin(h, p, "geo_head")


def ngp_color(d: Tensor, geo: Tensor, p: Dict[str, Tensor]) -> Tensor:
    """models/inr/meta_ngp.py:165-190 (direction is normalised before the SH
    encoder, which normalises again)."""
    d = d / d.norm(dim=-1, keepdim=True).clamp_min(1e-9)
    h = torch.cat([geo, sh16(d)], dim=-1)
    n = _mlp_depth(p, "color_mlp")
    for i in range(n):
        h = torch.relu(_lin(h, p, f"color_mlp.{i}.linear"))
    return torch.sigmoid(_lin(h, p, f"color_mlp.{n}"))


def ngp_forward(x6: Tensor, p: Dict[str, Tensor], box_min: Tensor, hash_cfg: dict) -> Tensor:
    """(P,6) -> (P,4) [rgb, sigma]; models/inr/meta_ngp.py:226-241."""
    sigma, geo = ngp_density(x6[:, :3], p, box_min, hash_cfg)
    return torch.cat([ngp_color(x6[:, 3:6], geo, p), sigma], dim=-1)


def subdict(state: Dict[str, Tensor], prefix: str) -> Dict[str, Tensor]:
    """Keys under ``prefix.`` with the prefix removed (metamodule.py:37-69)."""
    pre = prefix + "."
    return {k[len(pre):]: v for k, v in state.items() if k.startswith(pre)}


# --------------------------------------------------------------------------
# a15  positional-encoding MLP expert (MetaNeRF, stand-alone)
# --------------------------------------------------------------------------
def pe_mlp_forward(x: Tensor, d: Tensor, p: Dict[str, Tensor], skips: Sequence[int] = (4,),
                   pe_xyz: int = 10, pe_dir: int = 4) -> Tuple[Tensor, Tensor]:
    """sigma (P,1), rgb (P,3); models/inr/meta_vanilla.py:109-154 with
    encoding_dir="frequency" (the only constructible variant, SURVEY.md note V)."""
    enc = freq_encode(x, pe_xyz)
    h = enc
    depth = 0
    while f"trunk.{depth}.linear.weight" in p:
        depth += 1
    for i in range(depth):
        if i in skips and i != 0:
            h = torch.cat([h, enc], dim=-1)
        h = torch.relu(_lin(h, p, f"trunk.{i}.linear"))
    sigma = trunc_exp(_lin(h, p, "sigma_head"))
    c = torch.cat([_lin(h, p, "geo_head"), freq_encode(d, pe_dir)], dim=-1)
    l = 0
    while f"color_mlp.layer{l}.linear.weight" in p:
        c = torch.relu(_lin(c, p, f"color_mlp.layer{l}.linear"))
        l += 1
    return sigma, torch.sigmoid(_lin(c, p, "color_mlp.color_out"))


# --------------------------------------------------------------------------
# a7, a16  container dispatch/combine and the shared background head
# --------------------------------------------------------------------------
def container_forward(x6: Tensor, state: Dict[str, Tensor], expert_mins: Sequence[Tensor], hash_cfg: dict,
                      margin: float, cluster_2d: bool, active: Optional[int] = None,
                      direct_routing: bool = False) -> Tensor:
    """(P,6) -> (P,4); models/inr/meta_container.py:275-343.  Soft routing blends
    all four channels linearly with the routing weight, experts in index order."""
    K = state["centroids"].shape[0]
    if active is not None:
        return ngp_forward(x6, subdict(state, f"submodules.{active}"), expert_mins[active], hash_cfg)
    with torch.no_grad():
        w, hard = routing(x6[:, :3], state["centroids"], margin, cluster_2d, direct=direct_routing)
    out = x6.new_zeros(x6.shape[0], 4)
    for k in range(K):
        sel = ((w[:, k] > 0) if w is not None else (hard == k)).nonzero().squeeze(1)
        if sel.numel() == 0:
            continue
        y = ngp_forward(x6[sel], subdict(state, f"submodules.{k}"), expert_mins[k], hash_cfg)
        if w is not None:
            out = out.index_add(0, sel, y * w[sel, k][:, None])
        else:
            out = out.index_copy(0, sel, y)
    return out


def background_color(d: Tensor, state: Dict[str, Tensor]) -> Tensor:
    """sigmoid(W1 relu(W0 SH16(normalize(d)) + b0) + b1); meta_container.py:347-382."""
    e = sh16(F.normalize(d, dim=-1))
    h = torch.relu(e.matmul(state["bg_mlp.0.weight"].t()) + state["bg_mlp.0.bias"])
    return torch.sigmoid(h.matmul(state["bg_mlp.2.weight"].t()) + state["bg_mlp.2.bias"])


# --------------------------------------------------------------------------
# a17  alpha compositing
# --------------------------------------------------------------------------
def composite(rgb_sigma: Tensor, t: Tensor, bg: Optional[Tensor], sigma_scale: float = 1.0):
    """(N,S,4),(N,S),(N,3) -> rgb (N,3), depth (N,), weights (N,S), acc (N,).
    No early termination.  nerfs/ray_rendering.py:137-165."""
    c = rgb_sigma[..., :3].clamp(0.0, 1.0)
    sig = rgb_sigma[..., 3].clamp_min(0.0)
    if sigma_scale != 1.0:
        sig = sig * float(sigma_scale)
    dt = (t[:, 1:] - t[:, :-1]).clamp_min(1e-4)
    dt = torch.cat([dt, dt[:, -1:]], dim=1)
    alpha = (1.0 - torch.exp(-sig * dt)).clamp(0.0, 1.0 - 1e-7)
    trans = torch.cumprod(torch.cat([torch.ones_like(alpha[:, :1]), 1.0 - alpha + 1e-10], dim=1), dim=1)[:, :-1]
    w = alpha * trans
    rgb = (w[..., None] * c).sum(dim=1)
    depth = (w * t).sum(dim=1)
    acc = w.sum(dim=1)
    if bg is not None:
        rgb = rgb + (1.0 - acc[:, None]) * bg
    return rgb, depth, w, acc


# --------------------------------------------------------------------------
# a18  photometric loss
# --------------------------------------------------------------------------
def srgb_to_linear(x: Tensor) -> Tensor:
    return torch.where(x <= 0.04045, x / 12.92, ((x + 0.055) / 1.055).pow(2.4))


def linear_to_srgb(x: Tensor) -> Tensor:
    x = x.clamp(0, 1)
    return torch.where(x <= 0.0031308, 12.92 * x, 1.055 * x.pow(1 / 2.4) - 0.055)


def mse_loss(pred_linear: Tensor, gt_srgb: Tensor, color_space: str = "linear") -> Tensor:
    """nerfs/losses.py:10-32 + nerfs/color_space.py:22-66."""
    gt = gt_srgb.float().clamp(0, 1)
    pred = pred_linear.float()
    if color_space == "linear":
        pred, gt = pred.clamp(0, 1), srgb_to_linear(gt).clamp(0, 1)
    elif color_space == "srgb":
        pred, gt = linear_to_srgb(pred).clamp(0, 1), gt.clamp(0, 1)
    elif color_space != "identity":
        raise ValueError(color_space)
    return F.mse_loss(pred, gt)


# --------------------------------------------------------------------------
# whole path: render_rays (stratified renderer)
# --------------------------------------------------------------------------
def render_rays(rays: Tensor, state: Dict[str, Tensor], expert_mins: Sequence[Tensor], hash_cfg: dict, *,
                S: int, margin: float, cluster_2d: bool = True, u: Optional[Tensor] = None,
                active: Optional[int] = None, use_bg: bool = True, bg_default: str = "white",
                sigma_scale: float = 1.0, direct_routing: bool = False):
    """nerfs/ray_rendering.py:290-345: sample -> points -> (routed) experts ->
    background -> composite.  Returns (rgb, depth, weights(N,S), acc)."""
    N = rays.shape[0]
    with torch.no_grad():
        t = stratified_t(rays[:, 6], rays[:, 7], S, u)
    x6 = sample_points(rays, t)
    y = container_forward(x6, state, expert_mins, hash_cfg, margin, cluster_2d, active,
                          direct_routing=direct_routing).view(N, S, 4)
    if use_bg:
        bg = background_color(rays[:, 3:6], state)
    elif bg_default == "white":
        bg = torch.ones(N, 3)
    elif bg_default == "black":
        bg = torch.zeros(N, 3)
    elif bg_default == "last_sample":
        bg = y[:, -1, :3]
    elif bg_default == "none":
        bg = None
    else:
        raise ValueError(bg_default)
    return composite(y, t, bg, sigma_scale)


def render_rays_pe(rays: Tensor, p: Dict[str, Tensor], *, S: int, u: Optional[Tensor] = None,
                   bg: Optional[Tensor] = None):
    """Config-1 path: stand-alone PE-MLP expert + the same sampler/compositor."""
    N = rays.shape[0]
    with torch.no_grad():
        t = stratified_t(rays[:, 6], rays[:, 7], S, u)
    x6 = sample_points(rays, t)
    sigma, rgb = pe_mlp_forward(x6[:, :3], x6[:, 3:6], p)
    return composite(torch.cat([rgb, sigma], dim=-1).view(N, S, 4), t, bg)


# --------------------------------------------------------------------------
# synthetic scenes: host-side scene construction lives in nerf_sys_b200/synthetic.py (bench.py
# uses it directly, so no timed region calls into this file); re-exported here for the tests.
# --------------------------------------------------------------------------
from nerf_sys_b200.synthetic import GLOBAL_AABB, HASH_CFG, aerial_rays, grid_scene, init_ngp_state, ring_pose  # noqa: E402,F401


# --------------------------------------------------------------------------
# proposal (hierarchical) resampling - ADDITIVE: the reference has none (SURVEY headline 4); the canonical NeRF sample_pdf
# --------------------------------------------------------------------------
def sample_pdf(t: Tensor, weights: Tensor, u: Tensor) -> Tuple[Tensor, Tensor]:
    """(N,S) coarse distances + weights, (N,S2) ascending u in [0,1) -> t_new (N,S2), merged ascending (N,S+S2)."""
    bins = 0.5 * (t[:, 1:] + t[:, :-1])
    w = weights[:, 1:-1] + 1e-5
    pdf = w / torch.cumsum(w, -1)[:, -1:]  # a sequential fp32 sum, as the kernel's (a pairwise sum moves the cdf by an ulp, and a
    # sample in a near-empty bin by (1 ulp / bin mass) of the bin width)
    cdf = torch.cat([torch.zeros_like(pdf[:, :1]), torch.cumsum(pdf, -1)], -1)
    inds = torch.searchsorted(cdf.contiguous(), u.contiguous(), right=True)
    below, above = (inds - 1).clamp_min(0), inds.clamp_max(cdf.shape[-1] - 1)
    c0, c1 = torch.gather(cdf, 1, below), torch.gather(cdf, 1, above)
    b0, b1 = torch.gather(bins, 1, below), torch.gather(bins, 1, above)
    denom = c1 - c0
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t_new = b0 + (u - c0) / denom * (b1 - b0)
    return t_new, torch.sort(torch.cat([t, t_new], -1), -1).values
