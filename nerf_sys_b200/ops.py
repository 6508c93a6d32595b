"""Tensor-level wrappers and autograd Functions over the C ABI (include/acnerf.h).

PyTorch is plumbing here: it owns device memory (caching allocator), the current
stream and the autograd graph.  Every computation is a call into
libacnerf_sm100a.so; inputs must be CUDA fp32 tensors - CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
import functools
import os
from typing import List, Optional, Sequence, Tuple

import torch


def first_order(backward):
    """Decorator for the `backward` of every op in this package.  The engine evaluates a backward pass with grad mode ON
    only when the caller asked for a differentiable gradient (`create_graph=True`: algo='maml', meta_core.py:54-59); the
    sm_100a kernels are first-order, so that request fails here instead of silently returning gradients without a graph.
    (`torch.autograd.function.once_differentiable` does not do this: it only raises when such a gradient is differentiated
    AND an incoming gradient already carried a graph.)"""

    @functools.wraps(backward)
    def wrapper(ctx, *grads):
        if torch.is_grad_enabled():
            raise RuntimeError("the sm_100a ops are first-order: autograd.grad(..., create_graph=True) (algo='maml') is not supported - "
                               "use algo='fomaml' or 'reptile'")
        return backward(ctx, *grads)

    return wrapper


from . import _lib
from ._lib import (ACN_MAX_EXPERTS, ACN_NGP_LAYERS, ACN_TILE, AcnA2aRoute, AcnGridDesc, AcnNgpExpert, AcnNgpExpertGrads, AcnPairs,
                   AcnPoints, check, host_floats, lib)

Tensor = torch.Tensor

# per-expert MLP tensors in ABI layer order (include/acnerf.h: AcnNgpExpert)
NGP_LAYER_KEYS = ("sigma_trunk.0.linear", "sigma_trunk.1.linear", "sigma_head", "geo_head", "color_mlp.0.linear",
                  "color_mlp.1.linear", "color_mlp.2")
NGP_LAYER_SHAPES = ((64, 32), (64, 64), (1, 64), (15, 64), (64, 31), (64, 64), (3, 64))
INTERP = {"linear": 0, "smoothstep": 1, "nearest": 2}

# Arithmetic of the fused expert MLPs: "f16tc" = tcgen05 tensor cores, fp16 operands / fp32 accumulate (the
# precision class of the reference's own GPU path); "fp32" = exact CUDA-core mode used by the tight parity tests.
_mlp_math = "f16tc"


def set_mlp_math(mode: str) -> str:
    """Select the expert-MLP arithmetic for subsequent forward calls; returns the previous mode."""
    global _mlp_math
    if mode not in ("f16tc", "fp32"):
        raise ValueError(f"mlp math must be 'f16tc' or 'fp32', got {mode!r}")
    prev, _mlp_math = _mlp_math, mode
    return prev


def get_mlp_math() -> str:
    return _mlp_math


class mlp_math:
    """Context manager: `with ops.mlp_math("f16tc"): ...`"""

    def __init__(self, mode: str):
        self.mode = mode

    def __enter__(self):
        self.prev = set_mlp_math(self.mode)
        return self

    def __exit__(self, *exc):
        set_mlp_math(self.prev)
        return False


_launches = 0  # kernels launched through this module (bench.py's gpu_launches claim)


def launches() -> int:
    return _launches


def _count(n: int):
    global _launches
    _launches += n


# ---- optional per-call CUDA-event timing (bench.py's live roofline) ----------------------
_timing = None  # None = off; else {name: [(start_event, end_event), ...]}


def kernel_timing(enable: bool):
    """Turn per-call event recording on (returns the dict that fills up) or off."""
    global _timing
    _timing = {} if enable else None
    return _timing


def kernel_timing_enabled() -> bool:
    return _timing is not None


class _timed:
    """Records a CUDA event pair on the current stream around one C-ABI call."""

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if _timing is not None:
            self.a = torch.cuda.Event(enable_timing=True)
            self.b = torch.cuda.Event(enable_timing=True)
            self.a.record()
        return self

    def __exit__(self, *exc):
        if _timing is not None:
            self.b.record()
            _timing.setdefault(self.name, []).append((self.a, self.b))
        return False


def _stream() -> C.c_void_p:
    """The current stream of the current device (raw handle; the Python Stream object costs ~10 us per call)."""
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))


class _NoGuard:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


_no_guard = _NoGuard()


def _on(dev):
    """Device guard for a launch: nothing when `dev` is already current (the one-process-per-GPU case)."""
    idx = dev.index if isinstance(dev, torch.device) else int(dev)
    if idx is None or idx == torch._C._cuda_getDevice():
        return _no_guard
    return torch.cuda.device(idx)


def _ptr(t: Optional[Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _need(t: Tensor, name: str, dtype=torch.float32) -> Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name} is on {t.device}: nerf_sys_b200 runs on CUDA (sm_100a) only, there is no CPU path")
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def _f32(t: Tensor, name: str) -> Tensor:
    """CUDA check + cast to contiguous fp32 (callers may hand fp16 under autocast)."""
    if not t.is_cuda:
        raise RuntimeError(f"{name} is on {t.device}: nerf_sys_b200 runs on CUDA (sm_100a) only, there is no CPU path")
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def device_ok(index: Optional[int] = None):
    idx = torch.cuda.current_device() if index is None else index
    check(lib().acn_device_ok(idx), "acn_device_ok")


# --------------------------------------------------------------------------
# descriptors
# --------------------------------------------------------------------------
def grid_desc(levels: int, log2_T: int, interp: str, res: Sequence[int], features: int = 2) -> AcnGridDesc:
    g = AcnGridDesc()
    g.levels, g.features, g.log2_T = int(levels), int(features), int(log2_T)
    g.interp = INTERP[str(interp).lower()]
    for i in range(_lib.ACN_MAX_LEVELS):
        g.res[i] = int(res[i]) if i < len(res) else 0
    return g


class PointSet:
    """Sample points: explicit (P,6) rows, or implicit (rays (N,8), t (N,S))."""

    def __init__(self, x6: Optional[Tensor] = None, rays: Optional[Tensor] = None, t: Optional[Tensor] = None, coherent: bool = False):
        # coherent: consecutive rays are neighbouring pixels of an image block (a renderer's hint, see sample_major())
        self.coherent = bool(coherent) and x6 is None
        self.tT = None
        # a renderer's list of (seg, capacity) to verify later: lets the container route without its host read (no_grad only)
        self.route_checks: Optional[list] = None
        if x6 is not None:
            self.x6 = _f32(x6, "x6")
            self.rays = self.t = None
            self.S, self.P = 1, self.x6.shape[0]
            self.device = self.x6.device
        else:
            self.x6 = None
            self.rays, self.t = _f32(rays, "rays"), _f32(t, "t")
            self.S, self.P = int(self.t.shape[1]), int(self.t.numel())
            self.device = self.rays.device

    def sample_major(self) -> "PointSet":
        """The same points with SAMPLE-MAJOR ids (AcnPoints.t_T): point p = sample p // N of ray p % N.  Consecutive points -
        the lanes of every kernel - are then the same sample of adjacent rays; for a block of image rows those are
        neighbouring pixels that share grid cells on most levels.  Outputs indexed by point id come back in (S,N) order
        (`to_ray_major`)."""
        if self.x6 is not None:
            raise ValueError("sample_major() needs implicit points (rays, t)")
        q = PointSet(rays=self.rays, t=self.t)
        q.tT = self.t.t().contiguous()
        q.route_checks = self.route_checks
        return q

    def to_ray_major(self, out: Tensor) -> Tensor:
        """(P, C) in this set's point order -> (P, C) in ray-major order (p = ray * S + sample)."""
        if self.tT is None:
            return out
        N = self.P // self.S
        return out.view(self.S, N, -1).transpose(0, 1).reshape(self.P, -1)

    def c(self) -> AcnPoints:
        p = AcnPoints()
        p.x6, p.rays, p.t = _ptr(self.x6), _ptr(self.rays), _ptr(self.t)
        p.S, p.P = self.S, self.P
        p.t_T = _ptr(self.tT)
        return p


class PairList:
    """Routed (point, expert) pairs grouped by expert; identity when point is None."""

    def __init__(self, M: int, point: Optional[Tensor] = None, weight: Optional[Tensor] = None, seg: Optional[Tensor] = None,
                 seg_host: Optional[List[int]] = None, store_out: Optional[Tensor] = None):
        """``seg_host`` None with a routed list = the host never read the segment table (speculative capacity; every expert
        is treated as live and the device-side flags of `route_status` / `acn_p2p_plan` tell the optimiser otherwise).
        ``weight`` None with ``point`` given = unit weights, unique point ids: the expert kernels STORE their rows, into
        ``store_out`` (P,4) when given (an owner's result rows inside its exchange slab)."""
        self.M, self.point, self.weight, self.seg, self.seg_host = int(M), point, weight, seg, seg_host
        self.store_out = store_out

    @property
    def identity(self) -> bool:
        return self.point is None

    def c(self) -> AcnPairs:
        p = AcnPairs()
        p.point, p.weight, p.seg, p.Mcap = _ptr(self.point), _ptr(self.weight), _ptr(self.seg), self.M
        return p


# --------------------------------------------------------------------------
# a1-a5 rays and sampling
# --------------------------------------------------------------------------
def ray_directions(H: int, W: int, fx: float, fy: float, cx: float, cy: float, center_pixels: bool, device) -> Tensor:
    out = torch.empty(H, W, 3, device=device, dtype=torch.float32)
    with _on(out.device):
        check(lib().acn_ray_directions(H, W, fx, fy, cx, cy, int(bool(center_pixels)), _ptr(out), _stream()), "acn_ray_directions")
    _count(1)
    return out


def _c2w_host(c2w: Tensor):
    m = c2w.detach().to("cpu", torch.float32)
    return host_floats(m[:3, :4].reshape(-1).tolist())


def _aabb_host(aabb: Optional[Tensor]):
    if aabb is None:
        return None
    return host_floats(aabb.detach().to("cpu", torch.float32).reshape(-1).tolist())


def get_rays(dirs: Tensor, c2w: Tensor, aabb: Optional[Tensor], near: Optional[float], far: Optional[float],
             max_bound: float = 1e10, invalid_value: float = 1e10) -> Tensor:
    d = _f32(dirs, "directions").reshape(-1, 3)
    out = torch.empty(d.shape[0], 8, device=d.device, dtype=torch.float32)
    with _on(d.device):
        check(lib().acn_get_rays(_ptr(d), d.shape[0], _c2w_host(c2w), _aabb_host(aabb), float(near or 0.0), float(far or 0.0),
                                 float(max_bound), float(invalid_value), _ptr(out), _stream()), "acn_get_rays")
    _count(1)
    return out.reshape(*dirs.shape[:-1], 8)


def raygen(H: int, W: int, fx, fy, cx, cy, center_pixels: bool, c2w: Tensor, aabb: Optional[Tensor], device, near=0.0, far=0.0,
           row0: int = 0, rows: Optional[int] = None) -> Tensor:
    rows = H - row0 if rows is None else rows
    out = torch.empty(rows * W, 8, device=device, dtype=torch.float32)
    with _on(out.device), _timed("raygen"):
        check(lib().acn_raygen(H, W, row0, rows, fx, fy, cx, cy, int(bool(center_pixels)), _c2w_host(c2w), _aabb_host(aabb),
                               float(near), float(far), _ptr(out), _stream()), "acn_raygen")
    _count(1)
    return out


def ray_aabb(o: Tensor, d: Tensor, aabb: Tensor, max_bound: float = 1e10, invalid_value: float = 1e10) -> Tuple[Tensor, Tensor]:
    o, d = _f32(o, "origins"), _f32(d, "directions")
    n = o.shape[0]
    tn, tf = torch.empty(n, device=o.device), torch.empty(n, device=o.device)
    with _on(o.device):
        check(lib().acn_ray_aabb(_ptr(o), _ptr(d), n, _aabb_host(aabb), float(max_bound), float(invalid_value), _ptr(tn), _ptr(tf),
                                 _stream()), "acn_ray_aabb")
    _count(1)
    return tn, tf


def clamp_near_far(rays: Tensor, override, eps: float = 1e-6, invalid_value: float = float("inf")) -> Tuple[Tensor, Tensor]:
    r = _f32(rays, "rays")
    n = r.shape[0]
    valid = torch.empty(n, device=r.device, dtype=torch.uint8)
    with _on(r.device):
        if override is None:
            check(lib().acn_clamp_near_far(_ptr(r), n, 1, 0, 0.0, 0, 0.0, float(eps), float(invalid_value), None, _ptr(valid),
                                           _stream()), "acn_clamp_near_far")
            out = rays
        else:
            lo, hi = override
            out = torch.empty_like(r)
            check(lib().acn_clamp_near_far(_ptr(r), n, 0, int(lo is not None), float(lo or 0.0), int(hi is not None),
                                           float(hi or 0.0), float(eps), float(invalid_value), _ptr(out), _ptr(valid), _stream()),
                  "acn_clamp_near_far")
    _count(1)
    return out, valid.bool()


_tau_cache = {}


def _tau(S: int, device) -> Tensor:
    """torch.linspace(0,1,S) evaluated by the CPU kernel (the reference's values), cached on the device."""
    key = (S, str(device))
    if key not in _tau_cache:
        _tau_cache[key] = torch.linspace(0.0, 1.0, S).to(device)
    return _tau_cache[key]


def stratified_t(rays: Tensor, S: int, u: Optional[Tensor]) -> Tensor:
    r = _f32(rays, "rays")
    N = r.shape[0]
    t = torch.empty(N, S, device=r.device, dtype=torch.float32)
    if u is not None:
        u = _f32(u, "u")
        if tuple(u.shape) != (N, S):
            raise ValueError(f"u must be ({N},{S}), got {tuple(u.shape)}")
    with _on(r.device), _timed("stratified_t"):
        check(lib().acn_stratified_t(_ptr(r), N, S, _ptr(_tau(S, r.device)), _ptr(u), _ptr(t), _stream()), "acn_stratified_t")
    _count(1)
    return t


def sample_points(rays: Tensor, t: Tensor) -> Tensor:
    r, tt = _f32(rays, "rays"), _f32(t, "t")
    N, S = tt.shape
    x6 = torch.empty(N * S, 6, device=r.device, dtype=torch.float32)
    with _on(r.device):
        check(lib().acn_sample_points(_ptr(r), _ptr(tt), N, S, _ptr(x6), _stream()), "acn_sample_points")
    _count(1)
    return x6


# --------------------------------------------------------------------------
# a6-a7 routing
# --------------------------------------------------------------------------
def routing(pts: PointSet, centroids_host: Sequence[float], K: int, margin: float, cluster_2d: bool):
    if K > ACN_MAX_EXPERTS:
        raise ValueError(f"at most {ACN_MAX_EXPERTS} experts per routing call (got {K})")
    ch = host_floats(centroids_host)
    cp = pts.c()
    with _on(pts.device):
        if margin > 1.0:
            w = torch.empty(pts.P, K, device=pts.device, dtype=torch.float32)
            check(lib().acn_routing(C.byref(cp), ch, K, float(margin), int(cluster_2d), _ptr(w), None, _stream()), "acn_routing")
            _count(1)
            return w, None
        hard = torch.empty(pts.P, device=pts.device, dtype=torch.int64)
        check(lib().acn_routing(C.byref(cp), ch, K, float(margin), int(cluster_2d), None, _ptr(hard), _stream()), "acn_routing")
    _count(1)
    return None, hard


def route_pack(pts: PointSet, centroids_host: Sequence[float], K: int, margin: float, cluster_2d: bool, sync: bool = True) -> PairList:
    """Routing + stable pack by expert.  One 4*(K+1)-byte device->host read fixes the exact
    pair count (the reference syncs K times through nonzero(), meta_container.py:309).
    ``sync=False`` skips that read: the list keeps its speculative capacity (``seg_host`` is None, the kernels take the
    segment table from the device) and the CALLER must check ``seg[K] <= pairs.M`` once it reads the counts."""
    if K > ACN_MAX_EXPERTS:
        raise ValueError(f"at most {ACN_MAX_EXPERTS} experts per routing call (got {K})")
    L = lib()
    ch = host_floats(centroids_host)
    cp = pts.c()
    dev = pts.device
    with _on(dev):
        cap = int(L.acn_route_pack_capacity(pts.P, K))
        ws = torch.empty(int(L.acn_route_pack_ws_bytes(pts.P, K)) // 4 + 1, device=dev, dtype=torch.int32)
        seg = torch.empty(K + 1, device=dev, dtype=torch.int32)
        while True:
            point = torch.empty(cap, device=dev, dtype=torch.int32)
            weight = torch.empty(cap, device=dev, dtype=torch.float32)
            with _timed("route_pack"):
                check(L.acn_route_pack(C.byref(cp), ch, K, float(margin), int(cluster_2d), _ptr(point), _ptr(weight), _ptr(seg), cap,
                                       _ptr(ws), _stream()), "acn_route_pack")
            _count(6)
            if not sync:
                return PairList(cap, point, weight, seg, None)
            seg_host = seg.tolist()
            if seg_host[K] <= cap:
                break
            cap = seg_host[K]
    M = seg_host[K]
    return PairList(M, point[:M], weight[:M], seg, seg_host)


@torch.no_grad()
def sample_pdf(t: Tensor, weights: Tensor, u: Tensor, merge: bool = True):
    """Proposal resampling (additive, see acn_sample_pdf): (N,S) coarse distances + weights, (N,S2) ascending u in [0,1) ->
    t_new (N,S2) and, with ``merge``, the ascending union (N,S+S2)."""
    tt, ww, uu = _need(t, "t"), _need(weights.detach(), "weights"), _need(u, "u")
    N, S = tt.shape
    S2 = uu.shape[1]
    t_new = torch.empty(N, S2, device=tt.device)
    merged = torch.empty(N, S + S2, device=tt.device) if merge else None
    with _on(tt.device), _timed("sample_pdf"):
        check(lib().acn_sample_pdf(_ptr(tt), _ptr(ww), _ptr(uu), N, S, S2, _ptr(t_new), _ptr(merged), _stream()), "acn_sample_pdf")
    _count(1)
    return t_new, merged


def route_capacity_ok(checks: Sequence[Tuple[Tensor, int]]) -> bool:
    """One device->host read for a whole frame: did every deferred route_pack(sync=False) fit its speculative capacity?"""
    if not checks:
        return True
    used = torch.stack([seg[-1] for seg, _ in checks]).tolist()
    return all(u <= cap for u, (_, cap) in zip(used, checks))


# --------------------------------------------------------------------------
# (e) expert-parallel dispatch / combine around the sample all-to-all
# --------------------------------------------------------------------------
def a2a_pack(pts: PointSet, pairs: PairList, K: int):
    """Compacts a routed pair list into all-to-all send rows (meta_container.py:305-313 across GPUs): returns
    rows6 (M,6) [xyz,dir] grouped by expert, point (M,) int32, weight (M,), counts (K host ints).  One
    4*(K+1)-byte device->host read (the reference reads K nonzero() sizes)."""
    if pairs.identity:
        raise ValueError("a2a_pack needs a routed pair list")
    dev = pts.device
    cap = pairs.M
    rows = torch.empty(cap, 6, device=dev, dtype=torch.float32)
    point = torch.empty(cap, device=dev, dtype=torch.int32)
    weight = torch.empty(cap, device=dev, dtype=torch.float32)
    counts = torch.empty(K + 1, device=dev, dtype=torch.int32)
    offs = torch.empty(K + 1, device=dev, dtype=torch.int32)
    cp, cq = pts.c(), pairs.c()
    with _on(dev), _timed("a2a_pack"):
        check(lib().acn_a2a_pack(C.byref(cp), C.byref(cq), K, _ptr(rows), _ptr(point), _ptr(weight), _ptr(counts), _ptr(offs),
                                 _stream()), "acn_a2a_pack")
    _count(2)
    ch = counts.tolist()
    M = ch[K]
    return rows[:M], point[:M], weight[:M], ch[:K]


def a2a_count(pairs: PairList, K: int) -> Tuple[Tensor, Tensor]:
    """Device-side (counts (K+1,), offs (K+1,)) of a routed pair list; counts[K] = offs[K] = live pairs."""
    dev = pairs.point.device
    counts = torch.empty(K + 1, device=dev, dtype=torch.int32)
    offs = torch.empty(K + 1, device=dev, dtype=torch.int32)
    cq = pairs.c()
    with _on(dev):
        check(lib().acn_a2a_count(C.byref(cq), K, _ptr(counts), _ptr(offs), _stream()), "acn_a2a_count")
    _count(1)
    return counts, offs


def make_route(bases: Sequence[int], row0: Sequence[int]) -> AcnA2aRoute:
    r = AcnA2aRoute()
    for e, (b, o) in enumerate(zip(bases, row0)):
        r.base[e], r.row0[e] = int(b), int(o)
    return r


def a2a_push_rows(pts: PointSet, pairs: PairList, K: int, offs: Tensor, route: AcnA2aRoute, M: int, row0_dev: Optional[Tensor] = None,
                  cap_rows: int = 0) -> Tuple[Tensor, Tensor]:
    """Writes the [xyz,dir] rows of the live pairs into their experts' owners (route); returns the compact
    (point (M,) int32, weight (M,)) of the pairs for the combine.  ``row0_dev`` (K int64 on the device, from
    `p2p_plan`) replaces the host's route rows; ``M`` is then an upper bound (the pair capacity)."""
    dev = pts.device
    point = torch.empty(M, device=dev, dtype=torch.int32)
    weight = torch.empty(M, device=dev, dtype=torch.float32)
    cp, cq = pts.c(), pairs.c()
    with _on(dev), _timed("a2a_push_rows"):
        check(lib().acn_a2a_push_rows(C.byref(cp), C.byref(cq), K, _ptr(offs), C.byref(route), _ptr(row0_dev), int(cap_rows), _ptr(point),
                                      _ptr(weight), _stream()), "acn_a2a_push_rows")
    _count(1)
    return point, weight


def a2a_pull_combine(route: AcnA2aRoute, point: Tensor, weight: Tensor, offs: Tensor, K: int, P: int, row0_dev: Optional[Tensor] = None) -> Tensor:
    dev = point.device
    out = torch.zeros(P, 4, device=dev, dtype=torch.float32)
    with _on(dev), _timed("a2a_pull_combine"):
        check(lib().acn_a2a_pull_combine(C.byref(route), _ptr(row0_dev), _ptr(point), _ptr(weight), _ptr(offs), K, point.numel(), _ptr(out),
                                         _stream()), "acn_a2a_pull_combine")
    _count(1)
    return out


def a2a_push_grad(d_out: Tensor, point: Tensor, weight: Tensor, offs: Tensor, K: int, route: AcnA2aRoute, row0_dev: Optional[Tensor] = None,
                  cap_rows: int = 0) -> None:
    g = _f32(d_out, "d_out")
    with _on(g.device), _timed("a2a_push_grad"):
        check(lib().acn_a2a_push_grad(_ptr(g), _ptr(point), _ptr(weight), _ptr(offs), K, point.numel(), C.byref(route), _ptr(row0_dev),
                                      int(cap_rows), _stream()), "acn_a2a_push_grad")
    _count(1)


def route_status(seg: Tensor, K: int, Mcap: int, live: Optional[Tensor], status: Optional[Tensor]) -> None:
    """Device-side reading of a segment table nobody reads back: live[k] = expert k received a sample; status |= 1 when the
    pair list overflowed its speculative capacity (sticky)."""
    with _on(seg.device):
        check(lib().acn_route_status(_ptr(seg), int(K), int(Mcap), _ptr(live), _ptr(status), _stream()), "acn_route_status")
    _count(1)


class A2aCombineFn(torch.autograd.Function):
    """out[point[j]] += weight[j] * y[j] (meta_container.py:322-337); backward d_y[j] = weight[j] * d_out[point[j]]."""

    @staticmethod
    def forward(ctx, y: Tensor, point: Tensor, weight: Tensor, P: int):
        yy = _f32(y.detach(), "y")
        M = yy.shape[0]
        if yy.dim() != 2 or yy.shape[1] != 4 or point.numel() != M or weight.numel() != M:
            raise ValueError("a2a_combine: y must be (M,4) with M point ids and weights")
        pt, wt = _need(point, "point", torch.int32), _f32(weight, "weight")
        out = torch.zeros(P, 4, device=yy.device, dtype=torch.float32)
        with _on(yy.device), _timed("a2a_combine"):
            check(lib().acn_a2a_combine(_ptr(yy), _ptr(pt), _ptr(wt), M, _ptr(out), _stream()), "acn_a2a_combine")
        _count(1)
        ctx.save_for_backward(pt, wt)
        return out

    @staticmethod
    @first_order
    def backward(ctx, d_out):
        pt, wt = ctx.saved_tensors
        g = _f32(d_out, "d_out")
        d_y = torch.empty(pt.numel(), 4, device=g.device, dtype=torch.float32)
        with _on(g.device), _timed("a2a_combine_bwd"):
            check(lib().acn_a2a_combine_bwd(_ptr(g), _ptr(pt), _ptr(wt), pt.numel(), _ptr(d_y), _stream()), "acn_a2a_combine_bwd")
        _count(1)
        return d_y, None, None, None


def a2a_combine(y: Tensor, point: Tensor, weight: Tensor, P: int) -> Tensor:
    return A2aCombineFn.apply(y, point, weight, int(P))


# --------------------------------------------------------------------------
# a9-a11 stand-alone encoders
# --------------------------------------------------------------------------
def hashgrid_indices(x01: Tensor, g: AcnGridDesc) -> Tensor:
    x = _need(x01, "x01")
    idx = torch.empty(x.shape[0], g.levels, 8, device=x.device, dtype=torch.int32)
    with _on(x.device):
        check(lib().acn_hashgrid_indices(_ptr(x), x.shape[0], C.byref(g), _ptr(idx), _stream()), "acn_hashgrid_indices")
    _count(1)
    return idx


class HashGridFn(torch.autograd.Function):
    """HashGridEncoder.forward and its autograd backward (models/encodings.py:293-381)."""

    @staticmethod
    def forward(ctx, x01: Tensor, table: Tensor, g: AcnGridDesc):
        x = _need(x01.detach().reshape(-1, 3), "x01")
        tb = _need(table.detach(), "hash_table")
        y = torch.empty(x.shape[0], g.levels * g.features, device=x.device, dtype=torch.float32)
        with _on(x.device):
            check(lib().acn_hashgrid_fwd(_ptr(x), x.shape[0], _ptr(tb), C.byref(g), _ptr(y), _stream()), "acn_hashgrid_fwd")
        _count(1)
        ctx.g, ctx.shape = g, table.shape
        ctx.save_for_backward(x)
        return y.reshape(*x01.shape[:-1], g.levels * g.features)

    @staticmethod
    @first_order
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        g = ctx.g
        dy = _f32(dy, "dy").reshape(x.shape[0], -1)
        grad = torch.zeros(ctx.shape, device=x.device, dtype=torch.float32)
        with _on(x.device):
            check(lib().acn_hashgrid_bwd(_ptr(x), x.shape[0], _ptr(dy), C.byref(g), _ptr(grad), _stream()), "acn_hashgrid_bwd")
        _count(1)
        return None, grad, None


def sh16(d: Tensor) -> Tensor:
    x = _f32(d, "d").reshape(-1, 3)
    out = torch.empty(x.shape[0], 16, device=x.device, dtype=torch.float32)
    with _on(x.device):
        check(lib().acn_sh16(_ptr(x), x.shape[0], _ptr(out), _stream()), "acn_sh16")
    _count(1)
    return out.reshape(*d.shape[:-1], 16)


def freq_encode(x: Tensor, n_freq: int, include_input: bool = True) -> Tensor:
    xx = _f32(x, "x").reshape(-1, 3)
    width = (3 if include_input else 0) + 6 * n_freq
    out = torch.empty(xx.shape[0], width, device=xx.device, dtype=torch.float32)
    with _on(xx.device):
        check(lib().acn_freq_encode(_ptr(xx), xx.shape[0], int(n_freq), int(include_input), _ptr(out), _stream()), "acn_freq_encode")
    _count(1)
    return out.reshape(*x.shape[:-1], width)


# --------------------------------------------------------------------------
# a8-a14 hash-grid experts over routed pairs
# --------------------------------------------------------------------------
class ExpertSpec:
    """Static (non-tensor) description of E experts for one fused call."""

    def __init__(self, grid: AcnGridDesc, box_min: Sequence[Sequence[float]], extent: Sequence[Sequence[float]]):
        self.grid, self.box_min, self.extent = grid, [list(map(float, b)) for b in box_min], [list(map(float, e)) for e in extent]
        self.E = len(self.box_min)
        if self.E > ACN_MAX_EXPERTS:
            raise ValueError(f"at most {ACN_MAX_EXPERTS} experts per fused call (got {self.E})")


def _expert_array(spec: ExpertSpec, tables: Optional[Sequence[Tensor]], mlp: Optional[Sequence[Tensor]]):
    arr = (AcnNgpExpert * spec.E)()
    for e in range(spec.E):
        if tables is not None:
            arr[e].table = _ptr(tables[e])
        if mlp is not None:
            for l in range(ACN_NGP_LAYERS):
                arr[e].w[l] = _ptr(mlp[e * 14 + 2 * l])
                arr[e].b[l] = _ptr(mlp[e * 14 + 2 * l + 1])
        for a in range(3):
            arr[e].box_min[a] = spec.box_min[e][a]
            arr[e].extent[a] = spec.extent[e][a]
    return arr


def _check_mlp(mlp: Sequence[Tensor], E: int) -> List[Tensor]:
    if len(mlp) != 14 * E:
        raise ValueError(f"expected {14 * E} MLP tensors, got {len(mlp)}")
    out = []
    for i, t in enumerate(mlp):
        l = (i % 14) // 2
        shape = NGP_LAYER_SHAPES[l] if i % 2 == 0 else (NGP_LAYER_SHAPES[l][0],)
        if tuple(t.shape) != shape:
            raise ValueError(f"expert {i // 14} {NGP_LAYER_KEYS[l]}.{'weight' if i % 2 == 0 else 'bias'} must be {shape}, got "
                             f"{tuple(t.shape)}: the fused sm_100a expert is specialised on the production shape "
                             "hidden=64, sigma_depth=2, color_depth=2, geo_feat_dim=15, SH16")
        out.append(_need(t.detach(), "mlp tensor"))
    return out


class NgpEncodeFn(torch.autograd.Function):
    """pairs -> enc (32, ld) feature-major; backward scatters into each expert's table grad."""

    @staticmethod
    def forward(ctx, pts: PointSet, pairs: PairList, spec: ExpertSpec, *tables: Tensor):
        L = lib()
        tb = [_need(t.detach(), "hash_table") for t in tables]
        rows = spec.grid.levels << spec.grid.log2_T
        for t in tb:
            if tuple(t.shape) != (rows, 2):
                raise ValueError(f"hash_table must be ({rows},2), got {tuple(t.shape)}")
        ld = int(L.acn_ngp_ld(pairs.M))
        enc = torch.empty(32, ld, device=pts.device, dtype=torch.float32)
        ex = _expert_array(spec, tb, None)
        cp, cq = pts.c(), pairs.c()
        with _on(pts.device), _timed("hash_fwd"):
            check(L.acn_ngp_encode_fwd(C.byref(cp), C.byref(cq), ex, spec.E, C.byref(spec.grid), _ptr(enc), _stream()),
                  "acn_ngp_encode_fwd")
        _count(1)
        ctx.pts, ctx.pairs, ctx.spec = pts, pairs, spec
        ctx.save_for_backward(*tb)
        return enc

    @staticmethod
    @first_order
    def backward(ctx, d_enc):
        L = lib()
        tb = ctx.saved_tensors
        pts, pairs, spec = ctx.pts, ctx.pairs, ctx.spec
        d_enc = _f32(d_enc, "d_enc")
        grads = (AcnNgpExpertGrads * spec.E)()
        outs = []
        for e in range(spec.E):
            if ctx.needs_input_grad[3 + e] and _segment_len(pairs, e) > 0:
                g = _table_grad(tb[e])
                grads[e].table = _ptr(g)
                outs.append(g)
            else:
                outs.append(None)
        ex = _expert_array(spec, tb, None)
        cp, cq = pts.c(), pairs.c()
        with _on(pts.device), _timed("hash_bwd"):
            check(L.acn_ngp_encode_bwd(C.byref(cp), C.byref(cq), ex, spec.E, C.byref(spec.grid), _ptr(d_enc), grads, _stream()),
                  "acn_ngp_encode_bwd")
        _count(1)
        return (None, None, None, *outs)


def _segment_len(pairs: PairList, e: int) -> int:
    if pairs.identity:
        return pairs.M
    if pairs.seg_host is None:  # never read back: live as far as the host knows
        return 1
    return pairs.seg_host[e + 1] - pairs.seg_host[e]


def _out_buffer(pts: "PointSet", pairs: PairList) -> Tensor:
    """(P,4) result buffer: plain stores for identity / unique-point lists, zero-filled for the weighted blend."""
    if pairs.store_out is not None:
        if not (pairs.store_out.is_cuda and pairs.store_out.dtype == torch.float32 and pairs.store_out.is_contiguous()
                and tuple(pairs.store_out.shape) == (pts.P, 4)):
            raise ValueError(f"store_out must be a contiguous CUDA fp32 ({pts.P},4) tensor")
        return pairs.store_out.detach()  # a fresh alias per call: the caller's tensor object never acquires a grad_fn
    if pairs.identity or pairs.weight is None:
        return torch.empty(pts.P, 4, device=pts.device, dtype=torch.float32)
    return torch.zeros(pts.P, 4, device=pts.device, dtype=torch.float32)


# ---- persistent gradient buffers for the hash tables ---------------------------------------------------------------
# A 128 MiB table gradient costs a 128 MiB zero-fill per step when the backward allocates it.  An optimiser that zeroes
# the gradient while it reads it (FusedAdam(persistent_grads=True): ACN_ADAM_ZERO_GRAD) registers one buffer per table here;
# the backward then scatters into that buffer, which is all zeros between an optimiser step and the next backward.
_grad_arena = {}  # table data_ptr -> [buffer, in_use, weakref to the parameter]


def register_grad_buffer(table: Tensor, buf: Tensor) -> None:
    import weakref
    _grad_arena[table.data_ptr()] = [buf, False, weakref.ref(table)]


def unregister_grad_buffer(table: Tensor) -> None:
    _grad_arena.pop(table.data_ptr(), None)


def release_grad_buffer(table: Tensor, clean: bool) -> None:
    """The optimiser is done with the buffer (``clean``: it zeroed it on the device; otherwise it is zero-filled here)."""
    ent = _grad_arena.get(table.data_ptr())
    if ent is not None and ent[1]:
        if not clean:
            ent[0].zero_()
        ent[1] = False


def _table_grad(table: Tensor) -> Tensor:
    ent = _grad_arena.get(table.data_ptr())
    if ent is not None and not ent[1] and ent[0].shape == table.shape and ent[0].device == table.device:
        owner = ent[2]()
        # a parameter that still holds a .grad (zero_grad(set_to_none=False), gradient accumulation) gets a fresh tensor:
        # autograd would otherwise add the buffer to itself
        if owner is not None and owner.grad is None:
            ent[1] = True
            return ent[0].detach()  # a fresh alias: autograd adopts it as .grad without a copy
    return torch.zeros_like(table)


class NgpMlpFn(torch.autograd.Function):
    """enc (32, ld) + directions -> out (P,4) = combine_k w_k * [rgb, sigma]_k."""

    @staticmethod
    def forward(ctx, enc: Tensor, pts: PointSet, pairs: PairList, spec: ExpertSpec, *mlp: Tensor):
        L = lib()
        ws = _check_mlp(mlp, spec.E)
        enc_c = _need(enc.detach(), "enc")
        need_grad = any(ctx.needs_input_grad)  # False under no_grad: inference keeps no activations
        dev = pts.device
        out = _out_buffer(pts, pairs)
        tc = _mlp_math == "f16tc"
        save = None
        if need_grad and not tc:  # the tensor-core backward recomputes the forward per tile instead
            save = torch.empty(int(L.acn_ngp_save_bytes(pairs.M)) // 4, device=dev, dtype=torch.float32)
        ex = _expert_array(spec, None, ws)
        cp, cq = pts.c(), pairs.c()
        with _on(dev), _timed("mlp_fwd"):
            if tc:
                check(L.acn_ngp_mlp_fwd_tc(C.byref(cp), C.byref(cq), ex, spec.E, _ptr(enc_c), _ptr(out), _stream()), "acn_ngp_mlp_fwd_tc")
            else:
                check(L.acn_ngp_mlp_fwd(C.byref(cp), C.byref(cq), ex, spec.E, _ptr(enc_c), _ptr(out), _ptr(save), _stream()),
                      "acn_ngp_mlp_fwd")
        _count(1)
        ctx.pts, ctx.pairs, ctx.spec, ctx.save, ctx.tc = pts, pairs, spec, save, tc
        ctx.save_for_backward(enc_c, *ws)
        return out

    @staticmethod
    @first_order
    def backward(ctx, d_out):
        L = lib()
        enc, *ws = ctx.saved_tensors
        pts, pairs, spec = ctx.pts, ctx.pairs, ctx.spec
        if ctx.save is None and not ctx.tc:
            raise RuntimeError("NgpMlpFn.backward without saved activations (forward ran under no_grad)")
        dev = pts.device
        d_out = _f32(d_out, "d_out")
        grads = (AcnNgpExpertGrads * spec.E)()
        outs: List[Optional[Tensor]] = []
        for e in range(spec.E):
            live = _segment_len(pairs, e) > 0
            for l in range(ACN_NGP_LAYERS):
                for j in range(2):
                    i = e * 14 + 2 * l + j
                    if ctx.needs_input_grad[4 + i] and live:
                        g = torch.empty_like(ws[i])
                        if j == 0:
                            grads[e].w[l] = _ptr(g)
                        else:
                            grads[e].b[l] = _ptr(g)
                        outs.append(g)
                    else:
                        outs.append(None)
        d_enc = torch.empty_like(enc)
        ex = _expert_array(spec, None, ws)
        cq = pairs.c()
        if ctx.tc:
            wsb = torch.empty(int(L.acn_ngp_bwd_tc_ws_bytes()) // 4, device=dev, dtype=torch.float32)
            cp = pts.c()
            with _on(dev), _timed("mlp_bwd"):
                check(L.acn_ngp_mlp_bwd_tc(C.byref(cp), C.byref(cq), ex, spec.E, _ptr(enc), _ptr(d_out), grads, _ptr(d_enc), _ptr(wsb),
                                           _stream()), "acn_ngp_mlp_bwd_tc")
            _count(3)
        else:
            wsb = torch.empty(int(L.acn_ngp_bwd_ws_bytes(pairs.M, spec.E)) // 4, device=dev, dtype=torch.float32)
            with _on(dev), _timed("mlp_bwd"):
                check(L.acn_ngp_mlp_bwd(C.byref(cq), ex, spec.E, _ptr(enc), _ptr(d_out), _ptr(ctx.save), grads, _ptr(d_enc), _ptr(wsb),
                                        _stream()), "acn_ngp_mlp_bwd")
            _count(1 + 14)
        ctx.save = None
        return (d_enc if ctx.needs_input_grad[0] else None, None, None, None, *outs)


def tc_selftest(X: Tensor, Y: Tensor, G: Tensor, mixed: bool):
    """acn_tc_selftest: (D1 = X Y^T, D2 = G^T X, D3 = G Y) through the three tcgen05 operand layouts."""
    L = lib()
    X, Y, G = _need(X, "X"), _need(Y, "Y"), _need(G, "G")
    if tuple(X.shape) != (128, 64) or tuple(Y.shape) != (64, 64) or tuple(G.shape) != (128, 64):
        raise ValueError("tc_selftest shapes: X (128,64), Y (64,64), G (128,64)")
    D1 = torch.empty(128, 64, device=X.device)
    D2 = torch.empty(64, 64, device=X.device)
    D3 = torch.empty(128, 64, device=X.device)
    with _on(X.device):
        check(L.acn_tc_selftest(_ptr(X), _ptr(Y), _ptr(G), int(bool(mixed)), _ptr(D1), _ptr(D2), _ptr(D3), _stream()), "acn_tc_selftest")
    _count(1)
    return D1, D2, D3


# Chunks of the pipelined tensor-core backward (MLP backward of chunk c+1 beside the table scatter of chunk c on a side stream).
# Measured on B200: the two kernels do not overlap usefully - the scatter saturates the SM's LSU queue - and the
# half-register-file MLP variant is slower, so the default is 1 (two back-to-back launches); the path stays for experiments.
_bwd_chunks = 1
_side_streams = {}


def set_bwd_chunks(n: int) -> int:
    """1 = no pipelining (two back-to-back launches), 2..8 = software pipeline over that many chunks."""
    global _bwd_chunks
    if not 1 <= int(n) <= 8:
        raise ValueError("bwd chunks must be in [1, 8]")
    prev, _bwd_chunks = _bwd_chunks, int(n)
    return prev


def _side_stream(dev) -> torch.cuda.Stream:
    key = torch.device(dev).index
    if key not in _side_streams:
        _side_streams[key] = torch.cuda.Stream(device=dev)
    return _side_streams[key]


class NgpExpertsTcFn(torch.autograd.Function):
    """Encode + tensor-core MLPs as ONE autograd node, so that the backward can interleave the MLP backward with the
    hash-table scatter (acn_ngp_bwd_tc_pipelined).  Inputs: E tables then 14*E MLP tensors."""

    @staticmethod
    def forward(ctx, pts: PointSet, pairs: PairList, spec: ExpertSpec, *tensors: Tensor):
        L = lib()
        E = spec.E
        tb = [_need(t.detach(), "hash_table") for t in tensors[:E]]
        ws = _check_mlp(tensors[E:], E)
        rows = spec.grid.levels << spec.grid.log2_T
        for t in tb:
            if tuple(t.shape) != (rows, 2):
                raise ValueError(f"hash_table must be ({rows},2), got {tuple(t.shape)}")
        dev = pts.device
        ld = int(L.acn_ngp_ld(pairs.M))
        enc = torch.empty(32, ld, device=dev, dtype=torch.float32)
        cp, cq = pts.c(), pairs.c()
        with _on(dev):
            with _timed("hash_fwd"):
                check(L.acn_ngp_encode_fwd(C.byref(cp), C.byref(cq), _expert_array(spec, tb, None), E, C.byref(spec.grid), _ptr(enc), _stream()),
                      "acn_ngp_encode_fwd")
            out = _out_buffer(pts, pairs)
            with _timed("mlp_fwd"):
                check(L.acn_ngp_mlp_fwd_tc(C.byref(cp), C.byref(cq), _expert_array(spec, None, ws), E, _ptr(enc), _ptr(out), _stream()),
                      "acn_ngp_mlp_fwd_tc")
        _count(2)
        if any(ctx.needs_input_grad):
            ctx.pts, ctx.pairs, ctx.spec = pts, pairs, spec
            ctx.save_for_backward(enc, *tb, *ws)
        return out

    @staticmethod
    @first_order
    def backward(ctx, d_out):
        L = lib()
        pts, pairs, spec = ctx.pts, ctx.pairs, ctx.spec
        E = spec.E
        enc, *rest = ctx.saved_tensors
        tb, ws = rest[:E], rest[E:]
        dev = pts.device
        d_out = _f32(d_out, "d_out")
        grads = (AcnNgpExpertGrads * E)()
        t_out: List[Optional[Tensor]] = []
        m_out: List[Optional[Tensor]] = []
        want_table = False
        for e in range(E):
            live = _segment_len(pairs, e) > 0
            if ctx.needs_input_grad[3 + e] and live:
                g = _table_grad(tb[e])
                grads[e].table = _ptr(g)
                t_out.append(g)
                want_table = True
            else:
                t_out.append(None)
            for l in range(ACN_NGP_LAYERS):
                for j in range(2):
                    i = e * 14 + 2 * l + j
                    if ctx.needs_input_grad[3 + E + i] and live:
                        g = torch.empty_like(ws[i])
                        if j == 0:
                            grads[e].w[l] = _ptr(g)
                        else:
                            grads[e].b[l] = _ptr(g)
                        m_out.append(g)
                    else:
                        m_out.append(None)
        d_enc = torch.empty_like(enc)
        ex = _expert_array(spec, tb, ws)
        cp, cq = pts.c(), pairs.c()
        with _on(dev):
            if want_table and _bwd_chunks > 1:
                wsb = torch.empty(int(L.acn_ngp_bwd_tc_pipelined_ws_bytes(_bwd_chunks)) // 4, device=dev, dtype=torch.float32)
                side = _side_stream(dev)
                with _timed("mlp_hash_bwd"):
                    check(L.acn_ngp_bwd_tc_pipelined(C.byref(cp), C.byref(cq), ex, E, C.byref(spec.grid), _ptr(enc), _ptr(d_out), grads, _ptr(d_enc),
                                                     _ptr(wsb), _bwd_chunks, _stream(), C.c_void_p(side.cuda_stream)), "acn_ngp_bwd_tc_pipelined")
                _count(2 + 2 * _bwd_chunks + 1)
            else:
                wsb = torch.empty(int(L.acn_ngp_bwd_tc_ws_bytes()) // 4, device=dev, dtype=torch.float32)
                with _timed("mlp_bwd"):
                    check(L.acn_ngp_mlp_bwd_tc(C.byref(cp), C.byref(cq), ex, E, _ptr(enc), _ptr(d_out), grads, _ptr(d_enc), _ptr(wsb), _stream()),
                          "acn_ngp_mlp_bwd_tc")
                _count(3)
                if want_table:
                    with _timed("hash_bwd"):
                        check(L.acn_ngp_encode_bwd(C.byref(cp), C.byref(cq), ex, E, C.byref(spec.grid), _ptr(d_enc), grads, _stream()),
                              "acn_ngp_encode_bwd")
                    _count(1)
        return (None, None, None, *t_out, *m_out)


_fused_inference = True  # no_grad + tensor-core mode: gather and MLPs in one kernel (acn_ngp_fused_fwd_tc)


def set_fused_inference(enabled: bool) -> bool:
    global _fused_inference
    prev, _fused_inference = _fused_inference, bool(enabled)
    return prev


@torch.no_grad()
def ngp_experts_fused(pts: PointSet, pairs: PairList, spec: ExpertSpec, tables: Sequence[Tensor], mlp: Sequence[Tensor]) -> Tensor:
    """Inference form of `ngp_experts`: one launch, no (32, ld) encoding in HBM."""
    L = lib()
    E = spec.E
    tb = [_need(t.detach(), "hash_table") for t in tables]
    ws = _check_mlp(mlp, E)
    rows = spec.grid.levels << spec.grid.log2_T
    for t in tb:
        if tuple(t.shape) != (rows, 2):
            raise ValueError(f"hash_table must be ({rows},2), got {tuple(t.shape)}")
    out = _out_buffer(pts, pairs)
    cp, cq = pts.c(), pairs.c()
    with _on(pts.device), _timed("fused_fwd"):
        check(L.acn_ngp_fused_fwd_tc(C.byref(cp), C.byref(cq), _expert_array(spec, tb, ws), E, C.byref(spec.grid), _ptr(out), _stream()),
              "acn_ngp_fused_fwd_tc")
    _count(1)
    return out


def ngp_experts(pts: PointSet, pairs: PairList, spec: ExpertSpec, tables: Sequence[Tensor], mlp: Sequence[Tensor]) -> Tensor:
    """(P,4) [rgb, sigma] of E hash-grid experts over `pairs` (MetaNGP.forward / MetaContainer.forward)."""
    if _mlp_math == "f16tc":
        if _fused_inference and not torch.is_grad_enabled() and spec.grid.interp != INTERP["nearest"] and spec.grid.levels == 16:
            return ngp_experts_fused(pts, pairs, spec, tables, mlp)
        return NgpExpertsTcFn.apply(pts, pairs, spec, *tables, *mlp)
    enc = NgpEncodeFn.apply(pts, pairs, spec, *tables)
    return NgpMlpFn.apply(enc, pts, pairs, spec, *mlp)


# --------------------------------------------------------------------------
# a16 background head
# --------------------------------------------------------------------------
class BgFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, d: Tensor, stride: int, N: int, w0: Tensor, b0: Tensor, w1: Tensor, b1: Tensor):
        L = lib()
        dd = d.detach()  # strided view (e.g. rays[:, 3:6]); `stride` floats between rows, 3 adjacent components
        if not (dd.is_cuda and dd.dtype == torch.float32 and dd.stride(1) == 1 and dd.stride(0) == stride):
            raise RuntimeError("BgFn: directions must be a CUDA fp32 (N,3) view with unit inner stride")
        p = [_need(t.detach(), "bg parameter") for t in (w0, b0, w1, b1)]
        if tuple(p[0].shape) != (32, 16) or tuple(p[2].shape) != (3, 32):
            raise ValueError("background head must be SH16 -> 32 -> 3 (bg_hidden=32, spherical encoding)")
        need_grad = any(ctx.needs_input_grad)
        rgb = torch.empty(N, 3, device=dd.device, dtype=torch.float32)
        save = torch.empty(int(L.acn_bg_save_bytes(N)) // 4, device=dd.device, dtype=torch.float32) if need_grad else None
        with _on(dd.device), _timed("bg_fwd"):
            check(L.acn_bg_forward(_ptr(dd), stride, N, _ptr(p[0]), _ptr(p[1]), _ptr(p[2]), _ptr(p[3]), _ptr(rgb), _ptr(save),
                                   _stream()), "acn_bg_forward")
        _count(1)
        ctx.N, ctx.save = N, save
        ctx.save_for_backward(*p)
        return rgb

    @staticmethod
    @first_order
    def backward(ctx, d_rgb):
        L = lib()
        w0, b0, w1, b1 = ctx.saved_tensors
        if ctx.save is None:
            raise RuntimeError("BgFn.backward without saved activations")
        d_rgb = _f32(d_rgb, "d_rgb")
        g = [torch.empty_like(t) for t in (w0, b0, w1, b1)]
        ws = torch.empty(int(L.acn_bg_bwd_ws_bytes(ctx.N)) // 4, device=w0.device, dtype=torch.float32)
        with _on(w0.device), _timed("bg_bwd"):
            check(L.acn_bg_backward(_ptr(d_rgb), ctx.N, _ptr(w0), _ptr(w1), _ptr(ctx.save), _ptr(g[0]), _ptr(g[1]), _ptr(g[2]),
                                    _ptr(g[3]), _ptr(ws), _stream()), "acn_bg_backward")
        _count(2)
        ctx.save = None
        return (None, None, None, *g)


def background(d: Tensor, w0: Tensor, b0: Tensor, w1: Tensor, b1: Tensor) -> Tensor:
    """(N,3) directions (any row stride that keeps the 3 components adjacent) -> (N,3) rgb."""
    if d.dim() != 2 or d.shape[1] != 3 or d.stride(1) != 1:
        d = d.reshape(-1, 3).contiguous()
    if not d.is_cuda:
        raise RuntimeError("background(): CUDA tensors only")
    if d.dtype != torch.float32:
        d = d.float().contiguous()
    return BgFn.apply(d, int(d.stride(0)), int(d.shape[0]), w0, b0, w1, b1)


# --------------------------------------------------------------------------
# a17 compositing
# --------------------------------------------------------------------------
class CompositeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rgb_sigma: Tensor, t: Tensor, bg: Optional[Tensor], sigma_scale: float):
        rs = _f32(rgb_sigma.detach(), "rgb_sigma")
        tt = _f32(t.detach(), "t_vals")
        N, S = tt.shape
        if tuple(rs.shape) != (N, S, 4):
            raise ValueError(f"rgb_sigma must be ({N},{S},4), got {tuple(rs.shape)}")
        b = None if bg is None else _f32(bg.detach(), "bg_rgb")
        dev = rs.device
        rgb, depth = torch.empty(N, 3, device=dev), torch.empty(N, device=dev)
        w, acc = torch.empty(N, S, device=dev), torch.empty(N, device=dev)
        with _on(dev), _timed("composite_fwd"):
            check(lib().acn_composite_fwd(_ptr(rs), _ptr(tt), _ptr(b), N, S, float(sigma_scale), _ptr(rgb), _ptr(depth), _ptr(w),
                                          _ptr(acc), _stream()), "acn_composite_fwd")
        _count(1)
        ctx.scale, ctx.has_bg = float(sigma_scale), b is not None
        ctx.save_for_backward(rs, tt, b if b is not None else rs.new_empty(0))
        return rgb, depth, w, acc

    @staticmethod
    @first_order
    def backward(ctx, g_rgb, g_depth, g_w, g_acc):
        rs, tt, b = ctx.saved_tensors
        N, S = tt.shape
        dev = rs.device
        b = b if ctx.has_bg else None
        g_rgb = torch.zeros(N, 3, device=dev) if g_rgb is None else _f32(g_rgb, "g_rgb")
        g_depth = None if g_depth is None else _f32(g_depth, "g_depth")
        g_acc = None if g_acc is None else _f32(g_acc, "g_acc")
        g_w = None if g_w is None else _f32(g_w, "g_weights")
        d_rs = torch.empty_like(rs)
        d_bg = torch.empty_like(b) if (b is not None and ctx.needs_input_grad[2]) else None
        with _on(dev), _timed("composite_bwd"):
            check(lib().acn_composite_bwd(_ptr(rs), _ptr(tt), _ptr(b), N, S, ctx.scale, _ptr(g_rgb), _ptr(g_depth), _ptr(g_acc),
                                          _ptr(g_w), _ptr(d_rs), _ptr(d_bg), _stream()), "acn_composite_bwd")
        _count(1)
        return d_rs, None, d_bg, None


def composite(rgb_sigma: Tensor, t: Tensor, bg: Optional[Tensor], sigma_scale: float = 1.0):
    return CompositeFn.apply(rgb_sigma, t, bg, float(sigma_scale))


@torch.no_grad()
def composite_sample_major(rgb_sigma: Tensor, t_T: Tensor, bg: Optional[Tensor], sigma_scale: float = 1.0):
    """`composite` for the (S*N, 4) output of a sample-major point set (PointSet.sample_major) and its (S,N) sample
    distances: one thread per ray over coalesced rows (acn_composite_fwd_sm), no transposed copies.  Returns
    rgb (N,3), depth (N,), weights (N,S) (a transposed view of the (S,N) buffer), acc (N,).  Inference only."""
    rs, tt = _f32(rgb_sigma, "rgb_sigma"), _f32(t_T, "t_T")
    S, N = tt.shape
    if rs.numel() != N * S * 4:
        raise ValueError(f"rgb_sigma must hold {S}x{N} rows of 4, got {tuple(rs.shape)}")
    b = None if bg is None else _f32(bg, "bg_rgb")
    dev = rs.device
    rgb, depth = torch.empty(N, 3, device=dev), torch.empty(N, device=dev)
    wT, acc = torch.empty(S, N, device=dev), torch.empty(N, device=dev)
    with _on(dev), _timed("composite_fwd"):
        check(lib().acn_composite_fwd_sm(_ptr(rs), _ptr(tt), _ptr(b), N, S, float(sigma_scale), _ptr(rgb), _ptr(depth), _ptr(wT),
                                         _ptr(acc), _stream()), "acn_composite_fwd_sm")
    _count(1)
    return rgb, depth, wT.t(), acc


# --------------------------------------------------------------------------
# a18 loss, (f)2 optimiser pieces
# --------------------------------------------------------------------------
class MseFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred: Tensor, gt: Tensor, space: int):
        p, g = _f32(pred.detach(), "pred"), _f32(gt.detach(), "gt")
        N = p.shape[0]
        loss = torch.empty(1, device=p.device)
        d_pred = torch.empty_like(p)
        ws = torch.empty(2048, device=p.device)
        with _on(p.device), _timed("mse_loss"):
            check(lib().acn_mse_loss(_ptr(p), _ptr(g), N, int(space), _ptr(loss), _ptr(d_pred), _ptr(ws), _stream()), "acn_mse_loss")
        _count(2)
        ctx.save_for_backward(d_pred)
        return loss.reshape(())

    @staticmethod
    @first_order
    def backward(ctx, g):
        (d_pred,) = ctx.saved_tensors
        return d_pred * g, None, None


def mse_loss(pred: Tensor, gt: Tensor, color_space: str = "linear") -> Tensor:
    return MseFn.apply(pred, gt, {"linear": 0, "srgb": 1, "identity": 2}[str(color_space).lower()])


def adam_step(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, lr: float, beta1: float, beta2: float, eps: float,
              weight_decay: float, step: int, grad_scale: Optional[Tensor] = None):
    for t, n in ((param, "param"), (grad, "grad"), (exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq")):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError(f"adam_step: {n} must be a contiguous CUDA fp32 tensor")
    with _on(param.device), _timed("adam"):
        check(lib().acn_adam_step(_ptr(param), _ptr(grad), _ptr(exp_avg), _ptr(exp_avg_sq), param.numel(), lr, beta1, beta2, eps,
                                  weight_decay, int(step), _ptr(grad_scale), _stream()), "acn_adam_step")
    _count(1)


def sgd_axpy(w: Tensor, g: Tensor, lr: float) -> Tensor:
    if isinstance(g, torch.Tensor) and g.requires_grad:
        raise RuntimeError("sgd_axpy: the gradient carries an autograd graph (create_graph=True / algo='maml'); the sm_100a ops are "
                           "first-order - use algo='fomaml' or 'reptile'")
    ww, gg = _need(w.detach(), "w"), _need(g.detach(), "g")
    out = torch.empty_like(ww)
    with _on(ww.device):
        check(lib().acn_sgd_axpy(_ptr(ww), _ptr(gg), float(lr), ww.numel(), _ptr(out), _stream()), "acn_sgd_axpy")
    _count(1)
    return out


def sumsq_accum(x: Tensor, acc: Tensor):
    xx = _need(x.detach(), "x")
    with _on(xx.device), _timed("sumsq"):
        check(lib().acn_sumsq_accum(_ptr(xx), xx.numel(), _ptr(acc), _stream()), "acn_sumsq_accum")
    _count(1)


# --------------------------------------------------------------------------
# a21 / (f)-1: occupancy-guided rendering (csrc/occ.cu)
# --------------------------------------------------------------------------
ACN_OCC_MAX_LEVELS = _lib.ACN_OCC_MAX_LEVELS


class PackedSegments:
    """Per-ray lists of [t0, t1) segments, packed ray by ray: ray_idx (M,) int32, t0 / t1 (M,), offsets (N,) int64 (first
    segment of each ray), counts (N,) int32.  (nerfacc's ray_indices + packed_info.)"""

    def __init__(self, N: int, ray_idx: Tensor, t0: Tensor, t1: Tensor, offsets: Tensor, counts: Tensor, cells: Optional[Tensor] = None):
        self.N, self.ray_idx, self.t0, self.t1, self.offsets, self.counts, self.cells = int(N), ray_idx, t0, t1, offsets, counts, cells

    @property
    def M(self) -> int:
        return int(self.t0.shape[0])


def _exclusive_offsets(counts: Tensor) -> Tuple[Tensor, int]:
    """(offsets int64, total): the one host read of a packed-list construction (nerfacc reads the same total)."""
    inc = torch.cumsum(counts, 0, dtype=torch.int64)
    total = int(inc[-1].item()) if counts.numel() else 0
    return inc - counts, total


@torch.no_grad()
def occ_march(rays: Tensor, near_jitter: Optional[Tensor], binaries: Tensor, aabbs_host: Sequence[float], step: float, cone_angle: float,
              max_samples: int = 4096, return_cells: bool = False) -> PackedSegments:
    """Multi-level occupancy-grid traversal of (N,8) rays [o, d, near, far] (OccGridEstimator.sampling before its visibility
    filter): count pass, exclusive scan, write pass."""
    rays = _need(rays, "rays")
    if binaries.dtype not in (torch.bool, torch.uint8) or binaries.dim() != 4 or not binaries.is_cuda:
        raise TypeError("binaries must be a CUDA bool/uint8 (levels, R, R, R) tensor")
    b8 = binaries.contiguous().view(torch.uint8)
    levels, R = int(b8.shape[0]), int(b8.shape[1])
    jit = None if near_jitter is None else _need(near_jitter, "near_jitter")
    N = rays.shape[0]
    dev = rays.device
    ah = host_floats(aabbs_host)
    counts = torch.zeros(N, dtype=torch.int32, device=dev)
    L = lib()
    with _on(dev), _timed("occ_march"):
        check(L.acn_occ_march_count(_ptr(rays), _ptr(jit), N, _ptr(b8), ah, levels, R, float(step), float(cone_angle), int(max_samples), _ptr(counts),
                                    _stream()), "acn_occ_march_count")
    offsets, M = _exclusive_offsets(counts)
    ray_idx = torch.empty(M, dtype=torch.int32, device=dev)
    t0, t1 = torch.empty(M, device=dev), torch.empty(M, device=dev)
    cells = torch.empty(M, dtype=torch.int32, device=dev) if return_cells else None
    if M:
        with _on(dev), _timed("occ_march"):
            check(L.acn_occ_march_write(_ptr(rays), _ptr(jit), N, _ptr(b8), ah, levels, R, float(step), float(cone_angle), int(max_samples), _ptr(offsets),
                                        _ptr(ray_idx), _ptr(t0), _ptr(t1), _ptr(cells), _stream()), "acn_occ_march_write")
    _count(2)
    return PackedSegments(N, ray_idx, t0, t1, offsets, counts, cells)


@torch.no_grad()
def occ_visibility(seg: PackedSegments, sigma: Tensor, early_stop_eps: float, alpha_thre: float) -> Tensor:
    """bool (M,): nerfacc.render_visibility_from_density."""
    sg = _need(sigma.detach().reshape(-1), "sigma")
    if sg.numel() != seg.M:
        raise ValueError(f"sigma has {sg.numel()} entries for {seg.M} segments")
    vis = torch.empty(seg.M, dtype=torch.uint8, device=sg.device)
    with _on(sg.device), _timed("occ_visibility"):
        check(lib().acn_occ_visibility(_ptr(seg.offsets), _ptr(seg.counts), seg.N, _ptr(seg.t0), _ptr(seg.t1), _ptr(sg), float(early_stop_eps),
                                       float(alpha_thre), _ptr(vis), _stream()), "acn_occ_visibility")
    _count(1)
    return vis.bool()


def pack_segments(ray_idx: Tensor, t0: Tensor, t1: Tensor, N: int) -> PackedSegments:
    """Packed lists from sorted ray ids (nerfacc.pack_info): counts by bincount, offsets by scan."""
    counts = torch.bincount(ray_idx.long(), minlength=N).to(torch.int32)
    offsets = torch.cumsum(counts, 0, dtype=torch.int64) - counts
    return PackedSegments(N, ray_idx.to(torch.int32).contiguous(), _f32(t0, "t0"), _f32(t1, "t1"), offsets, counts)


@torch.no_grad()
def seg_union(lists: Sequence[PackedSegments], N: int) -> PackedSegments:
    """_merge_segments_union (ray_rendering.py:196-256) as one kernel: per ray, the sorted distinct boundaries of every
    list's segments; consecutive boundaries are the merged segments.  Every list covers the same N rays."""
    from ._lib import AcnSegLists
    if not 1 <= len(lists) <= ACN_MAX_EXPERTS:
        raise ValueError(f"1..{ACN_MAX_EXPERTS} lists")
    dev = lists[0].t0.device
    sl = AcnSegLists()
    sl.K = len(lists)
    for k, s in enumerate(lists):
        if s.N != N:
            raise ValueError("every list must cover the same rays")
        sl.offsets[k], sl.counts[k], sl.t0[k], sl.t1[k] = s.offsets.data_ptr(), s.counts.data_ptr(), s.t0.data_ptr(), s.t1.data_ptr()
    counts = torch.zeros(N, dtype=torch.int32, device=dev)
    L = lib()
    with _on(dev), _timed("seg_union"):
        check(L.acn_seg_union(C.byref(sl), N, None, _ptr(counts), None, None, None, _stream()), "acn_seg_union")
    offsets, M = _exclusive_offsets(counts)
    ray_idx = torch.empty(M, dtype=torch.int32, device=dev)
    t0, t1 = torch.empty(M, device=dev), torch.empty(M, device=dev)
    if M:
        with _on(dev), _timed("seg_union"):
            check(L.acn_seg_union(C.byref(sl), N, _ptr(offsets), None, _ptr(ray_idx), _ptr(t0), _ptr(t1), _stream()), "acn_seg_union")
    _count(2)
    return PackedSegments(N, ray_idx, t0, t1, offsets, counts)


class PackedRenderFn(torch.autograd.Function):
    """render_weight_from_density + the three accumulate_along_rays of the occupancy renderer (ray_rendering.py:477-487)."""

    @staticmethod
    def forward(ctx, sigma: Tensor, rgb: Tensor, seg: PackedSegments):
        sg, cl = _need(sigma.detach().reshape(-1), "sigma"), _need(rgb.detach(), "rgb")
        if sg.numel() != seg.M or tuple(cl.shape) != (seg.M, 3):
            raise ValueError("sigma (M,) / rgb (M,3) do not match the packed segments")
        dev, N = sg.device, seg.N
        w = torch.empty(seg.M, device=dev)
        rgb_map, depth, acc = torch.empty(N, 3, device=dev), torch.empty(N, device=dev), torch.empty(N, device=dev)
        with _on(dev), _timed("packed_render_fwd"):
            check(lib().acn_packed_render_fwd(_ptr(seg.offsets), _ptr(seg.counts), N, _ptr(seg.t0), _ptr(seg.t1), _ptr(sg), _ptr(cl), _ptr(w),
                                              _ptr(rgb_map), _ptr(depth), _ptr(acc), _stream()), "acn_packed_render_fwd")
        _count(1)
        ctx.seg = seg
        ctx.save_for_backward(sg, cl, w)
        ctx.sigma_shape = sigma.shape
        return rgb_map, depth, w, acc

    @staticmethod
    @first_order
    def backward(ctx, d_rgb, d_depth, d_w, d_acc):
        sg, cl, w = ctx.saved_tensors
        seg = ctx.seg
        g_sigma, g_rgb = torch.empty_like(sg), torch.empty_like(cl)
        # contiguous copies of expanded incoming gradients must outlive the launch (a temporary freed inside the argument list
        # is handed to the next temporary by the caching allocator)
        gr, gd, ga, gw = (None if t is None else _f32(t, "gradient") for t in (d_rgb, d_depth, d_acc, d_w))
        with _on(sg.device), _timed("packed_render_bwd"):
            check(lib().acn_packed_render_bwd(_ptr(seg.offsets), _ptr(seg.counts), seg.N, _ptr(seg.t0), _ptr(seg.t1), _ptr(sg), _ptr(cl), _ptr(w),
                                              _ptr(gr), _ptr(gd), _ptr(ga), _ptr(gw), _ptr(g_sigma), _ptr(g_rgb), _stream()), "acn_packed_render_bwd")
        _count(1)
        return g_sigma.view(ctx.sigma_shape), g_rgb, None


def packed_render(sigma: Tensor, rgb: Tensor, seg: PackedSegments):
    """rgb_map (N,3), depth (N,), weights (M,), acc (N,) - without background."""
    return PackedRenderFn.apply(sigma, rgb, seg)
