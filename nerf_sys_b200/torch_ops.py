"""``torch.ops.acnerf.*``: the tensor-signature ops of the path registered with ``torch.library.custom_op`` (+ ``register_fake``
for shape propagation under FakeTensor / torch.compile / export, + ``register_autograd``), over the same C-ABI calls the
``ops.py`` wrappers make ("torch custom-op extension").

Registered here: volume rendering (a17), packed occupancy rendering (a21), the MSE loss (a18), SH16 / frequency encoders
(a10-a11), the stand-alone hash-grid encoder forward (a9).  The routed expert evaluation (`ops.ngp_experts`) and the sample
exchange keep ``torch.autograd.Function``: their inputs are host-side descriptors (speculative pair-list capacities, per-expert
pointer blocks in the kernel-parameter array, device-resident segment tables the host never reads, NVLink slab routes) whose
lifetime and aliasing rules - a result stored in place in a peer-mapped slab, a table gradient adopted from a persistent
buffer - are exactly what the functional custom-op schema forbids (no aliasing outputs, no hidden mutation)."""
from __future__ import annotations

from typing import List, Tuple

import torch
from torch import Tensor

from . import ops
from ._lib import check, lib

_C = ops.C


# ---- a17 volume rendering ----------------------------------------------------------------------------------------------
@torch.library.custom_op("acnerf::volume_render", mutates_args=())
def volume_render(rgb_sigma: Tensor, t_vals: Tensor, bg_rgb: Tensor, sigma_scale: float) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """(N,S,4),(N,S),(N,3) -> rgb (N,3), depth (N,), weights (N,S), acc (N,)  (nerfs/ray_rendering.py:114-165)"""
    rs, tt, b = ops._f32(rgb_sigma, "rgb_sigma"), ops._f32(t_vals, "t_vals"), ops._f32(bg_rgb, "bg_rgb")
    N, S = tt.shape
    dev = rs.device
    rgb, depth = torch.empty(N, 3, device=dev), torch.empty(N, device=dev)
    w, acc = torch.empty(N, S, device=dev), torch.empty(N, device=dev)
    with ops._on(dev):
        check(lib().acn_composite_fwd(ops._ptr(rs), ops._ptr(tt), ops._ptr(b), N, S, float(sigma_scale), ops._ptr(rgb), ops._ptr(depth), ops._ptr(w),
                                      ops._ptr(acc), ops._stream()), "acn_composite_fwd")
    ops._count(1)
    return rgb, depth, w, acc


@volume_render.register_fake
def _(rgb_sigma, t_vals, bg_rgb, sigma_scale):
    N, S = t_vals.shape
    f = rgb_sigma.new_empty
    return f((N, 3)), f((N,)), f((N, S)), f((N,))


@torch.library.custom_op("acnerf::volume_render_backward", mutates_args=())
def volume_render_backward(rgb_sigma: Tensor, t_vals: Tensor, bg_rgb: Tensor, sigma_scale: float, g_rgb: Tensor, g_depth: Tensor, g_w: Tensor,
                           g_acc: Tensor) -> Tuple[Tensor, Tensor]:
    rs, tt, b = ops._f32(rgb_sigma, "rgb_sigma"), ops._f32(t_vals, "t_vals"), ops._f32(bg_rgb, "bg_rgb")
    N, S = tt.shape
    d_rs, d_bg = torch.empty_like(rs), torch.empty_like(b)
    # incoming gradients may be expanded views (e.g. of a .sum()): the contiguous copies must stay alive until the launch
    gr, gd, ga, gw = (ops._f32(g, "gradient") for g in (g_rgb, g_depth, g_acc, g_w))
    with ops._on(rs.device):
        check(lib().acn_composite_bwd(ops._ptr(rs), ops._ptr(tt), ops._ptr(b), N, S, float(sigma_scale), ops._ptr(gr), ops._ptr(gd), ops._ptr(ga),
                                      ops._ptr(gw), ops._ptr(d_rs), ops._ptr(d_bg), ops._stream()), "acn_composite_bwd")
    ops._count(1)
    return d_rs, d_bg


@volume_render_backward.register_fake
def _(rgb_sigma, t_vals, bg_rgb, sigma_scale, g_rgb, g_depth, g_w, g_acc):
    return torch.empty_like(rgb_sigma), torch.empty_like(bg_rgb)


def _vr_setup(ctx, inputs, output):
    ctx.save_for_backward(inputs[0], inputs[1], inputs[2])
    ctx.scale = inputs[3]


def _vr_backward(ctx, g_rgb, g_depth, g_w, g_acc):
    rs, tt, b = ctx.saved_tensors
    d_rs, d_bg = torch.ops.acnerf.volume_render_backward(rs, tt, b, ctx.scale, g_rgb, g_depth, g_w, g_acc)
    return d_rs, None, d_bg, None


volume_render.register_autograd(_vr_backward, setup_context=_vr_setup)


# ---- a18 loss --------------------------------------------------------------------------------------------------------------
@torch.library.custom_op("acnerf::mse_loss", mutates_args=())
def mse_loss(pred: Tensor, gt: Tensor, color_space: int) -> Tuple[Tensor, Tensor]:
    """-> (loss (), d loss / d pred (N,3)); color_space 0 linear, 1 srgb, 2 identity  (nerfs/losses.py:10-32)"""
    p, g = ops._f32(pred, "pred"), ops._f32(gt, "gt")
    loss, d_pred, ws = torch.empty(1, device=p.device), torch.empty_like(p), torch.empty(2048, device=p.device)
    with ops._on(p.device):
        check(lib().acn_mse_loss(ops._ptr(p), ops._ptr(g), p.shape[0], int(color_space), ops._ptr(loss), ops._ptr(d_pred), ops._ptr(ws),
                                 ops._stream()), "acn_mse_loss")
    ops._count(2)
    return loss.reshape(()), d_pred


@mse_loss.register_fake
def _(pred, gt, color_space):
    return pred.new_empty(()), torch.empty_like(pred)


def _mse_setup(ctx, inputs, output):
    ctx.save_for_backward(output[1])


def _mse_backward(ctx, g_loss, _g_dpred):
    (d_pred,) = ctx.saved_tensors
    return d_pred * g_loss, None, None


mse_loss.register_autograd(_mse_backward, setup_context=_mse_setup)


# ---- a10 / a11 encoders (no gradient w.r.t. positions / directions is ever taken: they are not learned) ------------------------
@torch.library.custom_op("acnerf::sh16", mutates_args=())
def sh16(dirs: Tensor) -> Tensor:
    return ops.sh16(dirs)


@sh16.register_fake
def _(dirs):
    return dirs.new_empty((*dirs.shape[:-1], 16))


@torch.library.custom_op("acnerf::freq_encode", mutates_args=())
def freq_encode(x: Tensor, n_freq: int, include_input: bool) -> Tensor:
    return ops.freq_encode(x, n_freq, include_input)


@freq_encode.register_fake
def _(x, n_freq, include_input):
    return x.new_empty((*x.shape[:-1], (3 if include_input else 0) + 6 * n_freq))


# ---- a21 packed rendering -----------------------------------------------------------------------------------------------------
@torch.library.custom_op("acnerf::packed_render", mutates_args=())
def packed_render(sigma: Tensor, rgb: Tensor, t0: Tensor, t1: Tensor, offsets: Tensor, counts: Tensor) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """segments packed ray by ray (offsets (N,) int64, counts (N,) int32) -> rgb (N,3), depth (N,), weights (M,), acc (N,)"""
    seg = ops.PackedSegments(offsets.shape[0], torch.empty(0, dtype=torch.int32, device=sigma.device), t0, t1, offsets, counts)
    N, M, dev = seg.N, sigma.numel(), sigma.device
    sg, cl = ops._f32(sigma.reshape(-1), "sigma"), ops._f32(rgb, "rgb")
    w = torch.empty(M, device=dev)
    rgb_map, depth, acc = torch.empty(N, 3, device=dev), torch.empty(N, device=dev), torch.empty(N, device=dev)
    with ops._on(dev):
        check(lib().acn_packed_render_fwd(ops._ptr(offsets), ops._ptr(counts), N, ops._ptr(t0), ops._ptr(t1), ops._ptr(sg), ops._ptr(cl), ops._ptr(w),
                                          ops._ptr(rgb_map), ops._ptr(depth), ops._ptr(acc), ops._stream()), "acn_packed_render_fwd")
    ops._count(1)
    return rgb_map, depth, w, acc


@packed_render.register_fake
def _(sigma, rgb, t0, t1, offsets, counts):
    N, M = offsets.shape[0], sigma.numel()
    f = sigma.new_empty
    return f((N, 3)), f((N,)), f((M,)), f((N,))


@torch.library.custom_op("acnerf::packed_render_backward", mutates_args=())
def packed_render_backward(sigma: Tensor, rgb: Tensor, t0: Tensor, t1: Tensor, offsets: Tensor, counts: Tensor, weights: Tensor, g_rgb: Tensor,
                           g_depth: Tensor, g_w: Tensor, g_acc: Tensor) -> Tuple[Tensor, Tensor]:
    sg, cl = ops._f32(sigma.reshape(-1), "sigma"), ops._f32(rgb, "rgb")
    g_sigma, g_c = torch.empty_like(sg), torch.empty_like(cl)
    gr, gd, ga, gw = (ops._f32(g, "gradient") for g in (g_rgb, g_depth, g_acc, g_w))  # kept alive until the launch
    with ops._on(sg.device):
        check(lib().acn_packed_render_bwd(ops._ptr(offsets), ops._ptr(counts), offsets.shape[0], ops._ptr(t0), ops._ptr(t1), ops._ptr(sg), ops._ptr(cl),
                                          ops._ptr(weights), ops._ptr(gr), ops._ptr(gd), ops._ptr(ga), ops._ptr(gw), ops._ptr(g_sigma), ops._ptr(g_c),
                                          ops._stream()), "acn_packed_render_bwd")
    ops._count(1)
    return g_sigma.view_as(sigma), g_c


@packed_render_backward.register_fake
def _(sigma, rgb, t0, t1, offsets, counts, weights, g_rgb, g_depth, g_w, g_acc):
    return torch.empty_like(sigma), torch.empty_like(rgb)


def _pr_setup(ctx, inputs, output):
    ctx.save_for_backward(*inputs, output[2])


def _pr_backward(ctx, g_rgb, g_depth, g_w, g_acc):
    sigma, rgb, t0, t1, offsets, counts, w = ctx.saved_tensors
    g_sigma, g_c = torch.ops.acnerf.packed_render_backward(sigma, rgb, t0, t1, offsets, counts, w, g_rgb, g_depth, g_w, g_acc)
    return g_sigma, g_c, None, None, None, None


packed_render.register_autograd(_pr_backward, setup_context=_pr_setup)

OPS: List[str] = ["volume_render", "volume_render_backward", "mse_loss", "sh16", "freq_encode", "packed_render", "packed_render_backward"]
