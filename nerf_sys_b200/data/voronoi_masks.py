"""Voronoi mask builder on the GPU: scripts/create_clusters.py ``compute_voronoi_opt`` (:386-556) /
``compute_voronoi_orig`` (:559-634) with the per-ray work in one kernel (csrc/voronoi.cu) - no (rays x samples x C) distance
matrices, no per-expert Python loops for the bounding boxes.  Same keyword arguments; the chunk sizes are accepted and unused
(nothing is staged).  The per-expert boxes / counts are updated in place on the device; `reduce_boxes` is the script's
cross-rank reduction (:928-932)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
from torch import Tensor

from .. import ops
from .._lib import check, host_floats, lib


@torch.no_grad()
def compute_voronoi_opt(rays: Tensor, *, ray_samples: int, centroids: Tensor, cluster_2d: bool, boundary_margin: float, device=None,
                        ray_chunk_size: int = 0, sample_chunk_size: int = 0, fp16: bool = False, update_aabbs: bool = False,
                        mins_out: Optional[Tensor] = None, maxs_out: Optional[Tensor] = None, counts_out: Optional[Tensor] = None,
                        to_cpu: bool = True) -> Tensor:
    """(N,8) rays -> (N,C) bool membership mask (on the CPU like the reference unless ``to_cpu=False``).  With ``update_aabbs``
    the (C,3) fp32 ``mins_out`` / ``maxs_out`` and the (C,) int64 ``counts_out`` on the device are updated with the positions
    of the member samples.  Distances are evaluated directly in fp32 (``fp16`` is accepted and ignored: the decision is
    an integer output)."""
    dev = torch.device(device) if device is not None else (rays.device if rays.is_cuda else torch.device("cuda", torch.cuda.current_device()))
    r = ops._f32(rays.to(dev), "rays")
    if r.dim() != 2 or r.shape[1] < 8:
        raise ValueError("rays must be (N, >=8) [o, d, near, far]")
    if r.shape[1] != 8:
        r = r[:, :8].contiguous()
    N, Cn = r.shape[0], int(centroids.shape[0])
    mask = torch.empty(N, Cn, dtype=torch.uint8, device=dev)
    ch = host_floats(centroids.detach().float().cpu().reshape(-1).tolist())
    ws = mins = maxs = cnts = None
    if update_aabbs:
        if mins_out is None or maxs_out is None or counts_out is None:
            raise ValueError("update_aabbs needs mins_out, maxs_out and counts_out")
        for t, dt, shape in ((mins_out, torch.float32, (Cn, 3)), (maxs_out, torch.float32, (Cn, 3)), (counts_out, torch.int64, (Cn,))):
            if not (t.is_cuda and t.dtype == dt and tuple(t.shape) == shape and t.is_contiguous()):
                raise ValueError("mins_out / maxs_out must be contiguous CUDA fp32 (C,3), counts_out CUDA int64 (C,)")
        mins, maxs, cnts = mins_out, maxs_out, counts_out
        ws = torch.empty(int(lib().acn_voronoi_ws_bytes()) // 4, dtype=torch.int32, device=dev)
    with ops._on(dev), ops._timed("voronoi_masks"):
        check(lib().acn_voronoi_masks(ops._ptr(r), N, int(ray_samples), ch, Cn, int(bool(cluster_2d)), float(boundary_margin), ops._ptr(mask),
                                      ops._ptr(mins), ops._ptr(maxs), ops._ptr(cnts), ops._ptr(ws), ops._stream()), "acn_voronoi_masks")
    ops._count(3 if update_aabbs else 1)
    out = mask.bool()
    return out.cpu() if to_cpu else out


compute_voronoi_orig = compute_voronoi_opt  # one GPU implementation behind both of the script's entry points


def reduce_boxes(mins: Tensor, maxs: Tensor, cnts: Tensor, group=None) -> None:
    """scripts/create_clusters.py:928-932: every rank routed its own images; the boxes and counts are combined in place."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(mins, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(maxs, op=dist.ReduceOp.MAX, group=group)
        dist.all_reduce(cnts, op=dist.ReduceOp.SUM, group=group)
