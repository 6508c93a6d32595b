"""Photometric loss over the CUDA renderer; signature of the reference's
``nerfs/losses.py`` compute_mse_loss (:10-32) and compute_loss (:154-166).  The Fisher-weighted
branch of the reference is unreachable there (no ``fisher_store`` is ever created, losses.py:76-80),
so ``compute_loss`` is the MSE loss."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import ops
from .color_space import color_space_transformer
from .ray_rendering import render_rays


def compute_mse_loss(P, model, data, params=None, active_module=None, reduction="mean"):
    gt_rgb, rays = data["rgbs"], data["rays"]
    pred_rgb, *_ = render_rays(model, rays, ray_samples=P.ray_samples, params=params, active_module=active_module,
                               chunk=P.chunk_points)
    if reduction == "mean" and pred_rgb.is_cuda:
        return ops.mse_loss(pred_rgb, gt_rgb.to(pred_rgb.device), getattr(P, "color_space", "linear"))
    pred_rgb, gt_rgb = color_space_transformer(pred_rgb, gt_rgb, color_space=P.color_space)
    return F.mse_loss(pred_rgb, gt_rgb, reduction=reduction)


def compute_loss(P, model, data, params=None, active_module=None, **kwargs):
    return compute_mse_loss(P, model, data, params=params, active_module=active_module)
