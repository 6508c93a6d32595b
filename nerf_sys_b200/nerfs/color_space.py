"""Colour-space helpers for the photometric loss (reference ``nerfs/color_space.py``:13-66).
Elementwise work on (N,3) tensors; the fused loss kernel is ops.mse_loss."""
import torch


def linear_to_srgb(x: torch.Tensor) -> torch.Tensor:
    x = x.clamp(0, 1)
    return torch.where(x <= 0.0031308, 12.92 * x, 1.055 * x.pow(1 / 2.4) - 0.055)


def srgb_to_linear(x: torch.Tensor) -> torch.Tensor:
    return torch.where(x <= 0.04045, x / 12.92, ((x + 0.055) / 1.055).pow(2.4))


def color_space_transformer(pred_linear: torch.Tensor, gt_tensor: torch.Tensor, color_space: str):
    cs = str(color_space).lower()
    pred = pred_linear.to(torch.float32)
    gt = gt_tensor.to(torch.float32).clamp(0, 1)
    if cs == "linear":
        pred, gt = pred.clamp(0, 1), srgb_to_linear(gt).clamp(0, 1)
    elif cs == "srgb":
        pred, gt = linear_to_srgb(pred).clamp(0, 1), gt.clamp(0, 1)
    elif cs == "identity":
        if (gt.max() > 1) or (gt.min() < 0):
            raise ValueError("GT out of [0,1]; identity mode assumes normalized linear GT.")
    else:
        raise ValueError(f"Invalid color_space={color_space!r}; use 'linear'|'srgb'|'identity'")
    return pred.to(pred_linear.dtype), gt.to(pred_linear.dtype).to(pred_linear.device)
