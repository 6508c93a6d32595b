"""Fused optimiser step: torch.optim.Adam semantics and state_dict layout
(``step``, ``exp_avg``, ``exp_avg_sq`` per parameter; the reference builds torch Adam/AdamW with
per-group learning rates, common/utils.py:16-76) with the arithmetic in csrc/optim.cu, plus
clip_grad_norm_ (meta_core.py:181-190, runtime_adapt.py:261-268) folded into the same pass: the
clip coefficient stays on the device and multiplies the gradient inside the Adam kernel, so the
128 MiB hash-table gradients are read once for the norm and once for the update, and there is no
host synchronisation."""
from __future__ import annotations

from typing import Optional

import torch

from . import ops


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm: Optional[float] = None):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.max_grad_norm = max_grad_norm
        self._scale = None

    @torch.no_grad()
    def clip_coef(self, max_norm: float) -> torch.Tensor:
        """Device scalar min(1, max_norm / (||g|| + 1e-6)) over every gradient (torch.nn.utils.clip_grad_norm_)."""
        dev = None
        acc = None
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if acc is None:
                    dev = p.device
                    acc = torch.zeros(1, device=dev, dtype=torch.float32)
                ops.sumsq_accum(p.grad, acc)
        if acc is None:
            return None
        return (max_norm / (acc.sqrt() + 1e-6)).clamp_(max=1.0)

    @torch.no_grad()
    def step(self, closure=None, grad_scale: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        scale = grad_scale
        if self.max_grad_norm is not None:
            c = self.clip_coef(self.max_grad_norm)
            if c is not None:
                scale = c if scale is None else c * scale
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                ops.adam_step(p.data, g, st["exp_avg"], st["exp_avg_sq"], float(group["lr"]), float(b1), float(b2), float(group["eps"]),
                              float(group["weight_decay"]), int(st["step"].item()), scale)
        return loss


def build_optimizer(model, lr_encoding=1e-2, lr_sigma=2e-3, lr_color=2e-3, lr_background=1e-3, max_grad_norm: Optional[float] = 1.0,
                    eps: float = 1e-8):
    """Per-group learning rates of the reference (common/args.py:115-125; common/utils.py:23-47)."""
    groups = model.get_param_groups()
    lrs = {"encoding": lr_encoding, "sigma": lr_sigma, "color": lr_color, "background": lr_background}
    pg = [{"params": g["params"], "lr": lrs[name]} for name, g in groups.items()]
    return FusedAdam(pg, betas=(0.9, 0.999), eps=eps, max_grad_norm=max_grad_norm)  # torch AdamW defaults, weight_decay 0
