"""exp(clamp(x, +-88.72)) whose backward is g*exp(clamp(x)) - the reference's
``models/trunc_exp.py``:30-57.  Inside the fused expert kernel this is computed in place
(csrc/ngp.cu); this module exists for callers that import ``trunc_exp`` directly."""
import torch

_EXP_MAX = {torch.float16: 11.089866488, torch.bfloat16: 88.722839111, torch.float32: 88.722839111, torch.float64: 709.782712893}


class _TruncExpFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        m = _EXP_MAX.get(x.dtype, _EXP_MAX[torch.float32])
        y = torch.exp(x.clamp(-m, m))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        return g * y


def trunc_exp(x: torch.Tensor) -> torch.Tensor:
    return _TruncExpFn.apply(x)
