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
import torch.distributed as dist

import ctypes as C

from . import ops
from ._lib import ACN_ADAM_MAX_TENSORS, AcnAdamTensor, check, lib


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam over the C ABI's multi-tensor kernels: per step one sum-of-squares launch (clip_grad_norm_),
    one prep launch (device step counters + clip coefficient) and one Adam launch per 64 tensors.  Nothing in
    ``step()`` reads device memory from the host, so a whole training step can be captured in a CUDA graph
    (nerf_sys_b200.engine.GraphedTrainStep).  ``state[p]["step"]`` is refreshed from the device counters by
    ``state_dict()`` (the checkpoint layout is torch Adam's)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm: Optional[float] = None):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.max_grad_norm = max_grad_norm
        self._steps_dev = None
        self._index = {}
        self._shard = None

    def shard_grad_norm(self, group, shared_params) -> "FusedAdam":
        """Expert parallelism: this optimiser holds one rank's experts plus the replicated ``shared_params``.  The clip
        of runtime_adapt.py:265 is over the norm of ALL parameters, so the sum of squares of the sharded tensors is
        all-reduced over ``group`` and the replicated ones (identical on every rank after their gradient all-reduce)
        are counted once."""
        world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self._shard = (group, {id(p) for p in shared_params}, world)
        return self

    # ---- device-side bookkeeping -------------------------------------------------------------
    def _ensure(self, dev):
        if self._steps_dev is not None and self._steps_dev.device == dev:
            return
        plist = [p for g in self.param_groups for p in g["params"]]
        self._index = {id(p): i for i, p in enumerate(plist)}
        steps = [int(float(self.state[p]["step"])) if (p in self.state and "step" in self.state[p]) else 0 for p in plist]
        self._steps_dev = torch.tensor(steps, dtype=torch.int32, device=dev)
        self._acc = torch.zeros(1, dtype=torch.float32, device=dev)
        self._coef = torch.ones(1, dtype=torch.float32, device=dev)

    def sync_steps(self):
        """Copy the device step counters into state[p]['step'] (host sync; checkpoint time only)."""
        if self._steps_dev is None:
            return
        host = self._steps_dev.cpu().tolist()
        for g in self.param_groups:
            for p in g["params"]:
                if p in self.state and len(self.state[p]):
                    self.state[p]["step"] = torch.tensor(float(host[self._index[id(p)]]), dtype=torch.float32)

    def state_dict(self):
        self.sync_steps()
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._steps_dev = None  # re-seeded from the loaded state on the next step

    def _descs(self):
        descs = []
        for group in self.param_groups:
            if tuple(group["betas"]) != tuple(self.param_groups[0]["betas"]) or group["eps"] != self.param_groups[0]["eps"]:
                raise NotImplementedError("FusedAdam: betas / eps must be shared by all groups")
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if "exp_avg" not in st:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                    raise RuntimeError("FusedAdam: parameters must be contiguous CUDA fp32 tensors (no CPU path)")
                descs.append((p, g, st, float(group["lr"]), float(group["weight_decay"])))
        return descs

    @torch.no_grad()
    def step(self, closure=None, grad_scale: Optional[torch.Tensor] = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        descs = self._descs()
        if not descs:
            return loss
        dev = descs[0][0].device
        self._ensure(dev)
        L = lib()
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        b1, b2 = self.param_groups[0]["betas"]
        eps = float(self.param_groups[0]["eps"])
        # descriptor arrays are rebuilt only when the set of live tensors (or their storage) changes; otherwise just the
        # gradient pointers and learning rates are refreshed (eager steps are host-bound: ~64 tensors x 9 fields per step)
        key = tuple((id(p), p.data_ptr(), st["exp_avg"].data_ptr()) for p, _, st, _, _ in descs)
        if getattr(self, "_arr_key", None) != key:
            chunks = []
            for c0 in range(0, len(descs), ACN_ADAM_MAX_TENSORS):
                part = descs[c0:c0 + ACN_ADAM_MAX_TENSORS]
                arr = (AcnAdamTensor * len(part))()
                for i, (p, g, st, lr, wd) in enumerate(part):
                    arr[i].param = p.data_ptr()
                    arr[i].exp_avg, arr[i].exp_avg_sq = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                    arr[i].n, arr[i].weight_decay, arr[i].step_index = p.numel(), wd, self._index[id(p)]
                chunks.append((arr, len(part)))
            self._arr_key, self._arr_chunks = key, chunks
        chunks = self._arr_chunks
        for ci, (arr, n) in enumerate(chunks):
            base = ci * ACN_ADAM_MAX_TENSORS
            for i in range(n):
                d = descs[base + i]
                arr[i].grad, arr[i].lr, arr[i].weight_decay = d[1].data_ptr(), d[3], d[4]
        clip = self.max_grad_norm is not None
        with torch.cuda.device(dev):
            if clip:
                self._acc.zero_()
                with ops._timed("sumsq"):
                    for arr, n in chunks:
                        check(L.acn_sumsq_multi(arr, n, self._acc.data_ptr(), stream), "acn_sumsq_multi")
                ops._count(len(chunks) + 1)
                if self._shard is not None and self._shard[2] > 1:
                    sh = [d for d in descs if id(d[0]) in self._shard[1]]
                    rep = torch.zeros_like(self._acc)
                    if sh:
                        arr = (AcnAdamTensor * len(sh))()
                        for i, (p, g, st, lr, wd) in enumerate(sh):
                            arr[i].param, arr[i].grad, arr[i].n = p.data_ptr(), g.data_ptr(), p.numel()
                        check(L.acn_sumsq_multi(arr, len(sh), rep.data_ptr(), stream), "acn_sumsq_multi")
                        ops._count(1)
                    part = self._acc - rep
                    dist.all_reduce(part, group=self._shard[0])
                    self._acc.copy_(part + rep)
            with ops._timed("adam"):
                for ci, (arr, n) in enumerate(chunks):
                    check(L.acn_adam_multi(arr, n, float(b1), float(b2), eps, self._steps_dev.data_ptr(),
                                           self._acc.data_ptr() if clip else None, float(self.max_grad_norm or 0.0),
                                           grad_scale.data_ptr() if grad_scale is not None else None, self._coef.data_ptr(),
                                           1 if ci == 0 else 0, stream), "acn_adam_multi")
            ops._count(2 * len(chunks))
        return loss


def build_optimizer(model, lr_encoding=1e-2, lr_sigma=2e-3, lr_color=2e-3, lr_background=1e-3, max_grad_norm: Optional[float] = 1.0,
                    eps: float = 1e-8):
    """Per-group learning rates of the reference (common/args.py:115-125; common/utils.py:23-47)."""
    groups = model.get_param_groups()
    lrs = {"encoding": lr_encoding, "sigma": lr_sigma, "color": lr_color, "background": lr_background}
    pg = [{"params": g["params"], "lr": lrs[name]} for name, g in groups.items()]
    return FusedAdam(pg, betas=(0.9, 0.999), eps=eps, max_grad_norm=max_grad_norm)  # torch AdamW defaults, weight_decay 0
