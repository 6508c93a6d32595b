"""Fused optimiser step: torch.optim.Adam semantics and state_dict layout
(``step``, ``exp_avg``, ``exp_avg_sq`` per parameter; the reference builds torch Adam/AdamW with
per-group learning rates, common/utils.py:16-76) with the arithmetic in csrc/optim.cu, plus
clip_grad_norm_ (meta_core.py:181-190, runtime_adapt.py:261-268) and GradScaler's unscale / skip
folded into the same pass: the gradient multiplier stays on the device and is applied inside the
Adam kernel, so the 128 MiB hash-table gradients are read once for the norm and once for the
update, and there is no host synchronisation."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import ops
from ._lib import ACN_ADAM_MAX_TENSORS, ACN_ADAM_ZERO_GRAD, AcnAdamTensor, check, lib

PERSISTENT_MIN_NUMEL = 1 << 20  # tensors at least this large (the hash tables) get a persistent gradient buffer


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam over the C ABI's multi-tensor kernels: per step one sum-of-squares launch (clip_grad_norm_),
    one prep launch (device step counters + gradient multiplier) and one Adam launch per 64 tensors.  Nothing in
    ``step()`` reads device memory from the host, so a whole training step can be captured in a CUDA graph
    (nerf_sys_b200.engine.GraphedTrainStep).  ``state[p]["step"]`` is refreshed from the device counters by
    ``state_dict()`` (the checkpoint layout is torch Adam's).

    ``persistent_grads``: every hash table gets ONE gradient buffer that the Adam kernel zeroes while it reads it
    (ACN_ADAM_ZERO_GRAD); the backward scatters into it instead of allocating and zero-filling 128 MiB per table per step.
    ``bind_live_flags``: device-side "this expert received a sample" flags (MetaContainer.live_flags); a tensor whose
    flag is 0 is skipped exactly as torch.optim.Adam skips a parameter whose ``.grad`` is None - without the host
    knowing, which is what lets the soft-routed container step live in a CUDA graph.
    AMP: ``scaler.step(opt)`` hands the loss scale and the found-inf flag over as tensors
    (``_step_supports_amp_scaling``); the gradients are unscaled inside the Adam kernel, the clip acts on the unscaled
    norm, and a non-finite step is skipped on the device (runtime_adapt.py:261-268)."""

    _step_supports_amp_scaling = True

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm: Optional[float] = None,
                 persistent_grads: bool = False):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.max_grad_norm = max_grad_norm
        self._steps_dev = None
        self._index = {}
        self._shard = None
        self._live = None       # (flags tensor int32, {id(param): slot})
        self._arena: Dict[int, tuple] = {}  # id(param) -> (param, buffer)
        if persistent_grads:
            self.enable_persistent_grads()

    # ---- persistent table gradients ---------------------------------------------------------------
    def enable_persistent_grads(self, min_numel: int = PERSISTENT_MIN_NUMEL) -> "FusedAdam":
        for g in self.param_groups:
            for p in g["params"]:
                if p.is_cuda and p.requires_grad and p.numel() >= min_numel and id(p) not in self._arena:
                    buf = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    self._arena[id(p)] = (p, buf)
                    ops.register_grad_buffer(p, buf)
        return self

    def release_persistent_grads(self) -> None:
        for p, _ in self._arena.values():
            ops.unregister_grad_buffer(p)
        self._arena = {}

    def zero_grad(self, set_to_none: bool = True) -> None:
        # a buffer handed to a backward whose step never ran (skipped by the caller) still holds that gradient
        for p, _ in self._arena.values():
            ops.release_grad_buffer(p, clean=False)
        super().zero_grad(set_to_none=set_to_none)

    # ---- expert liveness ------------------------------------------------------------------------
    def bind_live_flags(self, flags: torch.Tensor, slot_of: Dict[int, int]) -> "FusedAdam":
        """``flags`` (int32 on the device), ``slot_of[id(param)]`` = index of the flag that gates the parameter."""
        if flags.dtype != torch.int32 or not flags.is_cuda:
            raise TypeError("live flags must be a CUDA int32 tensor")
        if any(not (0 <= s < min(255, flags.numel())) for s in slot_of.values()):
            raise ValueError("liveness slot outside the flag tensor (at most 255 slots)")
        self._live = (flags, dict(slot_of))
        self._arr_key = None
        return self

    def shard_grad_norm(self, group, shared_params, exchange=None) -> "FusedAdam":
        """Expert parallelism: this optimiser holds one rank's experts plus the replicated ``shared_params``.  The clip
        of runtime_adapt.py:265 is over the norm of ALL parameters, so the sum of squares of the sharded tensors is
        all-reduced over ``group`` and the replicated ones (identical on every rank after their gradient all-reduce)
        are counted once.  With ``exchange`` (parallel.PeerExchange) the one-float reduction rides on a slab barrier
        instead of an NCCL all-reduce (graph-capturable, no NCCL kernel on the step)."""
        if exchange is not None:
            world = exchange.world
        else:
            world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self._shard = (group, {id(p) for p in shared_params}, world, exchange)
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
        self._rep = torch.zeros(1, dtype=torch.float32, device=dev)
        self._coef = torch.tensor([1.0, 0.0], dtype=torch.float32, device=dev)  # [gradient multiplier, step skipped]

    @property
    def clip_coef(self) -> torch.Tensor:
        """Device scalar: the multiplier applied to the gradients in the last step (clip coefficient / loss scale)."""
        return self._coef[0]

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
                if g.dtype != torch.float32:
                    raise RuntimeError("FusedAdam: gradients must be fp32 (the sm_100a ops return fp32 gradients under autocast too)")
                descs.append((p, g, st, float(group["lr"]), float(group["weight_decay"])))
        return descs

    def _flags(self, p, g) -> int:
        f = 0
        if self._live is not None:
            slot = self._live[1].get(id(p))
            if slot is not None:
                f |= slot + 1
        ent = self._arena.get(id(p))
        if ent is not None and g.data_ptr() == ent[1].data_ptr():
            f |= ACN_ADAM_ZERO_GRAD
        return f

    @torch.no_grad()
    def step(self, closure=None, inv_scale: Optional[torch.Tensor] = None):
        """``inv_scale`` (device float tensor): multiplies every gradient first (1 / loss scale).  Under
        ``GradScaler.step(opt)`` the scale and the found-inf flag arrive as ``self.grad_scale`` / ``self.found_inf``."""
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
        amp_scale = getattr(self, "grad_scale", None)   # torch's convention: gradients are DIVIDED by it
        found_inf = getattr(self, "found_inf", None)
        if amp_scale is not None:
            inv = amp_scale.to(dev, torch.float32).reciprocal()
            inv_scale = inv if inv_scale is None else inv_scale * inv
        if found_inf is not None:
            found_inf = found_inf.to(dev, torch.float32).reshape(-1)
        if inv_scale is not None:
            inv_scale = inv_scale.to(dev, torch.float32).reshape(-1)
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
                arr[i].grad, arr[i].lr, arr[i].weight_decay, arr[i].flags = d[1].data_ptr(), d[3], d[4], self._flags(d[0], d[1])
        live = self._live[0].data_ptr() if self._live is not None else None
        clip = self.max_grad_norm is not None
        with torch.cuda.device(dev):
            if clip:
                self._acc.zero_()
                with ops._timed("sumsq"):
                    for arr, n in chunks:
                        check(L.acn_sumsq_multi(arr, n, live, self._acc.data_ptr(), stream), "acn_sumsq_multi")
                ops._count(len(chunks) + 1)
                if self._shard is not None and self._shard[2] > 1:
                    sh = [d for d in descs if id(d[0]) in self._shard[1]]
                    self._rep.zero_()
                    if sh:
                        arr = (AcnAdamTensor * len(sh))()
                        for i, (p, g, st, lr, wd) in enumerate(sh):
                            arr[i].param, arr[i].grad, arr[i].n = p.data_ptr(), g.data_ptr(), p.numel()
                        check(L.acn_sumsq_multi(arr, len(sh), None, self._rep.data_ptr(), stream), "acn_sumsq_multi")
                        ops._count(1)
                    part = self._acc - self._rep
                    if self._shard[3] is not None:
                        part = self._shard[3].allreduce_small(part)  # one float on a slab barrier (fixed summation order)
                    else:
                        dist.all_reduce(part, group=self._shard[0])
                    self._acc.copy_(part + self._rep)
            with ops._timed("adam"):
                for ci, (arr, n) in enumerate(chunks):
                    check(L.acn_adam_multi(arr, n, float(b1), float(b2), eps, self._steps_dev.data_ptr(),
                                           self._acc.data_ptr() if clip else None, float(self.max_grad_norm or 0.0),
                                           inv_scale.data_ptr() if inv_scale is not None else None,
                                           found_inf.data_ptr() if found_inf is not None else None, live, self._coef.data_ptr(),
                                           1 if ci == 0 else 0, stream), "acn_adam_multi")
            ops._count(2 * len(chunks))
        for p, g, *_ in descs:  # the Adam kernel left the persistent buffers zeroed
            ent = self._arena.get(id(p))
            if ent is not None:
                ops.release_grad_buffer(p, clean=g.data_ptr() == ent[1].data_ptr())
        return loss


def build_optimizer(model, lr_encoding=1e-2, lr_sigma=2e-3, lr_color=2e-3, lr_background=1e-3, max_grad_norm: Optional[float] = 1.0,
                    eps: float = 1e-8, persistent_grads: bool = False):
    """Per-group learning rates of the reference (common/args.py:115-125; common/utils.py:23-47)."""
    groups = model.get_param_groups()
    lrs = {"encoding": lr_encoding, "sigma": lr_sigma, "color": lr_color, "background": lr_background}
    pg = [{"params": g["params"], "lr": lrs[name]} for name, g in groups.items()]
    # torch AdamW defaults, weight_decay 0
    return FusedAdam(pg, betas=(0.9, 0.999), eps=eps, max_grad_norm=max_grad_norm, persistent_grads=persistent_grads)


def bind_expert_liveness(opt: FusedAdam, model, experts=None) -> FusedAdam:
    """Gate every expert's tensors in ``opt`` on the container's device-side live flags (``model.live_flags``, written by
    the deferred routing of `MetaContainer.evaluate` / `parallel.container_evaluate_p2p`)."""
    flags, slots = model.live_flag_slots(experts)
    slot_of = {}
    for k, slot in slots.items():
        for p in model.submodules[k].parameters():
            slot_of[id(p)] = slot
    return opt.bind_live_flags(flags, slot_of)
