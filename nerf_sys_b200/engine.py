"""CUDA-graph execution of a training step.

The reference's hot loops (pipelines/offline_stage/meta_core.py:123-141, runtime_stage/runtime_adapt.py:249-310)
run ``zero_grad -> loss -> backward -> clip -> optimizer.step`` as ~60 separate launches per step driven from
Python.  On B200 the kernels of one step add up to ~2 ms, so the step is launch-bound unless the launches are
replayed from a graph.  ``GraphedTrainStep`` captures that exact sequence (the same Python code, run once under
stream capture) and replays it; every kernel in it is one of this library's, launched on the capturing stream.

Requirements (checked, not worked around): fixed batch shapes, an optimiser whose step reads nothing back to
the host (``nerf_sys_b200.optim.FusedAdam``), and a loss whose routing needs no host round trip (a single
active expert per step: ``active_module`` / K = 1 - the meta-training and per-expert adaptation steps).
"""
from __future__ import annotations

from typing import Callable, Optional, Sequence

import torch

from . import ops


class GraphedTrainStep:
    def __init__(self, optimizer, loss_fn: Callable[..., torch.Tensor], example_inputs: Sequence[torch.Tensor], warmup: int = 3,
                 post_backward: Optional[Callable[[], None]] = None):
        """``loss_fn(*inputs)`` returns the scalar loss.  ``warmup`` REAL optimisation steps are run eagerly on
        ``example_inputs`` before capture (lazy state, allocator pools, kernel attributes)."""
        if not all(t.is_cuda for t in example_inputs):
            raise RuntimeError("GraphedTrainStep: inputs must be CUDA tensors (no CPU path)")
        self.opt, self.loss_fn, self.post = optimizer, loss_fn, post_backward
        self.static_in = [t.detach().clone() for t in example_inputs]
        if ops.kernel_timing_enabled():
            raise RuntimeError("GraphedTrainStep: turn ops.kernel_timing off before capture (events cannot be captured)")
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._body()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        n0 = ops.launches()
        with torch.cuda.graph(self.graph):
            self.static_loss = self._body()
        self.launches_per_step = ops.launches() - n0

    def _body(self) -> torch.Tensor:
        self.opt.zero_grad(set_to_none=True)
        loss = self.loss_fn(*self.static_in)
        loss.backward()
        if self.post is not None:
            self.post()
        self.opt.step()
        return loss.detach()

    def __call__(self, *inputs: torch.Tensor) -> torch.Tensor:
        """Copies the batch into the captured buffers (async, current stream), replays the step and returns the
        loss tensor of the captured graph (valid until the next call)."""
        if len(inputs) != len(self.static_in):
            raise ValueError(f"expected {len(self.static_in)} inputs, got {len(inputs)}")
        for s, t in zip(self.static_in, inputs):
            if t.shape != s.shape:
                raise ValueError(f"GraphedTrainStep: batch shape {tuple(t.shape)} differs from the captured {tuple(s.shape)}")
            s.copy_(t, non_blocking=True)
        self.graph.replay()
        ops._count(self.launches_per_step)
        return self.static_loss
