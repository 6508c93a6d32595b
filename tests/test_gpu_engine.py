"""GPU tests of the step machinery around the kernels: multi-tensor Adam + clip against torch.optim.Adam +
clip_grad_norm_ (common/utils.py:16-76, meta_core.py:181-190 in the reference), and CUDA-graph replay of a
training step against the same step driven eagerly."""
import copy

import pytest
import torch

from oracle import acnerf_oracle as O
from test_gpu_parity import _big_scene, close

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_fused_adam_matches_torch_adam_with_clipping():
    from nerf_sys_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(3)
    shapes = [(64, 32), (64,), (1, 64), (15, 64), (3,), (4099, 2), (1,), (70001,)]  # odd sizes: scalar tails, unaligned counts
    ps = [torch.randn(*s, generator=g) for s in shapes]
    mine = [p.clone().to(DEV).requires_grad_(True) for p in ps]
    ref = [p.clone().double().requires_grad_(True) for p in ps]
    groups = lambda xs: [{"params": xs[:4], "lr": 1e-2}, {"params": xs[4:], "lr": 2e-3}]
    opt = FusedAdam(groups(mine), betas=(0.9, 0.999), eps=1e-8, max_grad_norm=1.0)
    opt_ref = torch.optim.Adam(groups(ref), betas=(0.9, 0.999), eps=1e-8)
    for step in range(12):
        grads = [torch.randn(*s, generator=g) * (10.0 if step % 3 == 0 else 0.01) for s in shapes]
        skip = 6 if step % 4 == 1 else -1  # a parameter without gradient keeps its state and step count (torch semantics)
        for i, (a, b, gr) in enumerate(zip(mine, ref, grads)):
            a.grad = None if i == skip else gr.to(DEV)
            b.grad = None if i == skip else gr.double()
        torch.nn.utils.clip_grad_norm_([b for b in ref if b.grad is not None], 1.0)
        opt_ref.step()
        opt.step()
    for a, b, s in zip(mine, ref, shapes):
        close(a, b.float(), 2e-6, f"param {s}")
    sd = opt.state_dict()
    steps = [int(v["step"]) for v in sd["state"].values()]
    assert steps == [12, 12, 12, 12, 12, 12, 9, 12], steps
    assert sorted(sd["state"][0]) == ["exp_avg", "exp_avg_sq", "step"]  # torch Adam's checkpoint layout
    # reload: step counters are re-seeded from the checkpoint
    opt2 = FusedAdam(groups(mine), betas=(0.9, 0.999), eps=1e-8, max_grad_norm=1.0)
    opt2.load_state_dict(copy.deepcopy(sd))
    for a in mine:
        a.grad = torch.ones_like(a)
    opt2.step()
    assert [int(v["step"]) for v in opt2.state_dict()["state"].values()] == [13] * 6 + [10, 13]


def _setup(N, S, n_batches):
    from nerf_sys_b200 import ops
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    from nerf_sys_b200.optim import build_optimizer
    m, st, cfg, mins, aabb = _big_scene(1, 1, 1, 14, seed=0)
    m.train()
    opt = build_optimizer(m, max_grad_norm=1.0)
    gen = torch.Generator().manual_seed(9)
    batches = [(O.aerial_rays(N, aabb, seed=20 + i).to(DEV), torch.rand(N, 3, generator=gen).to(DEV), torch.rand(N, S, generator=gen).to(DEV))
               for i in range(n_batches)]

    def loss_fn(rays, gt, u):
        rgb, *_ = render_rays(m, rays, ray_samples=S, u=u)
        return ops.mse_loss(rgb, gt)

    return m, opt, batches, loss_fn


def test_graph_replay_equals_eager_training():
    """Same seed, same batches (jitter injected): N steps replayed from a CUDA graph leave the parameters where
    N eager steps do.  Both arms use the tensor-core MLPs; only the hash-table scatter order differs run to run."""
    from nerf_sys_b200.engine import GraphedTrainStep
    N, S, steps = 1024, 32, 8
    m, opt, batches, loss_fn = _setup(N, S, steps)
    ref_losses = []
    for b in batches:
        opt.zero_grad(set_to_none=True)
        l = loss_fn(*b)
        l.backward()
        opt.step()
        ref_losses.append(l.item())
    ref_params = {n: p.detach().clone() for n, p in m.named_parameters()}

    m, opt, batches, loss_fn = _setup(N, S, steps)
    snap = {n: p.detach().clone() for n, p in m.named_parameters()}
    step = GraphedTrainStep(opt, loss_fn, batches[0], warmup=1)
    # capture ran real optimisation steps on batches[0]: rewind parameters, moments and step counters
    with torch.no_grad():
        for n, p in m.named_parameters():
            p.copy_(snap[n])
    for stt in opt.state.values():
        stt["exp_avg"].zero_()
        stt["exp_avg_sq"].zero_()
    opt._steps_dev.zero_()
    losses = [step(*b).item() for b in batches]
    for a, b in zip(losses, ref_losses):
        assert abs(a - b) <= 1e-4 * max(1e-3, abs(b)), (losses, ref_losses)
    for n, p in m.named_parameters():
        # the table gradient is an atomic scatter (summation order differs run to run) and Adam divides by sqrt(v) + eps
        close(p, ref_params[n], 1e-3 if n.endswith("hash_table") else 1e-4, n)
    assert step.launches_per_step >= 10
    assert [int(v["step"]) for v in opt.state_dict()["state"].values()] == [steps] * len(opt.state)
