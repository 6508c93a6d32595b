"""Multi-GPU correctness check (run under torch.distributed.run with one rank per GPU, NCCL):
  1. expert-parallel container forward/backward (sample all-to-all) == the same container evaluated on one GPU;
  2. ray-sharded full-frame render == the single-GPU frame;
prints one line per check from rank 0 and exits non-zero on a mismatch."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = __file__.rsplit("/", 2)[0]
sys.path.insert(0, ROOT)
sys.path.insert(0, ROOT + "/tests")
from nerf_sys_b200 import ops, parallel as PL  # noqa: E402
from nerf_sys_b200.nerfs.ray_rendering import render_image, render_rays  # noqa: E402
from nerf_sys_b200.nerfs.scene_box import SceneBox  # noqa: E402
from oracle import acnerf_oracle as O  # noqa: E402
from test_gpu_parity import _big_scene  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
ex = PL.PeerExchange(rows_cap=2 * 1024 * 32 * world + 8 * 128)
for mode, path in (("fp32", "nccl"), ("f16tc", "nccl"), ("fp32", "p2p"), ("f16tc", "p2p")):
    ops.set_mlp_math(mode)
    m, st, cfg, mins, aabb = _big_scene(8, 2, 4, 14)
    m = m.to(dev).train()
    N, S = 1024, 32
    rays = O.aerial_rays(N, aabb, seed=50 + rank).to(dev)
    g = torch.Generator().manual_seed(7 + rank)
    u, cot = torch.rand(N, S, generator=g).to(dev), torch.randn(N * S, 4, generator=g).to(dev)
    t = ops.stratified_t(rays, S, u)
    pts = ops.PointSet(rays=rays, t=t)
    # reference: the whole container on this GPU, this rank's rays
    m.zero_grad(set_to_none=True)
    y_ref = m.evaluate(pts)
    (y_ref * cot).sum().backward()
    g_ref = {n: (p.grad.clone() if p.grad is not None else torch.zeros_like(p)) for n, p in m.named_parameters() if "bg_mlp" not in n}
    for v in g_ref.values():  # an expert's gradient is the sum of every rank's contribution
        dist.all_reduce(v)
    m.zero_grad(set_to_none=True)
    y = PL.container_evaluate_ep(m, pts) if path == "nccl" else PL.container_evaluate_p2p(m, pts, ex)
    (y * cot).sum().backward()
    tol = 1e-5 if mode == "fp32" else 2e-3
    err = (y - y_ref).abs().max().item()
    owner = PL.expert_owner(8, world)
    gerr = 0.0
    for n, p in m.named_parameters():
        if not n.startswith("submodules."):
            continue
        e = int(n.split(".")[1])
        if owner[e] == rank:
            ref = g_ref[n]
            got = p.grad if p.grad is not None else torch.zeros_like(p)
            gerr = max(gerr, ((got - ref).norm() / ref.norm().clamp_min(1e-30)).item())
        else:
            assert p.grad is None or p.grad.abs().max().item() == 0, f"rank {rank} touched expert {e}"
    stats = torch.tensor([err, gerr], device=dev)
    dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    gtol = 1e-4 if mode == "fp32" else 3e-2
    good = stats[0].item() <= tol and stats[1].item() <= gtol
    ok = ok and good
    if rank == 0:
        print(f"[{mode}] expert-parallel a2a ({path}), world {world}: out max-abs err {stats[0].item():.2e} (tol {tol}), "
              f"owner-gradient rel-L2 {stats[1].item():.2e} (tol {gtol}) -> {'ok' if good else 'MISMATCH'}", flush=True)

# one expert-parallel optimisation step == the same step on one GPU over the concatenated batch: loss, the global clip
# coefficient (sharded sum of squares all-reduced, replicated head counted once) and the owner's updated parameters
from nerf_sys_b200.optim import FusedAdam, build_optimizer  # noqa: E402
ops.set_mlp_math("fp32")
N, S = 512, 32
rays = O.aerial_rays(N, aabb, seed=70 + rank).to(dev)
g = torch.Generator().manual_seed(17 + rank)
u, gt = torch.rand(N, S, generator=g).to(dev), torch.rand(N, 3, generator=g).to(dev)
all_rays, all_u, all_gt = [torch.empty(world * N, *x.shape[1:], device=dev) for x in (rays, u, gt)]
for dst, src in ((all_rays, rays), (all_u, u), (all_gt, gt)):
    dist.all_gather_into_tensor(dst, src.contiguous())
ref_m = _big_scene(8, 2, 4, 14)[0].to(dev).train()
ref_opt = build_optimizer(ref_m, max_grad_norm=0.05)  # small enough that the clip is active
rgb, *_ = render_rays(ref_m, all_rays, ray_samples=S, u=all_u)
ref_loss = ops.mse_loss(rgb, all_gt)
ref_loss.backward()
ref_opt.step()
ep_m = _big_scene(8, 2, 4, 14)[0].to(dev).train()
ep_m.set_expert_parallel(True, exchange=ex)
groups, shared = PL.owned_param_groups(ep_m)
lrs = {"encoding": 1e-2, "sigma": 2e-3, "color": 2e-3, "background": 1e-3}
ep_opt = FusedAdam([{"params": gr["params"], "lr": lrs[n]} for n, gr in groups.items()], max_grad_norm=0.05).shard_grad_norm(None, shared)
rgb, *_ = render_rays(ep_m, rays, ray_samples=S, u=u)
loss = ops.mse_loss(rgb, gt)
(loss / world).backward()
PL.allreduce_shared_grads(shared, average=False, world=world)
ep_opt.step()
mean_loss = loss.detach().clone()
dist.all_reduce(mean_loss)
mean_loss /= world
owner = PL.expert_owner(8, world)
ref_p = dict(ref_m.named_parameters())
perr = 0.0
for n, p in ep_m.named_parameters():
    mine = (n.startswith("submodules.") and owner[int(n.split(".")[1])] == rank) or n.startswith("bg_mlp")
    if mine:
        perr = max(perr, (p.detach() - ref_p[n].detach()).abs().max().item())
stats = torch.tensor([abs(mean_loss.item() - ref_loss.item()), abs(ep_opt.clip_coef.item() - ref_opt.clip_coef.item()) / ref_opt.clip_coef.item(), perr],
                     device=dev)
dist.all_reduce(stats, op=dist.ReduceOp.MAX)
good = stats[0].item() <= 1e-6 and stats[1].item() <= 1e-4 and ref_opt.clip_coef.item() < 0.999 and stats[2].item() <= 2e-4
ok = ok and good
if rank == 0:
    print(f"expert-parallel optimisation step, world {world}: loss diff {stats[0].item():.1e}, clip coefficient {ep_opt.clip_coef.item():.6f} vs "
          f"{ref_opt.clip_coef.item():.6f} (rel {stats[1].item():.1e}), owner parameters max-abs diff {stats[2].item():.1e} -> {'ok' if good else 'MISMATCH'}",
          flush=True)
ep_m.set_expert_parallel(False)

# early-posting shared-gradient reducer (config 2 at N ranks): hooks post inside backward(), wait() reduces; vs dist.all_reduce
qs = [torch.nn.Parameter(torch.randn(32, 16, device=dev)), torch.nn.Parameter(torch.randn(3, device=dev))]
red = PL.SlabGradReducer(ex, qs, scale=1.0 / world, early=True)
worst = 0.0
for it in range(3):
    for q in qs:
        q.grad = None
    ((qs[0] ** 2).sum() * (rank + 1 + it) + (qs[1] * torch.arange(3.0, device=dev)).sum() * (rank + 2)).backward()
    loc = torch.cat([q.grad.reshape(-1) for q in qs])
    allg = [torch.zeros_like(loc) for _ in range(world)]
    dist.all_gather(allg, loc)
    terms = torch.stack(allg).double()
    exact, mag = terms.sum(0) / world, terms.abs().sum(0) / world  # the ranks' terms have mixed signs: errors scale with sum |term|
    red.wait()
    mine = torch.cat([q.grad.reshape(-1) for q in qs]).double()
    worst = max(worst, ((mine - exact).abs() / mag.clamp_min(1e-30)).max().item())
red.remove()
flag = torch.tensor([worst], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MAX)
good = flag.item() <= 1e-6  # fp32 sum of `world` terms in rank order against the fp64 sum, relative to sum |term|
ok = ok and good
if rank == 0:
    print(f"early-posting slab reducer (arrive in backward, wait before the optimiser), world {world}: max error vs the fp64 sum, relative to sum |term| "
          f"{flag.item():.1e} -> {'ok' if good else 'MISMATCH'}", flush=True)

# ray-sharded render
ops.set_mlp_math("f16tc")
m, st, cfg, mins, aabb = _big_scene(8, 2, 4, 14)
m = m.to(dev).eval()
H, W = 270, 480
c2w = torch.tensor([[0.0, 0.0, -1.0, 0.2], [0.0, 1.0, 0.0, 0.0], [1.0, 0.0, 0.0, 0.0]])
kw = dict(H=H, W=W, fx=300.0, fy=300.0, cx=W / 2, cy=H / 2, c2w=c2w, scene_box=SceneBox(aabb.to(dev)), ray_samples=32)
full = render_image(m, **kw)
sharded = PL.render_image_sharded(lambda r0, r1: render_image(m, row_range=(r0, r1), **kw), H, W)
e = max((a.reshape(-1) - b.reshape(-1)).abs().max().item() for a, b in zip(full, sharded))
good = e <= 1e-6
ok = ok and good
if rank == 0:
    print(f"ray-sharded render {W}x{H}, world {world}: max-abs err vs single-GPU frame {e:.2e} -> {'ok' if good else 'MISMATCH'}", flush=True)
ex.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
