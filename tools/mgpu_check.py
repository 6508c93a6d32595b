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
