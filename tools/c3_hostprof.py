"""Host-side profile (cProfile, rank 0) of the expert-parallel config-3 step over peer memory.
torchrun --nproc-per-node 2 tools/c3_hostprof.py"""
import cProfile
import os
import pstats
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from nerf_sys_b200 import ops, parallel as PL  # noqa: E402
from nerf_sys_b200.nerfs.losses import compute_mse_loss  # noqa: E402
from nerf_sys_b200.optim import FusedAdam  # noqa: E402
from oracle import acnerf_oracle as O  # noqa: E402

rank, world, local = bench.dist_env()
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ops.set_mlp_math("f16tc")
model, aabb = bench.build_c5(dev)
model.train()
ex = PL.PeerExchange(rows_cap=2 * bench.N_RAYS * bench.S_TRAIN * world + 16 * 128)
model.set_expert_parallel(True, exchange=ex)
groups, shared = PL.owned_param_groups(model)
lrs = {"encoding": 1e-2, "sigma": 2e-3, "color": 2e-3, "background": 1e-3}
opt = FusedAdam([{"params": g["params"], "lr": lrs[n]} for n, g in groups.items()], max_grad_norm=1.0).shard_grad_norm(None, shared)
red = PL.SharedGradReducer(shared, average=False)
rays = [O.aerial_rays(bench.N_RAYS, aabb, seed=1000 + 16 * rank + b).to(dev) for b in range(4)]
gts = [torch.rand(bench.N_RAYS, 3, device=dev) for _ in range(4)]


def step(i):
    opt.zero_grad(set_to_none=True)
    loss = compute_mse_loss(bench.P, model, {"rays": rays[i % 4], "rgbs": gts[i % 4]})
    (loss / world).backward()
    red.wait()
    opt.step()


for i in range(5):
    step(i)
torch.cuda.synchronize()
dist.barrier()
n = 30
pr = cProfile.Profile()
t0 = time.perf_counter()
pr.enable()
for i in range(n):
    step(i)
pr.disable()
t_host = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
if rank == 0:
    print(f"host enqueue {t_host / n * 1e3:.3f} ms/step, with final sync {t_all / n * 1e3:.3f} ms/step")
    pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
dist.barrier()
red.remove()
model.set_expert_parallel(False)
ex.close()
dist.destroy_process_group()
