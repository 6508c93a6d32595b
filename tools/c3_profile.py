"""Per-kernel device time of the config-3 training step (8-expert container, one GPU): the same step bench.py's train_c3
times, driven with a CUDA-event pair around every C-ABI call.  python tools/c3_profile.py [steps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from nerf_sys_b200 import ops, parallel as PL  # noqa: E402
from nerf_sys_b200.nerfs.losses import compute_mse_loss  # noqa: E402
from nerf_sys_b200.optim import FusedAdam  # noqa: E402
from oracle import acnerf_oracle as O  # noqa: E402  (synthetic rays only)

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
ops.set_mlp_math("f16tc")
model, aabb = bench.build_c5(dev)
model.train()
groups, shared = PL.owned_param_groups(model)
lrs = {"encoding": 1e-2, "sigma": 2e-3, "color": 2e-3, "background": 1e-3}
opt = FusedAdam([{"params": g["params"], "lr": lrs[n]} for n, g in groups.items()], max_grad_norm=1.0)
rays = [O.aerial_rays(bench.N_RAYS, aabb, seed=1000 + b).to(dev) for b in range(4)]
gts = [torch.rand(bench.N_RAYS, 3, device=dev) for _ in range(4)]


def step(i):
    opt.zero_grad(set_to_none=True)
    loss = compute_mse_loss(bench.P, model, {"rays": rays[i % 4], "rgbs": gts[i % 4]})
    loss.backward()
    opt.step()
    return loss


for i in range(3):
    step(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(steps):
    step(i)
e1.record()
torch.cuda.synchronize()
print(f"step {e0.elapsed_time(e1) / steps:.3f} ms")
timing = ops.kernel_timing(True)
for i in range(steps):
    step(i)
torch.cuda.synchronize()
ops.kernel_timing(False)
tot = 0.0
for name, evs in timing.items():
    t = sum(a.elapsed_time(b) for a, b in evs) / steps
    tot += t
    print(f"  {name:18s} {t * 1e3:9.1f} us/step  ({len(evs) / steps:.0f} calls)")
print(f"  sum of timed calls {tot:.3f} ms")

# ---- config-5 frame (1920x1080 x 64, eval) through the same container --------------------------------------------
from nerf_sys_b200.nerfs.ray_rendering import render_image  # noqa: E402
from nerf_sys_b200.nerfs.scene_box import SceneBox  # noqa: E402
del opt
model.zero_grad(set_to_none=True)
model.eval()
torch.cuda.empty_cache()
sb = SceneBox(aabb.to(dev))
kw = dict(H=1080, W=1920, fx=1480.0, fy=1480.0, cx=960.0, cy=540.0, scene_box=sb, ray_samples=64)
render_image(model, c2w=O.ring_pose(0, 20, radius=0.9), **kw)
torch.cuda.synchronize()
e0.record()
for i in range(3):
    render_image(model, c2w=O.ring_pose(1 + i, 20, radius=0.9), **kw)
e1.record()
torch.cuda.synchronize()
print(f"frame {e0.elapsed_time(e1) / 3:.2f} ms")
timing = ops.kernel_timing(True)
for i in range(3):
    render_image(model, c2w=O.ring_pose(1 + i, 20, radius=0.9), **kw)
torch.cuda.synchronize()
ops.kernel_timing(False)
tot = 0.0
for name, evs in timing.items():
    t = sum(a.elapsed_time(b) for a, b in evs) / 3
    tot += t
    print(f"  {name:18s} {t:9.3f} ms/frame  ({len(evs) / 3:.0f} calls)")
print(f"  sum of timed calls {tot:.3f} ms")
