"""Does the order of the samples matter to the hash-grid gather?  One config-5 image block (136 rows x 1920 pixels x 64
samples) through the stand-alone encoder, level by level, with the points listed (a) ray-major (consecutive samples of a
ray in adjacent lanes) and (b) sample-major (the same sample of adjacent pixels in adjacent lanes)."""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from nerf_sys_b200 import ops  # noqa: E402
from oracle import acnerf_oracle as O  # noqa: E402

dev = torch.device("cuda", 0)
aabb, cent, boxes = O.grid_scene(2, 4)
H, W, S = 136, 1920, 64
rays = ops.raygen(1080, W, 1480.0, 1480.0, W / 2, 540.0, True, O.ring_pose(1, 20, radius=0.9), aabb.to(dev), dev, row0=400, rows=H)
t = ops.stratified_t(rays, S, None)
x6 = ops.sample_points(rays, t)                      # (N*S, 6) ray-major
lo, hi = aabb[0].to(dev), aabb[1].to(dev)
x01 = ((x6[:, :3] - lo) / (hi - lo)).clamp(1e-6, 1 - 1e-6).contiguous()
N = rays.shape[0]
orders = {"ray-major": x01, "sample-major": x01.view(N, S, 3).transpose(0, 1).reshape(-1, 3).contiguous()}
res = [16, 23, 33, 48, 70, 101, 147, 212, 307, 445, 645, 933, 1351, 1955, 2830, 4095]
ev = lambda: torch.cuda.Event(enable_timing=True)
tot = {k: 0.0 for k in orders}
for r in res:
    gd = ops.grid_desc(1, 20, "linear", [r])
    table = torch.randn(1 << 20, 2, device=dev) * 0.1
    line = f"res {r:5d}:"
    for name, x in orders.items():
        tf = 0.0
        with torch.no_grad():
            for it in range(4):
                a, b = ev(), ev()
                a.record()
                ops.HashGridFn.apply(x, table, gd)
                b.record()
                torch.cuda.synchronize()
                if it >= 1:
                    tf += a.elapsed_time(b) / 3
        tot[name] += tf
        line += f"  {name} {tf * 1e3:8.1f} us"
    print(line)
print("total: " + "  ".join(f"{k} {v:.3f} ms" for k, v in tot.items()))
