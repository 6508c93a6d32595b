"""route_pack at the config-5 chunk size (136 rows x 1920 pixels x 64 samples, 8 experts): total device time per call, and
(under `ncu --metrics gpu__time_duration.sum`) the per-kernel split.  python tools/route_bench.py [reps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_sys_b200 import ops  # noqa: E402
from oracle import acnerf_oracle as O  # noqa: E402  (scene constants only)

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = torch.device("cuda", 0)
aabb, cent, boxes = O.grid_scene(2, 4)
H, W, S = 136, 1920, 64
rays = ops.raygen(1080, W, 1480.0, 1480.0, W / 2, 540.0, True, O.ring_pose(1, 20, radius=0.9), aabb.to(dev), dev, row0=400, rows=H)
t = ops.stratified_t(rays, S, None)
pts = ops.PointSet(rays=rays, t=t)
ch = cent.reshape(-1).tolist()
pairs = ops.route_pack(pts, ch, 8, 1.05, True)
print(f"P = {pts.P}, pairs = {pairs.seg_host}")
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    ops.route_pack(pts, ch, 8, 1.05, True)
e1.record()
torch.cuda.synchronize()
print(f"route_pack {e0.elapsed_time(e1) / reps:.3f} ms per call (incl. the host read of seg)")
