"""One config-5 image block (136 rows x 1920 px x 64 samples, 8 experts) rendered with ray-major and with sample-major point ids;
run under ncu to compare the gather kernel's sector counts (profiles/r1_v6_hash_fwd_order.csv)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from nerf_sys_b200 import ops  # noqa: E402
from nerf_sys_b200.nerfs.ray_rendering import render_rays  # noqa: E402
from nerf_sys_b200 import synthetic as SY  # noqa: E402

dev = torch.device("cuda", 0)
ops.set_mlp_math("f16tc")
model, aabb = bench.build_container(dev)
model.eval()
rays = ops.raygen(1080, 1920, 1480.0, 1480.0, 960.0, 540.0, True, SY.ring_pose(1, 20, radius=0.9), aabb.to(dev), dev, row0=400, rows=136)
with torch.no_grad():
    for coherent in (False, True, False, True):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rgb, *_ = render_rays(model, rays, ray_samples=64, coherent=coherent)
        e1.record()
        torch.cuda.synchronize()
        print(f"coherent={coherent}: {e0.elapsed_time(e1):.3f} ms, mean rgb {rgb.mean().item():.6f}")
