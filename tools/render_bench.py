"""1080p frame time of the 8-expert container for every levels-per-thread setting of the gather kernel, and the per-kernel
breakdown of one frame (GPU)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nerf_sys_b200 import ops, synthetic as SY  # noqa: E402
from nerf_sys_b200.nerfs.ray_rendering import render_image  # noqa: E402
from nerf_sys_b200.nerfs.scene_box import SceneBox  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
model, aabb = bench.build_container(dev)
model.eval()
H, W, f = 1080, 1920, 1480.0
sb = SceneBox(aabb.to(dev))
poses = [SY.ring_pose(i, 20, radius=0.9) for i in range(12)]


def run(tag, **kw):
    def frame(i):
        return render_image(model, H=H, W=W, fx=f, fy=f, cx=W / 2, cy=H / 2, c2w=poses[i], scene_box=sb, ray_samples=64, **kw)
    for i in range(2):
        frame(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 6
    for i in range(n):
        out = frame(2 + i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{tag:40s} {ms:7.2f} ms/frame  {H * W / ms / 1e3:7.1f} M rays/s", flush=True)
    return out


prev = ops.set_fused_inference(False)
two = run("two kernels (gather, then MLP)")
ops.set_fused_inference(True)
one = run("fused gather + MLP kernel")
print("fused vs two-kernel: max |diff| rgb/depth/acc:", [float((x - y).abs().max()) for x, y in zip(two, one)])
for eps in (1e-2, 1e-3):
    cut = run(f"fused kernel + early termination, eps = {eps}", early_stop_eps=eps)
    print(f"   max |rgb diff| vs full evaluation: {float((cut[0] - one[0]).abs().max()):.2e}; mean acc {float(one[2].mean()):.3f}")
timing = ops.kernel_timing(True)
render_image(model, H=H, W=W, fx=f, fy=f, cx=W / 2, cy=H / 2, c2w=poses[0], scene_box=sb, ray_samples=64)
torch.cuda.synchronize()
ops.kernel_timing(False)
for k, evs in timing.items():
    print(f"  {k:16s} {sum(a.elapsed_time(b) for a, b in evs):8.2f} ms / frame ({len(evs)} calls)")
