"""Occupancy-guided renderer vs the stratified renderer on the 8-expert bench scene (GPU): rays/s, samples per ray, and the
per-kernel time of one occupancy frame.  The grids are built from the experts' own densities (3 EMA updates over all cells)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nerf_sys_b200 import ops, synthetic as SY  # noqa: E402
from nerf_sys_b200.models.inr import MetaContainer  # noqa: E402
from nerf_sys_b200.nerfs.ray_rendering import render_rays  # noqa: E402
from nerf_sys_b200.nerfs.scene_box import SceneBox  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
aabb, cent, boxes = SY.grid_scene(2, 4)
cfg = dict(SY.HASH_CFG)
cfg["log2_hashmap_size"] = 20
st = SY.init_ngp_state(8, aabb, cent, boxes, cfg, seed=3, table_std=0.1)
occ = {"use_occ": True, "resolution": 128, "levels": 4, "warmup_steps": 3, "update_interval": 1}
model = MetaContainer(8, cent, aabb, "instant", boundary_margin=1.05, cluster_2d=True, use_bg_nerf=True, occ_conf=occ,
                      expert_box_list=[SceneBox(b) for b in boxes], hidden=64, sigma_depth=2, color_depth=2, color_hidden=64,
                      dir_encoding="spherical", use_sigmoid_rgb=True, hash_enc_conf=cfg)
model.load_state_dict(st, strict=False)
model = model.to(dev).train()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for step in range(4):
    model.maybe_update_expert_occupancies(step)
e1.record()
torch.cuda.synchronize()
print(f"4 grid updates (8 experts x 4 levels x 128^3 cells each): {e0.elapsed_time(e1):.1f} ms; occupied fraction per expert (level 0): "
      f"{[round(float(s.occ_grid.binaries[0].float().mean()), 3) for s in model.submodules]}", flush=True)
model.eval()
H, W, f = 1080, 1920, 1480.0
rays = ops.raygen(H, W, f, f, W / 2, H / 2, True, SY.ring_pose(2, 20, radius=0.9), aabb.to(dev), dev, row0=400, rows=64)  # 122 880 rays


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, out


with torch.no_grad():
    assert model.occ_ready
    ms_o, out_o = timed(lambda: render_rays(model, rays, ray_samples=64))
    model.use_occ = False  # the viewer's switch (viewer.py:725)
    ms_s, out_s = timed(lambda: render_rays(model, rays, ray_samples=64, coherent=True))
    model.use_occ = True
    N = rays.shape[0]
    print(f"occupancy renderer : {ms_o:8.2f} ms for {N} rays = {N / ms_o / 1e3:6.2f} M rays/s, {out_o[2].shape[0] / N:.1f} merged segments per ray, mean acc {float(out_o[3].mean()):.3f}")
    print(f"stratified (64/ray): {ms_s:8.2f} ms for {N} rays = {N / ms_s / 1e3:6.2f} M rays/s, mean acc {float(out_s[3].mean()):.3f}; mean |rgb diff| {float((out_o[0] - out_s[0]).abs().mean()):.4f}")
    timing = ops.kernel_timing(True)
    render_rays(model, rays, ray_samples=64)
    torch.cuda.synchronize()
    ops.kernel_timing(False)
    print({k: (round(sum(a.elapsed_time(b) for a, b in v), 2), len(v)) for k, v in timing.items()})
