"""Where a row-sharded 1080p frame spends its time at N ranks (torchrun): rows only / + uint8 conversion / + gather on rank 0,
with and without the per-frame routing-capacity read."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nerf_sys_b200 import ops, parallel as PL, synthetic as SY  # noqa: E402
from nerf_sys_b200.nerfs.ray_rendering import render_image  # noqa: E402
from nerf_sys_b200.nerfs.scene_box import SceneBox  # noqa: E402

rank, world, local = bench.dist_env()
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ctx = bench.Ctx(dev, rank, world)
model, aabb = bench.build_container(dev)
model.eval()
H, W, f = 1080, 1920, 1480.0
sb = SceneBox(aabb.to(dev))
poses = [SY.ring_pose(i, 20, radius=0.9) for i in range(12)]
r0, r1 = PL.shard_range(H, world, rank)


def rows(i):
    return render_image(model, H=H, W=W, fx=f, fy=f, cx=W / 2, cy=H / 2, c2w=poses[i % 12], scene_box=sb, ray_samples=64, row_range=(r0, r1))


def run(tag, fn, n=6):
    for i in range(2):
        fn(i)
    ms, _ = bench.timed(ctx, lambda i: fn(2 + i), n)
    # per-rank time of the same loop (no max): load balance
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.barrier()
    e0.record()
    for i in range(n):
        fn(2 + i)
    e1.record()
    torch.cuda.synchronize()
    mine = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
    allr = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allr, mine)
    if rank == 0:
        print(f"{tag:46s} {ms / n:7.3f} ms/frame (max over ranks); per rank: {[round(float(a), 2) for a in allr]}", flush=True)


with torch.no_grad():
    run("rows only (render_image row_range)", rows)
    run("rows + uint8 conversion", lambda i: PL.to_uint8(rows(i)[0]))
    run("rows + uint8 + gather on rank 0 (bench path)", lambda i: PL.render_image_sharded(lambda a, b: rows(i), H, W, as_uint8=True, dst=0))
    timing = ops.kernel_timing(True)
    rows(3)
    torch.cuda.synchronize()
    ops.kernel_timing(False)
    if rank == 0:
        print({k: round(sum(a.elapsed_time(b) for a, b in v), 3) for k, v in timing.items()})
dist.barrier()
dist.destroy_process_group()
