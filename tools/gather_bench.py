"""acn_ngp_encode_fwd over the training-step pair lists (config 2: identity pairs of one expert; config 3: 8 experts, routed)
for every levels-per-thread setting of the gather kernel (GPU)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nerf_sys_b200 import _lib, ops, synthetic as SY  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
L = _lib.lib()


def time_encode(model, rays, S, tag):
    u = torch.rand(rays.shape[0], S, device=dev)
    t = ops.stratified_t(rays, S, u)
    pts = ops.PointSet(rays=rays, t=t)
    K = len(model.submodules)
    tables = [sub.xyz_encoder.hash_table for sub in model.submodules]
    spec = model.expert_spec()
    if K > 1:
        pairs = ops.route_pack(pts, model.centroids_host(), K, model.boundary_margin, model.cluster_2d)
    else:
        pairs = ops.PairList(pts.P)
    ref = None
    for lg in (1, 2, 4, 8, 16):
        L.acn_launch_hint(_lib.ACN_HINT_GATHER_LEVELS, lg)
        with torch.no_grad():
            enc = ops.NgpEncodeFn.apply(pts, pairs, spec, *tables)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                enc = ops.NgpEncodeFn.apply(pts, pairs, spec, *tables)
            e1.record()
            torch.cuda.synchronize()
        if ref is None:
            ref = enc.clone()
        print(f"{tag}: levels/thread {lg:2d}: {e0.elapsed_time(e1) / 20 * 1e3:8.1f} us   max|diff| vs 1: {float((enc - ref).abs().max()):.1e}", flush=True)
    L.acn_launch_hint(_lib.ACN_HINT_GATHER_LEVELS, 0)


model, aabb = bench.build_container(dev)
time_encode(model, SY.aerial_rays(bench.N_RAYS, aabb, seed=1000).to(dev), 96, "config 3 (8 experts, 8192 x 96 aerial)")
del model
torch.cuda.empty_cache()
m1, st, cfg, box = bench.build_scene(0, 1, dev)
pool = bench.make_pool(box, dev, 1, seed=10)
time_encode(m1, pool[0], 96, "config 2 (1 expert, 8192 x 96 ring views)")
