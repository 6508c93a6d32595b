"""Generate tests/golden/voronoi.npz with the UNMODIFIED reference's compute_voronoi_orig (scripts/create_clusters.py:559-634).

    python tests/golden/make_golden_voronoi.py     (build container only)"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
for name in ("nerfacc", "viser", "viser.transforms"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["nerfacc"].OccGridEstimator = object
sys.modules["viser"].transforms = sys.modules["viser.transforms"]
sys.path.insert(0, "/root/reference/adaptive_nerf")
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from scripts.create_clusters import compute_voronoi_orig  # noqa: E402
from nerf_sys_b200 import synthetic as SY  # noqa: E402  (scene constants and rays: inputs only)


def main():
    out = {}
    aabb, cent, _ = SY.grid_scene(2, 4)
    rays = SY.aerial_rays(3000, aabb, seed=5)
    out["rays"], out["centroids"] = rays.numpy(), cent.numpy()
    for tag, c2d, margin, S in (("soft2d", True, 1.05, 64), ("hard2d", True, 1.0, 64), ("soft3d", False, 1.15, 33)):
        m = compute_voronoi_orig(rays, ray_samples=S, ray_chunk_size=1024, sample_chunk_size=1 << 16, centroids=cent, cluster_2d=c2d,
                                 device=torch.device("cpu"), boundary_margin=margin)
        out[f"{tag}_mask"], out[f"{tag}_cfg"] = m.numpy(), np.array([float(c2d), margin, S])
    np.savez_compressed(os.path.join(HERE, "voronoi.npz"), **out)
    print({k: v.shape for k, v in out.items()}, {k: out[k].mean() for k in out if k.endswith("mask")})


if __name__ == "__main__":
    main()
