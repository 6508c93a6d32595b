"""Generate tests/golden/episode.npz by running the UNMODIFIED reference's TaskDataset routing methods on seeded rays.

    python tests/golden/make_golden_episode.py     (build container only: /root/reference must exist)
The methods are called on an instance created without __init__ (which wants a full RamRaysDataset); only the attributes they
read are set (aabb, cells, cell_bounds, cell_sizes, routing_policy, rays, device)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/adaptive_nerf")
from data.task_dataset import TaskDataset  # noqa: E402


def make_rays(N, aabb, seed):
    g = torch.Generator().manual_seed(seed)
    lo, hi = aabb[0], aabb[1]
    ctr, half = (lo + hi) / 2, (hi - lo) / 2
    o = ctr + (torch.rand(N, 3, generator=g) - 0.5) * 2 * half * 2.5          # inside and around the region
    o[: N // 3, 0] = lo[0] - 0.3 - torch.rand(N // 3, generator=g) * 0.5     # aerial: above the region, looking down
    target = ctr + (torch.rand(N, 3, generator=g) - 0.5) * 2 * half * 1.1
    d = torch.nn.functional.normalize(target - o, dim=-1)
    d[N // 2: N // 2 + 16, 1] = 0.0                                           # axis-parallel components
    d[N // 2 + 16: N // 2 + 24] = torch.tensor([1.0, 0.0, 0.0])
    near = torch.rand(N, generator=g) * 0.1
    far = near + torch.rand(N, generator=g) * 4.0
    far[-8:] = near[-8:] - 0.1                                                # empty [near, far]
    return torch.cat([o, d, near[:, None], far[:, None]], 1)


def main():
    out = {}
    for tag, aabb, cells, N in (("a", [[-0.05, -0.6, -0.55], [0.5, 0.05, 0.0]], (1, 6, 6), 6000),
                                ("b", [[0.0, 0.0, 0.0], [1.0, 2.0, 3.0]], (3, 5, 4), 6000)):
        aabb_t = torch.tensor(aabb, dtype=torch.float32)
        rays = make_rays(N, aabb_t, seed=len(tag) + N)
        out[f"{tag}_aabb"], out[f"{tag}_cells"], out[f"{tag}_rays"] = aabb_t.numpy(), np.array(cells), rays.numpy()
        for policy in ("dda", "alpha"):
            ds = TaskDataset.__new__(TaskDataset)
            ds.device = torch.device("cpu")
            ds.aabb, ds.cells, ds.routing_policy, ds.rays = aabb_t, cells, policy, rays
            ds.cell_bounds, ds.cell_sizes = TaskDataset._build_cell_bounds(aabb_t, cells, ds.device)
            valid, t0, t1, seg = TaskDataset._region_segment(rays, aabb_t)
            iv = valid.nonzero().reshape(-1)
            if policy == "dda":
                cid, blen = ds._dda_maxoverlap(rays[iv], t0[iv], t1[iv], max_steps=64)
            else:
                cid, blen = ds._alpha_local_max_overlap(rays[iv], aabb_t, cells, 0.7, t0[iv], t1[iv], seg[iv])
            bins = ds._route_and_bin(rays, aabb_t, cells, 0.7)
            out[f"{tag}_{policy}_valid"] = valid.numpy()
            out[f"{tag}_{policy}_cid"] = cid.numpy()
            out[f"{tag}_{policy}_len"] = blen.numpy()
            out[f"{tag}_{policy}_bin_sizes"] = np.array([b.numel() for b in bins])
            out[f"{tag}_{policy}_bins"] = torch.cat(bins).numpy()
        out[f"{tag}_cell_bounds"] = ds.cell_bounds.numpy()
    np.savez_compressed(os.path.join(HERE, "episode.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
