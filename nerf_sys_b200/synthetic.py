"""Synthetic scenes, initial weights, camera poses and ray batches for the benchmarks, the
smoke test and the parity tests.  Host-side scene construction only - plain CPU tensors that are then loaded into the
CUDA modules (and, in the tests, into the CPU oracle, which re-exports these names); none of the path's arithmetic
lives here."""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

GLOBAL_AABB = [[-0.05, -1.1, -1.1], [0.5, 1.1, 1.1]]
HASH_CFG = {"levels": 16, "features_per_level": 2, "log2_hashmap_size": 20, "max_res": 4096,
            "min_res": 16, "interpolation": "Linear"}


def grid_scene(gy: int, gz: int, overlap: float = 0.06):
    """K = gy*gz experts on a 1 x gy x gz YZ grid: centroids at tile centres
    (x = mid), expert boxes = tile grown by ``overlap`` of the tile size, clipped
    to the global box."""
    aabb = torch.tensor(GLOBAL_AABB, dtype=torch.float32)
    lo, hi = aabb[0], aabb[1]
    sy, sz = (hi[1] - lo[1]) / gy, (hi[2] - lo[2]) / gz
    cents, boxes = [], []
    for iy in range(gy):
        for iz in range(gz):
            y0, z0 = lo[1] + iy * sy, lo[2] + iz * sz
            cents.append(torch.stack([(lo[0] + hi[0]) * 0.5, y0 + 0.5 * sy, z0 + 0.5 * sz]))
            bmin = torch.stack([lo[0], torch.maximum(y0 - overlap * sy, lo[1]), torch.maximum(z0 - overlap * sz, lo[2])])
            bmax = torch.stack([hi[0], torch.minimum(y0 + sy + overlap * sy, hi[1]),
                                torch.minimum(z0 + sz + overlap * sz, hi[2])])
            boxes.append(torch.stack([bmin, bmax]))
    return aabb, torch.stack(cents).float(), [b.float() for b in boxes]


def init_ngp_state(K: int, aabb: Tensor, centroids: Tensor, boxes: List[Tensor], hash_cfg: dict,
                   seed: int = 0, table_std: Optional[float] = None, hidden: int = 64,
                   geo: int = 15, bg_hidden: int = 32) -> Dict[str, Tensor]:
    """A state dict with the reference's keys and shapes, filled
    from a seeded generator: nn.Linear-style U(-1/sqrt(in), 1/sqrt(in)) weights,
    sigma-head bias -1, hash tables U(-1e-3,1e-3) or N(0,table_std)."""
    g = torch.Generator().manual_seed(seed)
    T = hash_cfg["levels"] << hash_cfg["log2_hashmap_size"]
    Fd = hash_cfg["features_per_level"]

    def lin(out_f, in_f):
        b = 1.0 / math.sqrt(in_f)
        return ((torch.rand(out_f, in_f, generator=g) * 2 - 1) * b, (torch.rand(out_f, generator=g) * 2 - 1) * b)

    st: Dict[str, Tensor] = {"scene_aabb_vec": torch.cat([aabb[0], aabb[1]]).float(), "centroids": centroids.float()}
    enc_dim = hash_cfg["levels"] * Fd
    for k in range(K):
        pre = f"submodules.{k}."
        st[pre + "aabb_extent"] = boxes[k][1] - boxes[k][0]
        if table_std is None:
            st[pre + "xyz_encoder.hash_table"] = (torch.rand(T, Fd, generator=g) * 2 - 1) * 1e-3
        else:
            st[pre + "xyz_encoder.hash_table"] = torch.randn(T, Fd, generator=g) * table_std
        for name, (o, i) in (("sigma_trunk.0.linear", (hidden, enc_dim)), ("sigma_trunk.1.linear", (hidden, hidden)),
                             ("sigma_head", (1, hidden)), ("geo_head", (geo, hidden)),
                             ("color_mlp.0.linear", (hidden, geo + 16)), ("color_mlp.1.linear", (hidden, hidden)),
                             ("color_mlp.2", (3, hidden))):
            w, b = lin(o, i)
            if name == "sigma_head":
                b = torch.full_like(b, -1.0)
            st[pre + name + ".weight"], st[pre + name + ".bias"] = w, b
    st["bg_mlp.0.weight"], st["bg_mlp.0.bias"] = lin(bg_hidden, 16)
    st["bg_mlp.2.weight"], st["bg_mlp.2.bias"] = lin(3, bg_hidden)
    return st


def ring_pose(i: int, n: int, radius: float = 0.9, height: float = -0.02) -> Tensor:
    """c2w (3,4) of the i-th of n cameras on a ring in the YZ plane at
    x = ``height`` (x is Down in DRB), looking at the scene centre."""
    ang = 2 * math.pi * i / n
    pos = torch.tensor([height, radius * math.cos(ang), radius * math.sin(ang)])
    target = torch.tensor([0.3, 0.0, 0.0])
    back = F.normalize(pos - target, dim=0)  # camera +z (Back)
    down = torch.tensor([1.0, 0.0, 0.0])
    right = F.normalize(torch.linalg.cross(-down, back), dim=0)
    up = torch.linalg.cross(back, right)
    return torch.stack([right, up, back, pos], dim=1).float()


def aerial_rays(N: int, aabb: Tensor, seed: int = 0, cone_deg: float = 35.0) -> Tensor:
    """Aerial rays: origins uniform over YZ in the top 10 %
    of the box's Down range, directions in a cone about +Down; near/far from
    the global box."""
    g = torch.Generator().manual_seed(seed)
    lo, hi = aabb[0], aabb[1]
    o = torch.rand(N, 3, generator=g) * (hi - lo) + lo
    o[:, 0] = lo[0] + torch.rand(N, generator=g) * 0.1 * (hi[0] - lo[0])
    th = torch.rand(N, generator=g) * math.radians(cone_deg)
    ph = torch.rand(N, generator=g) * 2 * math.pi
    d = torch.stack([torch.cos(th), torch.sin(th) * torch.cos(ph), torch.sin(th) * torch.sin(ph)], dim=-1)
    n, f = _slab_near_far(o, d, aabb)
    return torch.cat([o, d, n[:, None], f[:, None]], dim=-1)


def _slab_near_far(o: Tensor, d: Tensor, aabb: Tensor, eps: float = 1e-8, bound: float = 1e10):
    """Entry / exit distances of rays through a box (used only to give the synthetic rays their near / far)."""
    tiny = torch.full_like(d, eps)
    d_safe = torch.where(d.abs() < eps, torch.where(d >= 0, tiny, -tiny), d)
    inv = 1.0 / d_safe
    ta, tb = (aabb[0][None, :] - o) * inv, (aabb[1][None, :] - o) * inv
    t_in = torch.minimum(ta, tb).amax(dim=-1).clamp(min=0.0, max=bound)
    t_out = torch.maximum(ta, tb).amin(dim=-1).clamp(min=0.0, max=bound)
    miss = t_out <= t_in
    bad = torch.tensor(bound, dtype=o.dtype)
    return torch.where(miss, bad, t_in), torch.where(miss, bad, t_out)
