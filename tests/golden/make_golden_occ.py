"""Generate tests/golden/occ_assembly.npz with the UNMODIFIED reference's own part of the occupancy renderer
(nerfs/ray_rendering.py: _intersect_rays_aabb :172-193, _merge_segments_union :196-256, render_rays_occ :349-491,
render_expert_occ :494-559) and the real MetaContainer._routing (models/inr/meta_container.py:97-134).

    python tests/golden/make_golden_occ.py     (build container only)

nerfacc (0.5.3, requirements.txt:8) is not installed here.  The reference module imports it at the top, so it is replaced
by a stub whose three packed primitives (pack_info, render_weight_from_density, accumulate_along_rays) are the oracle's
restatements; the experts are duck types (scene_box, occupancy_marching -> the oracle's traversal over fixed random grids,
__call__ -> a closed-form field).  What the vectors pin is therefore the reference's assembly: prefilter, per-expert lists in
global ray ids, union, midpoints, routing at the midpoints, evaluation where the weight is positive, the sigma / rgb blend,
the single integration's inputs, background.  nerfacc's own arithmetic stays unpinned.

One adaptation, because the reference's container path cannot run as shipped: render_rays_occ hands _routing a (1,M,3)
view (ray_rendering.py:437) and _routing asserts a 2-D input (meta_container.py:111), so the full-container occupancy
render raises AssertionError on its first call.  The script reshapes that one argument to (M,3) and calls the real
_routing; the lines after it (W.squeeze(0), the hard-routing scatter) work unchanged on the (M,K) / (M,) results."""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import acnerf_oracle as O  # noqa: E402
from oracle import occ_oracle as OO  # noqa: E402

na = types.ModuleType("nerfacc")
na.OccGridEstimator = object
na.pack_info = lambda ray_indices, n_rays: OO.pack_info(ray_indices, n_rays)
na.render_weight_from_density = lambda t_starts, t_ends, sigmas, packed_info=None, ray_indices=None, n_rays=None: OO.weights_from_density(
    t_starts, t_ends, sigmas, _RI[0])
na.accumulate_along_rays = lambda weights, values, ray_indices, n_rays: OO.accumulate(weights, values, ray_indices, n_rays)
_RI = [None]  # render_weight_from_density receives packed_info only; the sample -> ray map of the same call is kept here
for name in ("viser", "viser.transforms"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["viser"].transforms = sys.modules["viser.transforms"]
sys.modules["nerfacc"] = na
sys.path.insert(0, "/root/reference/adaptive_nerf")
import nerfs.ray_rendering as RR  # noqa: E402
from models.inr.meta_container import MetaContainer  # noqa: E402
from nerfs.scene_box import SceneBox  # noqa: E402

_pack = na.pack_info


def _pack_and_remember(ray_indices, n_rays):
    _RI[0] = ray_indices
    return _pack(ray_indices, n_rays)


na.pack_info = _pack_and_remember

STEP, CONE = 0.01, 0.004


class Expert:
    use_bg_nerf = True

    def __init__(self, k, box, binaries, aabbs):
        self.k, self.scene_box, self.binaries, self.aabbs, self.occ_grid = k, SceneBox(torch.stack([box[0], box[1]])), binaries, aabbs, True
        self.calls = []

    def occupancy_marching(self, rays, params=None, render_step_size=None, alpha_thre=None, cone_angle=None):
        r, a, b, _, _ = OO.march(rays, self.binaries, self.aabbs, STEP, CONE)
        self.calls.append((r, a, b))
        return r, a, b

    def __call__(self, xd, params=None):
        return OO.toy_field(xd, self.k)

    def background_color(self, d):
        return OO.toy_background(d)


def main():
    g = torch.Generator().manual_seed(31)
    aabb, cent, boxes = O.grid_scene(2, 2)
    K, N, R = 4, 160, 16
    rays = O.aerial_rays(N, aabb, seed=9)
    rays[:6, 7] = rays[:6, 6]  # empty [near, far]: prefiltered out of every expert
    out = {"rays": rays.numpy(), "centroids": cent.numpy(), "boxes": torch.stack([torch.stack([b[0], b[1]]) for b in boxes]).numpy(),
           "step_cone": np.array([STEP, CONE], dtype=np.float32)}
    experts = []
    for k in range(K):
        roi = torch.cat([boxes[k][0], boxes[k][1]])
        aabbs = torch.stack([roi, OO.enlarge_aabb(roi, 2.0)])
        binaries = torch.rand(2, R, R, R, generator=g) < (0.30 if k != 3 else 0.0)  # expert 3 proposes nothing
        experts.append(Expert(k, boxes[k], binaries, aabbs))
        out[f"binaries{k}"], out[f"aabbs{k}"] = binaries.numpy(), aabbs.numpy()
        hit = RR._intersect_rays_aabb(rays, experts[k].scene_box)
        out[f"hit{k}"] = hit.numpy()
    for tag, margin, c2d in (("soft2d", 1.05, True), ("hard3d", 1.0, False)):
        m = MetaContainer.__new__(MetaContainer)
        torch.nn.Module.__init__(m)
        m.register_buffer("centroids", cent)
        m.boundary_margin, m.cluster_2d = margin, c2d
        m._coord_idx = (1, 2) if c2d else (0, 1, 2)
        m.submodules, m.use_bg_nerf = experts, True
        m.background_color = lambda d: OO.toy_background(d)
        real_routing = MetaContainer._routing.__get__(m)
        m._routing = lambda x: real_routing(x.reshape(-1, 3))  # see the module docstring
        for e in experts:
            e.calls.clear()
        rgb, depth, w, acc = RR.render_rays_occ(m, rays)
        out[f"{tag}_rgb"], out[f"{tag}_depth"], out[f"{tag}_w"], out[f"{tag}_acc"] = rgb.numpy(), depth.numpy(), w.numpy(), acc.numpy()
        out[f"{tag}_cfg"] = np.array([margin, float(c2d)])
        if tag == "soft2d":
            # the union the reference built inside that call, from the same per-expert lists
            lists = []
            for k, e in enumerate(experts):
                r, a, b = e.calls[0]
                hit_idx = torch.from_numpy(out[f"hit{k}"]).nonzero().squeeze(1)
                if a.numel():
                    lists.append((hit_idx[r], a, b))
                out[f"list{k}_ri"], out[f"list{k}_t0"], out[f"list{k}_t1"] = hit_idx[r].numpy(), a.numpy(), b.numpy()
            ri, t0, t1 = RR._merge_segments_union([l[0] for l in lists], [l[1] for l in lists], [l[2] for l in lists])
            out["union_ri"], out["union_t0"], out["union_t1"] = ri.numpy(), t0.numpy(), t1.numpy()
    # one expert alone (render_expert_occ), and the expert that proposes nothing (background only)
    for k in (1, 3):
        rgb, depth, w, acc = RR.render_expert_occ(experts[k], rays)
        out[f"expert{k}_rgb"], out[f"expert{k}_depth"], out[f"expert{k}_w"], out[f"expert{k}_acc"] = rgb.numpy(), depth.numpy(), w.numpy(), acc.numpy()
    np.savez_compressed(os.path.join(HERE, "occ_assembly.npz"), **out)
    print({k: v.shape for k, v in out.items() if not k.startswith(("binaries", "aabbs"))})
    print("merged segments", out["union_ri"].shape, "acc mean", float(out["soft2d_acc"].mean()), "hits", [int(out[f"hit{k}"].sum()) for k in range(K)])


if __name__ == "__main__":
    main()
