"""GPU parity of the Voronoi mask builder (csrc/voronoi.cu): masks against the golden vectors of the unmodified
compute_voronoi_orig and, at scale, against the oracle - identical except where a ray's best distance ratio sits within 1e-5
of the margin (the reference itself has two formulations, cdist ratio on the CPU and expanded squared distances on CUDA, that
differ there); per-expert boxes and counts against the oracle's restatement of the CUDA branch."""
import pytest
import torch

from conftest import T, load_golden
from oracle import acnerf_oracle as O
from oracle import voronoi_oracle as VO

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _borderline(rays, S, cent, c2d, margin):
    """(N,C) bool: the ray's smallest distance ratio to c is within 1e-5 of the margin (fp64)"""
    start = 1 if c2d else 0
    x = VO.sample_positions(rays, S).double()
    D = torch.cdist(x.reshape(-1, 3)[:, start:], cent[:, start:].double())
    if margin == 1.0:  # strict Voronoi: a sample within 1e-5 (relative) of a bisector
        two = D.topk(2, dim=1, largest=False).values
        tie = ((two[:, 1] - two[:, 0]) <= 1e-5 * two[:, 1]).view(rays.shape[0], S).any(1)
        return tie[:, None].expand(-1, cent.shape[0])
    ratio = (D / D.min(dim=1, keepdim=True).values.clamp_min(1e-30)).view(rays.shape[0], S, -1)
    return ((ratio - margin).abs() <= 1e-5 * margin).any(dim=1)


@pytest.mark.parametrize("tag", ["soft2d", "hard2d", "soft3d"])
def test_voronoi_masks_vs_reference_golden(tag):
    from nerf_sys_b200.data.voronoi_masks import compute_voronoi_opt
    z = load_golden("voronoi")
    c2d, margin, S = z[f"{tag}_cfg"]
    rays, cent = T(z["rays"]), T(z["centroids"])
    m = compute_voronoi_opt(rays.to(DEV), ray_samples=int(S), ray_chunk_size=1024, sample_chunk_size=1 << 16, centroids=cent, cluster_2d=bool(c2d),
                            device=torch.device(DEV), boundary_margin=float(margin))
    assert m.device.type == "cpu" and m.dtype == torch.bool and m.shape == (3000, 8)
    ref = T(z[f"{tag}_mask"])
    mism = m != ref
    assert not (mism & ~_borderline(rays, int(S), cent, bool(c2d), float(margin))).any()
    assert mism.sum().item() <= 3, mism.sum().item()
    assert (m.sum(1) >= 1).all()  # every ray belongs somewhere


def test_voronoi_masks_boxes_and_counts_at_scale_vs_oracle():
    from nerf_sys_b200.data.voronoi_masks import compute_voronoi_opt
    aabb, cent, _ = O.grid_scene(2, 4)
    N, S = 200_000, 256  # the script's ray_samples default
    rays = O.aerial_rays(N, aabb, seed=9)
    for c2d, margin in ((True, 1.05), (True, 1.0)):
        mins = torch.full((8, 3), float("inf"), device=DEV)
        maxs = torch.full((8, 3), float("-inf"), device=DEV)
        cnts = torch.zeros(8, dtype=torch.int64, device=DEV)
        m = compute_voronoi_opt(rays.to(DEV), ray_samples=S, centroids=cent, cluster_2d=c2d, boundary_margin=margin, update_aabbs=True,
                                mins_out=mins, maxs_out=maxs, counts_out=cnts)
        sub = slice(0, 20_000)  # the oracle materialises (N,S,C) distances
        ok, mins_o, maxs_o, cnts_o = VO.voronoi_boxes(rays[sub], S, cent, c2d, margin)
        mism = m[sub] != ok.any(dim=1)
        assert not (mism & ~_borderline(rays[sub], S, cent, c2d, margin)).any() and mism.sum().item() <= 20
        # boxes / counts of the same 20 000 rays, streamed in two calls (the running update must accumulate)
        mins2 = torch.full((8, 3), float("inf"), device=DEV)
        maxs2 = torch.full((8, 3), float("-inf"), device=DEV)
        cnts2 = torch.zeros(8, dtype=torch.int64, device=DEV)
        for lo, hi in ((0, 7_001), (7_001, 20_000)):
            compute_voronoi_opt(rays[lo:hi].to(DEV), ray_samples=S, centroids=cent, cluster_2d=c2d, boundary_margin=margin, update_aabbs=True,
                                mins_out=mins2, maxs_out=maxs2, counts_out=cnts2)
        assert ((cnts2.cpu() - cnts_o).abs() <= 1e-4 * cnts_o.clamp_min(1) + 2).all(), (cnts2.cpu(), cnts_o)
        assert (mins2.cpu() - mins_o).abs().max() <= 1e-4 and (maxs2.cpu() - maxs_o).abs().max() <= 1e-4
        assert (mins.cpu() <= mins2.cpu()).all() and (maxs.cpu() >= maxs2.cpu()).all() and (cnts.cpu() >= cnts2.cpu()).all()
        assert int(cnts.sum()) >= N * S
