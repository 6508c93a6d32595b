"""GPU parity of the episode-sampler routing kernel (csrc/episode.cu) against the golden vectors written by the
unmodified reference and, at scale, against oracle/episode_oracle.py: valid masks, cell ids, keep flags and bins exact."""
import pytest
import torch

from conftest import T, load_golden
from oracle import episode_oracle as EO

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("tag", ["a", "b"])
@pytest.mark.parametrize("policy", ["dda", "alpha"])
def test_episode_router_equals_the_reference_golden(tag, policy):
    from nerf_sys_b200.data.episode_routing import EpisodeRouter
    z = load_golden("episode")
    rays, aabb, cells = T(z[f"{tag}_rays"]), T(z[f"{tag}_aabb"]), tuple(int(c) for c in z[f"{tag}_cells"])
    router = EpisodeRouter(aabb, cells, routing_policy=policy, assignment_checkpoint=0.7)
    assert torch.equal(router.cell_bounds.cpu(), T(z[f"{tag}_cell_bounds"]))
    valid, cid, blen, keep = router.route(rays.to(DEV))
    v = T(z[f"{tag}_{policy}_valid"])
    assert torch.equal(valid.cpu(), v)
    assert torch.equal(cid.cpu()[v], T(z[f"{tag}_{policy}_cid"]))          # best_cid: bit-exact
    assert torch.equal(blen.cpu()[v], T(z[f"{tag}_{policy}_len"]))         # the in-cell lengths too: same fp32 operations
    assert (cid.cpu()[~v] == -1).all()
    bins = router.route_and_bin(rays.to(DEV))
    sizes = z[f"{tag}_{policy}_bin_sizes"].tolist()
    assert [b.numel() for b in bins] == sizes
    for mine, ref in zip(bins, torch.split(T(z[f"{tag}_{policy}_bins"]), sizes)):
        assert torch.equal(mine.cpu(), torch.sort(ref).values)


@pytest.mark.parametrize("policy", ["dda", "alpha"])
def test_episode_router_one_million_rays_vs_oracle(policy):
    """Every training ray of a region at once (the reference does this on the CPU before step 1); (N,6) rays without bounds too."""
    from nerf_sys_b200.data.episode_routing import EpisodeRouter
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    g = torch.Generator().manual_seed(3)
    aabb = torch.tensor([[-0.05, -0.6, -0.55], [0.5, 0.05, 0.0]])
    N = 1 << 20
    ctr, half = (aabb[0] + aabb[1]) / 2, (aabb[1] - aabb[0]) / 2
    o = ctr + (torch.rand(N, 3, generator=g) - 0.5) * 5 * half
    d = torch.nn.functional.normalize(ctr + (torch.rand(N, 3, generator=g) - 0.5) * 2.2 * half - o, dim=-1)
    near = torch.rand(N, generator=g) * 0.1
    rays = torch.cat([o, d, near[:, None], (near + torch.rand(N, generator=g) * 3)[:, None]], 1)
    for cols, cells in ((8, (1, 6, 6)), (6, (2, 7, 5))):
        r = rays[:, :cols].contiguous()
        router = EpisodeRouter(aabb, cells, routing_policy=policy)
        valid, cid, blen, keep = router.route(r.to(DEV))
        v_o, c_o, l_o, k_o = EO.route(r, aabb, cells, policy, alpha=0.7)
        assert torch.equal(valid.cpu(), v_o) and torch.equal(cid.cpu(), c_o) and torch.equal(keep.cpu(), k_o)
        assert torch.equal(blen.cpu(), l_o)
        assert 0.2 < float(v_o.float().mean()) < 1.0
