"""CPU checks of the occupancy grid's bookkeeping (nerf_sys_b200/models/occ_grid.py: buffers / checkpoint keys of
nerfacc.OccGridEstimator, EMA update, visibility pre-marking) and of the oracle's pieces against brute force."""
import math

import torch

from oracle import occ_oracle as OO


def test_occ_grid_buffers_update_and_premark_on_cpu():
    from nerf_sys_b200.models.occ_grid import OccGridEstimator
    roi = torch.tensor([-1.0, -0.5, -0.25, 1.0, 0.5, 0.25])
    g = OccGridEstimator(roi_aabb=roi, resolution=8, levels=3)
    sd = g.state_dict()
    assert set(sd) == {"resolution", "aabbs", "occs", "binaries"}  # grid_coords / grid_indices are not persistent (as in nerfacc)
    assert sd["resolution"].tolist() == [8, 8, 8] and sd["resolution"].dtype == torch.int32
    assert tuple(sd["aabbs"].shape) == (3, 6) and tuple(sd["binaries"].shape) == (3, 8, 8, 8) and sd["binaries"].dtype == torch.bool
    assert torch.allclose(sd["aabbs"][2], torch.tensor([-4.0, -2.0, -1.0, 4.0, 2.0, 1.0]))
    assert g.grid_coords[9].tolist() == [0, 1, 1]  # cell id = (ix * R + iy) * R + iz
    # EMA update: occupancy = 1 inside the unit ball, 0 outside
    g.train()
    fn = lambda x: (x.norm(dim=-1) < 0.45).float()  # noqa: E731
    torch.manual_seed(0)
    g.update_every_n_steps(step=0, occ_eval_fn=fn, occ_thre=0.5, ema_decay=0.9, warmup_steps=4, n=1)
    occ0 = g.occs.clone()
    assert g.binaries[0].any() and not g.binaries[0].all()
    assert float(g.binaries[0].float().mean()) > float(g.binaries[2].float().mean())  # the ball fills less of the coarser boxes
    g.update_every_n_steps(step=1, occ_eval_fn=lambda x: torch.zeros(x.shape[0]), occ_thre=0.5, ema_decay=0.9, warmup_steps=4, n=1)
    assert torch.allclose(g.occs, occ0 * 0.9)  # max(decayed, 0)
    g.update_every_n_steps(step=3, occ_eval_fn=fn, n=2)  # not a multiple of n: nothing happens
    assert torch.allclose(g.occs, occ0 * 0.9)
    g.update_every_n_steps(step=8, occ_eval_fn=fn, occ_thre=0.5, ema_decay=0.9, warmup_steps=4, n=1)  # after warm-up: sampled cells only
    assert (g.occs >= occ0 * 0.9 * 0.9 - 1e-6).all()
    g.eval()
    try:
        g.update_every_n_steps(step=0, occ_eval_fn=fn)
        raise AssertionError("update outside training must raise")
    except RuntimeError:
        pass
    # visibility pre-marking: one camera at x = -3 looking down +x (OpenCV axes: z forward), 90 degree field of view
    K = torch.tensor([[[32.0, 0.0, 32.0], [0.0, 32.0, 32.0], [0.0, 0.0, 1.0]]])
    c2w = torch.tensor([[[0.0, 0.0, 1.0, -3.0], [-1.0, 0.0, 0.0, 0.0], [0.0, -1.0, 0.0, 0.0]]])
    g.mark_invisible_cells(K, c2w, width=64, height=64, near_plane=0.5)
    occs = g.occs.view(3, 8, 8, 8)
    assert (occs[0] == 0).all()                     # the finest box lies inside the frustum
    assert (occs[2][0] == -1).all()                 # cells at x = -4 are behind the camera
    assert (occs[2][7, 3:5, 3:5] == 0).all()        # far cells on the axis are seen
    assert len(g._get_all_cells()[2]) == int((occs[2] >= 0).sum())


def test_anneal_schedule_matches_reference_formula():
    from nerf_sys_b200.models.inr import MetaNGP
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    box = SceneBox(torch.tensor([[0.0, 0.0, 0.0], [1.0, 2.0, 2.0]]))
    occ = {"use_occ": True, "resolution": 4, "levels": 1, "warmup_steps": 5, "alpha_thre": 0.02, "alpha_thre_start": 0.0}
    m = MetaNGP(occ_conf=occ, scene_box=box, color_depth=2, hash_enc_conf={"levels": 16, "log2_hashmap_size": 8})
    assert abs(m.render_step_size - 3.0 / 1000.0) < 1e-9 and m.cone_angle == 1.0 / 256.0
    vals = []
    for step in range(7):
        m._anneal_alpha_thre(step)
        vals.append(m.alpha_thre)
    want = [0.5 * (1 - math.cos(math.pi * s / 4)) * 0.02 for s in range(5)] + [0.02, 0.02]  # meta_ngp.py:245-259
    assert all(abs(a - b) < 1e-12 for a, b in zip(vals, want))
    assert "scene_aabb" in m.state_dict() and "occ_grid.occs" in m.state_dict()


def test_oracle_union_and_packed_render_against_brute_force():
    ri = [torch.tensor([0, 0, 2]), torch.tensor([0, 2, 2, 5])]
    t0 = [torch.tensor([1.0, 2.0, 0.5]), torch.tensor([1.5, 0.5, 3.0, 7.0])]
    t1 = [torch.tensor([2.0, 3.0, 1.0]), torch.tensor([2.5, 1.0, 4.0, 8.0])]
    r, a, b = OO.union(ri, t0, t1)
    assert r.tolist() == [0, 0, 0, 0, 2, 2, 2, 5]
    assert a.tolist() == [1.0, 1.5, 2.0, 2.5, 0.5, 1.0, 3.0, 7.0] and b.tolist() == [1.5, 2.0, 2.5, 3.0, 1.0, 3.0, 4.0, 8.0]  # the gap [1, 3] of ray 2 is a segment
    sigma, rgb = torch.tensor([1.0, 2.0, 0.0, 4.0, 1.0, 1.0, 1.0, 3.0]), torch.rand(8, 3)
    rgb_map, depth, w, acc = OO.packed_render(r, a, b, sigma, rgb, 6)
    al = 1 - torch.exp(-sigma * (b - a))
    assert torch.allclose(w[:4], al[:4] * torch.tensor([1.0, float(1 - al[0]), float((1 - al[0]) * (1 - al[1])), float((1 - al[0]) * (1 - al[1]) * (1 - al[2]))]))
    assert torch.allclose(acc[0], w[:4].sum()) and acc[1] == 0 and torch.allclose(depth[5], w[7] * 7.5)
