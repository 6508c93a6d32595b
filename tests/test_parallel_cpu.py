"""Host-side logic of the multi-GPU partitioning (nerf_sys_b200/parallel.py) on CPU with the gloo backend,
world_size 2 and 3: the expert-parallel all-to-all dispatch/combine (forward values and the gradients that travel
back to the owning rank), the ray-sharded frame assembly, the shared-gradient all-reduce, and the partitions.
No CUDA kernel is involved: the expert function is a stand-in torch function."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nerf_sys_b200 import parallel as PL


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _spawn(fn, world, *args):
    port = _free_port()
    mp.spawn(_entry, args=(fn, world, port, args), nprocs=world, join=True)


def _entry(rank, fn, world, port, args):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fn(rank, world, *args)
    finally:
        dist.destroy_process_group()


def test_partitions():
    for n in (0, 1, 7, 8, 1080):
        for world in (1, 2, 3, 8):
            blocks = [PL.shard_range(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert PL.expert_owner(8, 8) == list(range(8))
    assert PL.expert_owner(8, 2) == [0, 0, 0, 0, 1, 1, 1, 1]
    assert PL.expert_owner(8, 3) == [0, 0, 0, 1, 1, 1, 2, 2]
    with pytest.raises(ValueError):
        PL.shard_range(4, 2, 2)


# ---- expert-parallel dispatch / combine ---------------------------------------------------------
K, D = 5, 6


def _make_weights():
    g = torch.Generator().manual_seed(0)
    return [torch.randn(D, 4, generator=g) for _ in range(K)]


def _routing(P, seed):
    """Random soft routing: every point goes to 1 or 2 experts; pairs grouped by expert like route_pack's output."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(P, D, generator=g)
    first = torch.randint(0, K, (P,), generator=g)
    second = torch.randint(0, K, (P,), generator=g)
    two = (torch.rand(P, generator=g) < 0.3) & (second != first)
    w1 = torch.where(two, torch.rand(P, generator=g) * 0.6 + 0.2, torch.ones(P))
    pts, ws, counts = [], [], []
    for e in range(K):
        a = torch.nonzero(first == e).squeeze(1)
        b = torch.nonzero(two & (second == e)).squeeze(1)
        pts.append(torch.cat([a, b]))
        ws.append(torch.cat([w1[a], 1 - w1[b]]))
        counts.append(int(a.numel() + b.numel()))
    return x, torch.cat(pts), torch.cat(ws), counts


def _dense(x, sel_point, sel_w, counts, Ws):
    out = torch.zeros(x.shape[0], 4)
    o = 0
    for e in range(K):
        idx = sel_point[o:o + counts[e]]
        out = out.index_add(0, idx, torch.tanh(x[idx] @ Ws[e]) * sel_w[o:o + counts[e], None])
        o += counts[e]
    return out


def _ep_worker(rank, world, frozen):
    Ws = [w.clone().requires_grad_(not frozen or e == 0) for e, w in enumerate(_make_weights())]
    P = 300 + 17 * rank  # ragged: every rank holds a different number of points
    x, sel_point, sel_w, counts = _routing(P, seed=10 + rank)
    if rank == 1:  # an expert that receives nothing from this rank, and an empty expert overall is covered by K=5 > world
        keep = torch.ones(sel_point.numel(), dtype=torch.bool)
        o = sum(counts[:3])
        keep[o:o + counts[3]] = False
        sel_point, sel_w, counts = sel_point[keep], sel_w[keep], counts[:3] + [0] + counts[4:]

    def expert_fn(rows, local, index):
        y = rows.new_zeros(rows.shape[0], 4)
        for e, ix in zip(local, index):
            y = y.index_add(0, ix, torch.tanh(rows[ix] @ Ws[e]))
        return y

    out = PL.expert_parallel_apply(x.index_select(0, sel_point), counts, expert_fn,
                                   lambda y: x.new_zeros(P, 4).index_add(0, sel_point, y * sel_w.unsqueeze(1)))
    ref_W = [w.clone().requires_grad_(True) for w in _make_weights()]
    ref = _dense(x, sel_point, sel_w, counts, ref_W)
    assert torch.allclose(out, ref, atol=1e-6), (out - ref).abs().max()
    # gradients: every rank's loss contributes to the owner's copy of an expert
    g = torch.Generator().manual_seed(99 + rank)
    cot = torch.randn(P, 4, generator=g)
    (out * cot).sum().backward()
    (ref * cot).sum().backward()
    owner = PL.expert_owner(K, world)
    for e in range(K):
        total = ref_W[e].grad.clone()
        dist.all_reduce(total)  # sum over ranks of d(loss_r)/dW_e
        if owner[e] == rank and Ws[e].requires_grad:
            assert torch.allclose(Ws[e].grad, total, atol=1e-5), (e, (Ws[e].grad - total).abs().max())
        elif Ws[e].requires_grad:
            assert Ws[e].grad is None or Ws[e].grad.abs().max() == 0, f"rank {rank} must not touch expert {e}"


@pytest.mark.parametrize("world", [2, 3])
def test_expert_parallel_all_to_all_matches_dense(world):
    _spawn(_ep_worker, world, False)


def test_expert_parallel_with_frozen_experts_does_not_deadlock():
    """Config 4: only expert 0 trains; ranks whose experts are all frozen still join the backward exchange."""
    _spawn(_ep_worker, 2, True)


def test_expert_parallel_single_process_path():
    Ws = [w.requires_grad_(True) for w in _make_weights()]
    x, sel_point, sel_w, counts = _routing(200, seed=3)

    def expert_fn(rows, local, index):
        y = rows.new_zeros(rows.shape[0], 4)
        for e, ix in zip(local, index):
            y = y.index_add(0, ix, torch.tanh(rows[ix] @ Ws[e]))
        return y

    out = PL.expert_parallel_apply(x.index_select(0, sel_point), counts, expert_fn,
                                   lambda y: x.new_zeros(200, 4).index_add(0, sel_point, y * sel_w.unsqueeze(1)))
    assert torch.allclose(out, _dense(x, sel_point, sel_w, counts, Ws), atol=1e-6)


# ---- ray-sharded render and shared gradients -----------------------------------------------------
def _render_worker(rank, world):
    H, W = 11, 7  # rows do not divide evenly

    def render_rows(r0, r1):
        rows = torch.arange(r0, r1).float().repeat_interleave(W)
        cols = torch.arange(W).float().repeat(r1 - r0)
        rgb = torch.stack([rows, cols, rows * 100 + cols], 1)
        return rgb, rows + 0.5, cols + 0.25

    rgb, depth, acc = PL.render_image_sharded(render_rows, H, W)
    full_rgb, full_d, full_a = render_rows(0, H)
    assert torch.equal(rgb, full_rgb.view(H, W, 3)) and torch.equal(depth, full_d) and torch.equal(acc, full_a)
    (own, rng) = PL.render_image_sharded(render_rows, H, W, gather=False)
    assert rng == PL.shard_range(H, world, rank) and own[0].shape[0] == (rng[1] - rng[0]) * W
    # shared head: mean of the per-rank gradients, parameters without gradient are skipped
    ps = [torch.zeros(3, 2, requires_grad=True), torch.zeros(5, requires_grad=True), torch.zeros(2, requires_grad=True)]
    ps[0].grad = torch.full((3, 2), float(rank + 1))
    ps[1].grad = torch.arange(5.0) * (rank + 1)
    PL.allreduce_shared_grads(ps)
    mean = sum(range(1, world + 1)) / world
    assert torch.allclose(ps[0].grad, torch.full((3, 2), mean)) and torch.allclose(ps[1].grad, torch.arange(5.0) * mean)
    # the same reduction started from inside backward() as soon as the shared gradients exist
    qs = [torch.ones(3, 2, requires_grad=True), torch.ones(5, requires_grad=True)]
    red = PL.SharedGradReducer(qs)
    for _ in range(2):
        for q in qs:
            q.grad = None
        ((qs[0].sum() + (qs[1] * torch.arange(5.0)).sum()) * (rank + 1)).backward()
        red.wait()
        assert torch.allclose(qs[0].grad, torch.full((3, 2), mean)) and torch.allclose(qs[1].grad, torch.arange(5.0) * mean)
    red.remove()
    assert ps[2].grad is None


@pytest.mark.parametrize("world", [2, 3])
def test_ray_sharded_frame_and_shared_grad_allreduce(world):
    _spawn(_render_worker, world)


def test_owned_param_groups_partition_experts_and_keep_the_shared_head():
    """Expert parallelism: a rank optimises its own experts plus the replicated background head (world 1 = everything)."""
    from types import SimpleNamespace
    import torch.nn as nn

    class Sub(nn.Module):
        def __init__(self):
            super().__init__()
            self.t, self.s, self.c = nn.Parameter(torch.zeros(4, 2)), nn.Parameter(torch.zeros(3)), nn.Parameter(torch.zeros(2))

        def get_param_groups(self):
            return {"encoding": {"params": [self.t]}, "sigma": {"params": [self.s]}, "color": {"params": [self.c]}}

    subs = [Sub() for _ in range(8)]
    model = SimpleNamespace(submodules=subs, use_bg_nerf=True, bg_dir_enc=nn.Identity(), bg_mlp=nn.Linear(16, 3))
    groups, shared = PL.owned_param_groups(model)
    assert len(groups["encoding"]["params"]) == 8 and len(shared) == 2 and groups["background"]["params"] == shared
    import nerf_sys_b200.parallel as mod
    real = mod._world
    try:
        for world in (2, 3, 8):
            seen = []
            for rank in range(world):
                mod._world = lambda group, w=world, r=rank: (w, r)
                g, sh = PL.owned_param_groups(model)
                lo, hi = PL.shard_range(8, world, rank)
                assert [id(p) for p in g["encoding"]["params"]] == [id(subs[e].t) for e in range(lo, hi)]
                assert len(g["sigma"]["params"]) == hi - lo == len(g["color"]["params"]) and len(sh) == 2
                seen += g["encoding"]["params"]
            assert len(seen) == 8, "every expert is trained by exactly one rank"
    finally:
        mod._world = real


def test_peer_layout_tiles_every_slab_without_overlap():
    """Peer-memory exchange: every (source, expert) sub-block lands in its owner's slab exactly once, expert blocks are
    contiguous, start on tile boundaries and appear in expert order; empty experts and ragged counts included."""
    g = torch.Generator().manual_seed(5)
    for world in (1, 2, 3, 8):
        K = 8
        owner = PL.expert_owner(K, world)
        table = torch.randint(0, 700, (world, K), generator=g)
        table[:, 3] = 0          # an expert nobody routes to
        table[0, 5] = 0          # a source with nothing for expert 5
        table = table.tolist()
        cover = {r: [] for r in range(world)}
        for rank in range(world):
            start, tot, used, row0 = PL.peer_layout(table, owner, rank, 128)
            for e in range(K):
                assert tot[e] == sum(table[s][e] for s in range(world)) and start[e] % 128 == 0
                assert start[e] <= row0[e] and row0[e] + table[rank][e] <= start[e] + tot[e]
                cover[owner[e]].append((row0[e], row0[e] + table[rank][e], e))
        for r in range(world):
            spans = sorted(x for x in cover[r] if x[1] > x[0])
            for a, b in zip(spans, spans[1:]):
                assert a[1] <= b[0], "sub-blocks must not overlap"
                assert a[2] <= b[2], "experts appear in order"
            mine = [e for e in range(K) if owner[e] == r]
            assert used[r] == sum((tot[e] + 127) // 128 * 128 for e in mine)
            for e in mine:  # the sources tile the expert's block exactly
                rows = sum(b - a for a, b, ee in spans if ee == e)
                assert rows == tot[e]


# ---- owner -> everyone parameter sync ---------------------------------------------------------------
def _sync_worker(rank, world):
    import torch.nn as nn

    class Sub(nn.Module):
        def __init__(self, big):
            super().__init__()
            self.table = nn.Parameter(torch.zeros((1 << 20) + 8 if big else 64, 2))
            self.w, self.b = nn.Parameter(torch.zeros(5, 3)), nn.Parameter(torch.zeros(5))

    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.submodules = nn.ModuleList([Sub(big=(e == 1)) for e in range(5)])

    m = M()
    owner = PL.expert_owner(5, world)
    with torch.no_grad():  # every rank "trained" every copy differently; only the owner's counts
        for e, sub in enumerate(m.submodules):
            for j, p in enumerate(sub.parameters()):
                p.fill_(1000.0 * rank + 10.0 * e + j)
    PL.sync_expert_params(m)
    for e, sub in enumerate(m.submodules):
        for j, p in enumerate(sub.parameters()):
            assert float(p.min()) == float(p.max()) == 1000.0 * owner[e] + 10.0 * e + j, (rank, e, j)


@pytest.mark.parametrize("world", [2, 3])
def test_sync_expert_params_broadcasts_from_the_owner(world):
    _spawn(_sync_worker, world)


# ---- Voronoi box reduction (scripts/create_clusters.py:928-932) ------------------------------------
def _boxes_worker(rank, world):
    from nerf_sys_b200.data.voronoi_masks import reduce_boxes
    mins = torch.tensor([[float(rank), -float(rank), 0.5]] * 3)
    maxs = torch.tensor([[float(rank), 10.0 + rank, 0.5]] * 3)
    cnts = torch.tensor([rank + 1, 0, 5], dtype=torch.int64)
    reduce_boxes(mins, maxs, cnts)
    assert mins[0].tolist() == [0.0, -float(world - 1), 0.5] and maxs[0].tolist() == [float(world - 1), 10.0 + world - 1, 0.5]
    assert cnts.tolist() == [world * (world + 1) // 2, 0, 5 * world]


def test_voronoi_box_reduction_min_max_sum():
    _spawn(_boxes_worker, 2)
