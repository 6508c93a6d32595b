"""GPU tests of the host-read-free training step (round 2): deferred routing + device-side expert liveness +
persistent table gradients zeroed inside the Adam kernel, GradScaler / autocast contract, the device-side exchange plan,
the barrier's small all-reduce, and the routing of non-finite positions.  Reference semantics cited per test."""
import ctypes as C

import pytest
import torch

from oracle import acnerf_oracle as O
from test_gpu_parity import _big_scene, close

pytestmark = pytest.mark.gpu
DEV = "cuda"


# ----------------------------------------------------------------------------- routing of non-finite positions
@pytest.mark.parametrize("K", [3, 5, 7])
def test_routing_of_nan_and_inf_positions_matches_torch(K):
    """meta_container.py:119-126: clamp_min propagates NaN, `dist <= margin*mind` is then false and the point gets no
    expert.  K not in {2,4,8,16} exercises the +inf padding centroids of the templated kernels."""
    from nerf_sys_b200 import ops
    g = torch.Generator().manual_seed(K)
    P = 4096 + 37
    cent = torch.rand(K, 3, generator=g) * 2 - 1
    x = torch.rand(P, 3, generator=g) * 2 - 1
    bad = torch.randperm(P, generator=g)[:300]
    x[bad[:100], 1] = float("nan")
    x[bad[100:200], 2] = float("inf")
    x[bad[200:250], 1] = float("-inf")
    x[bad[250:], :] = float("nan")
    x6 = torch.cat([x, torch.zeros(P, 3)], 1)
    w_ref, _ = O.routing(x, cent, 1.05, True, direct=True)
    w, _ = ops.routing(ops.PointSet(x6=x6.to(DEV)), cent.reshape(-1).tolist(), K, 1.05, True)
    member_ref = torch.nan_to_num(w_ref, nan=0.0) > 0
    assert not member_ref[bad].any()  # the reference gives a non-finite point no expert
    member = torch.nan_to_num(w.cpu(), nan=0.0) > 0
    rows = (member != member_ref).any(1).nonzero().squeeze(1)
    assert rows.numel() == 0, (rows[:8].tolist(), x[rows[:8]].tolist(), w.cpu()[rows[:8]].tolist(), w_ref[rows[:8]].tolist())
    pairs = ops.route_pack(ops.PointSet(x6=x6.to(DEV)), cent.reshape(-1).tolist(), K, 1.05, True)
    pt = pairs.point.cpu()
    for k in range(K):
        seg = pt[pairs.seg_host[k]:pairs.seg_host[k + 1]]
        assert torch.equal(seg[seg >= 0], member_ref[:, k].nonzero().squeeze(1).int()), k  # stable order, no stray slot
        assert (seg[(seg >= 0).sum():] == -1).all()


# ----------------------------------------------------------------------------- optimiser: loss scale, clip, skip, liveness
def _rand_params(shapes, g):
    return [torch.randn(*s, generator=g) for s in shapes]


def test_fused_adam_unscale_clip_skip_match_gradscaler_flow():
    """runtime_adapt.py:261-268: scaler.unscale_(opt) -> clip_grad_norm_ -> scaler.step(opt).  The fused step takes the
    inverse scale as a device tensor, clips on the UNSCALED norm and skips a step whose gradients are not finite."""
    from nerf_sys_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(11)
    shapes = [(64, 32), (64,), (4099, 2), (70001,)]
    ps = _rand_params(shapes, g)
    mine = [p.clone().to(DEV).requires_grad_(True) for p in ps]
    ref = [p.clone().double().requires_grad_(True) for p in ps]
    opt = FusedAdam([{"params": mine, "lr": 1e-2}], max_grad_norm=1.0)
    opt_ref = torch.optim.Adam([{"params": ref, "lr": 1e-2}])
    scale = 65536.0
    inv = torch.tensor([1.0 / scale], device=DEV)
    for step in range(8):
        grads = [torch.randn(*s, generator=g) * (5.0 if step % 2 == 0 else 1e-3) for s in shapes]
        poison = step == 3
        for a, b, gr in zip(mine, ref, grads):
            a.grad = (gr * scale).to(DEV)
            b.grad = gr.double()
        if poison:
            mine[2].grad[17, 1] = float("inf")
        else:  # a GradScaler skips the whole step on inf
            torch.nn.utils.clip_grad_norm_(ref, 1.0)
            opt_ref.step()
        opt.step(inv_scale=inv)
        assert float(opt._coef[1]) == (1.0 if poison else 0.0)
    for a, b, s in zip(mine, ref, shapes):
        close(a, b.float(), 2e-6, f"param {s}")
    assert [int(v["step"]) for v in opt.state_dict()["state"].values()] == [7] * 4  # the skipped step does not count


def test_gradscaler_step_through_the_torch_amp_contract():
    """torch.amp.GradScaler.step(FusedAdam): scale / found_inf arrive as tensors (_step_supports_amp_scaling)."""
    from nerf_sys_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(12)
    shapes = [(33, 7), (1025,)]
    ps = _rand_params(shapes, g)
    mine = [p.clone().to(DEV).requires_grad_(True) for p in ps]
    ref = [p.clone().to(DEV).requires_grad_(True) for p in ps]
    opt = FusedAdam([{"params": mine, "lr": 1e-2}], max_grad_norm=0.5)
    opt_ref = torch.optim.Adam([{"params": ref, "lr": 1e-2}])
    scaler = torch.amp.GradScaler("cuda", init_scale=1024.0)
    for step in range(4):
        x = torch.randn(1025, generator=g).to(DEV)
        for params, o in ((mine, opt), (ref, opt_ref)):
            o.zero_grad(set_to_none=True)
            loss = (params[0].sum() * 0.3 + (params[1] * x).pow(2).sum()) * 10
            if o is opt:
                scaler.scale(loss).backward()
                scaler.step(o)
                scaler.update()
            else:
                loss.backward()
                torch.nn.utils.clip_grad_norm_(params, 0.5)
                o.step()
    for a, b in zip(mine, ref):
        close(a, b, 2e-6, "GradScaler.step(FusedAdam)")


def _container_steps(mode, steps=4, graph=False):
    from nerf_sys_b200 import ops
    from nerf_sys_b200.engine import GraphedTrainStep
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    from nerf_sys_b200.optim import bind_expert_liveness, build_optimizer
    m, st, cfg, mins, aabb = _big_scene(8, 2, 4, 14, seed=4)
    m.train()
    fast = mode != "eager"
    opt = build_optimizer(m, max_grad_norm=0.5)
    if fast:
        opt.enable_persistent_grads(min_numel=1 << 16)  # the test's 2^14-row tables are below the default threshold
        m.set_deferred_routing(True)
        bind_expert_liveness(opt, m)
    N, S = 1024, 32
    gen = torch.Generator().manual_seed(5)
    batches = []
    for i in range(steps):
        if i % 2 == 0:  # samples over every tile
            rays = O.aerial_rays(N, aabb, seed=60 + i)
        else:           # nadir rays over the tiles of experts 2 and 5 only: six experts receive nothing (config 4)
            rays = torch.zeros(N, 8)
            for j, k in enumerate((2, 5)):
                sl = slice(j * N // 2, (j + 1) * N // 2)
                rays[sl, 0] = -0.04
                rays[sl, 1:3] = st["centroids"][k, 1:3] + (torch.rand(N // 2, 2, generator=gen) - 0.5) * 0.2
                rays[sl, 3] = 1.0
                rays[sl, 6], rays[sl, 7] = 0.0, 0.5
        batches.append((rays.to(DEV), torch.rand(N, 3, generator=gen).to(DEV), torch.rand(N, S, generator=gen).to(DEV)))

    def loss_fn(rays, gt, u):
        rgb, *_ = render_rays(m, rays, ray_samples=S, u=u)
        return ops.mse_loss(rgb, gt)

    losses = []
    if graph:
        snap = {n: p.detach().clone() for n, p in m.named_parameters()}
        step = GraphedTrainStep(opt, loss_fn, batches[0], warmup=1)
        with torch.no_grad():
            for n, p in m.named_parameters():
                p.copy_(snap[n])
        for stt in opt.state.values():
            stt["exp_avg"].zero_()
            stt["exp_avg_sq"].zero_()
        opt._steps_dev.zero_()
        for b in batches:
            losses.append(step(*b).item())
    else:
        for b in batches:
            opt.zero_grad(set_to_none=True)
            l = loss_fn(*b)
            l.backward()
            opt.step()
            losses.append(l.item())
    m.check_routing()
    sd = opt.state_dict()
    names = [n for n, _ in m.named_parameters()]
    order = {id(p): i for i, gr in enumerate(opt.param_groups) for p in gr["params"]}
    steps_of = {}
    flat = [p for gr in opt.param_groups for p in gr["params"]]
    by_id = {id(p): n for n, p in m.named_parameters()}
    for idx, p in enumerate(flat):
        if idx in sd["state"]:
            steps_of[by_id[id(p)]] = int(sd["state"][idx]["step"])
    return {n: p.detach().clone() for n, p in m.named_parameters()}, losses, steps_of, opt, m


@pytest.mark.parametrize("graph", [False, True])
def test_deferred_routing_step_equals_the_host_read_step(graph):
    """The step without a host read (speculative pair capacity, gradients for every expert, device-side liveness flags,
    table gradients scattered into persistent buffers that the Adam kernel zeroes) leaves every parameter and every
    Adam step counter where the host-read step does - including the experts that received no sample in half of the
    steps (torch: p.grad is None -> skipped, meta_container.py:309-311) - eagerly and replayed from a CUDA graph."""
    ref_p, ref_l, ref_steps, _, _ = _container_steps("eager")
    got_p, got_l, got_steps, opt, m = _container_steps("deferred", graph=graph)
    for a, b in zip(got_l, ref_l):
        assert abs(a - b) <= 1e-5 * max(1e-3, abs(b)), (got_l, ref_l)
    for n in ref_p:
        close(got_p[n], ref_p[n], 1e-3 if n.endswith("hash_table") else 2e-4, n)
    # experts other than 2 and 5 stepped only on the even batches
    for n, s in ref_steps.items():
        assert got_steps.get(n, 0) == s, (n, got_steps.get(n), s)
    assert ref_steps["submodules.0.sigma_head.weight"] == 2 and ref_steps["submodules.2.sigma_head.weight"] == 4
    # persistent gradient buffers are all zero between steps
    for p, buf in opt._arena.values():
        assert float(buf.abs().max()) == 0.0
    assert len(opt._arena) == 8


def test_persistent_grad_buffer_is_adopted_without_a_copy_and_survives_accumulation():
    from nerf_sys_b200 import ops
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    from nerf_sys_b200.optim import build_optimizer
    m, st, cfg, mins, aabb = _big_scene(1, 1, 1, 14, seed=1)
    m.train()
    opt = build_optimizer(m, max_grad_norm=None).enable_persistent_grads(min_numel=1 << 16)  # 2^14-row table
    table = m.submodules[0].xyz_encoder.hash_table
    buf = opt._arena[id(table)][1]
    rays = O.aerial_rays(256, aabb, seed=3).to(DEV)
    u = torch.rand(256, 16, device=DEV)
    gt = torch.rand(256, 3, device=DEV)

    def bwd():
        rgb, *_ = render_rays(m, rays, ray_samples=16, u=u)
        ops.mse_loss(rgb, gt).backward()

    opt.zero_grad(set_to_none=True)
    bwd()
    assert table.grad.data_ptr() == buf.data_ptr()  # autograd adopted the alias: no 128 MiB copy, no zero-fill
    g1 = table.grad.clone()
    bwd()                                           # gradient accumulation without zero_grad: must not add the buffer to itself
    close(table.grad, 2 * g1, 1e-6, "accumulated gradient")
    opt.step()
    assert float(buf.abs().max()) == 0.0            # zeroed by the Adam kernel
    opt.zero_grad(set_to_none=True)
    bwd()
    opt.zero_grad(set_to_none=True)                 # a step that never ran: the buffer is cleaned on the next zero_grad
    assert float(buf.abs().max()) == 0.0
    opt.release_persistent_grads()


# ----------------------------------------------------------------------------- exchange plan, barrier payload
def test_p2p_plan_kernel_matches_peer_layout():
    """acn_p2p_plan (csrc/a2a.cu) against parallel.peer_layout for a fabricated 4-rank count table, every rank's view."""
    from nerf_sys_b200 import parallel as PL
    from nerf_sys_b200._lib import ACN_MAX_EXPERTS, check, i32, lib
    K, world, cap = 8, 4, 64 * 128
    g = torch.Generator().manual_seed(2)
    table = torch.randint(0, 700, (world, K + 1), generator=g).int()
    table[:, 3] = 0          # an expert nobody routes to
    table[1, 5] = 0
    table[:, K] = table[:, :K].sum(1)
    owner = PL.expert_owner(K, world)
    ex = PL.PeerExchange(rows_cap=cap)
    try:
        w = ACN_MAX_EXPERTS + 1
        for s in range(world):
            ex.sync[32 + s * w:32 + s * w + K + 1] = table[s].to(DEV)
        for rank in range(world):
            arr = (i32 * K)(*owner)
            ex.status.zero_()
            ex.g.fill_(7.0)
            check(lib().acn_p2p_plan(C.c_void_p(ex._ptr), world, rank, K, arr, ex.cap, C.c_void_p(ex.row0_dev.data_ptr()),
                                     C.c_void_p(ex.seg_own.data_ptr()), C.c_void_p(ex.point_own.data_ptr()), C.c_void_p(ex.live_own.data_ptr()),
                                     C.c_void_p(ex.g.data_ptr()), C.c_void_p(ex.status.data_ptr()),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream)), "acn_p2p_plan")
            start, tot, used, row0 = PL.peer_layout(table[:, :K].tolist(), owner, rank, 128)
            mine = [e for e in range(K) if owner[e] == rank]
            assert ex.row0_dev[:K].tolist() == row0
            segs = [start[e] for e in mine] + [used[rank]]
            assert ex.seg_own[:len(mine) + 1].tolist() == segs
            assert ex.live_own[:len(mine)].tolist() == [int(tot[e] > 0) for e in mine]
            pt = ex.point_own[:used[rank]].cpu()
            want = torch.full((used[rank],), -1, dtype=torch.int32)
            for e in mine:
                want[start[e]:start[e] + tot[e]] = torch.arange(start[e], start[e] + tot[e], dtype=torch.int32)
            assert torch.equal(pt, want)
            grow = ex.g[:used[rank]].cpu()
            assert (grow[want < 0] == 0).all() and (grow[want >= 0] == 7.0).all()  # padding rows of the gradient block zeroed
            assert int(ex.status.item()) == 0
        # overflow and truncation are reported, not overrun
        ex.sync[32 + K] = -1
        check(lib().acn_p2p_plan(C.c_void_p(ex._ptr), world, 0, K, arr, 1024, C.c_void_p(ex.row0_dev.data_ptr()),
                                 C.c_void_p(ex.seg_own.data_ptr()), C.c_void_p(ex.point_own.data_ptr()), C.c_void_p(ex.live_own.data_ptr()),
                                 None, C.c_void_p(ex.status.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)), "acn_p2p_plan")
        assert int(ex.status.item()) == 6 and int(ex.seg_own[2].item()) <= 1024
        with pytest.raises(RuntimeError):
            ex.check()
    finally:
        ex.close()


def test_barrier_epochs_and_payload_allreduce_on_one_rank_and_in_a_graph():
    from nerf_sys_b200 import parallel as PL
    ex = PL.PeerExchange(rows_cap=256)
    try:
        e0 = int(ex.sync[17].item())
        x = torch.arange(643, device=DEV, dtype=torch.float32)
        y = ex.allreduce_small(x, scale=0.5)
        close(y, x.cpu() * 0.5, 0, "payload all-reduce, world 1")
        assert int(ex.sync[17].item()) == e0 + 1 and int(ex.sync[0].item()) == e0 + 1  # device-side epoch, own flag raised
        out = torch.zeros_like(x)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            ex.barrier(payload=x, reduced=out)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            ex.barrier(payload=x, reduced=out, scale=2.0)
        e1 = int(ex.sync[17].item())
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize()
        assert int(ex.sync[17].item()) == e1 + 3  # a replayed barrier advances its own epoch
        close(out, x.cpu() * 2.0, 0, "graph-replayed payload")
    finally:
        ex.close()


def test_peer_exchange_step_in_a_cuda_graph_on_one_gpu():
    """The whole expert-parallel optimisation step (device-side plan, stores into the slab, liveness flags, slab
    reductions) captured in a CUDA graph with every slab local (world 1) == the eager host-read step."""
    from nerf_sys_b200 import ops, parallel as PL
    from nerf_sys_b200.engine import GraphedTrainStep
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    from nerf_sys_b200.optim import FusedAdam, bind_expert_liveness, build_optimizer
    N, S, steps = 768, 32, 3
    lrs = {"encoding": 1e-2, "sigma": 2e-3, "color": 2e-3, "background": 1e-3}
    gen = torch.Generator().manual_seed(15)
    _, _, _, _, aabb = _big_scene(8, 2, 4, 14, seed=6)
    batches = [(O.aerial_rays(N, aabb, seed=80 + i).to(DEV), torch.rand(N, 3, generator=gen).to(DEV), torch.rand(N, S, generator=gen).to(DEV))
               for i in range(steps)]

    def run(p2p):
        m, *_ = _big_scene(8, 2, 4, 14, seed=6)
        m.train()

        def loss_fn(rays, gt, u):
            rgb, *_ = render_rays(m, rays, ray_samples=S, u=u)
            return ops.mse_loss(rgb, gt)

        if not p2p:
            opt = build_optimizer(m, max_grad_norm=0.5)
            for b in batches:
                opt.zero_grad(set_to_none=True)
                loss_fn(*b).backward()
                opt.step()
            return {n: p.detach().clone() for n, p in m.named_parameters()}
        ex = PL.PeerExchange(rows_cap=3 * N * S)
        try:
            m.set_expert_parallel(True, exchange=ex)
            groups, shared = PL.owned_param_groups(m)
            opt = FusedAdam([{"params": gr["params"], "lr": lrs[n]} for n, gr in groups.items()], max_grad_norm=0.5,
                            persistent_grads=True).shard_grad_norm(None, shared, exchange=ex)
            bind_expert_liveness(opt, m)
            red = PL.SlabGradReducer(ex, shared)
            snap = {n: p.detach().clone() for n, p in m.named_parameters()}
            step = GraphedTrainStep(opt, loss_fn, batches[0], warmup=1, post_backward=red.wait)
            with torch.no_grad():
                for n, p in m.named_parameters():
                    p.copy_(snap[n])
            for stt in opt.state.values():
                stt["exp_avg"].zero_()
                stt["exp_avg_sq"].zero_()
            opt._steps_dev.zero_()
            for b in batches:
                step(*b)
            torch.cuda.synchronize()
            m.check_routing()
            out = {n: p.detach().clone() for n, p in m.named_parameters()}
            opt.release_persistent_grads()
            return out
        finally:
            m.set_expert_parallel(False)
            ex.close()

    ref, got = run(False), run(True)
    for n in ref:
        close(got[n], ref[n], 1e-3 if n.endswith("hash_table") else 2e-4, n)


# ----------------------------------------------------------------------------- first-order contract
def test_second_order_requests_fail_loudly():
    """algo='maml' asks autograd.grad(..., create_graph=True) (meta_core.py:54-59); the custom ops are first-order."""
    from nerf_sys_b200 import ops
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    m, st, cfg, mins, aabb = _big_scene(1, 1, 1, 12, seed=2)
    m.train()
    rays = O.aerial_rays(64, aabb, seed=5).to(DEV)
    rgb, *_ = render_rays(m, rays, ray_samples=8)
    loss = ops.mse_loss(rgb, torch.rand(64, 3, device=DEV))
    w = m.submodules[0].sigma_head.weight
    with pytest.raises(RuntimeError, match="first-order"):
        torch.autograd.grad(loss, [w], create_graph=True)
    rgb, *_ = render_rays(m, rays, ray_samples=8)
    loss = ops.mse_loss(rgb, torch.rand(64, 3, device=DEV))
    (g,) = torch.autograd.grad(loss, [w])  # the first-order request (fomaml / reptile) is served
    assert g.shape == w.shape and not g.requires_grad
    with pytest.raises(RuntimeError):
        ops.sgd_axpy(w, (g * w).sum() * w, 0.015)  # a gradient that still carries a graph


def test_split_barrier_arrive_then_wait_reduces_and_replays_in_a_graph():
    """acn_p2p_arrive / acn_p2p_wait (world 1: every slab local): one epoch per pair, the payload reduced at the wait, work
    enqueued between the two halves, replayable from a CUDA graph; SlabGradReducer(early=True) leaves the gradients scaled."""
    from nerf_sys_b200 import parallel as PL
    ex = PL.PeerExchange(rows_cap=256)
    try:
        e0 = int(ex.sync[17].item())
        x = torch.arange(643, device=DEV, dtype=torch.float32)
        out = torch.zeros_like(x)
        ex.arrive(x)
        busy = torch.randn(1 << 20, device=DEV).sin().sum()  # work between the two halves
        ex.wait(643, out, 0.25)
        close(out, x.cpu() * 0.25, 0, "split barrier payload")
        assert int(ex.sync[17].item()) == e0 + 1 and torch.isfinite(busy)
        ex.barrier()  # a full barrier afterwards keeps counting
        assert int(ex.sync[17].item()) == e0 + 2
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            ex.arrive(x)
            ex.wait(643, out, 1.0)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            ex.arrive(x)
            y = x * 2
            ex.wait(643, out, 2.0)
        e1 = int(ex.sync[17].item())
        for _ in range(3):
            gr.replay()
        torch.cuda.synchronize()
        assert int(ex.sync[17].item()) == e1 + 3
        close(out, x.cpu() * 2.0, 0, "graph-replayed split barrier")
    finally:
        ex.close()
