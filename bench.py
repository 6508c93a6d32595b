#!/usr/bin/env python
"""bench.py - train & render rays/s of the per-ray hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU PyTorch path

Workload at N=1 = BASELINE.json configs[1]: single-expert hash-grid NeRF (L16 F2 T2^20, 64-wide MLPs)
trained on synthetic 800x800 views, 8192-ray batch x 96 samples.  A step is one training pass of
the hot path over one batch: stratified sampling -> hash grid -> MLPs -> compositing -> MSE ->
backward -> clip(1.0) -> Adam on every tensor (incl. the dense 128 MiB table).  At N>1 every rank
owns one spatial expert and its own 8192-ray batch (the meta-training shard of SURVEY 8(e): tasks
never cross experts); the only collective is the NCCL all-reduce of the shared background head's
643 gradient floats - weak scaling.

`value` times K steps with inputs resident in HBM; `e2e` times the same step through the
reference-facing API (compute_mse_loss -> backward -> optimizer.step) with the batch copied from
pinned host memory and the loss read back every step.  L2: the per-step working set (table +
gradient + Adam moments = 512 MiB) exceeds the 126 MB L2, so there is no explicit flush.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_RAYS, S_TRAIN, LOG2_T = 8192, 96, 20
IMG, FOCAL, N_VIEWS = 800, 1111.0, 40
WORKLOAD = "config[1]: single-expert hash-grid NeRF training, synthetic 800x800 views, 8192-ray batch x 96 samples"

# algorithmic bytes / flops per sample point (DESIGN.md "Kernels")
ALGO = {
    "hash_fwd": ("hbm", 12 + 1024 + 128),       # xyz + 16 levels x 8 corners x 8 B + 32 features out
    "hash_bwd": ("hbm", 12 + 128 + 2 * 1024),   # xyz + d_enc + read-modify-write of the same rows
    "mlp_fwd": ("tensor", 26880),               # 13 440 MAC
    "mlp_bwd": ("tensor", 53760),               # dgrad + wgrad
    "composite_fwd": ("hbm", 24),
    "composite_bwd": ("hbm", 40),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "tensor": d["bf16_tflops_sustained"], "src": "MEASURED_PEAKS.json (hbm_gbs copy bandwidth; bf16_tflops_sustained for tensor-bound kernels)"}
    return {"hbm": 6650.0, "tensor": 1400.0, "src": "fallback (B200_PROFILING.md)"}


class Clocks:
    """nvidia-smi sampled every 200 ms during the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            uuid = str(torch.cuda.get_device_properties(device).uuid)
            sel = ["-i", uuid if uuid.startswith("GPU-") else "GPU-" + uuid]
        except Exception:
            sel = ["-i", str(device)]
        try:
            self.p = subprocess.Popen(["nvidia-smi", *sel, f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for n, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def dist_env():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    return rank, world, local


# ------------------------------------------------------------------------------------------
# scene, data
# ------------------------------------------------------------------------------------------
def build_scene(rank, world, device, seed=0, table_std=None):
    """One expert per rank: rank r owns tile r of a 1 x gy x gz grid over the global box."""
    from nerf_sys_b200.models.inr import MetaContainer
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    from oracle import acnerf_oracle as O  # scene constants / initial weights only (shared with the CPU baseline)
    gy, gz = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}.get(world, (1, world))
    aabb, cent, boxes = O.grid_scene(gy, gz)
    cfg = dict(O.HASH_CFG)
    cfg["log2_hashmap_size"] = LOG2_T
    box, c = boxes[rank], cent[rank:rank + 1]
    st = O.init_ngp_state(1, box, c, [box], cfg, seed=seed + 17 * rank, table_std=table_std)
    g = torch.Generator().manual_seed(1234)  # the shared background head starts identical on every rank
    for k in ("bg_mlp.0.weight", "bg_mlp.0.bias", "bg_mlp.2.weight", "bg_mlp.2.bias"):
        st[k] = (torch.rand(st[k].shape, generator=g) * 2 - 1) * (0.25 if "0." in k else 0.17)
    model = MetaContainer(1, c, box, "instant", boundary_margin=1.05, cluster_2d=True, use_bg_nerf=True, occ_conf={"use_occ": False},
                          expert_box_list=[SceneBox(box)], hidden=64, sigma_depth=2, color_depth=2, color_hidden=64,
                          dir_encoding="spherical", use_sigmoid_rgb=True, hash_enc_conf=cfg)
    model.load_state_dict(st, strict=True)
    return model.to(device), st, cfg, box


def ring_pose_in(box, i, n):
    from oracle import acnerf_oracle as O
    c2w = O.ring_pose(i, n, radius=0.4)
    ctr = (box[0] + box[1]) * 0.5
    half = (box[1] - box[0]) * 0.5
    c2w = c2w.clone()
    c2w[1, 3] = ctr[1] + c2w[1, 3] * half[1] / 0.5
    c2w[2, 3] = ctr[2] + c2w[2, 3] * half[2] / 0.5
    return c2w


def make_pool(box, device, batches, seed):
    """`batches` x 8192 rays drawn from random pixels of 40 synthetic 800x800 ring views of this rank's box."""
    from nerf_sys_b200 import ops
    g = torch.Generator().manual_seed(seed)
    per_view = (batches * N_RAYS + N_VIEWS - 1) // N_VIEWS
    out = []
    for v in range(N_VIEWS):
        rays = ops.raygen(IMG, IMG, FOCAL, FOCAL, IMG / 2, IMG / 2, True, ring_pose_in(box, v, N_VIEWS), box, device)
        _, valid = ops.clamp_near_far(rays, None)
        idx = torch.randint(0, IMG * IMG, (per_view * 2,), generator=g).to(device)
        idx = idx[valid[idx]][:per_view]
        out.append(rays[idx])
    rays = torch.cat(out)[torch.randperm(len(out) * per_view, generator=g).to(device)][:batches * N_RAYS]
    return rays.view(batches, N_RAYS, 8).contiguous()


@torch.no_grad()
def teacher_colors(pool, rank, world, device):
    """Targets = the same rays rendered once through a frozen teacher (seed 1, N(0,0.1) table): exactly representable."""
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    teacher, *_ = build_scene(rank, world, device, seed=1, table_std=0.1)
    teacher.eval()
    out = torch.empty(pool.shape[0], N_RAYS, 3, device=device)
    for b in range(pool.shape[0]):
        out[b] = render_rays(teacher, pool[b], ray_samples=S_TRAIN)[0].clamp(0, 1)
    del teacher
    torch.cuda.empty_cache()
    return out


def build_c5(device):
    """Config 5 scene: 8 experts on a 1x2x4 grid, 2^20-row tables (1 GiB), trained-like N(0,0.1) tables; replicated on every rank."""
    from nerf_sys_b200.models.inr import MetaContainer
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    from oracle import acnerf_oracle as O  # scene constants / initial weights only
    aabb, cent, boxes = O.grid_scene(2, 4)
    cfg = dict(O.HASH_CFG)
    cfg["log2_hashmap_size"] = LOG2_T
    st = O.init_ngp_state(8, aabb, cent, boxes, cfg, seed=3, table_std=0.1)
    model = MetaContainer(8, cent, aabb, "instant", boundary_margin=1.05, cluster_2d=True, use_bg_nerf=True, occ_conf={"use_occ": False},
                          expert_box_list=[SceneBox(b) for b in boxes], hidden=64, sigma_depth=2, color_depth=2, color_hidden=64,
                          dir_encoding="spherical", use_sigmoid_rgb=True, hash_enc_conf=cfg)
    model.load_state_dict(st, strict=True)
    return model.to(device).eval(), aabb


@torch.no_grad()
def render_c5(device, world, barrier, max_over_ranks, frames=4):
    """1920x1080 x 64 samples through the 8-expert container, image rows sharded over the ranks, frame gathered on every rank."""
    from nerf_sys_b200 import parallel as PL
    from nerf_sys_b200.nerfs.ray_rendering import render_image
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    from oracle import acnerf_oracle as O
    model, aabb = build_c5(device)
    H, W, f = 1080, 1920, 1480.0
    sb = SceneBox(aabb.to(device))

    def frame(i):
        c2w = O.ring_pose(i, 20, radius=0.9)
        kw = dict(H=H, W=W, fx=f, fy=f, cx=W / 2, cy=H / 2, c2w=c2w, scene_box=sb, ray_samples=64)
        return PL.render_image_sharded(lambda r0, r1: render_image(model, row_range=(r0, r1), **kw), H, W)

    frame(0)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(frames):
        rgb, depth, acc = frame(1 + i)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    cover = float((acc > 0.01).float().mean().item())
    del model
    torch.cuda.empty_cache()
    return {"value": frames * H * W / (ms / 1e3), "unit": "rays/s", "ms_per_frame": ms / frames, "coverage": cover,
            "config": f"config[4]: 1920x1080 x 64 samples, 8 experts (soft routing, margin 1.05), eval, rows sharded over {world} GPU(s), "
                      "frame all-gathered"}


def train_c3(device, rank, world, barrier, max_over_ranks, steps=10, warmup=3):
    """Config 3: the 8-expert container (soft routing, margin 1.05) trained on 8192 aerial rays per GPU x 96 samples: full-container
    forward + backward + global clip + Adam (the runtime_adapt step, runtime_adapt.py:249-268).  world > 1: experts sharded in blocks
    over the ranks; every routed sample travels to its expert's owner and its [rgb,sigma] back (and the gradient rows in the backward
    pass) - measured twice: through NVLink peer memory inside the pack/combine kernels (the product path) and through NCCL
    all-to-alls; background-head gradients and the clip norm are all-reduced."""
    from nerf_sys_b200 import parallel as PL
    from nerf_sys_b200.nerfs.losses import compute_mse_loss
    from nerf_sys_b200.optim import FusedAdam
    from oracle import acnerf_oracle as O  # scene constants / synthetic rays only
    nb = 4
    lrs = {"encoding": 1e-2, "sigma": 2e-3, "color": 2e-3, "background": 1e-3}

    def run(path):
        model, aabb = build_c5(device)
        model.train()
        ex = None
        if world > 1:
            if path == "p2p":
                ex = PL.PeerExchange(rows_cap=2 * N_RAYS * S_TRAIN * world + 16 * 128)
            model.set_expert_parallel(True, exchange=ex)
        groups, shared = PL.owned_param_groups(model)
        opt = FusedAdam([{"params": gr["params"], "lr": lrs[n]} for n, gr in groups.items()], max_grad_norm=1.0).shard_grad_norm(None, shared)
        rays = [O.aerial_rays(N_RAYS, aabb, seed=1000 + 16 * rank + b).to(device) for b in range(nb)]
        g = torch.Generator().manual_seed(900 + rank)
        gts = [torch.rand(N_RAYS, 3, generator=g).to(device) for _ in range(nb)]

        def step(i):
            opt.zero_grad(set_to_none=True)
            loss = compute_mse_loss(P, model, {"rays": rays[i % nb], "rgbs": gts[i % nb]})
            (loss / world).backward()  # the step's loss is the mean over the world x 8192 rays of all ranks
            PL.allreduce_shared_grads(shared, average=False, world=world)  # eager step: enqueued while the table scatter runs
            opt.step()
            return loss

        for i in range(max(warmup, 3)):
            step(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            loss = step(i)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1))
        lv = float(loss.item())
        if ex is not None:
            model.set_expert_parallel(False)
            ex.close()
        del model, opt
        torch.cuda.empty_cache()
        return ms, lv

    ms, lv = run("p2p")
    out = {"value": world * N_RAYS * steps / (ms / 1e3), "unit": "rays/s", "ms_per_step": ms / steps, "loss": lv,
           "config": f"config[2]: 8 experts (2^20-row tables), {N_RAYS} aerial rays/GPU x {S_TRAIN} samples, soft routing, fwd+bwd+clip+Adam, "
                     + ("all 8 experts on one GPU" if world == 1 else
                        f"experts sharded {8 / world:g} per GPU over {world} GPUs; sample rows (24 B), results (16 B) and gradient rows (16 B) "
                        "written / read in the owner's memory over NVLink by the pack and combine kernels, eager launches")}
    if world > 1:
        ms2, lv2 = run("nccl")
        out["nccl_all_to_all"] = {"value": world * N_RAYS * steps / (ms2 / 1e3), "unit": "rays/s", "ms_per_step": ms2 / steps, "loss": lv2}
    return out


def adapt_c4(device, steps=10, warmup=3):
    """Config 4 (one GPU): incremental adaptation of the 8-expert container on a new aerial batch whose samples fall in experts
    {2, 5} only (nadir rays over those two tiles); the other six experts get no gradient, so their parameters and Adam state are not
    touched (asserted bit-unchanged) and the step costs two experts' worth of scatter and optimiser traffic."""
    from nerf_sys_b200.nerfs.losses import compute_mse_loss
    from nerf_sys_b200.optim import build_optimizer
    model, aabb = build_c5(device)
    model.train()
    opt = build_optimizer(model, max_grad_norm=1.0)
    cent = model.centroids.detach().cpu()
    g = torch.Generator().manual_seed(77)
    nb, batches = 4, []
    for b in range(nb):
        rays = torch.zeros(N_RAYS, 8)
        for i, k in enumerate((2, 5)):
            sl = slice(i * N_RAYS // 2, (i + 1) * N_RAYS // 2)
            rays[sl, 0] = float(aabb[0, 0]) + 0.01
            rays[sl, 1:3] = cent[k, 1:3] + (torch.rand(N_RAYS // 2, 2, generator=g) - 0.5) * 0.3
            rays[sl, 3] = 1.0  # straight down: the YZ position (the routing key) never changes along the ray
            rays[sl, 6], rays[sl, 7] = 0.0, float(aabb[1, 0] - aabb[0, 0]) - 0.02
        batches.append((rays.to(device), torch.rand(N_RAYS, 3, generator=g).to(device)))
    frozen = {n: p.detach().clone() for n, p in model.named_parameters() if n.startswith("submodules.") and int(n.split(".")[1]) not in (2, 5)}

    def step(i):
        opt.zero_grad(set_to_none=True)
        loss = compute_mse_loss(P, model, {"rays": batches[i % nb][0], "rgbs": batches[i % nb][1]})
        loss.backward()
        opt.step()
        return loss

    for i in range(max(warmup, 3)):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        loss = step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    named = dict(model.named_parameters())
    unchanged = all(torch.equal(v, named[n].detach()) for n, v in frozen.items())
    no_state = all(len(opt.state.get(named[n], {})) == 0 for n in frozen)
    assert unchanged and no_state, "experts outside {2,5} must stay bit-unchanged, without optimiser state"
    out = {"value": N_RAYS * steps / (ms / 1e3), "unit": "rays/s", "ms_per_step": ms / steps, "loss": float(loss.item()),
           "frozen_experts_bit_unchanged": bool(unchanged and no_state),
           "config": f"config[3]: 8-expert container, {N_RAYS} nadir rays x {S_TRAIN} samples over the tiles of experts 2 and 5, fwd+bwd+clip+Adam; "
                     "six experts receive no gradient and no optimiser step"}
    del model, opt
    torch.cuda.empty_cache()
    return out


class P:  # the argparse namespace fields compute_mse_loss reads (common/args.py)
    ray_samples, chunk_points, color_space = S_TRAIN, 4456448, "identity"


def allreduce_bg(model, world):
    """Shared background head: NCCL all-reduce (mean) of its 643 gradient floats."""
    if world == 1:
        return
    import torch.distributed as dist
    gs = [p.grad for p in model.bg_mlp.parameters() if p.grad is not None]
    flat = torch.cat([g.reshape(-1) for g in gs])
    dist.all_reduce(flat)
    flat /= world
    o = 0
    for g in gs:
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()


def train_step(model, opt, rays, rgbs, world, reducer=None):
    from nerf_sys_b200.nerfs.losses import compute_mse_loss
    opt.zero_grad(set_to_none=True)
    loss = compute_mse_loss(P, model, {"rays": rays, "rgbs": rgbs})
    loss.backward()
    if reducer is not None:
        reducer.wait()
    else:
        allreduce_bg(model, world)
    opt.step()
    return loss


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's PyTorch path as restated by the oracle (bit-identical on the goldens)
# ------------------------------------------------------------------------------------------
def cpu_train_rays_per_s(n_rays, steps, warmup, seed=0):
    from oracle import acnerf_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    aabb, cent, boxes = O.grid_scene(1, 1)
    cfg = dict(O.HASH_CFG)
    cfg["log2_hashmap_size"] = LOG2_T
    st = O.init_ngp_state(1, aabb, cent, boxes, cfg, seed=seed)
    params = [v.requires_grad_(True) for k, v in st.items() if ("weight" in k or "bias" in k or "hash_table" in k)]
    opt = torch.optim.Adam(params, lr=2e-3)
    g = torch.Generator().manual_seed(7)
    times = []
    for i in range(warmup + steps):
        rays = O.aerial_rays(n_rays, aabb, seed=100 + i)
        u, gt = torch.rand(n_rays, S_TRAIN, generator=g), torch.rand(n_rays, 3, generator=g)
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        rgb, *_ = O.render_rays(rays, st, [boxes[0][0]], cfg, S=S_TRAIN, margin=1.05, u=u)
        loss = O.mse_loss(rgb, gt, "identity")
        loss.backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return n_rays * len(times) / sum(times), sum(times) / len(times)


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    n = 2048
    t0 = time.perf_counter()
    rps, sec = cpu_train_rays_per_s(n, args.steps, args.warmup)
    cores = os.cpu_count() or 1
    sample = f"{n}-ray x {S_TRAIN}-sample slices of the 8192-ray batch, fwd+bwd+clip+Adam, torch {torch.__version__} CPU fp32"
    print(json.dumps({
        "impl": "reference", "metric": "train_rays_per_sec", "value": rps, "unit": "rays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": WORKLOAD, "l2": "n/a (CPU)"},
        "cpu_baseline": {"value": rps, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rps, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = its pure-PyTorch path on the host cores via oracle/ (the reference is Python and does not exist on the GPU "
                "box; the oracle is bit-identical to it on tests/golden)", "wall_s": time.perf_counter() - t0}))


# ------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-render", action="store_true")
    ap.add_argument("--no-c3", action="store_true", help="skip the 8-expert container training measurement (config[2])")
    ap.add_argument("--no-graph", action="store_true", help="drive every step from Python instead of replaying a CUDA graph")
    ap.add_argument("--bwd-chunks", type=int, default=1, help="chunks of the pipelined backward (1 = MLP backward, then table scatter)")
    ap.add_argument("--mlp-math", default="f16tc", choices=["f16tc", "fp32"],
                    help="expert-MLP arithmetic: tcgen05 fp16/fp32-accumulate (default) or the exact fp32 CUDA-core mode")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    from nerf_sys_b200 import ops
    from nerf_sys_b200.optim import build_optimizer
    rank, world, local = dist_env()
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 under torch.distributed.run"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ops.device_ok(local)
    ops.set_mlp_math(args.mlp_math)
    ops.set_bwd_chunks(args.bwd_chunks)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    K, W = args.steps, max(args.warmup, 3)

    model, st, cfg, box = build_scene(rank, world, dev)
    model.train()
    opt = build_optimizer(model, max_grad_norm=1.0)
    nb = min(32, K + W)
    pool = make_pool(box, dev, nb, seed=10 + rank)
    rgbs = teacher_colors(pool, rank, world, dev)
    pool_h, rgbs_h = pool.cpu().pin_memory(), rgbs.cpu().pin_memory()

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return t.item()

    from nerf_sys_b200.nerfs.losses import compute_mse_loss
    from nerf_sys_b200 import parallel as PL
    # shared background head: its gradients are all-reduced (mean) on a side stream underneath the expert backward
    reducer = PL.SharedGradReducer(list(model.bg_mlp.parameters())) if world > 1 else None
    use_graph = not args.no_graph
    if use_graph:
        from nerf_sys_b200.engine import GraphedTrainStep
        gstep = GraphedTrainStep(opt, lambda r, c: compute_mse_loss(P, model, {"rays": r, "rgbs": c}), (pool[0], rgbs[0]), warmup=W,
                                 post_backward=(reducer.wait if reducer is not None else None))

        def step(r, c):
            return gstep(r, c)
    else:
        def step(r, c):
            return train_step(model, opt, r, c, world, reducer)

    # ---- resident-input arm ------------------------------------------------------------
    for i in range(W):
        step(pool[i % nb], rgbs[i % nb])
    barrier()
    clocks = Clocks(local)
    l0 = ops.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(K):
        loss = step(pool[(W + i) % nb], rgbs[(W + i) % nb])
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = ops.launches() - l0
    final_loss = float(loss.item())

    # ---- end-to-end arm: host batches in, loss out, every step -------------------------------
    # Every step: pinned host batch -> device (async), the step, the step's loss -> pinned host (async).  The host reads the
    # loss of step i-1 while step i runs (what a training loop that logs its loss does), so the device never waits for the
    # host; all K losses are read inside the timed region, the last one behind a final synchronisation.
    loss_h = torch.zeros(K, dtype=torch.float32).pin_memory()
    evs = [torch.cuda.Event() for _ in range(K)]
    losses_read = []
    barrier()
    e0.record()
    for i in range(K):
        if use_graph:  # pinned host batch straight into the captured step's input buffers (one H2D copy per tensor)
            l = gstep(pool_h[(W + i) % nb], rgbs_h[(W + i) % nb])
        else:
            l = step(pool_h[(W + i) % nb].to(dev, non_blocking=True), rgbs_h[(W + i) % nb].to(dev, non_blocking=True))
        loss_h[i:i + 1].copy_(l.detach().reshape(1), non_blocking=True)
        evs[i].record()
        if i > 0:
            evs[i - 1].synchronize()
            losses_read.append(float(loss_h[i - 1]))
    evs[K - 1].synchronize()
    losses_read.append(float(loss_h[K - 1]))
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    assert len(losses_read) == K and all(v == v for v in losses_read), "every step's loss must reach the host"
    clk = clocks.stop()

    # ---- per-kernel device time: the same K steps driven eagerly with an event pair around every C-ABI call
    # (events cannot be recorded between the nodes of a replayed graph)
    ops.set_bwd_chunks(1)  # unpipelined, so that the MLP backward and the table scatter are timed one at a time
    timing = ops.kernel_timing(True)
    for i in range(K):
        train_step(model, opt, pool[(W + i) % nb], rgbs[(W + i) % nb], world, reducer)
    barrier()
    ops.kernel_timing(False)
    ops.set_bwd_chunks(args.bwd_chunks)

    # ---- per-call device time -> roofline ------------------------------------------------
    pk = peaks()
    P_step = N_RAYS * S_TRAIN
    kern = {}
    for name, evs in timing.items():
        tot = sum(a.elapsed_time(b) for a, b in evs)
        kern[name] = {"ms_per_step": tot / K, "calls_per_step": len(evs) / K}
        if name in ALGO:
            bound, per_pt = ALGO[name]
            sec = tot / len(evs) / 1e3
            ach = per_pt * P_step / sec / (1e9 if bound == "hbm" else 1e12)
            kern[name].update({"bound": bound, "achieved": ach, "unit": "GB/s" if bound == "hbm" else "TFLOP/s", "frac": ach / pk[bound]})
    top = max((n for n in kern if n in ALGO), key=lambda n: kern[n]["ms_per_step"])
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(top)
    roofline = {"kernel": top, "bound": kern[top]["bound"], "achieved": kern[top]["achieved"], "peak": pk[kern[top]["bound"]],
                "unit": kern[top]["unit"], "frac": kern[top]["frac"], "traffic": traffic, "peak_source": pk["src"],
                "algorithmic_per_point": ALGO[top][1]}

    # ---- render throughput (no_grad, eval), one 800x800 view per measurement ---------------
    render = None
    if not args.no_render:
        from nerf_sys_b200.nerfs.ray_rendering import render_image
        from nerf_sys_b200.nerfs.scene_box import SceneBox
        model.eval()
        sb = SceneBox(box.to(dev))
        for v in range(2):
            render_image(model, H=IMG, W=IMG, fx=FOCAL, fy=FOCAL, cx=IMG / 2, cy=IMG / 2, c2w=ring_pose_in(box, v, N_VIEWS), scene_box=sb,
                         ray_samples=64)
        barrier()
        e0.record()
        nv = 5
        for v in range(nv):
            render_image(model, H=IMG, W=IMG, fx=FOCAL, fy=FOCAL, cx=IMG / 2, cy=IMG / 2, c2w=ring_pose_in(box, v + 2, N_VIEWS),
                         scene_box=sb, ray_samples=64)
        e1.record()
        barrier()
        rms = max_over_ranks(e0.elapsed_time(e1))
        render = {"value": world * nv * IMG * IMG / (rms / 1e3), "unit": "rays/s", "config": "800x800 x 64 samples, eval, no_grad, raygen on device",
                  "ms_per_frame": rms / nv}
        model.train()

    # secondary measurements: a failure here is recorded in its own field and must not cost the headline line (every rank runs
    # the same code, so an exception is raised on all of them alike)
    def guarded(fn, *a):
        try:
            return fn(*a)
        except Exception as e:  # noqa: BLE001
            torch.cuda.synchronize()
            return {"error": f"{type(e).__name__}: {e}"[:300]}

    c5 = guarded(render_c5, dev, world, barrier, max_over_ranks) if not args.no_render else None
    c3 = guarded(train_c3, dev, rank, world, barrier, max_over_ranks) if not args.no_c3 else None
    c4 = guarded(adapt_c4, dev) if (world == 1 and not args.no_c3) else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n = 2048
        rps, sec = cpu_train_rays_per_s(n, 6, 1)
        cpu = {"value": rps, "unit": "rays/s", "cores": os.cpu_count() or 1, "kind": "port",
               "sample": f"6 steps of a {n}-ray x {S_TRAIN}-sample slice of the same batch shape (fwd+bwd+clip+Adam, {sec:.2f} s/step), "
                         f"torch {torch.__version__} CPU fp32, all host threads"}
    if rank == 0:
        bytes_in = pool_h[0].numel() * 4 + rgbs_h[0].numel() * 4
        print(json.dumps({
            "metric": "train_rays_per_sec", "value": world * N_RAYS * K / (ms / 1e3), "unit": "rays/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16" if args.mlp_math == "f16tc" else "f32",
            "data": "synthetic", "config": {"workload": WORKLOAD, "rays_per_gpu": N_RAYS, "samples": S_TRAIN, "experts_per_gpu": 1,
                                            "table": "2^20 x 16 levels x 2 fp32 = 128 MiB", "mlp_math": ("tcgen05 kind::f16, fp16 operands, fp32 accumulate in TMEM (hash grid, compositing, loss, Adam in fp32)"
                                                         if args.mlp_math == "f16tc" else "fp32 FFMA (exact mode)"),
                                            "l2": "per-step working set 512 MiB (table+grad+Adam moments) > 126 MB L2, no flush",
                                            "launch": "CUDA graph replay of zero_grad+fwd+bwd+clip+Adam" if use_graph else "eager Python launches",
                                            "kernel_timing": "separate eager pass of the same K steps, CUDA-event pair per C-ABI call, backward unpipelined",
                                            "backward": (f"{args.bwd_chunks}-chunk software pipeline: table scatter of chunk c on a side stream beside the "
                                                         "MLP backward of chunk c+1") if args.bwd_chunks > 1 else "MLP backward, then table scatter",
                                            "collective": "none" if world == 1 else "nccl all-reduce of 643 bg-head grads on a side stream under the expert backward"},
            "clocks": clk, "e2e": {"value": world * N_RAYS * K / (ms_e2e / 1e3), "unit": "rays/s", "h2d_bytes_per_step": bytes_in,
                                   "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / K,
                                   "note": "loss of step i-1 read on the host while step i runs; all K losses read inside the timed region"},
            "gpu_launches": launches, "roofline": roofline, "kernels": kern, "render": render, "render_c5": c5, "train_c3": c3, "adapt_c4": c4, "cpu_baseline": cpu,
            "final_loss": final_loss}))
    if world > 1:
        # Leave without tearing NCCL down: the captured graph still references the communicator's kernels and
        # destroy_process_group() was seen to hang behind it.  All ranks meet, flush, and exit 0.
        sys.stdout.flush()
        sys.stderr.flush()
        torch.cuda.synchronize()
        torch.distributed.barrier()
        os._exit(0)


if __name__ == "__main__":
    main()
