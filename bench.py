#!/usr/bin/env python
"""bench.py - train & render rays/s of the per-ray hot path on the 8-expert scene (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W                    # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU PyTorch path, same config

Headline workload at every N = BASELINE.json configs[2]: the city-scale 8-expert container (2x4 YZ grid, one
2^20-row x 16-level fp32 hash table = 128 MiB per expert, 64-wide MLPs, soft routing with margin 1.05) trained on
8192 synthetic aerial rays PER GPU x 96 samples.  A step is one pass of the hot path over one batch: stratified
sampling -> routing + pack -> hash grids -> MLPs -> combine -> compositing -> MSE -> backward -> global clip(1.0) ->
Adam on every tensor that received a gradient (the runtime_adapt step, runtime_adapt.py:249-268).
  N = 1 : all 8 experts on the one GPU.
  N > 1 : experts sharded in blocks over the ranks, every rank routes its own 8192 rays; each routed sample travels
          to its expert's owner and its [rgb,sigma] back (gradient rows in the backward pass) through NVLink peer
          memory written / read inside the pack and combine kernels; background-head gradients and the clip norm are
          reduced on slab barriers.  No NCCL kernel inside the step.  Weak scaling (global batch = N x 8192 rays).
The step is a CUDA graph at every N: no host read, no Python between kernels.

`value` times K steps with inputs resident in HBM; `e2e` times the same step through the reference-facing API
(compute_mse_loss -> backward -> optimizer.step, graph-replayed) with every batch copied from pinned host memory and
every loss read back.  L2: the per-step working set (tables + gradients + Adam moments: 4 GiB at N=1, 512 MiB at N=8)
exceeds the 126 MB L2, so there is no explicit flush.
Secondary records in the same line: render_c5 (configs[4], 1080p, rows sharded), train_c2 (configs[1], one expert per
GPU), render_800, adapt_c4 (configs[3]), train_c1 (configs[0], PE-MLP), parity_check (N > 1).
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

N_RAYS, S_TRAIN, LOG2_T, K_EXPERTS = 8192, 96, 20, 8
IMG, FOCAL, N_VIEWS = 800, 1111.0, 40
LRS = {"encoding": 1e-2, "sigma": 2e-3, "color": 2e-3, "background": 1e-3}  # common/args.py:115-125
TABLE_ELEMS = 16 * (1 << LOG2_T) * 2
MLP_ELEMS = 13715


def workload_config(world):
    """The `config` object both arms print (the reference arm runs exactly this on the host cores)."""
    return {
        "workload": "config[2]: city-scale 8-expert decomposition (2x4 grid, soft routing margin 1.05), synthetic aerial rays, "
                    f"{N_RAYS} rays per GPU x {S_TRAIN} samples, fwd + bwd + global clip(1.0) + per-group Adam over every touched tensor",
        "experts": K_EXPERTS, "rays_per_gpu": N_RAYS, "samples": S_TRAIN, "gpus": world,
        "table": "2^20 rows x 16 levels x 2 fp32 = 128 MiB per expert (1 GiB for the scene)",
        "expert_placement": "all 8 experts on one GPU" if world == 1 else
                            f"{K_EXPERTS / world:g} expert(s) per GPU, routed samples exchanged with the owner (sample all-to-all)",
        "l2": "per-step working set (tables + gradients + Adam moments) >= 512 MiB per GPU > 126 MB L2, no flush",
    }


# per-unit algorithmic bytes / flops (DESIGN.md "Kernels"); unit = pair (routed sample), point, ray, elem
ALGO = {
    "hash_fwd": ("hbm", 12 + 1024 + 128, "pair"),       # xyz + 16 levels x 8 corners x 8 B + 32 features out
    "hash_bwd": ("hbm", 12 + 128 + 2 * 1024, "pair"),   # xyz + d_enc + read-modify-write of the same rows
    "mlp_fwd": ("tensor", 26880, "pair"),               # 13 440 MAC
    "mlp_bwd": ("tensor", 53760, "pair"),               # dgrad + wgrad
    "fused_fwd": ("hbm", 12 + 1024 + 16, "pair"),       # gather + MLP in one kernel: the encoding never exists in HBM
    "route_pack": ("hbm", 16 + 8, "point"),             # 16 B read per point, 8 B per pair written
    "composite_fwd": ("hbm", 24, "point"),
    "composite_bwd": ("hbm", 40, "point"),
    "adam": ("hbm", 32, "elem"),                        # p, m, v read + written, g read and zeroed
    "sumsq": ("hbm", 4, "elem"),
    "a2a_push_rows": ("hbm", 28 + 8, "pair"),
    "a2a_pull_combine": ("hbm", 16 + 8 + 16, "pair"),
    "a2a_push_grad": ("hbm", 16 + 8 + 16, "pair"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "tensor": d["bf16_tflops_sustained"],
                "src": "MEASURED_PEAKS.json (hbm_gbs copy bandwidth; bf16_tflops_sustained for tensor-bound kernels timed inside a step)"}
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


class P:  # the argparse namespace fields compute_mse_loss reads (common/args.py)
    ray_samples, chunk_points, color_space = S_TRAIN, 4456448, "identity"


# ------------------------------------------------------------------------------------------
# scenes, data
# ------------------------------------------------------------------------------------------
def build_container(device, seed=3, table_std=0.1, K=K_EXPERTS):
    """The 8-expert scene: experts on a 1x2x4 grid, 2^20-row tables (1 GiB), trained-like N(0,0.1) tables."""
    from nerf_sys_b200 import synthetic as SY
    from nerf_sys_b200.models.inr import MetaContainer
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    gy, gz = {8: (2, 4), 4: (2, 2), 2: (1, 2), 1: (1, 1)}[K]
    aabb, cent, boxes = SY.grid_scene(gy, gz)
    cfg = dict(SY.HASH_CFG)
    cfg["log2_hashmap_size"] = LOG2_T
    st = SY.init_ngp_state(K, aabb, cent, boxes, cfg, seed=seed, table_std=table_std)
    model = MetaContainer(K, cent, aabb, "instant", boundary_margin=1.05, cluster_2d=True, use_bg_nerf=True, occ_conf={"use_occ": False},
                          expert_box_list=[SceneBox(b) for b in boxes], hidden=64, sigma_depth=2, color_depth=2, color_hidden=64,
                          dir_encoding="spherical", use_sigmoid_rgb=True, hash_enc_conf=cfg)
    model.load_state_dict(st, strict=True)
    del st
    return model.to(device), aabb


def build_scene(rank, world, device, seed=0, table_std=None):
    """configs[1]: one expert per rank, rank r owns tile r of a 1 x gy x gz grid over the global box."""
    from nerf_sys_b200 import synthetic as SY
    from nerf_sys_b200.models.inr import MetaContainer
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    gy, gz = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}.get(world, (1, world))
    aabb, cent, boxes = SY.grid_scene(gy, gz)
    cfg = dict(SY.HASH_CFG)
    cfg["log2_hashmap_size"] = LOG2_T
    box, c = boxes[rank], cent[rank:rank + 1]
    st = SY.init_ngp_state(1, box, c, [box], cfg, seed=seed + 17 * rank, table_std=table_std)
    g = torch.Generator().manual_seed(1234)  # the shared background head starts identical on every rank
    for k in ("bg_mlp.0.weight", "bg_mlp.0.bias", "bg_mlp.2.weight", "bg_mlp.2.bias"):
        st[k] = (torch.rand(st[k].shape, generator=g) * 2 - 1) * (0.25 if "0." in k else 0.17)
    model = MetaContainer(1, c, box, "instant", boundary_margin=1.05, cluster_2d=True, use_bg_nerf=True, occ_conf={"use_occ": False},
                          expert_box_list=[SceneBox(box)], hidden=64, sigma_depth=2, color_depth=2, color_hidden=64,
                          dir_encoding="spherical", use_sigmoid_rgb=True, hash_enc_conf=cfg)
    model.load_state_dict(st, strict=True)
    return model.to(device), st, cfg, box


def ring_pose_in(box, i, n):
    from nerf_sys_b200 import synthetic as SY
    c2w = SY.ring_pose(i, n, radius=0.4)
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
def teacher_colors(teacher, pool):
    """Targets = the same rays rendered once (eval mode) through a frozen teacher: exactly representable."""
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    teacher.eval()
    out = torch.empty(pool.shape[0], pool.shape[1], 3, device=pool.device)
    for b in range(pool.shape[0]):
        out[b] = render_rays(teacher, pool[b], ray_samples=S_TRAIN)[0].clamp(0, 1)
    return out


def aerial_pool(aabb, rank, nb, device):
    from nerf_sys_b200 import synthetic as SY
    return torch.stack([SY.aerial_rays(N_RAYS, aabb, seed=1000 + 16 * rank + b) for b in range(nb)]).to(device)


class Ctx:
    """Per-process measurement context: device, ranks, collectives used around (never inside) timed regions."""

    def __init__(self, dev, rank, world):
        self.dev, self.rank, self.world = dev, rank, world

    def barrier(self):
        if self.world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.world == 1:
            return ms
        t = torch.tensor([ms], device=self.dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return t.item()


def timed(ctx, fn, n):
    """n calls of fn between a barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.barrier()
    e0.record()
    out = None
    for i in range(n):
        out = fn(i)
    e1.record()
    ctx.barrier()
    return ctx.max_over_ranks(e0.elapsed_time(e1)), out


# ------------------------------------------------------------------------------------------
# configs[2]: the 8-expert training step (headline)
# ------------------------------------------------------------------------------------------
class ContainerTrainer:
    """The 8-expert container, its optimiser, and - at N > 1 - the peer-memory exchange; `graph()` captures the step."""

    def __init__(self, ctx, active=None, table_std=0.1, max_grad_norm=1.0, rows_cap=None):
        from nerf_sys_b200 import parallel as PL
        from nerf_sys_b200.optim import FusedAdam, bind_expert_liveness
        self.ctx, self.PL = ctx, PL
        self.model, self.aabb = build_container(ctx.dev, table_std=table_std)
        self.model.train()
        self.ex = self.reducer = None
        if ctx.world > 1:
            cap = rows_cap or 3 * N_RAYS * S_TRAIN  # 3x the balanced load of one owner (aerial rays are uniform over the tiles)
            self.ex = PL.PeerExchange(rows_cap=cap)
            self.model.set_expert_parallel(True, exchange=self.ex)
            groups, shared = PL.owned_param_groups(self.model)
            self.opt = FusedAdam([{"params": gr["params"], "lr": LRS[n]} for n, gr in groups.items()], max_grad_norm=max_grad_norm,
                                 persistent_grads=True).shard_grad_norm(None, shared, exchange=self.ex)
            self.reducer = PL.SlabGradReducer(self.ex, shared, scale=1.0)
        else:
            self.model.set_deferred_routing(True)
            groups = self.model.get_param_groups()
            self.opt = FusedAdam([{"params": gr["params"], "lr": LRS[n]} for n, gr in groups.items()], max_grad_norm=max_grad_norm,
                                 persistent_grads=True)
        bind_expert_liveness(self.opt, self.model)
        self.gstep = None

    def loss_fn(self, rays, rgbs):
        from nerf_sys_b200.nerfs.losses import compute_mse_loss
        loss = compute_mse_loss(P, self.model, {"rays": rays, "rgbs": rgbs})
        # the step's loss is the mean over the world x 8192 rays of all ranks
        return loss / self.ctx.world if self.ctx.world > 1 else loss

    def eager_step(self, rays, rgbs):
        self.opt.zero_grad(set_to_none=True)
        loss = self.loss_fn(rays, rgbs)
        loss.backward()
        if self.reducer is not None:
            self.reducer.wait()
        self.opt.step()
        return loss.detach()

    def graph(self, rays, rgbs, warmup):
        from nerf_sys_b200.engine import GraphedTrainStep
        self.gstep = GraphedTrainStep(self.opt, self.loss_fn, (rays, rgbs), warmup=warmup,
                                      post_backward=(self.reducer.wait if self.reducer is not None else None))
        return self.gstep

    def check(self):
        self.model.check_routing()

    def close(self):
        self.gstep = None
        self.opt.release_persistent_grads()
        if self.ex is not None:
            self.model.set_expert_parallel(False)
            self.ex.close()
        self.model = self.opt = None
        torch.cuda.empty_cache()


def pair_stats(model, rays):
    """Routed (sample, expert) pairs of one batch (host read, outside every timed region)."""
    from nerf_sys_b200 import ops
    t = ops.stratified_t(rays, S_TRAIN, torch.rand(rays.shape[0], S_TRAIN, device=rays.device))
    pl = ops.route_pack(ops.PointSet(rays=rays, t=t), model.centroids_host(), len(model.submodules), model.boundary_margin, model.cluster_2d)
    counts, _ = ops.a2a_count(pl, len(model.submodules))
    return counts.tolist()


def kernel_table(timing, K, units, pk):
    kern = {}
    for name, evs in timing.items():
        tot = sum(a.elapsed_time(b) for a, b in evs)
        kern[name] = {"ms_per_step": tot / K, "calls_per_step": len(evs) / K}
        if name in ALGO and units.get(ALGO[name][2]):
            bound, per_unit, unit = ALGO[name]
            sec = tot / K / 1e3
            ach = per_unit * units[unit] / sec / (1e9 if bound == "hbm" else 1e12)
            kern[name].update({"bound": bound, "achieved": ach, "unit": "GB/s" if bound == "hbm" else "TFLOP/s", "frac": ach / pk[bound],
                               "algorithmic_per_" + unit: per_unit, unit + "s_per_step": units[unit]})
    return kern


def roofline_of(kern, pk, prefer=None):
    cands = [n for n in kern if "bound" in kern[n]]
    if not cands:
        return None
    top = prefer if prefer in cands else max(cands, key=lambda n: kern[n]["ms_per_step"])
    traffic, binding = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        ent = tj.get(top)
        if isinstance(ent, dict):
            traffic, binding = ent.get("dram_bytes_per_launch"), ent.get("binding")
        else:
            traffic = ent
    k = kern[top]
    unit = ALGO[top][2]
    return {"kernel": top, "bound": k["bound"], "achieved": k["achieved"], "peak": pk[k["bound"]], "unit": k["unit"], "frac": k["frac"],
            "traffic": traffic, "binding_resource": binding, "peak_source": pk["src"], "algorithmic_per_" + unit: ALGO[top][1],
            unit + "s_per_launch": k[unit + "s_per_step"] / max(k["calls_per_step"], 1e-9), "ms_per_launch": k["ms_per_step"] / max(k["calls_per_step"], 1e-9)}


def train_c3(ctx, K, W, args):
    """Headline: returns (line fields, trainer-independent extras)."""
    from nerf_sys_b200 import ops
    tr = ContainerTrainer(ctx)
    nb = min(16, K + W)
    pool = aerial_pool(tr.aabb, ctx.rank, nb, ctx.dev)
    teacher, _ = build_container(ctx.dev, seed=11, table_std=0.1)
    rgbs = teacher_colors(teacher, pool)
    del teacher
    torch.cuda.empty_cache()
    pool_h, rgbs_h = pool.cpu().pin_memory(), rgbs.cpu().pin_memory()
    counts = pair_stats(tr.model, pool[0])
    use_graph = not args.no_graph
    if use_graph:
        gstep = tr.graph(pool[0], rgbs[0], warmup=W)
        step = lambda r, c: gstep(r, c)  # noqa: E731
    else:
        step = tr.eager_step
        for i in range(W):
            step(pool[i % nb], rgbs[i % nb])
    for i in range(W):
        step(pool[i % nb], rgbs[i % nb])
    # ---- resident-input arm ------------------------------------------------------------
    clocks = Clocks(ctx.dev.index)
    l0 = ops.launches()
    ms, loss = timed(ctx, lambda i: step(pool[(W + i) % nb], rgbs[(W + i) % nb]), K)
    launches = ops.launches() - l0
    final_loss = float(loss.item()) * ctx.world
    tr.check()
    # ---- end-to-end arm: host batches in, loss out, every step -------------------------------
    # pinned host batch -> device (async), the step, the step's loss -> pinned host (async).  The host reads the loss of step
    # i-1 while step i runs (what a training loop that logs its loss does); all K losses are read inside the timed region.
    loss_h = torch.zeros(K, dtype=torch.float32).pin_memory()
    evs = [torch.cuda.Event() for _ in range(K)]
    losses_read = []

    def e2e_step(i):
        if use_graph:  # pinned host batch straight into the captured step's input buffers (one H2D copy per tensor)
            l = step(pool_h[(W + i) % nb], rgbs_h[(W + i) % nb])
        else:
            l = step(pool_h[(W + i) % nb].to(ctx.dev, non_blocking=True), rgbs_h[(W + i) % nb].to(ctx.dev, non_blocking=True))
        loss_h[i:i + 1].copy_(l.detach().reshape(1), non_blocking=True)
        evs[i].record()
        if i > 0:
            evs[i - 1].synchronize()
            losses_read.append(float(loss_h[i - 1]))
        if i == K - 1:
            evs[i].synchronize()
            losses_read.append(float(loss_h[i]))

    ms_e2e, _ = timed(ctx, e2e_step, K)
    assert len(losses_read) == K and all(v == v for v in losses_read), "every step's loss must reach the host"
    clk = clocks.stop()
    tr.check()
    # ---- per-kernel device time: the same steps driven eagerly with an event pair around every C-ABI call
    # (events cannot be recorded between the nodes of a replayed graph)
    KT = min(K, 10)
    timing = ops.kernel_timing(True)
    for i in range(KT):
        tr.eager_step(pool[(W + i) % nb], rgbs[(W + i) % nb])
    ctx.barrier()
    ops.kernel_timing(False)
    pk = peaks()
    owned = K_EXPERTS // ctx.world if ctx.world > 1 else K_EXPERTS
    n_pairs_local = sum(counts[:K_EXPERTS])            # pairs this rank routes (push / pull / route kernels)
    n_pairs_owner = n_pairs_local                      # balanced scene: an owner receives about what a rank sends
    units = {"pair": n_pairs_owner, "point": N_RAYS * S_TRAIN, "elem": owned * (TABLE_ELEMS + MLP_ELEMS) + 643}
    kern = kernel_table(timing, KT, units, pk)
    roof = roofline_of(kern, pk)
    fields = {
        "value": ctx.world * N_RAYS * K / (ms / 1e3), "ms_per_step": ms / K,
        "e2e": {"value": ctx.world * N_RAYS * K / (ms_e2e / 1e3), "unit": "rays/s",
                "h2d_bytes_per_step": pool_h[0].numel() * 4 + rgbs_h[0].numel() * 4, "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / K,
                "note": "per GPU: pinned-host rays + target colours copied in, loss copied out every step; loss of step i-1 read on the "
                        "host while step i runs, all K losses read inside the timed region"},
        "gpu_launches": launches, "roofline": roof, "kernels": kern, "clocks": clk, "final_loss": final_loss,
        "routed_pairs_per_gpu_step": n_pairs_local, "pairs_per_expert_batch0": counts[:K_EXPERTS],
        "impl_detail": {
            "launch": (f"CUDA graph replay of zero_grad + fwd + bwd + clip + Adam ({tr.gstep.launches_per_step} kernel launches per step, no host read)"
                       if use_graph else "eager Python launches"),
            "mlp_math": ("tcgen05 kind::f16, fp16 operands, fp32 accumulate in TMEM (hash grid, compositing, loss, Adam in fp32)"
                         if args.mlp_math == "f16tc" else "fp32 FFMA (exact mode)"),
            "exchange": "none (one GPU)" if ctx.world == 1 else
                        "NVLink peer memory: sample rows (24 B), results (16 B) and gradient rows (16 B) written / read in the owner's slab by "
                        "the pack and combine kernels; counts, slab layout, bg-head gradient sum and clip-norm sum on device-side slab barriers; "
                        "no NCCL kernel in the step",
            "optimizer": "FusedAdam: persistent table-gradient buffers zeroed inside the Adam kernel, device-side expert liveness flags, "
                         "global clip norm",
            "kernel_timing": f"separate eager pass of {KT} steps, CUDA-event pair per C-ABI call",
        },
    }
    return fields, tr


# ------------------------------------------------------------------------------------------
# multi-GPU parity on the bench scene (printed with every N > 1 line; a mismatch fails the run)
# ------------------------------------------------------------------------------------------
def parity_check(ctx, tr):
    """(1) the sharded peer-memory container == the same container evaluated whole on this GPU (outputs and the owner's
    gradients), on this rank's bench rays; (2) one expert-parallel optimisation step == the single-GPU step over the
    concatenated batch; (3) the row-sharded 1080p frame == the unsharded frame.  fp32-exact MLP mode for (1), (2)."""
    import torch.distributed as dist
    from nerf_sys_b200 import ops, parallel as PL
    from nerf_sys_b200.nerfs.ray_rendering import render_image, render_rays
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    from nerf_sys_b200.optim import FusedAdam, bind_expert_liveness, build_optimizer
    from nerf_sys_b200 import synthetic as SY
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    out = {}
    prev = ops.set_mlp_math("fp32")
    model, ex = tr.model, tr.ex
    owner = PL.expert_owner(K_EXPERTS, world)
    PL.sync_expert_params(model)  # the timed steps trained each expert on its owner only: every rank gets the owners' copies
    rays = SY.aerial_rays(N_RAYS, tr.aabb, seed=4000 + rank).to(dev)
    g = torch.Generator().manual_seed(40 + rank)
    u, cot = torch.rand(N_RAYS, S_TRAIN, generator=g).to(dev), torch.randn(N_RAYS * S_TRAIN, 4, generator=g).to(dev)
    pts = ops.PointSet(rays=rays, t=ops.stratified_t(rays, S_TRAIN, u))
    # (1) reference: the whole container on this GPU
    model.zero_grad(set_to_none=True)
    model.set_expert_parallel(False)
    y_ref = model.evaluate(pts)
    (y_ref * cot).sum().backward()
    mlp_ref = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None and "hash_table" not in n and "bg_mlp" not in n}
    tab_ref = {n: p.grad for n, p in model.named_parameters() if p.grad is not None and "hash_table" in n}
    # an expert's gradient is the sum of every rank's contribution: reduce the small tensors exactly, the tables by norm of the owner's
    for n in sorted(mlp_ref):
        dist.all_reduce(mlp_ref[n])
    tab_sum = {}
    for n in sorted(tab_ref):
        e = int(n.split(".")[1])
        t = tab_ref[n].clone()
        dist.reduce(t, dst=owner[e])
        if owner[e] == rank:
            tab_sum[n] = t
        del t
    model.zero_grad(set_to_none=True)
    tab_ref = None
    model.set_expert_parallel(True, exchange=ex)
    y = model.evaluate(pts)
    (y * cot).sum().backward()
    err = (y - y_ref).abs().max().item()
    gerr = terr = 0.0
    for n, p in model.named_parameters():
        if not n.startswith("submodules.") or owner[int(n.split(".")[1])] != rank:
            continue
        got = p.grad if p.grad is not None else torch.zeros_like(p)
        ref = tab_sum[n] if "hash_table" in n else mlp_ref.get(n)
        if ref is None:
            continue
        rel = ((got - ref).norm() / ref.norm().clamp_min(1e-30)).item()
        if "hash_table" in n:
            terr = max(terr, rel)
        else:
            gerr = max(gerr, rel)
    stats = torch.tensor([err, gerr, terr], device=dev)
    dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    out["container_p2p_vs_single_gpu"] = {"out_max_abs": stats[0].item(), "owner_mlp_grad_rel_l2": stats[1].item(),
                                          "owner_table_grad_rel_l2": stats[2].item(), "tol": [1e-5, 1e-4, 1e-4],
                                          "ok": bool(stats[0] <= 1e-5 and stats[1] <= 1e-4 and stats[2] <= 1e-4)}
    model.zero_grad(set_to_none=True)
    for p, _ in tr.opt._arena.values():
        ops.release_grad_buffer(p, clean=False)
    del tab_sum, mlp_ref, y, y_ref
    torch.cuda.empty_cache()
    # (2) one optimisation step, fresh models on both sides (clip active: max_norm far below the gradient norm)
    Nc, CLIP_CHECK = 2048, 1e-4
    rays = SY.aerial_rays(Nc, tr.aabb, seed=4100 + rank).to(dev)
    g = torch.Generator().manual_seed(60 + rank)
    u, gt = torch.rand(Nc, S_TRAIN, generator=g).to(dev), torch.rand(Nc, 3, generator=g).to(dev)
    all_rays, all_u, all_gt = [torch.empty(world * Nc, *x.shape[1:], device=dev) for x in (rays, u, gt)]
    for dst, src in ((all_rays, rays), (all_u, u), (all_gt, gt)):
        dist.all_gather_into_tensor(dst, src.contiguous())
    ref_m, _ = build_container(dev, seed=21)
    ref_m.train()
    ref_opt = build_optimizer(ref_m, max_grad_norm=CLIP_CHECK)
    rgb, *_ = render_rays(ref_m, all_rays, ray_samples=S_TRAIN, u=all_u)
    ref_loss = ops.mse_loss(rgb, all_gt, "identity")
    ref_loss.backward()
    ref_opt.step()
    ep_m, _ = build_container(dev, seed=21)
    ep_m.train()
    ep_m.set_expert_parallel(True, exchange=ex)
    groups, shared = PL.owned_param_groups(ep_m)
    ep_opt = FusedAdam([{"params": gr["params"], "lr": LRS[n]} for n, gr in groups.items()], max_grad_norm=CLIP_CHECK).shard_grad_norm(None, shared, exchange=ex)
    bind_expert_liveness(ep_opt, ep_m)
    red = PL.SlabGradReducer(ex, shared, scale=1.0)
    rgb, *_ = render_rays(ep_m, rays, ray_samples=S_TRAIN, u=u)
    loss = ops.mse_loss(rgb, gt, "identity")
    (loss / world).backward()
    red.wait()
    ep_opt.step()
    red.remove()
    ex.reducer = tr.reducer
    mean_loss = loss.detach().clone()
    dist.all_reduce(mean_loss)
    mean_loss /= world
    ref_p = dict(ref_m.named_parameters())
    perr = 0.0
    for n, p in ep_m.named_parameters():
        mine = (n.startswith("submodules.") and owner[int(n.split(".")[1])] == rank) or n.startswith("bg_mlp")
        if mine:
            perr = max(perr, (p.detach() - ref_p[n].detach()).abs().max().item())
    c_ep, c_ref = ep_opt.clip_coef.item(), ref_opt.clip_coef.item()
    stats = torch.tensor([abs(mean_loss.item() - ref_loss.item()), abs(c_ep - c_ref) / max(c_ref, 1e-30), perr], device=dev)
    dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    out["ep_optimizer_step_vs_concatenated_batch"] = {
        "loss_abs_diff": stats[0].item(), "clip_coef": c_ep, "clip_coef_ref": c_ref, "clip_coef_rel_diff": stats[1].item(),
        "owner_param_max_abs_diff": stats[2].item(), "tol": [1e-6, 1e-4, 2e-4],
        "ok": bool(stats[0] <= 1e-6 and stats[1] <= 1e-4 and c_ref < 0.999 and stats[2] <= 2e-4)}
    ep_m.set_expert_parallel(False)
    del ref_m, ep_m, ref_opt, ep_opt
    torch.cuda.empty_cache()
    ops.set_mlp_math(prev)
    # (3) row-sharded 1080p frame vs the unsharded frame (tensor-core mode, the product path)
    model.set_expert_parallel(False)
    model.eval()
    H, W, f = 1080, 1920, 1480.0
    kw = dict(H=H, W=W, fx=f, fy=f, cx=W / 2, cy=H / 2, c2w=SY.ring_pose(2, 20, radius=0.9), scene_box=SceneBox(tr.aabb.to(dev)), ray_samples=64)
    full = render_image(model, **kw)
    sharded = PL.render_image_sharded(lambda r0, r1: render_image(model, row_range=(r0, r1), **kw), H, W)
    e = max((a.reshape(-1).float() - b.reshape(-1).float()).abs().max().item() for a, b in zip(full, sharded))
    stats = torch.tensor([e], device=dev)
    dist.all_reduce(stats, op=dist.ReduceOp.MAX)
    out["sharded_1080p_frame_vs_unsharded"] = {"max_abs": stats[0].item(), "tol": 1e-6, "ok": bool(stats[0] <= 1e-6)}
    model.train()
    model.set_expert_parallel(True, exchange=ex)
    out["ok"] = all(v["ok"] for v in out.values())
    return out


# ------------------------------------------------------------------------------------------
# configs[4]: 1080p render, rows sharded
# ------------------------------------------------------------------------------------------
@torch.no_grad()
def render_c5(ctx, frames=8):
    """1920x1080 x 64 samples through the 8-expert container, image rows sharded over the ranks, uint8 tiles gathered on rank 0
    (the frame a viewer receives).  `value`: the frame stays on rank 0's GPU; `e2e`: pose from the host, the finished uint8 frame
    copied to pinned host memory."""
    from nerf_sys_b200 import ops, parallel as PL, synthetic as SY
    from nerf_sys_b200.nerfs.ray_rendering import render_image
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    model, aabb = build_container(ctx.dev)
    model.eval()
    H, W, f = 1080, 1920, 1480.0
    sb = SceneBox(aabb.to(ctx.dev))
    poses = [SY.ring_pose(i, 20, radius=0.9) for i in range(2 * frames + 2)]

    def frame(i, **kw):
        args = dict(H=H, W=W, fx=f, fy=f, cx=W / 2, cy=H / 2, c2w=poses[i], scene_box=sb, ray_samples=64)
        return PL.render_image_sharded(lambda r0, r1: render_image(model, row_range=(r0, r1), **args), H, W, **kw)

    own = frame(0, gather=False)
    acc = own[2] if ctx.world == 1 else own[0][2]
    cover = float((acc > 0.01).float().mean().item()) if acc.numel() else None
    for i in range(3):  # warm-up: allocator pools, NCCL gather connections (a frame is ~5 ms at 8 GPUs: one-off costs show)
        frame(i, as_uint8=True, dst=0)
    ms, _ = timed(ctx, lambda i: frame(1 + i, as_uint8=True, dst=0), frames)
    # end to end: the frame a viewer receives - uint8 rgb on rank 0's host
    host = torch.empty(H, W, 3, dtype=torch.uint8).pin_memory()
    ev = torch.cuda.Event()

    def e2e_frame(i):
        rgb8 = frame(1 + frames + i, as_uint8=True, dst=0)
        if ctx.rank == 0:
            host.copy_(rgb8[0], non_blocking=True)
            ev.record()
            ev.synchronize()

    e2e_frame(0)
    ms_e2e, _ = timed(ctx, e2e_frame, frames)
    # per-kernel pass on one frame
    timing = ops.kernel_timing(True)
    frame(1)
    ctx.barrier()
    ops.kernel_timing(False)
    rows = PL.shard_range(H, ctx.world, ctx.rank)
    pts = (rows[1] - rows[0]) * W * 64
    pk = peaks()
    kern = kernel_table(timing, 1, {"pair": int(pts * 1.03), "point": pts}, pk)
    roof = roofline_of(kern, pk, prefer="fused_fwd" if "fused_fwd" in kern else "hash_fwd")
    del model
    torch.cuda.empty_cache()
    return {"metric": "render_rays_per_sec", "value": frames * H * W / (ms / 1e3), "unit": "rays/s", "ms_per_frame": ms / frames, "coverage": cover,
            "e2e": {"value": frames * H * W / (ms_e2e / 1e3), "unit": "rays/s", "ms_per_frame": ms_e2e / frames, "h2d_bytes_per_step": 48,
                    "d2h_bytes_per_step": H * W * 3, "note": "camera pose from the host; uint8 tiles gathered on rank 0 and copied to pinned host memory"},
            "roofline": roof, "kernels": kern, "scaling": "strong",
            "config": f"config[4]: 1920x1080 x 64 samples, 8 experts (soft routing, margin 1.05), eval, rows sharded over {ctx.world} GPU(s), "
                      "experts replicated"}


# ------------------------------------------------------------------------------------------
# configs[1]: one expert per GPU (meta-training shard), configs[3]: adaptation, configs[0]: PE-MLP
# ------------------------------------------------------------------------------------------
def train_c2(ctx, K, W):
    """configs[1] per GPU: single-expert hash-grid NeRF on synthetic 800x800 views, 8192-ray batch x 96 samples; at N > 1 every rank
    owns one expert (tasks never cross experts, meta_train_step.py:60-96) and the only shared state is the background head, whose
    643 gradient floats are averaged on a slab barrier."""
    from nerf_sys_b200 import parallel as PL
    from nerf_sys_b200.engine import GraphedTrainStep
    from nerf_sys_b200.nerfs.losses import compute_mse_loss
    from nerf_sys_b200.nerfs.ray_rendering import render_image
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    from nerf_sys_b200.optim import build_optimizer
    model, st, cfg, box = build_scene(ctx.rank, ctx.world, ctx.dev)
    model.train()
    opt = build_optimizer(model, max_grad_norm=1.0, persistent_grads=True)
    nb = min(16, K + W)
    pool = make_pool(box, ctx.dev, nb, seed=10 + ctx.rank)
    teacher, *_ = build_scene(ctx.rank, ctx.world, ctx.dev, seed=1, table_std=0.1)
    rgbs = teacher_colors(teacher, pool)
    del teacher
    ex = red = None
    if ctx.world > 1:
        ex = PL.PeerExchange(rows_cap=128)
        red = PL.SlabGradReducer(ex, list(model.bg_mlp.parameters()), scale=1.0 / ctx.world, early=True)  # posted inside the backward
    gstep = GraphedTrainStep(opt, lambda r, c: compute_mse_loss(P, model, {"rays": r, "rgbs": c}), (pool[0], rgbs[0]), warmup=W,
                             post_backward=(red.wait if red is not None else None))
    for i in range(W):
        gstep(pool[i % nb], rgbs[i % nb])
    ms, loss = timed(ctx, lambda i: gstep(pool[(W + i) % nb], rgbs[(W + i) % nb]), K)
    out = {"value": ctx.world * N_RAYS * K / (ms / 1e3), "unit": "rays/s", "ms_per_step": ms / K, "loss": float(loss.item()), "scaling": "weak",
           "config": "config[1]: single-expert hash-grid NeRF training per GPU, synthetic 800x800 views, 8192-ray batch x 96 samples, CUDA graph; "
                     + ("one GPU" if ctx.world == 1 else "one expert per GPU, background-head gradients averaged on a slab barrier (no NCCL kernel)")}
    # render throughput (no_grad, eval), one 800x800 view per measurement
    model.eval()
    sb = SceneBox(box.to(ctx.dev))
    kw = dict(H=IMG, W=IMG, fx=FOCAL, fy=FOCAL, cx=IMG / 2, cy=IMG / 2, scene_box=sb, ray_samples=64)
    for v in range(2):
        render_image(model, c2w=ring_pose_in(box, v, N_VIEWS), **kw)
    nv = 5
    rms, _ = timed(ctx, lambda v: render_image(model, c2w=ring_pose_in(box, v + 2, N_VIEWS), **kw), nv)
    render = {"value": ctx.world * nv * IMG * IMG / (rms / 1e3), "unit": "rays/s", "ms_per_frame": rms / nv,
              "config": "800x800 x 64 samples, one expert per GPU, eval, no_grad, raygen on device"}
    gstep = None
    opt.release_persistent_grads()
    if red is not None:
        red.remove()
    if ex is not None:
        ex.close()
    del model, opt
    torch.cuda.empty_cache()
    return out, render


def adapt_c4(ctx, steps=10, warmup=3):
    """configs[3]: incremental adaptation of the 8-expert container on a new aerial batch whose samples fall in experts {2, 5} only
    (nadir rays over those two tiles); the other six experts get no sample, so their parameters and Adam state are not touched
    (asserted bit-unchanged: the device-side liveness flags stand in for `p.grad is None`) and the step costs two experts' worth of
    scatter and optimiser traffic.  Same graph-captured step as the headline; at N > 1 the experts are sharded."""
    # every ray of every rank lands in experts 2 and 5: their owners receive world x 4096 x 96 rows each (a deliberately
    # unbalanced exchange), so the slab is sized for that instead of the 3x-balanced default
    tr = ContainerTrainer(ctx, rows_cap=max(3, ctx.world + 1) * N_RAYS * S_TRAIN)
    model = tr.model
    cent = model.centroids.detach().cpu()
    aabb = tr.aabb
    g = torch.Generator().manual_seed(77 + ctx.rank)
    nb, batches = 4, []
    for b in range(nb):
        rays = torch.zeros(N_RAYS, 8)
        for i, k in enumerate((2, 5)):
            sl = slice(i * N_RAYS // 2, (i + 1) * N_RAYS // 2)
            rays[sl, 0] = float(aabb[0, 0]) + 0.01
            rays[sl, 1:3] = cent[k, 1:3] + (torch.rand(N_RAYS // 2, 2, generator=g) - 0.5) * 0.3
            rays[sl, 3] = 1.0  # straight down: the YZ position (the routing key) never changes along the ray
            rays[sl, 6], rays[sl, 7] = 0.0, float(aabb[1, 0] - aabb[0, 0]) - 0.02
        batches.append((rays.to(ctx.dev), torch.rand(N_RAYS, 3, generator=g).to(ctx.dev)))
    owned = {n for gr in tr.opt.param_groups for p in gr["params"] for n, q in model.named_parameters() if q is p}
    frozen = {n: p.detach().clone() for n, p in model.named_parameters()
              if n in owned and n.startswith("submodules.") and int(n.split(".")[1]) not in (2, 5)}
    gstep = tr.graph(batches[0][0], batches[0][1], warmup=warmup)
    ms, loss = timed(ctx, lambda i: gstep(*batches[i % nb]), steps)
    tr.check()
    named = dict(model.named_parameters())
    unchanged = all(torch.equal(v, named[n].detach()) for n, v in frozen.items())
    st = tr.opt.state
    moments_zero = all((not st.get(named[n])) or float(st[named[n]]["exp_avg"].abs().max()) == 0.0 for n in frozen)
    tr.opt.sync_steps()
    steps_zero = all((not st.get(named[n])) or float(st[named[n]]["step"]) == 0.0 for n in frozen)
    ok = bool(unchanged and moments_zero and steps_zero)
    out = {"value": ctx.world * N_RAYS * steps / (ms / 1e3), "unit": "rays/s", "ms_per_step": ms / steps, "loss": float(loss.item()) * ctx.world,
           "frozen_experts_bit_unchanged": ok, "scaling": "weak",
           "config": f"config[3]: 8-expert container, {N_RAYS} nadir rays per GPU x {S_TRAIN} samples over the tiles of experts 2 and 5, "
                     "fwd+bwd+clip+Adam (CUDA graph); six experts receive no sample: parameters bit-unchanged, moments and step counters stay 0"}
    tr.close()
    assert ok, "experts outside {2,5} must stay bit-unchanged, with untouched optimiser state"
    return out


def train_c1(ctx, steps=10, warmup=3):
    """configs[0]: tiny scene, 2 spatial experts, positional-encoding MLP NeRF (MetaNeRF stand-alone per expert, hard routing by
    centroid as the reference would dispatch it), 4096-ray batch x 64 samples, fwd + bwd + Adam.  Tensor-pipe-bound config."""
    from nerf_sys_b200 import ops, synthetic as SY
    from nerf_sys_b200.models.inr import MetaNeRF
    from nerf_sys_b200.optim import FusedAdam
    N, S = 4096, 64
    aabb, cent, boxes = SY.grid_scene(1, 2)
    torch.manual_seed(5)
    experts = [MetaNeRF().to(ctx.dev).train() for _ in range(2)]
    opt = FusedAdam([{"params": [p for e in experts for p in e.parameters()], "lr": 5e-4}], max_grad_norm=1.0)
    rays = SY.aerial_rays(N, aabb, seed=300 + ctx.rank).to(ctx.dev)
    g = torch.Generator().manual_seed(31)
    gt = torch.rand(N, 3, generator=g).to(ctx.dev)
    chost = cent.reshape(-1).tolist()
    pk = peaks()

    def step(i):
        opt.zero_grad(set_to_none=True)
        with torch.no_grad():
            t = ops.stratified_t(rays, S, torch.rand(N, S, device=ctx.dev))
            x6 = ops.sample_points(rays, t)
            _, hard = ops.routing(ops.PointSet(x6=x6), chost, 2, 1.0, True)
        y = torch.zeros(N * S, 4, device=ctx.dev)
        for k, ex in enumerate(experts):
            sel = (hard == k).nonzero().squeeze(1)  # the reference's per-expert host read (meta_container.py:309)
            if sel.numel():
                y = y.index_copy(0, sel, ex.evaluate(ops.PointSet(x6=x6.index_select(0, sel))))
        rgb, *_ = ops.composite(y.view(N, S, 4), t, torch.ones(N, 3, device=ctx.dev))
        loss = ops.mse_loss(rgb, gt, "identity")
        loss.backward()
        opt.step()
        return loss

    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        loss = step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    timing = ops.kernel_timing(True)
    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    ops.kernel_timing(False)
    # MACs per point of the default MetaNeRF (meta_vanilla.py:19-35): trunk 63x256 + 3x256x256 + (256+63)x256 + 3x256x256, heads 16x256,
    # colour (15+27)x128 + 128x3
    macs = 63 * 256 + 3 * 256 * 256 + (256 + 63) * 256 + 3 * 256 * 256 + 16 * 256 + 42 * 128 + 128 * 3
    kern = {}
    for name, evs in timing.items():
        tot = sum(a.elapsed_time(b) for a, b in evs)
        kern[name] = {"ms_per_step": tot / 3, "calls_per_step": len(evs) / 3}
    for name, mult in (("pe_mlp_fwd", 2), ("pe_mlp_bwd", 4)):
        if name in kern:
            tf = mult * macs * N * S / (kern[name]["ms_per_step"] / 1e3) / 1e12
            kern[name].update({"bound": "tensor", "achieved": tf, "unit": "TFLOP/s", "frac": tf / pk["tensor"], "flop_per_point": mult * macs})
    top = max((n for n in kern if "bound" in kern[n]), key=lambda n: kern[n]["ms_per_step"], default=None)
    roof = None if top is None else {"kernel": top + " (k_dense_tc / k_wgrad_tc launches, one per layer)", "bound": "tensor",
                                      "achieved": kern[top]["achieved"], "peak": pk["tensor"], "unit": "TFLOP/s", "frac": kern[top]["frac"],
                                      "traffic": None}
    return {"value": N * steps / (ms / 1e3), "unit": "rays/s", "ms_per_step": ms / steps, "loss": float(loss.item()), "roofline": roof, "kernels": kern,
            "config": "config[0]: 2 spatial experts (1x2 grid), PE-MLP MetaNeRF (8x256 trunk, skip at 4, 128-wide colour), hard routing, "
                      "4096 aerial rays x 64 samples, fwd+bwd+clip+Adam, eager launches (per-expert host read as in the reference's dispatch), one GPU"}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's PyTorch path as restated by the oracle (bit-identical on the goldens), SAME config:
# the 8-expert container step on one GPU's 8192-ray batch, all host threads
# ------------------------------------------------------------------------------------------
def cpu_c3(steps, warmup, budget_s=None):
    """Returns (rays/s, seconds per step, rays per step, steps timed).  One step = the oracle's 8-expert render_rays (soft routing
    1.05, cdist routing as the reference) + MSE + backward + clip_grad_norm_(1.0) + torch.optim.Adam with the reference's
    per-group learning rates (common/utils.py:16-76), optimizer.zero_grad() setting gradients to None as torch's default does.
    ``budget_s``: if the first step shows that steps+warmup full batches would not fit, the batch is cut to a slice (reported)."""
    from oracle import acnerf_oracle as O
    from nerf_sys_b200 import synthetic as SY
    torch.set_num_threads(os.cpu_count() or 1)
    aabb, cent, boxes = SY.grid_scene(2, 4)
    cfg = dict(SY.HASH_CFG)
    cfg["log2_hashmap_size"] = LOG2_T
    st = SY.init_ngp_state(K_EXPERTS, aabb, cent, boxes, cfg, seed=3, table_std=0.1)
    groups = {"encoding": [], "sigma": [], "color": [], "background": []}
    for k, v in st.items():
        if not ("weight" in k or "bias" in k or "hash_table" in k):
            continue
        v.requires_grad_(True)
        name = "encoding" if "hash_table" in k else "background" if k.startswith("bg_mlp") else "color" if "color_mlp" in k else "sigma"
        groups[name].append(v)
    params = [p for g in groups.values() for p in g]
    opt = torch.optim.Adam([{"params": ps, "lr": LRS[n]} for n, ps in groups.items()])
    mins = [b[0] for b in boxes]
    g = torch.Generator().manual_seed(7)
    n_rays, times, i = N_RAYS, [], 0
    while len(times) < steps:
        rays = SY.aerial_rays(n_rays, aabb, seed=1000 + i)
        u, gt = torch.rand(n_rays, S_TRAIN, generator=g), torch.rand(n_rays, 3, generator=g)
        t0 = time.perf_counter()
        opt.zero_grad()
        rgb, *_ = O.render_rays(rays, st, mins, cfg, S=S_TRAIN, margin=1.05, u=u)
        loss = O.mse_loss(rgb, gt, "identity")
        loss.backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        dt = time.perf_counter() - t0
        if i == 0 and budget_s is not None and dt * (steps + warmup) > budget_s and n_rays > 1024:
            n_rays = max(1024, int(n_rays * budget_s / (dt * (steps + warmup))) // 1024 * 1024)
        elif i >= warmup:
            times.append(dt)
        i += 1
    return n_rays * len(times) / sum(times), sum(times) / len(times), n_rays, len(times)


def run_reference(args):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    t0 = time.perf_counter()
    rps, sec, n_rays, n = cpu_c3(args.steps, args.warmup, budget_s=270.0)
    cores = os.cpu_count() or 1
    sample = (f"{n} steps of one GPU's share of the workload: the 8-expert container on a {n_rays}-ray x {S_TRAIN}-sample aerial batch "
              f"(fwd+bwd+clip+per-group Adam, {sec:.2f} s/step), torch {torch.__version__} CPU fp32, {cores} host threads"
              + ("" if n_rays == N_RAYS else f"; the batch was cut from {N_RAYS} rays to keep the run within a few minutes"))
    print(json.dumps({
        "impl": "reference", "metric": "train_rays_per_sec", "value": rps, "unit": "rays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus),
        "cpu_baseline": {"value": rps, "unit": "rays/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rps, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference = its pure-PyTorch path on the host cores via oracle/ (the reference is Python, has no installable package and "
                "does not exist on the GPU box; the oracle is bit-identical to it on tests/golden).  One process, the rays of one GPU.",
        "wall_s": time.perf_counter() - t0}))


# ------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-render", action="store_true", help="skip render_c5 / render_800")
    ap.add_argument("--no-secondary", action="store_true", help="headline only (no train_c2 / adapt_c4 / train_c1 / render / parity)")
    ap.add_argument("--no-parity", action="store_true", help="skip the multi-GPU parity check at N > 1")
    ap.add_argument("--no-graph", action="store_true", help="drive every step from Python instead of replaying a CUDA graph")
    ap.add_argument("--mlp-math", default="f16tc", choices=["f16tc", "fp32"],
                    help="expert-MLP arithmetic: tcgen05 fp16/fp32-accumulate (default) or the exact fp32 CUDA-core mode")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    from nerf_sys_b200 import ops
    rank, world, local = dist_env()
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 under torch.distributed.run"
    assert world in (1, 2, 4, 8), "the 8 experts are sharded in equal blocks: 1, 2, 4 or 8 GPUs"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ops.device_ok(local)
    ops.set_mlp_math(args.mlp_math)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)  # plumbing only: IPC-handle exchange, barriers around timed regions, parity checks
    ctx = Ctx(dev, rank, world)
    K, W = args.steps, max(args.warmup, 3)

    head, tr = train_c3(ctx, K, W, args)

    # secondary measurements: a failure here is recorded in its own field and must not cost the headline line (every rank runs
    # the same code, so an exception is raised on all of them alike)
    def guarded(fn, *a, **kw):
        try:
            return fn(*a, **kw)
        except Exception as e:  # noqa: BLE001
            torch.cuda.synchronize()
            return {"error": f"{type(e).__name__}: {e}"[:400]}

    parity = None
    if world > 1 and not args.no_parity and not args.no_secondary:
        parity = guarded(parity_check, ctx, tr)
    tr.close()
    sec = not args.no_secondary
    c5 = guarded(render_c5, ctx) if (sec and not args.no_render) else None
    c2 = guarded(train_c2, ctx, min(K, 20), W) if sec else None
    c2, r800 = c2 if isinstance(c2, tuple) else (c2, None)
    c4 = guarded(adapt_c4, ctx) if sec else None
    c1 = guarded(train_c1, ctx) if (sec and world == 1) else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rps, s, n_rays, n = cpu_c3(2, 1)
        cpu = {"value": rps, "unit": "rays/s", "cores": os.cpu_count() or 1, "kind": "port",
               "sample": f"{n} steps (after 1 warm-up) of the same workload: the 8-expert container on a {n_rays}-ray x {S_TRAIN}-sample aerial "
                         f"batch, fwd+bwd+clip+per-group Adam ({s:.2f} s/step), torch {torch.__version__} CPU fp32, all host threads"}
    if rank == 0:
        line = {"metric": "train_rays_per_sec", "value": head.pop("value"), "unit": "rays/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": head.pop("ms_per_step"), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f16" if args.mlp_math == "f16tc" else "f32", "data": "synthetic", "config": workload_config(world)}
        line.update(head)
        line.update({"parity_check": parity, "render_c5": c5, "train_c2": c2, "render_800": r800, "adapt_c4": c4, "train_c1": c1,
                     "cpu_baseline": cpu})
        print(json.dumps(line))
    ok = parity is None or bool(parity.get("ok", False))
    if world > 1:
        sys.stdout.flush()
        sys.stderr.flush()
        torch.cuda.synchronize()
        torch.distributed.barrier()
        os._exit(0 if ok else 1)
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
