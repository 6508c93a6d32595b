"""Single-node multi-GPU partitioning of the hot path; one process per GPU, torch.distributed
for the plumbing (NCCL over NVLink on the box, gloo in the CPU tests of the host logic).

The reference is single-GPU; the natural shards of its path are

* render (config 5): image ROWS are split across ranks, every expert is replicated, no collective on the data
  path, one gather of the finished tiles (`render_image_sharded`);
* meta-training (config 2/3 with ``active_module``): EXPERTS (regions) are split across ranks, tasks never cross
  experts (pipelines/offline_stage/meta_train_step.py:60-96); the only shared parameters are the 643 floats of
  the background head (`allreduce_shared_grads`);
* full-container training / adaptation (config 3/4, soft routing): rays are split across ranks AND experts are
  sharded, so every routed sample travels to the rank that owns its expert and its [rgb, sigma] travels back:
  one all-to-all per direction per step (`expert_parallel_apply`), the reverse pair in the backward pass.

Everything here is index bookkeeping around collectives; the arithmetic stays in the CUDA library.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

Tensor = torch.Tensor


def _first_order_backward(fn):
    from .ops import first_order  # deferred: ops loads the shared library
    return first_order(fn)


# ---------------------------------------------------------------------------------------------
# partitions
# ---------------------------------------------------------------------------------------------
def shard_range(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of ``n`` items for ``rank``; the first n % world ranks get one extra."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def expert_owner(K: int, world: int) -> List[int]:
    """owner[e] = rank holding expert e: contiguous blocks (experts of one rank are neighbours on the grid)."""
    owner = [0] * K
    for r in range(world):
        lo, hi = shard_range(K, world, r)
        for e in range(lo, hi):
            owner[e] = r
    return owner


def _world(group) -> Tuple[int, int]:
    if not dist.is_available() or not dist.is_initialized():
        return 1, 0
    return dist.get_world_size(group), dist.get_rank(group)


# ---------------------------------------------------------------------------------------------
# shared parameters
# ---------------------------------------------------------------------------------------------
def allreduce_shared_grads(params: Sequence[Tensor], group=None, average: bool = True, world: Optional[int] = None) -> None:
    """One flat all-reduce over the gradients of replicated parameters (the background head: 643 floats).
    ``world``: the group size when the caller already knows it (the c10d queries cost tens of microseconds per step)."""
    if world is None:
        world, _ = _world(group)
    if world == 1:
        return
    gs = [p.grad for p in params if p.grad is not None]
    if not gs:
        return
    flat = torch.cat([g.reshape(-1) for g in gs])
    dist.all_reduce(flat, group=group)
    if average:
        flat /= world
    o = 0
    for g in gs:
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()


class SharedGradReducer:
    """`allreduce_shared_grads` started from inside the backward pass.  The background head's gradients are complete
    right after the compositing backward - before the expert MLP backward and the table scatter, which are 60 % of a
    step - so their all-reduce runs on a side stream underneath those kernels; ``wait()`` joins it in front of the
    optimiser.  Works under CUDA-graph capture (the side stream forks from and re-joins the capturing stream)."""

    def __init__(self, params: Sequence[Tensor], group=None, average: bool = True):
        self.params = [p for p in params if p.requires_grad]
        self.group, self.average = group, average
        self._ready = 0
        self._launched = False
        self._world = _world(group)[0]
        self._stream = torch.cuda.Stream(self.params[0].device) if self.params and self.params[0].is_cuda else None
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]

    def _on_grad(self, _p) -> None:
        self._ready += 1
        if self._ready < len(self.params) or self._world == 1:
            return
        self._launched = True
        if self._stream is None:
            allreduce_shared_grads(self.params, self.group, self.average, self._world)
            return
        self._stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._stream):
            allreduce_shared_grads(self.params, self.group, self.average, self._world)

    def wait(self) -> None:
        """Call between backward() and optimizer.step()."""
        if self._launched:
            if self._stream is not None:
                torch.cuda.current_stream().wait_stream(self._stream)
        else:  # some parameter got no gradient this step: reduce what there is, in line
            allreduce_shared_grads(self.params, self.group, self.average, self._world)
        self._ready, self._launched = 0, False

    def remove(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []


# ---------------------------------------------------------------------------------------------
# ray-sharded full-frame render
# ---------------------------------------------------------------------------------------------
def to_uint8(rgb: Tensor) -> Tensor:
    """[0,1] floats -> round-to-nearest 8-bit (what a viewer / video writer receives)."""
    return (rgb.clamp(0, 1) * 255.0 + 0.5).to(torch.uint8)


def render_image_sharded(render_rows: Callable[[int, int], Tuple[Tensor, Tensor, Tensor]], H: int, W: int, group=None,
                         gather: bool = True, as_uint8: bool = False, dst: Optional[int] = None):
    """``render_rows(r0, r1)`` renders image rows [r0, r1) on this rank and returns (rgb (rows*W,3), depth (rows*W,),
    acc (rows*W,)).  Rows are split in contiguous blocks (`shard_range`).
      gather=False            -> ((rgb, depth, acc) of this rank's block, (r0, r1))
      as_uint8=True, dst=r    -> the viewer's frame: 8-bit rgb tiles (3 B / pixel, 6.2 MB at 1080p) gathered on rank ``r``
                                 only, received in place in the frame buffer; returns (rgb8 (H,W,3) uint8, None, None) there
                                 and (None, None, None) elsewhere
      default                 -> every rank gets the fp32 frame (rgb (H,W,3), depth (H*W,), acc (H*W,)): one
                                 all-gather of (rows*W, 5) tiles straight into the frame-ordered buffer."""
    world, rank = _world(group)
    r0, r1 = shard_range(H, world, rank)
    rgb, depth, acc = render_rows(r0, r1)
    if world == 1:
        if as_uint8:
            return to_uint8(rgb).view(H, W, 3), None, None
        return rgb.view(H, W, 3), depth.view(-1), acc.view(-1)
    if not gather:
        return (rgb, depth, acc), (r0, r1)
    even = H % world == 0
    rows_max = (H + world - 1) // world
    n = (r1 - r0) * W
    if as_uint8:
        dst = 0 if dst is None else int(dst)
        tile = to_uint8(rgb).reshape(-1)
        if not even and tile.numel() < rows_max * W * 3:
            tile = torch.cat([tile, tile.new_zeros(rows_max * W * 3 - tile.numel())])
        if rank != dst:
            dist.gather(tile, None, dst=dst, group=group)
            return None, None, None
        if even:  # tiles land in place: the gather list is the frame itself, viewed per rank
            frame = torch.empty(world, rows_max * W * 3, device=rgb.device, dtype=torch.uint8)
            dist.gather(tile, list(frame.unbind(0)), dst=dst, group=group)
            return frame.view(H, W, 3), None, None
        tiles = [torch.empty_like(tile) for _ in range(world)]
        dist.gather(tile, tiles, dst=dst, group=group)
        parts = []
        for r in range(world):
            a, b = shard_range(H, world, r)
            parts.append(tiles[r][:(b - a) * W * 3])
        return torch.cat(parts).view(H, W, 3), None, None
    tile = torch.empty(rows_max * W, 5, device=rgb.device, dtype=rgb.dtype)  # one message: rgb | depth | acc
    tile[:n, 0:3], tile[:n, 3], tile[:n, 4] = rgb.view(-1, 3), depth.view(-1), acc.view(-1)
    if n < rows_max * W:
        tile[n:].zero_()
    full = torch.empty(world, rows_max * W, 5, device=rgb.device, dtype=rgb.dtype)
    dist.all_gather_into_tensor(full.view(world * rows_max * W, 5), tile, group=group)
    if even:
        full = full.view(H * W, 5)
    else:
        parts = []
        for r in range(world):
            a, b = shard_range(H, world, r)
            parts.append(full[r, :(b - a) * W])
        full = torch.cat(parts, 0)
    return full[:, 0:3].reshape(H, W, 3), full[:, 3], full[:, 4]


# ---------------------------------------------------------------------------------------------
# expert-parallel dispatch / combine
# ---------------------------------------------------------------------------------------------
class _AllToAll(torch.autograd.Function):
    """all_to_all_single with per-peer row counts; the backward pass is the same exchange with the counts swapped."""

    @staticmethod
    def forward(ctx, x: Tensor, send: List[int], recv: List[int], group):
        ctx.send, ctx.recv, ctx.group = send, recv, group
        out = x.new_empty((sum(recv),) + tuple(x.shape[1:]))
        dist.all_to_all_single(out, x.contiguous(), output_split_sizes=recv, input_split_sizes=send, group=group)
        return out

    @staticmethod
    def backward(ctx, g: Tensor):
        out = g.new_empty((sum(ctx.send),) + tuple(g.shape[1:]))
        dist.all_to_all_single(out, g.contiguous(), output_split_sizes=ctx.send, input_split_sizes=ctx.recv, group=ctx.group)
        return out, None, None, None


def exchange_counts(counts: Sequence[int], group=None) -> List[List[int]]:
    """counts[e] = routed samples this rank holds for expert e  ->  table[s][e] of every rank s (host ints)."""
    world, rank = _world(group)
    if world == 1:
        return [list(map(int, counts))]
    table: List[Optional[List[int]]] = [None] * world
    dist.all_gather_object(table, list(map(int, counts)), group=group)
    return table  # type: ignore[return-value]


def expert_parallel_apply(rows: Tensor, counts: Sequence[int], expert_fn: Callable[[Tensor, List[int], List[Tensor]], Tensor],
                          combine: Callable[[Tensor], Tensor], group=None) -> Tensor:
    """Soft-routed container forward with experts sharded over ranks (meta_container.py:305-321 semantics).

    rows       (M, D) this rank's routed samples (xyz + direction), one row per (point, expert) pair, GROUPED BY
               EXPERT 0..K-1 (on the GPU: `ops.a2a_pack`, built straight from rays and t)
    counts     K host ints: rows per expert (sum = M)
    expert_fn(rows (R, D), local_experts, index) -> (R, out_dim): evaluates this rank's experts; ``index[j]`` holds the
               rows of ``rows`` that belong to ``local_experts[j]`` (every row belongs to exactly one).
    combine(y (M, out_dim)) -> out (P, out_dim): blends the returned rows into the points with the routing weights
               (on the GPU: `ops.a2a_combine`).
    The result is differentiable w.r.t. the parameters expert_fn uses (gradients reach the owning rank through the
    reversed all-to-all)."""
    world, rank = _world(group)
    K = len(counts)
    owner = expert_owner(K, world)
    mine = [e for e in range(K) if owner[e] == rank]
    if rows.shape[0] != sum(counts):
        raise ValueError(f"{rows.shape[0]} routed rows but counts sum to {sum(counts)}")
    dev = rows.device
    if world == 1:
        bounds = [0]
        for c in counts:
            bounds.append(bounds[-1] + int(c))
        index = [torch.arange(bounds[e], bounds[e + 1], device=dev) for e in range(K)]
        y = expert_fn(rows, list(range(K)), index)
    else:
        table = exchange_counts(counts, group)
        send = [sum(int(counts[e]) for e in range(K) if owner[e] == d) for d in range(world)]
        recv = [sum(table[s][e] for e in mine) for s in range(world)]
        got = _AllToAll.apply(rows, send, recv, group)
        # received layout: [source 0: expert mine[0], mine[1], ... | source 1: ...]; regroup by local expert
        index, off = [[] for _ in mine], 0
        for s in range(world):
            for j, e in enumerate(mine):
                n = table[s][e]
                if n:
                    index[j].append(torch.arange(off, off + n, device=dev))
                off += n
        index = [torch.cat(ix) if ix else torch.empty(0, dtype=torch.long, device=dev) for ix in index]
        y_owner = expert_fn(got, mine, index)
        if torch.is_grad_enabled() and not y_owner.requires_grad:
            # every rank must take part in the backward exchange even if its own experts are frozen (config 4)
            y_owner = y_owner + torch.zeros(1, device=dev, dtype=y_owner.dtype, requires_grad=True)
        y = _AllToAll.apply(y_owner, recv, send, group)
    return combine(y)


def _evaluate_local(model, params, rows: Tensor, local: List[int], point: Tensor, segs: List[int], seg_dev: Optional[Tensor] = None) -> Tensor:
    """This rank's experts ``local`` over received rows: ``point`` (segs[-1],) int32 lists, expert by expert (each group
    padded to 128 slots with -1), the rows of ``rows`` (R,6) that belong to it.  One fused launch per direction."""
    from . import ops
    spec_all = model.expert_spec()
    dev = rows.device
    if seg_dev is None:
        seg_dev = torch.tensor(segs, dtype=torch.int32, device=dev)
    pl = ops.PairList(segs[-1], point, torch.ones(segs[-1], device=dev), seg_dev, segs)
    spec = ops.ExpertSpec(spec_all.grid, [spec_all.box_min[e] for e in local], [spec_all.extent[e] for e in local])
    tables = [model.submodules[e].xyz_encoder.hash_table for e in local]
    mlp: List[Tensor] = []
    for e in local:
        mlp.extend(model.submodules[e].mlp_tensors(model._sub_params(params, e)))
    return ops.ngp_experts(ops.PointSet(x6=rows), pl, spec, tables, mlp)


def container_evaluate_ep(model, pts, params=None, group=None) -> Tensor:
    """MetaContainer.forward over a point set with the experts sharded across ``group`` (block placement,
    `expert_owner`): route on this rank, ship every routed sample to its expert's owner, evaluate there with the
    fused kernels, ship [rgb, sigma] back and blend with the routing weights.  Every rank holds the full container
    (same state_dict); only the owner's copy of an expert is evaluated and receives gradients."""
    from . import ops
    world, rank = _world(group)
    K = len(model.submodules)
    with torch.no_grad():
        pairs = ops.route_pack(pts, model.centroids_host(), K, model.boundary_margin, model.cluster_2d)
        # send rows [xyz, dir] built from (rays, t), padding dropped, grouped by expert = by destination rank;
        # counts: the one host read of the step (the reference reads K nonzero() sizes, meta_container.py:309)
        rows6, sel_point, sel_weight, counts = ops.a2a_pack(pts, pairs, K)

    def expert_fn(rows: Tensor, local: List[int], index: List[Tensor]) -> Tensor:
        R = rows.shape[0]
        if R == 0 or not local:
            return rows.new_zeros(R, 4)
        point, segs = [], [0]
        for ix in index:
            n = int(ix.numel())
            pad = (-n) % ops.ACN_TILE
            point.append(ix.to(torch.int32))
            if pad:
                point.append(torch.full((pad,), -1, dtype=torch.int32, device=rows.device))
            segs.append(segs[-1] + n + pad)
        return _evaluate_local(model, params, rows.contiguous(), local, torch.cat(point), segs)

    return expert_parallel_apply(rows6, counts, expert_fn, lambda y: ops.a2a_combine(y, sel_point, sel_weight, pts.P), group)


def owned_param_groups(model, group=None):
    """`MetaContainer.get_param_groups()` restricted to what this rank trains under expert parallelism: its own
    experts' tensors plus the replicated background head (whose gradients `allreduce_shared_grads` averages).
    Returns (groups, shared_params)."""
    world, rank = _world(group)
    owner = expert_owner(len(model.submodules), world)
    enc, sig, col = [], [], []
    for e, sub in enumerate(model.submodules):
        if owner[e] != rank:
            continue
        g = sub.get_param_groups()
        enc.extend(g["encoding"]["params"])
        sig.extend(g["sigma"]["params"])
        col.extend(g["color"]["params"])
    shared = []
    if getattr(model, "use_bg_nerf", False):
        shared = list(model.bg_dir_enc.parameters()) + list(model.bg_mlp.parameters())
    groups = {}
    for name, ps in (("encoding", enc), ("sigma", sig), ("color", col), ("background", shared)):
        if ps:
            groups[name] = {"params": ps}
    return groups, shared


@torch.no_grad()
def sync_expert_params(model, group=None) -> None:
    """Under expert parallelism only the owner's copy of an expert is trained; every other rank's copy goes stale.  Before
    anything that needs the whole container on one rank - a checkpoint written by rank 0, replicated-expert rendering
    (`render_image_sharded`), evaluation - broadcast each expert's tensors from its owner (one flat message per expert)."""
    world, rank = _world(group)
    if world == 1:
        return
    owner = expert_owner(len(model.submodules), world)
    for e, sub in enumerate(model.submodules):
        ps = [p.data for p in sub.parameters()]
        small = [p for p in ps if p.numel() < (1 << 20)]
        for p in ps:
            if p.numel() >= (1 << 20):  # hash table: in place, no staging copy
                dist.broadcast(p, src=dist.get_global_rank(group, owner[e]) if group is not None else owner[e], group=group)
        if small:
            flat = torch.cat([p.reshape(-1) for p in small])
            dist.broadcast(flat, src=dist.get_global_rank(group, owner[e]) if group is not None else owner[e], group=group)
            o = 0
            for p in small:
                p.copy_(flat[o:o + p.numel()].view_as(p))
                o += p.numel()


# ---------------------------------------------------------------------------------------------
# the same exchange through NVLink peer memory: the kernels write / read the owner's buffers themselves
# ---------------------------------------------------------------------------------------------
class _Slab:
    """A raw device allocation exposed through __cuda_array_interface__ so torch can view it without owning it."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


class PeerExchange:
    """One exchange slab per rank, mapped by every peer (CUDA IPC over NVLink / NVSwitch), holding up to ``rows_cap``
    received rows: sync area | rows (cap,6) | y (cap,4) | g (cap,4).  Senders write sample rows and result gradients
    straight into the owner's slab and read the owner's [rgb,sigma] rows out of it (csrc/a2a.cu); no collective carries
    data.  Ordering between the phases is `barrier()`: one small kernel per rank that raises this rank's flag in every
    slab and waits for every flag in its own (release / acquire at system scope), stream-ordered like any other launch.
    The barrier number lives in the slab and is advanced on the device, the routed-sample counts and the slab layout
    (`plan`) never visit the host, and small all-reduces (background-head gradients, the clip norm) ride on a barrier:
    a whole expert-parallel training step is a fixed launch sequence and can be captured in a CUDA graph."""

    ROW_FLOATS = (6, 4, 4)

    def __init__(self, rows_cap: int, group=None, device=None, timeout_s: Optional[float] = None):
        import ctypes as C
        import os
        from ._lib import ACN_MAX_EXPERTS, ACN_MAX_PEERS, ACN_P2P_SYNC_BYTES, ACN_TILE, check, lib
        self.group = group
        self.world, self.rank = _world(group)
        if self.world > ACN_MAX_PEERS:
            raise ValueError(f"PeerExchange spans one box: at most {ACN_MAX_PEERS} ranks")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.cap = (int(rows_cap) + ACN_TILE - 1) // ACN_TILE * ACN_TILE
        self.timeout_s = float(os.environ.get("ACN_P2P_TIMEOUT_S", 120.0) if timeout_s is None else timeout_s)
        self._lib, self._check, self._C = lib(), check, C
        head = ACN_P2P_SYNC_BYTES // 4
        n_floats = head + self.cap * sum(self.ROW_FLOATS)
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            check(self._lib.acn_p2p_alloc(n_floats * 4, C.byref(ptr), handle), "acn_p2p_alloc")
        self._ptr = int(ptr.value)
        self._opened: List[int] = []
        handles: List[Optional[bytes]] = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(handles, handle.raw, group=group)
        self.peer = [0] * self.world  # slab base of every rank as mapped into this process
        for r in range(self.world):
            if r == self.rank:
                self.peer[r] = self._ptr
                continue
            q = C.c_void_p()
            with torch.cuda.device(self.device):
                check(self._lib.acn_p2p_open(C.c_char_p(handles[r]), C.byref(q)), "acn_p2p_open")
            self.peer[r] = int(q.value)
            self._opened.append(int(q.value))
        self._bases = (C.c_void_p * self.world)(*self.peer)
        flat = torch.as_tensor(_Slab(self._ptr, n_floats), device=self.device)
        self.sync = flat[:head].view(torch.int32)  # flag[16] | status | epoch | pad | table[16][ACN_MAX_EXPERTS+1] at int 32 | payload
        o0 = head
        o1, o2 = o0 + self.cap * 6, o0 + self.cap * 10
        self.rows, self.y, self.g = flat[o0:o1].view(self.cap, 6), flat[o1:o2].view(self.cap, 4), flat[o2:].view(self.cap, 4)
        self._byte_off = {"rows": o0 * 4, "y": o1 * 4, "g": o2 * 4}
        # device-side plan of one exchange (acn_p2p_plan)
        dev = self.device
        self.row0_dev = torch.zeros(ACN_MAX_EXPERTS, dtype=torch.int64, device=dev)
        self.seg_own = torch.zeros(ACN_MAX_EXPERTS + 1, dtype=torch.int32, device=dev)
        self.point_own = torch.full((self.cap,), -1, dtype=torch.int32, device=dev)
        self.live_own = torch.zeros(ACN_MAX_EXPERTS, dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self._routes = {}
        self._owner_arr = {}
        self.reducer = None  # SlabGradReducer: shared-parameter gradients that ride on the backward's barrier
        if self.world > 1:
            dist.barrier(group=group)  # every slab is mapped (and zeroed) before the first flag is written
        self.barrier()

    def base(self, kind: str, r: int) -> int:
        return self.peer[r] + self._byte_off[kind]

    def route(self, kind: str, K: int, owner: Sequence[int]):
        """Static half of an exchange route: the owners' buffers as mapped here (the rows come from `plan`)."""
        from . import ops
        key = (kind, K, tuple(owner))
        if key not in self._routes:
            self._routes[key] = ops.make_route([self.base(kind, owner[e]) for e in range(K)], [0] * K)
        return self._routes[key]

    def barrier(self, counts: Optional[Tensor] = None, payload: Optional[Tensor] = None, reduced: Optional[Tensor] = None,
                scale: float = 1.0) -> None:
        """Every rank's earlier work on its stream (including stores into peers) is complete and visible before any rank's
        later work.  ``counts`` (int32, <= 17 entries): the ranks also exchange them (`plan` reads the table on the
        device).  ``payload`` (fp32, <= 1024 entries): all-reduced (sum in rank order, times ``scale``) into ``reduced``."""
        C = self._C
        n = 0 if counts is None else int(counts.numel())
        npay = 0 if payload is None else int(payload.numel())
        if npay and (reduced is None or reduced.numel() != npay or not payload.is_contiguous() or not reduced.is_contiguous()):
            raise ValueError("barrier payload needs a contiguous `reduced` tensor of the same size")
        from . import ops
        with torch.cuda.device(self.device), ops._timed("p2p_barrier"):
            self._check(self._lib.acn_p2p_barrier(self._bases, self.world, self.rank,
                                                  None if counts is None else C.c_void_p(counts.data_ptr()), n,
                                                  None if not npay else C.c_void_p(payload.data_ptr()), npay,
                                                  None if not npay else C.c_void_p(reduced.data_ptr()), float(scale), self.timeout_s,
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)), "acn_p2p_barrier")
        ops._count(1)

    def arrive(self, payload: Optional[Tensor] = None) -> None:
        """First half of a barrier: publish ``payload`` (fp32, <= 1024 entries) and raise this rank's flag, without waiting."""
        C = self._C
        npay = 0 if payload is None else int(payload.numel())
        from . import ops
        with torch.cuda.device(self.device), ops._timed("p2p_arrive"):
            self._check(self._lib.acn_p2p_arrive(self._bases, self.world, self.rank, None if not npay else C.c_void_p(payload.data_ptr()), npay,
                                                 C.c_void_p(torch.cuda.current_stream().cuda_stream)), "acn_p2p_arrive")
        ops._count(1)

    def wait(self, n_payload: int = 0, reduced: Optional[Tensor] = None, scale: float = 1.0) -> None:
        """Second half: wait for every rank's `arrive`, then sum the payload slots (rank order, times ``scale``) into ``reduced``."""
        C = self._C
        from . import ops
        with torch.cuda.device(self.device), ops._timed("p2p_wait"):
            self._check(self._lib.acn_p2p_wait(self._bases, self.world, self.rank, int(n_payload), None if reduced is None else C.c_void_p(reduced.data_ptr()),
                                               float(scale), self.timeout_s, C.c_void_p(torch.cuda.current_stream().cuda_stream)), "acn_p2p_wait")
        ops._count(1)

    def allreduce_small(self, x: Tensor, scale: float = 1.0) -> Tensor:
        """Sum of a small fp32 tensor (<= 1024 elements) over the ranks, identical bits on every rank."""
        flat = x.reshape(-1).contiguous()
        out = torch.empty_like(flat)
        self.barrier(payload=flat, reduced=out, scale=scale)
        return out.view_as(x)

    def plan(self, K: int, owner: Sequence[int]) -> None:
        """Slab layout of the exchange whose counts the last `barrier(counts=...)` gathered: fills row0_dev, seg_own,
        point_own, live_own on the device (see `peer_layout` for the same arithmetic on the host)."""
        from ._lib import i32
        C = self._C
        key = (K, tuple(owner))
        if key not in self._owner_arr:
            self._owner_arr[key] = (i32 * K)(*[int(o) for o in owner])
        from . import ops
        with torch.cuda.device(self.device), ops._timed("p2p_plan"):
            self._check(self._lib.acn_p2p_plan(C.c_void_p(self._ptr), self.world, self.rank, K, self._owner_arr[key], self.cap,
                                               C.c_void_p(self.row0_dev.data_ptr()), C.c_void_p(self.seg_own.data_ptr()),
                                               C.c_void_p(self.point_own.data_ptr()), C.c_void_p(self.live_own.data_ptr()),
                                               C.c_void_p(self.g.data_ptr()), C.c_void_p(self.status.data_ptr()),
                                               C.c_void_p(torch.cuda.current_stream().cuda_stream)), "acn_p2p_plan")
        ops._count(1)

    def table(self, n: int) -> List[List[int]]:
        """Host copy of the last exchanged counts, one row of ``n`` ints per rank (diagnostics / tests; a host sync)."""
        from ._lib import ACN_MAX_EXPERTS
        w = ACN_MAX_EXPERTS + 1
        host = self.sync[32:32 + self.world * w].tolist()
        return [host[r * w:r * w + n] for r in range(self.world)]

    def check(self) -> None:
        """One host read: raises if an exchange overflowed a slab or a pair list since the last check (the steps in
        between then computed on truncated lists).  Call it every so often, and before trusting a checkpoint."""
        st = int(self.status.item())
        if st & 2:
            raise RuntimeError(f"PeerExchange: an owner received more rows than the slab capacity {self.cap}; allocate a larger rows_cap")
        if st & 4:
            raise RuntimeError("PeerExchange: a rank's routed pair list overflowed its speculative capacity (acn_route_pack_capacity)")

    def close(self) -> None:
        if self._lib is None:
            return
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)
        self.rows = self.y = self.g = self.sync = None
        for q in self._opened:
            self._lib.acn_p2p_close(q)
        self._lib.acn_p2p_free(self._ptr)
        self._lib = None


class SlabGradReducer:
    """Sum of the replicated parameters' gradients (the background head: 643 floats) over the ranks WITHOUT a collective
    kernel: the flat gradient rides as the payload of the barrier the expert-parallel backward needs anyway (between
    "gradient rows pushed" and "owner backward"), so nothing runs beside the persistent tensor-core backward kernel.
    ``wait()`` goes between backward() and optimizer.step(): it copies the reduced values into the .grad tensors, or - if
    the backward did not pass that barrier with every gradient ready - performs the reduction on a barrier of its own."""

    def __init__(self, ex: PeerExchange, params: Sequence[Tensor], scale: float = 1.0, early: bool = False):
        """``early``: post the gradients from inside the backward pass, as soon as the last of them exists (`PeerExchange.arrive`
        from a post-accumulate hook), and only wait for the peers in ``wait()`` - the expert backward between the two absorbs the
        skew between the ranks.  For steps with NO other barrier between the background head's backward and the optimiser (one
        expert per rank, config 2); the expert-parallel container step carries the payload on its own backward barrier instead."""
        self.ex, self.scale = ex, float(scale)
        self.params = [p for p in params if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(max(n, 1), device=ex.device)
        self.reduced = torch.zeros(max(n, 1), device=ex.device)
        self._done = False
        self._posted = False
        self._ready = 0
        self._hooks = []
        if early and ex.world > 1:
            self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        else:
            ex.reducer = self

    def _on_grad(self, _p) -> None:
        self._ready += 1
        if self._ready == len(self.params):
            self.ex.arrive(self.pack())
            self._posted = True

    def ready(self) -> bool:
        return bool(self.params) and all(p.grad is not None for p in self.params)

    def pack(self) -> Tensor:
        torch.cat([p.grad.reshape(-1) for p in self.params], out=self.flat)
        return self.flat

    def mark_done(self) -> None:
        self._done = True

    def wait(self) -> None:
        if not self.params or self.ex.world == 1:
            self._done = False
            return
        if self._hooks:
            if not self._posted:  # some parameter got no gradient this step: post what there is now
                for p in self.params:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
                self.ex.arrive(self.pack())
            self.ex.wait(self.flat.numel(), self.reduced, self.scale)
            self._posted, self._ready = False, 0
            self._done = True
        if not self._done:
            if not self.ready():
                for p in self.params:  # a parameter without a gradient this step contributes zeros
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
            self.ex.barrier(payload=self.pack(), reduced=self.reduced, scale=self.scale)
        # the reduced values become the gradients in place: .grad is re-pointed at slices of the (static) reduction buffer - no copy
        # kernels, and under CUDA-graph capture the optimiser's launch records these very addresses
        o = 0
        for p in self.params:
            p.grad = self.reduced[o:o + p.numel()].view_as(p)
            o += p.numel()
        self._done = False

    def remove(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []
        if self.ex.reducer is self:
            self.ex.reducer = None


class _PeerCombine(torch.autograd.Function):
    """forward: barrier (every owner's results are in its slab), pull the rows of this rank's pairs from their owners and
    blend (out[point] += w*y).  backward: push w*d_out[point] into the owners' gradient rows, barrier (carrying the shared
    gradients when a SlabGradReducer is attached), hand this rank's gradient rows to the expert backward.  ``anchor`` is a
    dummy input that requires grad, so that every rank takes part in the backward exchange even if its own experts are
    frozen (config 4)."""

    @staticmethod
    def forward(ctx, y_owner: Tensor, anchor: Tensor, ex: PeerExchange, route_y, route_g, sel_point: Tensor, sel_weight: Tensor, offs: Tensor,
                K: int, P: int):
        from . import ops
        ex.barrier()
        out = ops.a2a_pull_combine(route_y, sel_point, sel_weight, offs, K, P, row0_dev=ex.row0_dev)
        ctx.ex, ctx.route_g, ctx.K = ex, route_g, K
        ctx.save_for_backward(sel_point, sel_weight, offs)
        return out

    @staticmethod
    @_first_order_backward
    def backward(ctx, d_out):
        from . import ops
        ex = ctx.ex
        sel_point, sel_weight, offs = ctx.saved_tensors
        ops.a2a_push_grad(d_out, sel_point, sel_weight, offs, ctx.K, ctx.route_g, row0_dev=ex.row0_dev, cap_rows=ex.cap)
        red = ex.reducer
        if red is not None and ex.world > 1 and red.ready():
            ex.barrier(payload=red.pack(), reduced=red.reduced, scale=red.scale)
            red.mark_done()
        else:
            ex.barrier()
        return ex.g, torch.zeros(1, device=d_out.device), None, None, None, None, None, None, None, None


def peer_layout(table: Sequence[Sequence[int]], owner: Sequence[int], rank: int, tile: int = 128):
    """Row layout of the owners' slabs for one exchange (the host statement of csrc/a2a.cu k_p2p_plan; the CPU tests and
    the GPU test of the kernel both use it).  table[s][e] = routed samples rank s holds for expert e.
    An owner's slab holds its experts in order, each expert's block = the sub-blocks of source ranks 0..world-1 back to
    back, blocks padded to ``tile`` rows.  Returns (start[e] first row of e's block in its owner's slab, tot[e] rows of e
    over all sources, used[r] rows of rank r's slab, row0[e] first row of THIS rank's sub-block of e)."""
    world, K = len(table), len(owner)
    start, tot, used = [0] * K, [0] * K, [0] * world
    for e in range(K):
        tot[e] = sum(int(table[s][e]) for s in range(world))
        start[e] = used[owner[e]]
        used[owner[e]] += (tot[e] + tile - 1) // tile * tile
    row0 = [start[e] + sum(int(table[s][e]) for s in range(rank)) for e in range(K)]
    return start, tot, used, row0


def container_evaluate_p2p(model, pts, ex: PeerExchange, params=None) -> Tensor:
    """`container_evaluate_ep` with the data path in our own kernels: the routed samples of expert e are written by the
    routing rank directly into the slab of e's owner (24 B rows over NVLink), the owner evaluates its experts in place
    (results stored straight into the slab), and the routing rank pulls the [rgb,sigma] rows it needs (16 B) while
    blending.  Layout of an owner's slab: its experts in order, each expert's block = the sub-blocks of source ranks
    0..world-1, blocks padded to 128 rows: an expert's rows are contiguous, so the received rows need no regrouping.
    Nothing is read back by the host: the counts travel with the opening barrier, `acn_p2p_plan` turns them into rows
    and segment tables on the device, the pair lists keep their speculative capacity (`ex.check()` reports an overflow
    after the fact).  Per call: three barrier kernels, one more in the backward pass.  No NCCL call on the path."""
    from . import ops
    world, rank = ex.world, ex.rank
    K = len(model.submodules)
    owner = expert_owner(K, world)
    mine = [e for e in range(K) if owner[e] == rank]
    dev = pts.device
    with torch.no_grad():
        pairs = ops.route_pack(pts, model.centroids_host(), K, model.boundary_margin, model.cluster_2d, sync=False)
        counts, offs = ops.a2a_count(pairs, K)
        # the counts travel with the step's opening barrier: nobody writes into a slab before its owner is done with the
        # previous step, and every rank ends up with the whole table in its own slab
        ex.barrier(counts=counts)
        ex.plan(K, owner)
        sel_point, sel_weight = ops.a2a_push_rows(pts, pairs, K, offs, ex.route("rows", K, owner), pairs.M, row0_dev=ex.row0_dev,
                                                  cap_rows=ex.cap)
        ex.barrier()
    if mine:
        spec_all = model.expert_spec()
        spec = ops.ExpertSpec(spec_all.grid, [spec_all.box_min[e] for e in mine], [spec_all.extent[e] for e in mine])
        tables = [model.submodules[e].xyz_encoder.hash_table for e in mine]
        mlp: List[Tensor] = []
        for e in mine:
            mlp.extend(model.submodules[e].mlp_tensors(model._sub_params(params, e)))
        # the owner's rows are its slab, row id == point id: results are stored in place (no blend, no copy)
        pl = ops.PairList(ex.cap, ex.point_own, None, ex.seg_own[:len(mine) + 1], None, store_out=ex.y)
        y_owner = ops.ngp_experts(ops.PointSet(x6=ex.rows), pl, spec, tables, mlp)
    else:
        y_owner = ex.y
    anchor = torch.zeros(1, device=dev, requires_grad=torch.is_grad_enabled())
    return _PeerCombine.apply(y_owner, anchor, ex, ex.route("y", K, owner), ex.route("g", K, owner), sel_point, sel_weight, offs, K, pts.P)
