"""K-expert container with the reference's interface (``models/inr/meta_container.py``
MetaContainer :21-503): same constructor, buffers / sub-module names (state_dict keys),
``forward``, ``_routing``, ``background_color``, ``get_param_groups``.

Dispatch differs by design: instead of K x (nonzero -> index_select -> expert -> index_add_),
one routing+pack kernel groups the (point, expert) pairs by expert and ONE fused launch
evaluates every expert over its own segment (csrc/routing.cu, csrc/ngp.cu)."""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from ... import ops
from ..encodings import FrequencyEncoder, SHEncoder
from ..metamodule import MetaModule
from .meta_ngp import MetaNGP


def build_expert(nerf_variant: str, **nerf_kwargs) -> nn.Module:
    if nerf_variant == "instant":
        return MetaNGP(**nerf_kwargs)
    raise NotImplementedError(
        f"nerf_variant={nerf_variant!r}: the reference's 'vanilla' expert cannot be built through its own container "
        "(SURVEY.md note V); the PE-MLP expert is driven stand-alone (nerf_sys_b200.models.inr.meta_vanilla)")


class MetaContainer(MetaModule):
    def __init__(self, num_submodules: int, centroids: torch.Tensor, aabb: torch.Tensor, nerf_variant: str = "instant",
                 boundary_margin: float = 1.0, cluster_2d: bool = True, joint_training: bool = False, use_bg_nerf: bool = True,
                 bg_hidden: int = 32, bg_encoding: str = "spherical", occ_conf: Optional[Dict] = None, **nerf_kwargs):
        super().__init__()
        assert num_submodules > 0
        assert centroids.ndim == 2 and centroids.size(0) == num_submodules
        assert boundary_margin >= 1.0
        occ_conf = occ_conf or {}
        self.register_buffer("scene_aabb_vec", torch.cat([aabb[0], aabb[1]], dim=0).float(), persistent=True)
        self.register_buffer("centroids", centroids.to(torch.float32), persistent=True)
        self.use_occ = bool(occ_conf.get("use_occ", False))
        self.boundary_margin, self.cluster_2d = float(boundary_margin), bool(cluster_2d)
        self.joint_training = bool(joint_training)
        self._coord_idx = (1, 2) if self.cluster_2d else (0, 1, 2)
        self.nerf_variant, self.dim_out = nerf_variant, 4

        expert_box_list = nerf_kwargs.pop("expert_box_list")
        self.submodules = nn.ModuleList()
        for box in expert_box_list:
            self.submodules.append(build_expert(nerf_variant, **{**nerf_kwargs, "occ_conf": occ_conf, "scene_box": box}))

        self.use_bg_nerf = bool(use_bg_nerf)
        if self.use_bg_nerf:
            if bg_encoding == "spherical":
                self.bg_dir_enc = SHEncoder(levels=4)
            else:
                self.bg_dir_enc = FrequencyEncoder(pe_dim=4, include_input=True, use_pi=False)
            self.bg_hidden_dim = int(bg_hidden)
            self.bg_mlp = nn.Sequential(nn.Linear(self.bg_dir_enc.out_dim, self.bg_hidden_dim, bias=True), nn.ReLU(),
                                        nn.Linear(self.bg_hidden_dim, 3, bias=True), nn.Sigmoid())
        self._cent_host = None
        self._ep = None  # (process group or None for the default group) when the experts are sharded over ranks
        self._deferred = False  # training-time routing without the host read (set_deferred_routing)
        self._live = None       # (live flags int32 (K,), status int32 (1,)) on the parameters' device

    # ---- expert parallelism (new: the reference is single-GPU) ----------
    def set_expert_parallel(self, enabled: bool = True, group=None, exchange=None) -> "MetaContainer":
        """Shard the experts over the ranks of ``group`` (block placement, parallel.expert_owner): `evaluate` /
        `forward` then route locally, ship every routed sample to its expert's owner with one all-to-all, evaluate
        there and ship [rgb, sigma] back.  Every rank keeps the full state_dict; only the owner's copy of an expert
        is evaluated, receives gradients and needs optimiser state (parallel.owned_param_groups).  With
        ``exchange`` (a parallel.PeerExchange) the samples and results move through NVLink peer memory inside the
        pack / combine kernels instead of NCCL all-to-alls."""
        object.__setattr__(self, "_ep", (group, exchange) if enabled else None)
        return self

    # ---- routing without a host read (CUDA-graph capturable training steps) -------------
    def set_deferred_routing(self, enabled: bool = True) -> "MetaContainer":
        """Training-mode `evaluate` normally reads the (K+1)-int segment table back to size its pair lists exactly and to
        learn which experts received samples (the reference reads K nonzero() sizes, meta_container.py:309).  Deferred:
        the pair lists keep their speculative capacity, every expert returns gradients, and the device-side
        ``live_flags`` tell the optimiser which experts to skip (optim.bind_expert_liveness) - the step makes no host
        read and can be captured in a CUDA graph.  ``check_routing()`` reports a capacity overflow after the fact."""
        object.__setattr__(self, "_deferred", bool(enabled))
        return self

    def _live_state(self):
        dev = self.centroids.device
        if self._live is None or self._live[0].device != dev:
            K = len(self.submodules)
            object.__setattr__(self, "_live", (torch.ones(K, dtype=torch.int32, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)))
        return self._live

    @property
    def live_flags(self) -> torch.Tensor:
        """int32 (K,) on the device: expert k received a sample in the last deferred `evaluate`."""
        return self._live_state()[0]

    def live_flag_slots(self, experts=None):
        """(flags tensor, {expert index: slot}) for optim.bind_expert_liveness; under peer-memory expert parallelism the
        flags are the exchange's (one slot per OWNED expert, written by acn_p2p_plan)."""
        K = len(self.submodules)
        if self._ep is not None and self._ep[1] is not None:
            from ... import parallel
            ex = self._ep[1]
            mine = [e for e, o in enumerate(parallel.expert_owner(K, ex.world)) if o == ex.rank]
            return ex.live_own, {e: j for j, e in enumerate(mine)}
        ks = range(K) if experts is None else experts
        return self.live_flags, {int(k): int(k) for k in ks}

    def check_routing(self) -> None:
        """One host read: raises if a deferred routing overflowed its speculative pair capacity since the last check."""
        if self._ep is not None and self._ep[1] is not None:
            self._ep[1].check()
        if self._live is not None and int(self._live[1].item()) != 0:
            self._live[1].zero_()
            raise RuntimeError("deferred routing produced more (sample, expert) pairs than the speculative pair-list capacity "
                               f"(boundary_margin {self.boundary_margin}); results since the last check are truncated")

    # ---- host copy of the centroids (they travel in the kernel-parameter block) ----
    def centroids_host(self) -> List[float]:
        key = (self.centroids.data_ptr(), self.centroids._version)
        if self._cent_host is None or self._cent_host[0] != key:
            self._cent_host = (key, self.centroids.detach().float().cpu().reshape(-1).tolist())
        return self._cent_host[1]

    # ---- routing ------------------------------------------------------------------
    def _routing(self, pts: torch.Tensor) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
        """(N,3) world points -> (weights (N,K), None) for soft routing or (None, argmin (N,))."""
        assert pts.dim() == 2 and pts.shape[-1] == 3, "pts must be (N,3)"
        x6 = torch.cat([pts.float(), torch.zeros_like(pts, dtype=torch.float32)], dim=-1)
        with torch.no_grad():
            return ops.routing(ops.PointSet(x6=x6), self.centroids_host(), len(self.submodules), self.boundary_margin, self.cluster_2d)

    def expert_spec(self) -> ops.ExpertSpec:
        boxes = [sub.box_host() for sub in self.submodules]
        return ops.ExpertSpec(self.submodules[0].xyz_encoder.grid_desc(), [b[0] for b in boxes], [b[1] for b in boxes])

    def _sub_params(self, params, k: int):
        return None if params is None else self.get_subdict(params, f"submodules.{k}")

    def evaluate(self, pts: ops.PointSet, params=None, active_module: Optional[int] = None) -> torch.Tensor:
        """(P,4) [rgb, sigma] for a point set (explicit rows or implicit ray samples), in the set's own point order."""
        K = len(self.submodules)
        if active_module is not None:
            return self.submodules[active_module].evaluate(pts, self._sub_params(params, active_module))
        for sub in self.submodules:
            sub._require_fused()
        if self._ep is not None:
            from ... import parallel
            if self._ep[1] is not None:
                return parallel.container_evaluate_p2p(self, pts, self._ep[1], params=params)
            if parallel._world(self._ep[0])[0] > 1:
                return parallel.container_evaluate_ep(self, pts, params=params, group=self._ep[0])
        if K == 1 and self.boundary_margin > 1.0:
            # a single centroid always gets weight 1: skip the pack, keep the fused launch
            return self.submodules[0].evaluate(pts, self._sub_params(params, 0))
        # a renderer that verifies the capacity once per frame (render_image) lets the pack skip its host read
        defer = getattr(pts, "route_checks", None) is not None and not torch.is_grad_enabled()
        defer_train = self._deferred and torch.is_grad_enabled()
        with torch.no_grad():
            pairs = ops.route_pack(pts, self.centroids_host(), K, self.boundary_margin, self.cluster_2d, sync=not (defer or defer_train))
            if defer_train:
                live, status = self._live_state()
                ops.route_status(pairs.seg, K, pairs.M, live, status)
        if defer:
            pts.route_checks.append((pairs.seg, pairs.M))
        tables = [sub.xyz_encoder.hash_table for sub in self.submodules]
        mlp: List[torch.Tensor] = []
        for k, sub in enumerate(self.submodules):
            mlp.extend(sub.mlp_tensors(self._sub_params(params, k)))
        return ops.ngp_experts(pts, pairs, self.expert_spec(), tables, mlp)

    def forward(self, x: torch.Tensor, params=None, active_module: Optional[int] = None) -> torch.Tensor:
        assert x.dim() == 2 and x.shape[-1] >= 6, "x must be (N,D>=6)"
        x6 = x if x.shape[-1] == 6 else x[:, :6]
        return self.evaluate(ops.PointSet(x6=x6), params=params, active_module=active_module)

    def _evaluate_expert_keys(self, x6: torch.Tensor, params, active_module):
        # with active_module the reference's density()/color() hand the expert its CONTAINER-level sub-dict
        # (meta_container.py:170-177, 236-239), unlike render_rays, whose caller passes expert-level keys
        if active_module is not None:
            sub = self.submodules[active_module]
            return sub.evaluate(ops.PointSet(x6=x6), self._sub_params(params, active_module))
        return self.evaluate(ops.PointSet(x6=x6), params=params)

    def density(self, xyz: torch.Tensor, params=None, active_module: Optional[int] = None) -> torch.Tensor:
        """Routed density query, sigma (N,) (meta_container.py:217-273): the route -> dispatch -> combine of `forward`
        with the sigma channel kept (soft routing blends sigma linearly with the routing weight, as `forward` does)."""
        assert xyz.dim() == 2 and xyz.shape[-1] == 3, "xyz must be (N,3)"
        d = torch.zeros_like(xyz, dtype=torch.float32)
        d[:, 2] = 1.0  # the colour branch is evaluated and dropped; any unit direction does
        out = self._evaluate_expert_keys(torch.cat([xyz.float(), d], dim=-1), params, active_module)
        return out[:, 3].to(dtype=xyz.dtype)

    def color(self, xyz: torch.Tensor, dirs: torch.Tensor, params=None, active_module: Optional[int] = None) -> torch.Tensor:
        """Routed RGB query, (N,3) (meta_container.py:137-215): rgb_mix = sum_k w_k rgb_k."""
        assert xyz.dim() == 2 and xyz.shape[-1] == 3, "xyz must be (N,3)"
        assert dirs.dim() == 2 and dirs.shape[-1] == 3, "dirs must be (N,3)"
        dirs = torch.nn.functional.normalize(dirs.to(xyz.device).float(), dim=-1)
        out = self._evaluate_expert_keys(torch.cat([xyz.float(), dirs], dim=-1), params, active_module)
        return out[:, :3].to(dtype=xyz.dtype)

    # ---- background ---------------------------------------------------------------
    def background_color(self, d: torch.Tensor) -> torch.Tensor:
        if not self.use_bg_nerf:
            raise RuntimeError("background_color called but use_bg_nerf=False")
        if d.dim() not in (2, 3):
            raise ValueError(f"background_color expects (N,3) or (B,N,3), got {tuple(d.shape)}")
        if not isinstance(self.bg_dir_enc, SHEncoder) or self.bg_hidden_dim != 32:
            raise NotImplementedError("the sm_100a background kernel implements SH16 -> 32 -> 3")
        l0, l1 = self.bg_mlp[0], self.bg_mlp[2]
        rgb = ops.background(d.reshape(-1, 3) if d.dim() == 3 else d, l0.weight, l0.bias, l1.weight, l1.bias)
        return rgb.view(*d.shape[:-1], 3)

    # ---- occupancy hooks (meta_container.py:385-470) -----------------------------------
    def maybe_update_expert_occupancies(self, step: int, params=None) -> None:
        for sub in self.submodules:  # `params` is handed over as it is, container-level keys included (meta_container.py:385-388)
            sub.maybe_update_occ_grid(step, params)

    def freeze_expert_occupancies(self, flag: bool) -> None:
        for sub in self.submodules:
            sub.occ_frozen = flag

    @torch.no_grad()
    def premark_invisible_expert_cells(self, metas, near_plane: float = 0.0, chunk: int = 32 ** 3) -> List[int]:
        """One-time visibility pruning of every expert's grid from the camera metadata; returns the marked cells per expert."""
        if self.cells_premarked or not self.use_occ:
            return [0] * len(self.submodules)
        marked = []
        for sub in self.submodules:
            sub.premark_invisible_cells(metas, near_plane=float(near_plane), chunk=chunk)
            marked.append(int((sub.occ_grid.occs < 0).sum().item()) if hasattr(sub.occ_grid, "occs") else 0)
        return marked

    @property
    def occ_ready(self) -> bool:
        return all(getattr(sub, "occ_ready", False) for sub in self.submodules)

    @property
    def cells_premarked(self) -> bool:
        return all(getattr(sub, "occ_premarked", False) for sub in self.submodules)

    # ---- optimiser groups ----------------------------------------------------------
    def get_param_groups(self) -> Dict[str, Dict]:
        enc, sig, col, bg = [], [], [], []
        for sub in self.submodules:
            g = sub.get_param_groups()
            enc.extend(g["encoding"]["params"])
            sig.extend(g["sigma"]["params"])
            col.extend(g["color"]["params"])
        if self.use_bg_nerf:
            bg.extend(list(self.bg_dir_enc.parameters()))
            bg.extend(list(self.bg_mlp.parameters()))
        groups: Dict[str, Dict] = {}
        for name, ps in (("encoding", enc), ("sigma", sig), ("color", col), ("background", bg)):
            if ps:
                groups[name] = {"params": ps}
        return groups
