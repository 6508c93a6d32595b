"""One hash-grid expert with the reference's interface (``models/inr/meta_ngp.py`` MetaNGP :15-469):
same constructor arguments, sub-module names (hence state_dict keys) and ``forward / density /
color`` signatures, evaluated by the fused sm_100a kernels (csrc/ngp.cu)."""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple, Union

import torch
from torch import Tensor

from ... import ops
from ...nerfs.scene_box import SceneBox
from ..encodings import FrequencyEncoder, HashGridEncoder, SHEncoder
from ..metamodule import MetaLayerBlock, MetaLinear, MetaModule, MetaSequential
from ..occ_grid import OccGridEstimator
from ..trunc_exp import trunc_exp


class MetaNGP(MetaModule):
    def __init__(self, *, occ_conf: Optional[Dict] = None, scene_box: SceneBox, hidden: int = 64, sigma_depth: int = 2,
                 color_hidden: int = 64, geo_feat_dim: int = 15, color_depth: int = 3, use_sigmoid_rgb: bool = True,
                 hash_enc_conf=None, dir_encoding: str = "spherical", **kwargs) -> None:
        super().__init__()
        hash_enc_conf = hash_enc_conf or {}
        occ_conf = occ_conf or {}
        self.use_occ = bool(occ_conf.get("use_occ", False))
        self.occ_ready = False
        self.occ_premarked = False
        self.register_buffer("aabb_extent", scene_box.extent.clone().float())
        self.register_buffer("enc_eps", torch.tensor(1e-6, dtype=torch.float32), persistent=False)
        self.geo_feat_dim, self.use_sigmoid_rgb, self.scene_box = int(geo_feat_dim), bool(use_sigmoid_rgb), scene_box
        assert isinstance(scene_box.aabb, torch.Tensor) and scene_box.aabb.shape == (2, 3)

        self.xyz_encoder = HashGridEncoder(levels=hash_enc_conf.get("levels", 4), min_res=hash_enc_conf.get("min_res", 16),
                                           max_res=hash_enc_conf.get("max_res", 4096),
                                           log2_hashmap_size=hash_enc_conf.get("log2_hashmap_size", 19),
                                           features_per_level=hash_enc_conf.get("features_per_level", 2),
                                           interpolation=hash_enc_conf.get("interpolation", "Linear"))
        dir_encoding = dir_encoding.lower()
        if dir_encoding == "spherical":
            self.dir_encoder = SHEncoder(levels=4)
        elif dir_encoding == "frequency":
            self.dir_encoder = FrequencyEncoder(in_dim=3, pe_dim=4, include_input=True, use_pi=False)
        else:
            raise ValueError(f"Unsupported dir_encoding: {dir_encoding}")
        in_ch_dir = self.dir_encoder.out_dim

        trunk, last = [], self.xyz_encoder.out_dim
        for _ in range(max(int(sigma_depth), 0)):
            trunk.append(MetaLayerBlock(last, hidden, activation="relu"))
            last = hidden
        self.sigma_trunk = MetaSequential(*trunk)
        self.sigma_head = MetaLinear(last, 1)
        with torch.no_grad():
            self.sigma_head.bias.fill_(-1.0)
        self.geo_head = MetaLinear(last, self.geo_feat_dim)
        self.sigma_act = trunc_exp

        color, last = [], self.geo_feat_dim + in_ch_dir
        for _ in range(max(int(color_depth), 0)):
            color.append(MetaLayerBlock(last, color_hidden, activation="relu"))
            last = color_hidden
        color.append(MetaLinear(last, 3))
        self.color_mlp = MetaSequential(*color)
        self.rgb_act = torch.nn.Sigmoid() if self.use_sigmoid_rgb else torch.nn.Identity()

        # occupancy state, as meta_ngp.py:103-142 (same attribute and buffer names: `scene_aabb`, `occ_grid.*` are checkpointed)
        if self.use_occ:
            self.render_step_size = (float(occ_conf["render_step_size"]) if occ_conf.get("render_step_size") is not None
                                     else float(scene_box.get_diagonal_length()) / 1000.0)
            self.occ_thre = float(occ_conf.get("occ_thre", 1e-2))
            self.alpha_thre = float(occ_conf.get("alpha_thre", 1e-2))
            self.cone_angle = float(occ_conf.get("cone_angle", 1.0 / 256.0))
            self.near_plane = float(occ_conf.get("near_plane", 0.05))
            self.far_plane = float(occ_conf.get("far_plane", 1e3))
            self.occ_update_interval = int(occ_conf.get("update_interval", 16))
            self.occ_warmup_steps = int(occ_conf.get("warmup_steps", 256))
            self.occ_cosine_anneal = bool(occ_conf.get("cosine_anneal", True))
            self.occ_alpha_thre_start = float(occ_conf.get("alpha_thre_start", 0.0))
            self.occ_alpha_thre_end = float(occ_conf.get("alpha_thre_end", self.alpha_thre))
            self.occ_ema_decay = float(occ_conf.get("ema_decay", 0.95))
            self.occ_resolution = int(occ_conf.get("resolution", 128))
            self.occ_levels = int(occ_conf.get("levels", 4))
            self.register_buffer("scene_aabb", torch.cat([scene_box.min, scene_box.max]).flatten().float())
            self.occ_grid = OccGridEstimator(roi_aabb=self.scene_aabb, resolution=self.occ_resolution, levels=self.occ_levels)
            self._check_aabb()
            self.occ_frozen = bool(occ_conf.get("occ_frozen", False))
            self.occ_ready = bool(occ_conf.get("occ_ready", False))
            self.num_occ_updates = 0

        # the fused kernel is specialised on the shape nerf_runner.py actually builds
        self._fused_ok = (self.xyz_encoder.levels == 16 and hidden == 64 and sigma_depth == 2 and color_hidden == 64
                          and color_depth == 2 and self.geo_feat_dim == 15 and dir_encoding == "spherical" and self.use_sigmoid_rgb)
        self._box_host = None

    # ---- fused-path plumbing ------------------------------------------------
    def _require_fused(self):
        if not self._fused_ok:
            raise NotImplementedError(
                "the sm_100a expert kernel implements the production architecture only: 16-level hash grid, hidden=64, "
                "sigma_depth=2, color_depth=2, color_hidden=64, geo_feat_dim=15, spherical directions, sigmoid rgb")

    def box_host(self):
        """(min, extent) as python floats for the kernel-parameter block (cached; boxes are constants)."""
        if self._box_host is None:
            self._box_host = (self.scene_box.min.detach().float().cpu().tolist(), self.aabb_extent.detach().float().cpu().tolist())
        return self._box_host

    def _apply(self, fn, *a, **k):
        self._box_host = None
        object.__setattr__(self, "_mlp_cache", None)
        return super()._apply(fn, *a, **k)

    def mlp_tensors(self, params: Optional[Dict[str, Tensor]] = None) -> List[Tensor]:
        """The 14 MLP tensors in ABI order, each taken from ``params`` when present (fast weights)."""
        own = getattr(self, "_mlp_cache", None)
        first = f"{ops.NGP_LAYER_KEYS[0]}.weight"
        if own is None or own[0][1] is not self.get_parameter(first):  # walking the module tree costs ~0.1 ms per expert per call
            named = dict(self.meta_named_parameters())
            own = [(f"{key}.{leaf}", named[f"{key}.{leaf}"]) for key in ops.NGP_LAYER_KEYS for leaf in ("weight", "bias")]
            object.__setattr__(self, "_mlp_cache", own)
        if params is None:
            return [p for _, p in own]
        return [params[n] if n in params else p for n, p in own]

    def expert_spec(self) -> ops.ExpertSpec:
        mn, ext = self.box_host()
        return ops.ExpertSpec(self.xyz_encoder.grid_desc(), [mn], [ext])

    # ---- reference interface --------------------------------------------------
    def _world_to_unit(self, x: Tensor) -> Tensor:
        x01 = (x - self.scene_box.min.to(x.device)) / self.aabb_extent
        return x01.clamp(self.enc_eps, 1.0 - self.enc_eps)

    def evaluate(self, pts: ops.PointSet, params=None) -> Tensor:
        """(P,4) [rgb, sigma] for a point set (explicit rows or implicit ray samples); expert-level ``params``."""
        self._require_fused()
        return ops.ngp_experts(pts, ops.PairList(pts.P), self.expert_spec(), [self.xyz_encoder.hash_table], self.mlp_tensors(params))

    def forward(self, x_d: Tensor, params=None) -> Tensor:
        """(...,6) [xyz, dir] -> (...,4) [rgb, sigma]."""
        assert x_d.shape[-1] == 6, f"Expected (...,6) [xyz,dir], got {x_d.shape}"
        return self.evaluate(ops.PointSet(x6=x_d.reshape(-1, 6)), params).reshape(*x_d.shape[:-1], 4)

    def density(self, x: Tensor, params=None, return_feats: bool = False) -> Union[Tensor, Dict[str, Tensor]]:
        """sigma (...,1) [and geo features] at world coordinates.  Layer-wise path (API parity for
        callers that query density alone, e.g. occupancy updates); the renderer uses forward()."""
        h = self.xyz_encoder(self._world_to_unit(x))
        h = self.sigma_trunk(h, params=self.get_subdict(params, "sigma_trunk"))
        sigma = self.sigma_act(self.sigma_head(h, params=self.get_subdict(params, "sigma_head")))
        if not return_feats:
            return sigma
        return {"sigma": sigma, "geo_feat": self.geo_head(h, params=self.get_subdict(params, "geo_head"))}

    def color(self, d: Tensor, geo_feat: Tensor, params=None) -> Tensor:
        d = d / d.norm(dim=-1, keepdim=True).clamp_min(1e-9)
        h = torch.cat([geo_feat, self.dir_encoder(d)], dim=-1)
        return self.rgb_act(self.color_mlp(h, params=self.get_subdict(params, "color_mlp")))

    # ---- occupancy (meta_ngp.py:244-443) ----------------------------------------------
    def _check_aabb(self) -> None:
        aabb = self.scene_aabb
        assert aabb.dtype == torch.float32 and aabb.isfinite().all() and aabb.numel() == 6
        if not (aabb[:3] < aabb[3:]).all():
            raise ValueError(f"AABB invalid: min>=max ({aabb[:3]} vs {aabb[3:]})")

    def _anneal_alpha_thre(self, step: int) -> None:
        """alpha threshold from start to end over the warm-up (cosine or linear), then held (meta_ngp.py:245-259)."""
        if step < self.occ_warmup_steps:
            t = step / max(1, self.occ_warmup_steps - 1)
            if self.occ_cosine_anneal:
                t = 0.5 * (1 - math.cos(math.pi * t))
            self.alpha_thre = (1 - t) * self.occ_alpha_thre_start + t * self.occ_alpha_thre_end
        else:
            self.alpha_thre = self.occ_alpha_thre_end

    @torch.no_grad()
    def density_at(self, x: Tensor, params=None) -> Tensor:
        """sigma (N,) at world points through the fused kernels (the occupancy update evaluates 4 x 128^3 points and the
        training-time visibility filter every marched sample: not a job for the layer-wise `density`)."""
        x = x.reshape(-1, 3).float()
        d = torch.zeros_like(x)
        d[:, 2] = 1.0  # the colour branch is evaluated and dropped; any unit direction does
        return self.evaluate(ops.PointSet(x6=torch.cat([x, d], dim=-1)), params)[:, 3]

    @torch.no_grad()
    def premark_invisible_cells(self, mds, near_plane: float = 0.05, chunk: int = 32 ** 3) -> None:
        """One-time visibility pruning (meta_ngp.py:319-353): cells no camera sees get occupancy -1."""
        if not self.use_occ or self.occ_premarked:
            return
        mds = [md for md in mds if md is not None]
        if len(mds) == 0:
            self.occ_premarked = True
            return
        device = self.occ_grid.aabbs.device

        def make_K(md):
            k = torch.as_tensor(md.intrinsics, dtype=torch.float32)
            if k.numel() == 9:
                return k.view(3, 3)
            if k.numel() == 4:
                fx, fy, cx, cy = k.unbind()
                return torch.tensor([[fx, 0.0, cx], [0.0, fy, cy], [0.0, 0.0, 1.0]], dtype=torch.float32)
            raise ValueError(f"Unsupported intrinsics shape: {tuple(k.shape)}")

        K = torch.stack([make_K(md) for md in mds]).to(device)
        c2w = torch.stack([torch.as_tensor(md.c2w, dtype=torch.float32) for md in mds]).to(device).clone()
        c2w[:, :3, :3] = c2w[:, :3, :3] @ torch.diag(torch.tensor([1.0, -1.0, -1.0], device=device))  # RUB -> RDF camera axes
        self.occ_grid.mark_invisible_cells(K=K, c2w=c2w, width=int(mds[0].W), height=int(mds[0].H), near_plane=float(near_plane), chunk=chunk)
        self.occ_premarked = True

    @torch.no_grad()
    def maybe_update_occ_grid(self, step: int, params=None) -> None:
        """Periodic EMA update of the occupancy grid during training (meta_ngp.py:355-387)."""
        if not (self.training and self.use_occ and not self.occ_frozen):
            return
        self.occ_ready = step >= self.occ_warmup_steps
        self._anneal_alpha_thre(step)
        self.occ_grid.update_every_n_steps(step=step, occ_eval_fn=lambda x: self.density_at(x, params) * self.render_step_size,
                                           occ_thre=self.occ_thre, ema_decay=self.occ_ema_decay, warmup_steps=self.occ_warmup_steps,
                                           n=self.occ_update_interval)
        self.num_occ_updates += 1

    @torch.no_grad()
    def occupancy_marching(self, rays: Tensor, *, params=None, render_step_size: Optional[float] = None, alpha_thre: Optional[float] = None,
                           cone_angle: Optional[float] = None) -> Tuple[Tensor, Tensor, Tensor]:
        """(ray_indices, t_starts, t_ends) of this expert's occupancy grid along (N,8) rays (meta_ngp.py:389-443); in training
        the samples are jittered and filtered by the current density (transmittance early stop + alpha threshold)."""
        if getattr(self, "occ_grid", None) is None:
            raise RuntimeError("MetaNGP: occ_grid missing")
        rays = rays.contiguous()
        o, d = rays[:, :3], rays[:, 3:6]
        sigma_fn = None
        if self.training:
            def sigma_fn(t_starts, t_ends, ray_indices):
                mids = 0.5 * (t_starts + t_ends)
                return self.density_at(o[ray_indices] + d[ray_indices] * mids[:, None], params)
        return self.occ_grid.sampling(rays_o=o, rays_d=d, sigma_fn=sigma_fn, t_min=rays[:, 6], t_max=rays[:, 7],
                                      render_step_size=self.render_step_size if render_step_size is None else render_step_size,
                                      stratified=self.training, cone_angle=self.cone_angle if cone_angle is None else cone_angle,
                                      alpha_thre=self.alpha_thre if alpha_thre is None else alpha_thre)

    def get_param_groups(self) -> Dict[str, Dict]:
        return {"encoding": {"params": list(self.xyz_encoder.parameters())},
                "sigma": {"params": list(self.sigma_trunk.parameters()) + list(self.sigma_head.parameters())
                          + list(self.geo_head.parameters())},
                "color": {"params": list(self.color_mlp.parameters())}}
