"""One hash-grid expert with the reference's interface (``models/inr/meta_ngp.py`` MetaNGP :15-469):
same constructor arguments, sub-module names (hence state_dict keys) and ``forward / density /
color`` signatures, evaluated by the fused sm_100a kernels (csrc/ngp.cu)."""
from __future__ import annotations

from typing import Dict, List, Optional, Union

import torch
from torch import Tensor

from ... import ops
from ...nerfs.scene_box import SceneBox
from ..encodings import FrequencyEncoder, HashGridEncoder, SHEncoder
from ..metamodule import MetaLayerBlock, MetaLinear, MetaModule, MetaSequential
from ..trunc_exp import trunc_exp


class MetaNGP(MetaModule):
    def __init__(self, *, occ_conf: Optional[Dict] = None, scene_box: SceneBox, hidden: int = 64, sigma_depth: int = 2,
                 color_hidden: int = 64, geo_feat_dim: int = 15, color_depth: int = 3, use_sigmoid_rgb: bool = True,
                 hash_enc_conf=None, dir_encoding: str = "spherical", **kwargs) -> None:
        super().__init__()
        hash_enc_conf = hash_enc_conf or {}
        occ_conf = occ_conf or {}
        if bool(occ_conf.get("use_occ", False)):
            raise NotImplementedError("occupancy-grid marching (SURVEY 8(f)-1) is not built yet; construct with use_occ=False")
        self.use_occ = False
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

    def maybe_update_occ_grid(self, step: int, params=None) -> None:
        return None

    def get_param_groups(self) -> Dict[str, Dict]:
        return {"encoding": {"params": list(self.xyz_encoder.parameters())},
                "sigma": {"params": list(self.sigma_trunk.parameters()) + list(self.sigma_head.parameters())
                          + list(self.geo_head.parameters())},
                "color": {"params": list(self.color_mlp.parameters())}}
