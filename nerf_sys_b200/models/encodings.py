"""Encoders with the reference's class names, constructor arguments and parameter names
(``models/encodings.py``: SHEncoder :84-151, HashGridEncoder :154-381, FrequencyEncoder :387-444),
computed by the sm_100a library.  The ``implementation`` argument is accepted for call-site
compatibility and ignored: there is one backend.

The hash grid follows the reference's *torch* backend function (every level hashed,
res_l = floor(min_res * g^l), table ``hash_table`` of shape (L*T, F) fp32), which is also the
state_dict layout of torch-backend checkpoints.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn

from .. import ops

MAX_SH_DEGREE = 4
INTERPOLATIONS = ("Nearest", "Linear", "Smoothstep")


class SHEncoder(nn.Module):
    def __init__(self, levels: int = 4, implementation: str = "tcnn") -> None:
        super().__init__()
        if levels != 4:
            raise ValueError(f"the sm_100a SH kernel implements levels=4 (16 components), got {levels}")
        self.levels, self.degree, self._out_dim = 4, 3, 16

    @property
    def out_dim(self) -> int:
        return self._out_dim

    def get_out_dim(self) -> int:
        return self._out_dim

    def forward(self, d: torch.Tensor) -> torch.Tensor:
        assert d.shape[-1] == 3, f"Expected (...,3), got {tuple(d.shape)}"
        return ops.sh16(d)


class HashGridEncoder(nn.Module):
    def __init__(self, levels: int = 16, min_res: int = 16, max_res: int = 4096, log2_hashmap_size: int = 19,
                 features_per_level: int = 2, hash_init_scale: float = 1e-3, implementation: str = "tcnn",
                 interpolation: Optional[str] = None) -> None:
        super().__init__()
        self.levels, self.min_res, self.max_res = int(levels), int(min_res), int(max_res)
        self.features_per_level, self.log2_hashmap_size = int(features_per_level), int(log2_hashmap_size)
        self.hash_init_scale = float(hash_init_scale)
        self.hash_table_size = 2 ** self.log2_hashmap_size
        if interpolation is not None and interpolation not in INTERPOLATIONS:
            interpolation = "Linear"
        self.interpolation = interpolation
        if self.features_per_level != 2:
            raise ValueError("the sm_100a hash-grid kernels implement features_per_level=2")
        L = self.levels
        self.growth_factor = 1.0 if L <= 1 else float(math.exp((math.log(self.max_res) - math.log(self.min_res)) / (L - 1)))
        lv = torch.arange(L, dtype=torch.float32)
        # same fp32 expression as the reference (encodings.py:203-215) so the integer resolutions agree
        res = torch.floor(self.min_res * (self.growth_factor ** lv)).to(torch.int32)
        self.register_buffer("level_resolutions", res, persistent=False)
        self.register_buffer("level_offsets", torch.arange(L, dtype=torch.int64) * self.hash_table_size, persistent=False)
        self.register_buffer("hash_primes", torch.tensor([1, 2654435761, 805459861], dtype=torch.int64), persistent=False)
        self._out_dim = L * self.features_per_level
        self.hash_table = nn.Parameter((torch.rand(self.hash_table_size * L, self.features_per_level) * 2 - 1) * self.hash_init_scale)
        self._res_host = [int(v) for v in res.tolist()]
        self._desc = None

    @property
    def out_dim(self) -> int:
        return self._out_dim

    def get_out_dim(self) -> int:
        return self._out_dim

    def grid_desc(self):
        interp = self.interpolation or "Linear"
        if self._desc is None or self._desc[0] != interp:
            self._desc = (interp, ops.grid_desc(self.levels, self.log2_hashmap_size, interp, self._res_host, self.features_per_level))
        return self._desc[1]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[-1] == 3, f"Expected (...,3), got {tuple(x.shape)}"
        return ops.HashGridFn.apply(x, self.hash_table, self.grid_desc())

    @torch.no_grad()
    def corner_indices(self, x: torch.Tensor) -> torch.Tensor:
        """(P,L,8) int32 table rows of the 8 cell corners (bit-exact contract; tests)."""
        return ops.hashgrid_indices(x.reshape(-1, 3).float().contiguous(), self.grid_desc())


class FrequencyEncoder(nn.Module):
    def __init__(self, in_dim: int = 3, pe_dim: int = 10, include_input: bool = True, use_pi: bool = False):
        super().__init__()
        self.in_dim, self.pe_dim = int(in_dim), int(pe_dim)
        self.include_input, self.use_pi = bool(include_input), bool(use_pi)
        if self.in_dim != 3:
            raise ValueError("the sm_100a frequency kernel implements in_dim=3")
        self.register_buffer("bands", 2.0 ** torch.arange(self.pe_dim, dtype=torch.float32), persistent=False)

    @property
    def out_dim(self) -> int:
        return self.in_dim * (2 * self.pe_dim + (1 if self.include_input else 0))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[-1] == self.in_dim, f"Expected (...,{self.in_dim}), got {tuple(x.shape)}"
        return ops.freq_encode(x, self.pe_dim, self.include_input)
