"""Fast-weight plumbing with the reference's interface (``models/metamodule/metamodule.py``:
MetaModule :13-69, MetaSequential :72-85, MetaLinear :129-156, MetaLayerBlock :159-192).

These classes exist so that parameter names (state_dict keys, ``meta_named_parameters``) and
the ``params`` override protocol match the reference; the experts do not run layer by layer -
MetaNGP hands all of its tensors (module weights or overrides) to one fused kernel."""
from __future__ import annotations

import warnings
from collections import OrderedDict
from typing import Dict, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


class MetaModule(nn.Module):
    def __init__(self):
        super().__init__()
        self._subdict_names = {}

    def meta_named_parameters(self, prefix: str = "", recurse: bool = True):
        """Parameters owned by MetaModule instances only (encoders are plain nn.Modules)."""
        yield from self._named_members(lambda m: m._parameters.items() if isinstance(m, MetaModule) else [], prefix=prefix,
                                       recurse=recurse)

    def meta_parameters(self, recurse: bool = True):
        for _, p in self.meta_named_parameters(recurse=recurse):
            yield p

    def get_subdict(self, params: Optional[Dict[str, torch.Tensor]], key: Optional[str] = None):
        """Entries of ``params`` under ``key.``, prefix removed; None (with a warning) when there are none."""
        if params is None:
            return None
        names = tuple(params.keys())
        ck = (key, names)
        if ck not in self._subdict_names:
            if key is None:
                self._subdict_names[ck] = list(names)
            else:
                pre = key + "."
                self._subdict_names[ck] = [n[len(pre):] for n in names if n.startswith(pre)]
        sub = self._subdict_names[ck]
        if not sub:
            warnings.warn(f"Module `{self.__class__.__name__}` has no parameter for submodule `{key}` in `params`.\n"
                          f"Using default parameters. Provided keys: [{', '.join(names)}]", stacklevel=2)
            return None
        if key is None:
            return OrderedDict((n, params[n]) for n in sub)
        return OrderedDict((n, params[f"{key}.{n}"]) for n in sub)


class MetaSequential(nn.Sequential, MetaModule):
    def forward(self, input, params: Optional[Dict[str, torch.Tensor]] = None):
        for name, module in self._modules.items():
            if isinstance(module, MetaModule):
                input = module(input, params=self.get_subdict(params, name))
            else:
                input = module(input)
        return input


class MetaLinear(nn.Linear, MetaModule):
    def forward(self, inputs: torch.Tensor, params: Optional[Dict[str, torch.Tensor]] = None) -> torch.Tensor:
        weight = self.weight if params is None else params.get("weight", self.weight)
        bias = self.bias if params is None else params.get("bias", self.bias)
        if inputs.dtype != weight.dtype:
            inputs = inputs.to(weight.dtype)
        return F.linear(inputs, weight, bias)


class MetaLayerBlock(MetaModule):
    def __init__(self, dim_in: int, dim_out: int, activation: Optional[str] = None, batched: bool = False):
        super().__init__()
        if batched:
            raise NotImplementedError("batched fast weights are not used by the NeRF experts")
        self.linear = MetaLinear(dim_in, dim_out)
        act = None if activation is None else activation.lower()
        if act is None:
            self.act = nn.Identity()
        elif act == "relu":
            self.act = nn.ReLU()
        elif act == "sigmoid":
            self.act = nn.Sigmoid()
        elif act == "softplus":
            self.act = nn.Softplus()
        elif act == "trunc_exp":
            from ..trunc_exp import trunc_exp
            self.act = trunc_exp
        else:
            raise ValueError(f"Unsupported activation: {activation}")

    def forward(self, x: torch.Tensor, params=None) -> torch.Tensor:
        return self.act(self.linear(x, params=self.get_subdict(params, "linear")))
