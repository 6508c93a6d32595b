"""MetaNeRF - the positional-encoding MLP expert (config 1), interface of the reference's
``models/inr/meta_vanilla.py`` (:13-154): same constructor arguments, attribute names, state_dict keys
(``trunk.{i}.linear.*``, ``sigma_head.*``, ``geo_head.*``, ``color_mlp.layer{l}.linear.*``, ``color_mlp.color_out.*``)
and ``params`` fast-weight protocol.  The network runs as tcgen05 GEMM launches (csrc/dense_tc.cu): one per layer
forward, two (dgrad + wgrad) per layer backward, activations as fp16 chunk-major tensors, fp32 accumulation."""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn
from torch import Tensor

from ... import ops
from ..._lib import check, lib
from ..encodings import FrequencyEncoder
from ..metamodule import MetaLayerBlock, MetaLinear, MetaModule, MetaSequential


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t: Optional[Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _chunks(n: int) -> int:
    return (n + 7) // 8


class _Net:
    """Static description of one MetaNeRF: widths, depth, skips, encodings."""

    def __init__(self, hidden, depth, skips, geo, color_hidden, color_depth, pe_xyz, incl_xyz, pe_dir, incl_dir, sigmoid):
        self.H, self.D, self.skips, self.G = hidden, depth, tuple(skips), geo
        self.CH, self.CD = color_hidden, color_depth
        self.pe_xyz, self.incl_xyz, self.pe_dir, self.incl_dir, self.sigmoid = pe_xyz, incl_xyz, pe_dir, incl_dir, sigmoid
        self.w_xyz = (3 if incl_xyz else 0) + 6 * pe_xyz
        self.w_dir = (3 if incl_dir else 0) + 6 * pe_dir
        if hidden % 8 or (color_depth > 1 and color_hidden % 8):
            raise NotImplementedError("the sm_100a PE-MLP needs hidden widths that are multiples of 8")
        if geo > 15:
            raise NotImplementedError("geo_feat_dim <= 15 (heads are evaluated as one 16-row GEMM)")

    def trunk_in(self, i: int) -> int:
        if i == 0:
            return self.w_xyz
        return self.H + self.w_xyz if i in self.skips else self.H


def _dense(L, a, kc, M, ld, W, transposed, bias, n_out, k_valid, act, mask=None, mask_chunks=0, out_h=None, out_chunk0=0, out_chunks=0,
           out_f=None):
    check(L.acn_dense_tc(_ptr(a), kc, M, ld, _ptr(W), W.stride(0), int(transposed), _ptr(bias), n_out, k_valid, act, _ptr(mask), ld, mask_chunks,
                         _ptr(out_h), ld, out_chunk0, out_chunks, _ptr(out_f), 0 if out_f is None else out_f.stride(0), _stream()),
          "acn_dense_tc")
    ops._count(1)


class PeMlpFn(torch.autograd.Function):
    """(points, weights...) -> (P,4) [rgb, sigma].  Weight order: trunk (w,b) x depth, sigma_head, geo_head, colour layers."""

    @staticmethod
    def forward(ctx, pts: ops.PointSet, net: _Net, *weights: Tensor):
        L = lib()
        ws = [ops._need(w.detach(), "PE-MLP weight") for w in weights]
        dev, P = pts.device, pts.P
        ld = (P + 127) // 128 * 128
        Hc, Xc, half = _chunks(net.H), _chunks(net.w_xyz), torch.float16
        cp = pts.c()
        need_grad = any(ctx.needs_input_grad)
        with torch.cuda.device(dev), ops._timed("pe_mlp_fwd"):
            pe = torch.empty(Xc, ld, 8, device=dev, dtype=half)
            check(L.acn_pe_encode_h(C.byref(cp), 0, net.pe_xyz, int(net.incl_xyz), None, 0, 0, _ptr(pe), ld, 0, _stream()), "acn_pe_encode_h")
            acts: List[Tensor] = []
            a, kc = pe, Xc
            for i in range(net.D):
                tail = Xc if (i + 1) in net.skips and i + 1 < net.D else 0  # the next layer concatenates the encoding
                out = torch.empty(Hc + tail, ld, 8, device=dev, dtype=half)
                _dense(L, a, kc, P, ld, ws[2 * i], False, ws[2 * i + 1], net.H, net.trunk_in(i), 1, out_h=out, out_chunks=Hc)
                if tail:
                    check(L.acn_pe_encode_h(C.byref(cp), 0, net.pe_xyz, int(net.incl_xyz), None, 0, 0, _ptr(out), ld, Hc, _stream()),
                          "acn_pe_encode_h")
                    ops._count(1)
                acts.append(out)
                a, kc = out, Hc + tail
            w_sig, b_sig, w_geo, b_geo = ws[2 * net.D:2 * net.D + 4]
            w_head = torch.cat([w_geo, w_geo.new_zeros(15 - net.G, net.H), w_sig], 0)       # rows 0..G-1 geo, row 15 sigma
            b_head = torch.cat([b_geo, b_geo.new_zeros(15 - net.G), b_sig], 0)
            heads = torch.empty(P, 16, device=dev)
            _dense(L, a, Hc, P, ld, w_head, False, b_head, 16, net.H, 0, out_f=heads)
            cin_w = net.G + net.w_dir
            cin = torch.empty(_chunks(cin_w), ld, 8, device=dev, dtype=half)
            check(L.acn_pe_encode_h(C.byref(cp), 1, net.pe_dir, int(net.incl_dir), _ptr(heads), 16, net.G, _ptr(cin), ld, 0, _stream()),
                  "acn_pe_encode_h")
            cacts: List[Tensor] = []
            a, kc, kin = cin, _chunks(cin_w), cin_w
            cw = ws[2 * net.D + 4:]
            rgb_raw = torch.empty(P, 4, device=dev)
            for l in range(net.CD):
                if l == net.CD - 1:
                    _dense(L, a, kc, P, ld, cw[2 * l], False, cw[2 * l + 1], 3, kin, 0, out_f=rgb_raw)
                else:
                    out = torch.empty(_chunks(net.CH), ld, 8, device=dev, dtype=half)
                    _dense(L, a, kc, P, ld, cw[2 * l], False, cw[2 * l + 1], net.CH, kin, 1, out_h=out, out_chunks=_chunks(net.CH))
                    cacts.append(out)
                    a, kc, kin = out, _chunks(net.CH), net.CH
            y = torch.empty(P, 4, device=dev)
            check(L.acn_pe_final_fwd(_ptr(rgb_raw), 4, _ptr(heads), 16, 15, int(net.sigmoid), P, _ptr(y), _stream()), "acn_pe_final_fwd")
            ops._count(4)
        if need_grad:
            ctx.pts, ctx.net, ctx.ld = pts, net, ld
            ctx.bufs = (pe, acts, cin, cacts, w_head)
            ctx.save_for_backward(y, *ws)
        return y

    @staticmethod
    @ops.first_order
    def backward(ctx, d_out: Tensor):
        L = lib()
        y, *ws = ctx.saved_tensors
        net, ld, pts = ctx.net, ctx.ld, ctx.pts
        pe, acts, cin, cacts, w_head = ctx.bufs
        dev, P = y.device, y.shape[0]
        half, Hc, Xc = torch.float16, _chunks(net.H), _chunks(net.w_xyz)
        d_out = ops._f32(d_out, "d_out")
        grads: List[Optional[Tensor]] = [None] * len(ws)

        def wgrad(dz, dz_chunks, a, kc, n_out, k_valid, dW, db):
            wsb = torch.empty(int(L.acn_wgrad_tc_ws_bytes(n_out, kc)) // 4, device=dev)
            check(L.acn_wgrad_tc(_ptr(dz), dz_chunks, _ptr(a), kc, P, ld, n_out, k_valid, _ptr(dW), dW.stride(0), _ptr(db), _ptr(inv_scale),
                                 _ptr(wsb), _stream()), "acn_wgrad_tc")
            ops._count(2)

        with torch.cuda.device(dev), ops._timed("pe_mlp_bwd"):
            g = torch.empty(2, ld, 8, device=dev, dtype=half)
            dsraw = torch.empty(P, device=dev)
            scale = torch.empty(4, device=dev)
            check(L.acn_pe_final_bwd(_ptr(d_out), _ptr(y), int(net.sigmoid), P, _ptr(g), ld, _ptr(dsraw), _ptr(scale), _stream()),
                  "acn_pe_final_bwd")
            ops._count(2)
            inv_scale = scale[1:2]
            # ---- colour branch, last layer first ----
            cw = ws[2 * net.D + 4:]
            base = 2 * net.D + 4
            dz, dzc, n_out = g, 2, 3
            for l in range(net.CD - 1, -1, -1):
                a = cin if l == 0 else cacts[l - 1]
                kin = net.G + net.w_dir if l == 0 else net.CH
                kc = _chunks(kin)
                dW, db = torch.empty_like(cw[2 * l]), torch.empty_like(cw[2 * l + 1])
                wgrad(dz, dzc, a, kc, n_out, kin, dW, db)
                grads[base + 2 * l], grads[base + 2 * l + 1] = dW, db
                if l > 0:  # gradient w.r.t. the previous hidden layer's pre-activation
                    nxt = torch.empty(_chunks(net.CH), ld, 8, device=dev, dtype=half)
                    _dense(L, dz, dzc, P, ld, cw[2 * l], True, None, net.CH, n_out, 0, mask=cacts[l - 1], mask_chunks=_chunks(net.CH), out_h=nxt,
                           out_chunks=_chunks(net.CH))
                    dz, dzc, n_out = nxt, _chunks(net.CH), net.CH
                else:      # gradient w.r.t. [geo | dir encoding]: only the first 16 columns are needed
                    gh = torch.empty(2, ld, 8, device=dev, dtype=half)
                    _dense(L, dz, dzc, P, ld, cw[0], True, None, 16, n_out, 0, out_h=gh, out_chunks=2)
            # ---- heads: gh = [d geo 0..G-1 | junk | d sigma_raw at column 15] ----
            for f in range(net.G, 15):
                check(L.acn_set_feature_h(_ptr(gh), ld, f, _ptr(torch.zeros(P, device=dev)), P, _stream()), "acn_set_feature_h")
            check(L.acn_set_feature_h(_ptr(gh), ld, 15, _ptr(dsraw), P, _stream()), "acn_set_feature_h")
            ops._count(1)
            dWh, dbh = torch.empty(16, net.H, device=dev), torch.empty(16, device=dev)
            wgrad(gh, 2, acts[-1], Hc, 16, net.H, dWh, dbh)
            grads[2 * net.D], grads[2 * net.D + 1] = dWh[15:16].contiguous(), dbh[15:16].contiguous()
            grads[2 * net.D + 2], grads[2 * net.D + 3] = dWh[:net.G].contiguous(), dbh[:net.G].contiguous()
            dz = torch.empty(Hc, ld, 8, device=dev, dtype=half)
            _dense(L, gh, 2, P, ld, w_head, True, None, net.H, 16, 0, mask=acts[-1], mask_chunks=Hc, out_h=dz, out_chunks=Hc)
            # ---- trunk ----
            for i in range(net.D - 1, -1, -1):
                a = pe if i == 0 else acts[i - 1]
                kin = net.trunk_in(i)
                kc = _chunks(kin) if i == 0 else Hc + (Xc if i in net.skips else 0)
                dW, db = torch.empty_like(ws[2 * i]), torch.empty_like(ws[2 * i + 1])
                wgrad(dz, Hc, a, kc, net.H, kin, dW, db)
                grads[2 * i], grads[2 * i + 1] = dW, db
                if i > 0:
                    nxt = torch.empty(Hc, ld, 8, device=dev, dtype=half)
                    _dense(L, dz, Hc, P, ld, ws[2 * i], True, None, net.H, net.H, 0, mask=acts[i - 1], mask_chunks=Hc, out_h=nxt, out_chunks=Hc)
                    dz = nxt
        ctx.bufs = None
        out = [gr if ctx.needs_input_grad[2 + j] else None for j, gr in enumerate(grads)]
        return (None, None, *out)


class MetaNeRF(MetaModule):
    """Meta-learnable NeRF with Fourier xyz features, NeRF-style skips and view-dependent colour
    (reference: models/inr/meta_vanilla.py:13-154; only ``encoding_dir="frequency"`` is constructible there, SURVEY note V)."""

    def __init__(self, *, hidden: int = 256, sigma_depth: int = 8, skips: Tuple[int, ...] = (4,), geo_feat_dim: int = 15,
                 color_hidden: int = 128, color_depth: int = 2, use_sigmoid_rgb: bool = True, pe_dim_xyz: int = 10,
                 include_input_xyz: bool = True, encoding_dir: str = "frequency", include_input_dir: bool = True) -> None:
        super().__init__()
        if encoding_dir.lower() != "frequency":
            raise NotImplementedError("MetaNeRF: the sm_100a path implements encoding_dir='frequency' "
                                      "(the reference's 'spherical' branch passes an unknown keyword to SHEncoder and cannot be built)")
        self.hidden, self.sigma_depth, self.skips = int(hidden), int(sigma_depth), tuple(int(s) for s in skips)
        self.geo_feat_dim, self.color_hidden, self.color_depth = int(geo_feat_dim), int(color_hidden), int(color_depth)
        self.use_sigmoid_rgb = bool(use_sigmoid_rgb)
        self.xyz_encoder = FrequencyEncoder(in_dim=3, pe_dim=int(pe_dim_xyz), include_input=bool(include_input_xyz), use_pi=False)
        self.dir_encoder = FrequencyEncoder(in_dim=3, pe_dim=4, include_input=bool(include_input_dir), use_pi=False)
        in_xyz, in_dir = self.xyz_encoder.out_dim, self.dir_encoder.out_dim
        self.trunk = nn.ModuleList()
        for i in range(self.sigma_depth):
            in_dim = in_xyz if i == 0 else self.hidden
            if i in self.skips and i != 0:
                in_dim = self.hidden + in_xyz
            self.trunk.append(MetaLayerBlock(in_dim, self.hidden, activation="relu"))
        self.sigma_head = MetaLinear(self.hidden, 1)
        self.geo_head = MetaLinear(self.hidden, self.geo_feat_dim)
        layers = []
        for l in range(self.color_depth):
            in_dim = self.geo_feat_dim + in_dir if l == 0 else self.color_hidden
            if l == self.color_depth - 1:
                layers.append(("color_out", MetaLinear(in_dim, 3)))
            else:
                layers.append((f"layer{l}", MetaLayerBlock(in_dim, self.color_hidden, activation="relu")))
        self.color_mlp = MetaSequential(OrderedDict(layers))
        self._in_ch_xyz = in_xyz
        self._net = _Net(self.hidden, self.sigma_depth, tuple(s for s in self.skips if s != 0), self.geo_feat_dim, self.color_hidden,
                         self.color_depth, int(pe_dim_xyz), bool(include_input_xyz), 4, bool(include_input_dir), self.use_sigmoid_rgb)

    def weight_names(self) -> List[str]:
        names = []
        for i in range(self.sigma_depth):
            names += [f"trunk.{i}.linear.weight", f"trunk.{i}.linear.bias"]
        names += ["sigma_head.weight", "sigma_head.bias", "geo_head.weight", "geo_head.bias"]
        for l in range(self.color_depth):
            stem = "color_mlp.color_out" if l == self.color_depth - 1 else f"color_mlp.layer{l}.linear"
            names += [stem + ".weight", stem + ".bias"]
        return names

    def mlp_tensors(self, params: Optional[Dict[str, Tensor]] = None) -> List[Tensor]:
        """Module weights, or the caller's fast weights where ``params`` has the key (metamodule.py:60-67 semantics)."""
        own = dict(self.named_parameters())
        return [(params[n] if params is not None and n in params else own[n]) for n in self.weight_names()]

    def evaluate(self, pts: ops.PointSet, params: Optional[Dict[str, Tensor]] = None) -> Tensor:
        """(P,4) [rgb, sigma] for explicit (P,6) rows or implicit ray samples."""
        return PeMlpFn.apply(pts, self._net, *self.mlp_tensors(params))

    def forward(self, x: Tensor, d: Tensor, params: Optional[Dict[str, Tensor]] = None) -> Dict[str, Tensor]:
        shape = x.shape[:-1]
        x6 = torch.cat([x.reshape(-1, 3), d.reshape(-1, 3)], dim=-1).contiguous()
        y = self.evaluate(ops.PointSet(x6=ops._f32(x6, "x6")), params)
        return {"sigma": y[:, 3:4].reshape(*shape, 1), "rgb": y[:, 0:3].reshape(*shape, 3)}

    def density(self, x: Tensor, params: Optional[Dict[str, Tensor]] = None) -> Dict[str, Tensor]:
        raise NotImplementedError("MetaNeRF.density/color are fused on sm_100a: call forward(x, d) or evaluate(points)")

    color = density
