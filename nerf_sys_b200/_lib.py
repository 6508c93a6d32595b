"""ctypes binding of libacnerf_sm100a.so (the C ABI in include/acnerf.h).

The library is the product: there is no Python / CPU fallback.  Importing this
module never touches the GPU; `lib()` loads the shared object and raises
``RuntimeError`` if it is missing or a symbol is absent, and every wrapper in
``ops.py`` raises on a non-zero status with ``acn_last_error()``.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libacnerf_sm100a.so")

ACN_MAX_LEVELS = 32
ACN_MAX_EXPERTS = 16
ACN_TILE = 128
ACN_NGP_LAYERS = 7

vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float


class AcnGridDesc(C.Structure):
    _fields_ = [("levels", i32), ("features", i32), ("log2_T", i32), ("interp", i32), ("res", i32 * ACN_MAX_LEVELS)]


class AcnPoints(C.Structure):
    _fields_ = [("x6", vp), ("rays", vp), ("t", vp), ("S", i32), ("P", i64), ("t_T", vp)]


class AcnPairs(C.Structure):
    _fields_ = [("point", vp), ("weight", vp), ("seg", vp), ("Mcap", i64)]


class AcnNgpExpert(C.Structure):
    _fields_ = [("table", vp), ("w", vp * ACN_NGP_LAYERS), ("b", vp * ACN_NGP_LAYERS), ("box_min", f32 * 3), ("extent", f32 * 3)]


class AcnNgpExpertGrads(C.Structure):
    _fields_ = [("table", vp), ("w", vp * ACN_NGP_LAYERS), ("b", vp * ACN_NGP_LAYERS)]


class AcnAdamTensor(C.Structure):
    _fields_ = [("param", vp), ("grad", vp), ("exp_avg", vp), ("exp_avg_sq", vp), ("n", i64), ("lr", f32), ("weight_decay", f32),
                ("step_index", i32), ("flags", i32)]


class AcnSegLists(C.Structure):
    _fields_ = [("offsets", vp * ACN_MAX_EXPERTS), ("counts", vp * ACN_MAX_EXPERTS), ("t0", vp * ACN_MAX_EXPERTS), ("t1", vp * ACN_MAX_EXPERTS),
                ("K", i32)]


class AcnA2aRoute(C.Structure):
    _fields_ = [("base", vp * ACN_MAX_EXPERTS), ("row0", i64 * ACN_MAX_EXPERTS)]


ACN_ADAM_MAX_TENSORS = 64
ACN_MAX_PEERS = 16
ACN_P2P_PAYLOAD_FLOATS = 1024
ACN_P2P_SYNC_BYTES = 4096 + 2 * ACN_MAX_PEERS * ACN_P2P_PAYLOAD_FLOATS * 4
ACN_ADAM_ZERO_GRAD = 0x100
ACN_OCC_MAX_LEVELS = 8
ACN_HINT_GATHER_LEVELS = 1
P = C.POINTER
fp = P(f32)  # host float arrays

# name -> (restype, argtypes); must list every symbol include/acnerf.h declares
SIGNATURES = {
    "acn_version": (C.c_int, []),
    "acn_last_error": (C.c_char_p, []),
    "acn_device_ok": (C.c_int, [C.c_int]),
    "acn_num_sms": (C.c_int, [C.c_int]),
    "acn_launch_hint": (C.c_int, [C.c_int, C.c_int]),
    "acn_episode_route": (C.c_int, [vp, C.c_int, i64, fp, P(i32), vp, vp, C.c_int, f32, f32, C.c_int, vp, vp, vp, vp, vp]),
    "acn_voronoi_ws_bytes": (i64, []),
    "acn_voronoi_masks": (C.c_int, [vp, i64, C.c_int, fp, C.c_int, C.c_int, f32, vp, vp, vp, vp, vp, vp]),
    "acn_occ_march_count": (C.c_int, [vp, vp, i64, vp, fp, C.c_int, C.c_int, f32, f32, C.c_int, vp, vp]),
    "acn_occ_march_write": (C.c_int, [vp, vp, i64, vp, fp, C.c_int, C.c_int, f32, f32, C.c_int, vp, vp, vp, vp, vp, vp]),
    "acn_occ_visibility": (C.c_int, [vp, vp, i64, vp, vp, vp, f32, f32, vp, vp]),
    "acn_seg_union": (C.c_int, [P(AcnSegLists), i64, vp, vp, vp, vp, vp, vp]),
    "acn_packed_render_fwd": (C.c_int, [vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "acn_packed_render_bwd": (C.c_int, [vp, vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "acn_ray_directions": (C.c_int, [C.c_int, C.c_int, f32, f32, f32, f32, C.c_int, vp, vp]),
    "acn_get_rays": (C.c_int, [vp, i64, fp, fp, f32, f32, f32, f32, vp, vp]),
    "acn_raygen": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, f32, f32, f32, f32, C.c_int, fp, fp, f32, f32, vp, vp]),
    "acn_ray_aabb": (C.c_int, [vp, vp, i64, fp, f32, f32, vp, vp, vp]),
    "acn_clamp_near_far": (C.c_int, [vp, i64, C.c_int, C.c_int, f32, C.c_int, f32, f32, f32, vp, vp, vp]),
    "acn_stratified_t": (C.c_int, [vp, i64, C.c_int, vp, vp, vp, vp]),
    "acn_sample_pdf": (C.c_int, [vp, vp, vp, i64, C.c_int, C.c_int, vp, vp, vp]),
    "acn_sample_points": (C.c_int, [vp, vp, i64, C.c_int, vp, vp]),
    "acn_routing": (C.c_int, [P(AcnPoints), fp, C.c_int, f32, C.c_int, vp, vp, vp]),
    "acn_route_pack_capacity": (i64, [i64, C.c_int]),
    "acn_route_pack_ws_bytes": (i64, [i64, C.c_int]),
    "acn_route_pack": (C.c_int, [P(AcnPoints), fp, C.c_int, f32, C.c_int, vp, vp, vp, i64, vp, vp]),
    "acn_hashgrid_indices": (C.c_int, [vp, i64, P(AcnGridDesc), vp, vp]),
    "acn_hashgrid_fwd": (C.c_int, [vp, i64, vp, P(AcnGridDesc), vp, vp]),
    "acn_hashgrid_bwd": (C.c_int, [vp, i64, vp, P(AcnGridDesc), vp, vp]),
    "acn_sh16": (C.c_int, [vp, i64, vp, vp]),
    "acn_freq_encode": (C.c_int, [vp, i64, C.c_int, C.c_int, vp, vp]),
    "acn_ngp_ld": (i64, [i64]),
    "acn_ngp_save_bytes": (i64, [i64]),
    "acn_ngp_bwd_ws_bytes": (i64, [i64, C.c_int]),
    "acn_ngp_encode_fwd": (C.c_int, [P(AcnPoints), P(AcnPairs), P(AcnNgpExpert), C.c_int, P(AcnGridDesc), vp, vp]),
    "acn_ngp_encode_bwd": (C.c_int, [P(AcnPoints), P(AcnPairs), P(AcnNgpExpert), C.c_int, P(AcnGridDesc), vp,
                                     P(AcnNgpExpertGrads), vp]),
    "acn_ngp_mlp_fwd": (C.c_int, [P(AcnPoints), P(AcnPairs), P(AcnNgpExpert), C.c_int, vp, vp, vp, vp]),
    "acn_ngp_mlp_bwd": (C.c_int, [P(AcnPairs), P(AcnNgpExpert), C.c_int, vp, vp, vp, P(AcnNgpExpertGrads), vp, vp, vp]),
    "acn_ngp_mlp_fwd_tc": (C.c_int, [P(AcnPoints), P(AcnPairs), P(AcnNgpExpert), C.c_int, vp, vp, vp]),
    "acn_ngp_fused_fwd_tc": (C.c_int, [P(AcnPoints), P(AcnPairs), P(AcnNgpExpert), C.c_int, P(AcnGridDesc), vp, vp]),
    "acn_ngp_bwd_tc_ws_bytes": (i64, []),
    "acn_ngp_mlp_bwd_tc": (C.c_int, [P(AcnPoints), P(AcnPairs), P(AcnNgpExpert), C.c_int, vp, vp, P(AcnNgpExpertGrads), vp, vp, vp]),
    "acn_ngp_bwd_tc_pipelined_ws_bytes": (i64, [C.c_int]),
    "acn_ngp_bwd_tc_pipelined": (C.c_int, [P(AcnPoints), P(AcnPairs), P(AcnNgpExpert), C.c_int, P(AcnGridDesc), vp, vp, P(AcnNgpExpertGrads), vp,
                                           vp, C.c_int, vp, vp]),
    "acn_tc_selftest": (C.c_int, [vp, vp, vp, C.c_int, vp, vp, vp, vp]),
    "acn_pe_encode_h": (C.c_int, [P(AcnPoints), C.c_int, C.c_int, C.c_int, vp, i64, C.c_int, vp, i64, C.c_int, vp]),
    "acn_dense_tc": (C.c_int, [vp, C.c_int, i64, i64, vp, i64, C.c_int, vp, C.c_int, C.c_int, C.c_int, vp, i64, C.c_int, vp, i64, C.c_int,
                               C.c_int, vp, i64, vp]),
    "acn_wgrad_tc_ws_bytes": (i64, [C.c_int, C.c_int]),
    "acn_wgrad_tc": (C.c_int, [vp, C.c_int, vp, C.c_int, i64, i64, C.c_int, C.c_int, vp, i64, vp, vp, vp, vp]),
    "acn_pe_final_fwd": (C.c_int, [vp, i64, vp, i64, C.c_int, C.c_int, i64, vp, vp]),
    "acn_pe_final_bwd": (C.c_int, [vp, vp, C.c_int, i64, vp, i64, vp, vp, vp]),
    "acn_set_feature_h": (C.c_int, [vp, i64, C.c_int, vp, i64, vp]),
    "acn_bg_save_bytes": (i64, [i64]),
    "acn_bg_forward": (C.c_int, [vp, i64, i64, vp, vp, vp, vp, vp, vp, vp]),
    "acn_bg_bwd_ws_bytes": (i64, [i64]),
    "acn_bg_backward": (C.c_int, [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "acn_composite_fwd": (C.c_int, [vp, vp, vp, i64, C.c_int, f32, vp, vp, vp, vp, vp]),
    "acn_composite_fwd_sm": (C.c_int, [vp, vp, vp, i64, C.c_int, f32, vp, vp, vp, vp, vp]),
    "acn_composite_fwd_strided": (C.c_int, [vp, i64, i64, vp, vp, i64, C.c_int, f32, vp, vp, vp, vp, vp]),
    "acn_composite_bwd": (C.c_int, [vp, vp, vp, i64, C.c_int, f32, vp, vp, vp, vp, vp, vp, vp]),
    "acn_mse_loss": (C.c_int, [vp, vp, i64, C.c_int, vp, vp, vp, vp]),
    "acn_mse_loss_static": (C.c_int, [vp, vp, i64, C.c_int, vp, vp, vp]),
    "acn_adam_step": (C.c_int, [vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i64, vp, vp]),
    "acn_sgd_axpy": (C.c_int, [vp, vp, f32, i64, vp, vp]),
    "acn_sumsq_accum": (C.c_int, [vp, i64, vp, vp]),
    "acn_sumsq_multi": (C.c_int, [P(AcnAdamTensor), C.c_int, vp, vp, vp]),
    "acn_adam_multi": (C.c_int, [P(AcnAdamTensor), C.c_int, f32, f32, f32, vp, vp, f32, vp, vp, vp, vp, C.c_int, vp]),
    "acn_a2a_pack": (C.c_int, [P(AcnPoints), P(AcnPairs), C.c_int, vp, vp, vp, vp, vp, vp]),
    "acn_a2a_combine": (C.c_int, [vp, vp, vp, i64, vp, vp]),
    "acn_a2a_combine_bwd": (C.c_int, [vp, vp, vp, i64, vp, vp]),
    "acn_a2a_count": (C.c_int, [P(AcnPairs), C.c_int, vp, vp, vp]),
    "acn_a2a_push_rows": (C.c_int, [P(AcnPoints), P(AcnPairs), C.c_int, vp, P(AcnA2aRoute), vp, i64, vp, vp, vp]),
    "acn_a2a_pull_combine": (C.c_int, [P(AcnA2aRoute), vp, vp, vp, vp, C.c_int, i64, vp, vp]),
    "acn_a2a_push_grad": (C.c_int, [vp, vp, vp, vp, C.c_int, i64, P(AcnA2aRoute), vp, i64, vp]),
    "acn_p2p_alloc": (C.c_int, [i64, P(C.c_void_p), vp]),
    "acn_p2p_free": (C.c_int, [vp]),
    "acn_p2p_open": (C.c_int, [vp, P(C.c_void_p)]),
    "acn_p2p_close": (C.c_int, [vp]),
    "acn_p2p_barrier": (C.c_int, [P(C.c_void_p), C.c_int, C.c_int, vp, C.c_int, vp, C.c_int, vp, f32, f32, vp]),
    "acn_p2p_arrive": (C.c_int, [P(vp), C.c_int, C.c_int, vp, C.c_int, vp]),
    "acn_p2p_wait": (C.c_int, [P(vp), C.c_int, C.c_int, C.c_int, vp, f32, f32, vp]),
    "acn_p2p_plan": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, P(i32), i64, vp, vp, vp, vp, vp, vp, vp]),
    "acn_route_status": (C.c_int, [vp, C.c_int, i64, vp, vp, vp]),
}

_lock = threading.Lock()
_lib = None


def lib():
    """The loaded shared library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python nerf_sys_b200/csrc/build.py` "
                "(or __graft_entry__.build()).  nerf_sys_b200 has no CPU or PyTorch fallback."
            )
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(h, name)
            except AttributeError as e:
                raise RuntimeError(f"{LIB_PATH} does not export {name}; rebuild the extension") from e
            fn.restype = res
            fn.argtypes = args
        if h.acn_version() != 1:
            raise RuntimeError(f"ABI mismatch: library reports version {h.acn_version()}, binding expects 1")
        _lib = h
    return _lib


def check(status: int, what: str):
    if status != 0:
        msg = lib().acn_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (status {status}): {msg}")


def host_floats(values):
    arr = (f32 * len(values))(*[float(v) for v in values])
    return arr
