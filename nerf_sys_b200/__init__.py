"""nerf_sys_b200: the B200 (sm_100a) per-ray hot path of Adaptive City NeRF
(psklavos1/NeRF-Sys) behind the reference's own model / renderer interfaces.

Layout (mirrors the part of ``adaptive_nerf/`` that sits on the hot path):
  csrc/            hand-written CUDA kernels + the C ABI (include/acnerf.h)
  _lib.py, ops.py  ctypes binding, autograd Functions (plumbing only)
  nerfs/           ray_sampling, scene_box, ray_rendering, losses, color_space
  models/          encodings, trunc_exp, metamodule, inr/{meta_ngp, meta_container}
  optim.py         fused Adam / clip
  parallel.py      ray-sharded render, expert-parallel step
There is no CPU path, no Triton, no tiny-cuda-nn and no multi-backend dispatch.
"""
from ._lib import LIB_PATH, lib  # noqa: F401

__all__ = ["lib", "LIB_PATH"]
