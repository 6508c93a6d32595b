"""Microbenchmark of acn_adam_multi / acn_sumsq_multi over hash-table-sized tensors (GPU):
gradient-zeroing variants of the Adam kernel against the separate memset it replaces."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_sys_b200._lib import AcnAdamTensor, check, lib  # noqa: E402


def run(n_tables, flags, memset, iters=10):
    n = 16 * (1 << 20) * 2
    dev = "cuda"
    ps = [torch.randn(n, device=dev) for _ in range(n_tables)]
    gs = [torch.randn(n, device=dev) for _ in range(n_tables)]
    ms = [torch.zeros(n, device=dev) for _ in range(n_tables)]
    vs = [torch.zeros(n, device=dev) for _ in range(n_tables)]
    arr = (AcnAdamTensor * n_tables)()
    for i in range(n_tables):
        arr[i].param, arr[i].grad, arr[i].exp_avg, arr[i].exp_avg_sq = ps[i].data_ptr(), gs[i].data_ptr(), ms[i].data_ptr(), vs[i].data_ptr()
        arr[i].n, arr[i].lr, arr[i].weight_decay, arr[i].step_index, arr[i].flags = n, 1e-2, 0.0, i, flags
    steps = torch.zeros(n_tables, dtype=torch.int32, device=dev)
    acc = torch.ones(1, device=dev)
    coef = torch.zeros(2, device=dev)
    L = lib()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def once():
        check(L.acn_adam_multi(arr, n_tables, 0.9, 0.999, 1e-8, steps.data_ptr(), acc.data_ptr(), 1.0, None, None, None, coef.data_ptr(), 1, st), "adam")
        if memset:
            for g in gs:
                g.zero_()

    for _ in range(3):
        once()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        once()
    e1.record()
    torch.cuda.synchronize()
    ms_ = e0.elapsed_time(e1) / iters
    bytes_ = n_tables * n * (32 if (flags & 0x100) or memset else 28)
    return ms_, bytes_ / ms_ / 1e6


if __name__ == "__main__":
    for nt in (1, 8):
        for name, flags, memset in (("no zeroing", 0, False), ("separate memset", 0, True), ("zero store in the kernel", 0x100, False)):
            ms_, gbs = run(nt, flags, memset)
            print(f"tables {nt}  {name:18s} {ms_ * 1e3:9.1f} us  {gbs:7.0f} GB/s")
