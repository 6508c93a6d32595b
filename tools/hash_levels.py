"""Per-level device time of the hash-grid gather / scatter at the C2 point count (GPU only).
Points are 96 consecutive samples along each of 8192 random rays through the unit cube, like the render path."""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from nerf_sys_b200 import ops  # noqa: E402

N, S = 8192, 96
g = torch.Generator().manual_seed(0)
o = torch.rand(N, 1, 3, generator=g)
d = torch.randn(N, 1, 3, generator=g)
d = d / d.norm(dim=-1, keepdim=True)
t = (torch.arange(S).float() / S * 1.2).view(1, S, 1)
x = (o + d * (t - 0.6)).clamp(1e-6, 1 - 1e-6).reshape(-1, 3).cuda().contiguous()
res = [16, 23, 33, 48, 70, 101, 147, 212, 307, 445, 645, 933, 1351, 1955, 2830, 4095]
P = x.shape[0]
ev = lambda: torch.cuda.Event(enable_timing=True)
for r in res:
    gd = ops.grid_desc(1, 20, "linear", [r])
    table = (torch.randn(1 << 20, 2, device="cuda") * 0.1).requires_grad_(True)
    dy = torch.randn(P, 2, device="cuda")
    tf = tb = 0.0
    for it in range(6):
        table.grad = None
        a, b, c = ev(), ev(), ev()
        a.record()
        y = ops.HashGridFn.apply(x, table, gd)
        b.record()
        y.backward(dy)
        c.record()
        torch.cuda.synchronize()
        if it >= 2:
            tf += a.elapsed_time(b) / 4
            tb += b.elapsed_time(c) / 4
    print(f"res {r:5d}: fwd {tf * 1e3:7.1f} us   bwd (zero-fill 8 MiB + scatter) {tb * 1e3:7.1f} us")
