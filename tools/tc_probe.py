"""Prints the error of the tcgen05 descriptor self-test per operand-layout combination (GPU only)."""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from nerf_sys_b200 import ops  # noqa: E402

mixed = len(sys.argv) > 1 and sys.argv[1] == "mixed"
g = torch.Generator().manual_seed(11)
X, Y, G = torch.randn(128, 64, generator=g), torch.randn(64, 64, generator=g), torch.randn(128, 64, generator=g)
D1, D2, D3 = ops.tc_selftest(X.cuda(), Y.cuda(), G.cuda(), mixed)
torch.cuda.synchronize()
Xh, Yh = X.half().double(), Y.half().double()
Gh = (G.bfloat16() if mixed else G.half()).double()
for name, got, ref in (("D1=X Y^T", D1, Xh @ Yh.T), ("D3=G Y", D3, Gh @ Yh), ("D2=G^T X", D2, Gh.T @ Xh)):
    got = got.double().cpu()
    err = (got - ref).abs().max().item()
    print(f"mixed={mixed} {name}: max err {err:.3e} (ref max {ref.abs().max().item():.3e}); got[0,:4]={got[0,:4].tolist()} ref[0,:4]={ref[0,:4].tolist()}")
