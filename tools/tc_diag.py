"""Error statistics of the tensor-core expert MLPs against the library's fp32 exact mode (GPU only)."""
import sys

import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
sys.path.insert(0, __file__.rsplit("/", 2)[0] + "/tests")
from nerf_sys_b200 import ops  # noqa: E402
from nerf_sys_b200.nerfs.ray_rendering import render_rays  # noqa: E402
from oracle import acnerf_oracle as O  # noqa: E402
from test_gpu_parity import _big_scene  # noqa: E402


def run(m, rays, u, gt, S, mode):
    m.zero_grad(set_to_none=True)
    with ops.mlp_math(mode):
        rgb, depth, w, acc = render_rays(m, rays, ray_samples=S, u=u)
        loss = ops.mse_loss(rgb, gt)
        loss.backward()
    torch.cuda.synchronize()
    return rgb.detach(), depth.detach(), acc.detach(), loss.item(), {n: p.grad.clone() for n, p in m.named_parameters() if p.grad is not None}


for K, gy, gz, T, N, S in [(1, 1, 1, 16, 2048, 96), (8, 2, 4, 14, 1536, 64)]:
    m, st, cfg, mins, aabb = _big_scene(K, gy, gz, T)
    m.train()
    rays = O.aerial_rays(N, aabb, seed=1).cuda()
    g = torch.Generator().manual_seed(5)
    u, gt = torch.rand(N, S, generator=g).cuda(), torch.rand(N, 3, generator=g).cuda()
    a = run(m, rays, u, gt, S, "f16tc")
    b = run(m, rays, u, gt, S, "fp32")
    print(f"K={K} N={N} S={S}: loss tc {a[3]:.6f} fp32 {b[3]:.6f}")
    for name, x, y in (("rgb", a[0], b[0]), ("depth", a[1], b[1]), ("acc", a[2], b[2])):
        print(f"  {name}: max-abs err {(x - y).abs().max().item():.3e} (max {y.abs().max().item():.3e}) nan {torch.isnan(x).sum().item()}")
    worst = []
    for n in b[4]:
        if n not in a[4]:
            print("  missing grad", n)
            continue
        x, y = a[4][n], b[4][n]
        rel = ((x - y).abs().max() / y.abs().max().clamp_min(1e-30)).item()
        l2 = ((x - y).norm() / y.norm().clamp_min(1e-30)).item()
        worst.append((rel, l2, n))
    worst.sort(reverse=True)
    for rel, l2, n in worst[:20]:
        print(f"  grad {n}: rel-max {rel:.3e} rel-L2 {l2:.3e}")
