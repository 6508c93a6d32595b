"""GPU parity of the positional-encoding MLP expert (config 1, MetaNeRF; reference models/inr/meta_vanilla.py:109-154)
running as tcgen05 GEMM layers (csrc/dense_tc.cu), against the golden vectors produced by the real reference and
against the CPU oracle at the reference's default widths (256 x 8 with a skip, 128-wide colour layer).
Stated tolerances (fp16 operands / fp32 accumulation through 10-11 layers): sigma, rgb 5e-3 relative to range;
gradients 3e-2 of max and 3e-2 relative L2; rendered rgb / depth 2e-3."""
import pytest
import torch

from conftest import T, load_golden, state_from
from oracle import acnerf_oracle as O
from test_gpu_parity import close, rel_close

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _grad_close(got, ref, name):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    rel_close(got, ref, 3e-2, name)
    l2 = ((got - ref).norm() / ref.norm().clamp_min(1e-30)).item()
    assert l2 < 3e-2, f"{name}: relative L2 {l2:.3e}"


def test_pe_mlp_golden_forward_and_gradients():
    from nerf_sys_b200.models.inr import MetaNeRF
    z = load_golden("pe_mlp")
    st = state_from(z)
    m = MetaNeRF(hidden=64, sigma_depth=8, skips=(4,), color_hidden=32, color_depth=2, encoding_dir="frequency")
    assert sorted(m.state_dict()) == sorted(st)
    m.load_state_dict(st, strict=True)
    m = m.to(DEV)
    out = m(T(z["x"]).to(DEV), T(z["d"]).to(DEV))
    close(out["sigma"], z["sigma"], 5e-3, "sigma")
    close(out["rgb"], z["rgb"], 5e-3, "rgb")
    (out["sigma"] * T(z["g_sigma"]).to(DEV)).sum().add((out["rgb"] * T(z["g_rgb"]).to(DEV)).sum()).backward()
    for n, p in m.named_parameters():
        _grad_close(p.grad, T(z["grad/" + n]), n)


def test_pe_mlp_fast_weights_take_the_gradient():
    """params overrides (meta_core.py:54-59): gradients flow to the override tensors, not to the module."""
    from nerf_sys_b200.models.inr import MetaNeRF
    z = load_golden("pe_mlp")
    m = MetaNeRF(hidden=64, color_hidden=32, encoding_dir="frequency")
    m.load_state_dict(state_from(z))
    m = m.to(DEV)
    fast = {n: p.detach().clone().requires_grad_(True) for n, p in m.named_parameters()}
    out = m(T(z["x"]).to(DEV), T(z["d"]).to(DEV), params=fast)
    loss = (out["sigma"] * T(z["g_sigma"]).to(DEV)).sum() + (out["rgb"] * T(z["g_rgb"]).to(DEV)).sum()
    gs = torch.autograd.grad(loss, tuple(fast.values()), allow_unused=True)
    for (n, _), g in zip(fast.items(), gs):
        _grad_close(g, T(z["grad/" + n]), n)
    assert all(p.grad is None for p in m.parameters())


def test_config1_default_widths_render_vs_oracle():
    """C1 shape, reference default widths (hidden 256, colour 128): 512 rays x 64 samples rendered and differentiated."""
    from nerf_sys_b200 import ops
    from nerf_sys_b200.models.inr import MetaNeRF
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    torch.manual_seed(0)
    m = MetaNeRF(encoding_dir="frequency")
    p_cpu = {k: v.detach().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    m = m.to(DEV).train()
    aabb, _, _ = O.grid_scene(1, 2)
    N, S = 512, 64
    rays = O.aerial_rays(N, aabb, seed=4)
    g = torch.Generator().manual_seed(8)
    u, gt = torch.rand(N, S, generator=g), torch.rand(N, 3, generator=g)
    rgb, depth, w, acc = render_rays(m, rays.to(DEV), ray_samples=S, u=u.to(DEV))
    loss = ops.mse_loss(rgb, gt.to(DEV), "identity")
    loss.backward()
    rgb_o, depth_o, w_o, acc_o = O.render_rays_pe(rays, p_cpu, S=S, u=u, bg=torch.ones(N, 3))
    loss_o = O.mse_loss(rgb_o, gt, "identity")
    loss_o.backward()
    close(rgb, rgb_o, 2e-3, "rgb")
    close(depth, depth_o, 2e-3, "depth")
    close(acc, acc_o, 2e-3, "acc")
    close(loss, loss_o, 2e-3, "loss")
    for n, p in m.named_parameters():
        _grad_close(p.grad, p_cpu[n].grad, n)
    # ragged sizes: P not a multiple of 128, fewer tiles than SMs
    m.eval()
    with torch.no_grad():
        for n_r, s in [(3, 5), (77, 33)]:
            r = O.aerial_rays(n_r, aabb, seed=9)
            a = render_rays(m, r.to(DEV), ray_samples=s)
            b = O.render_rays_pe(r, {k: v.detach() for k, v in p_cpu.items()}, S=s, u=None, bg=torch.ones(n_r, 3))
            close(a[0], b[0], 2e-3, f"rgb {n_r}x{s}")
