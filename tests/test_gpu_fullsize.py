"""GPU parity at BASELINE.json's own sizes ("What's weak" 2-4): the 8-expert product configuration with
2^20-row tables, a full 136-row 1080p render block, 2^20-point routing against torch.cdist, the AMP contract of
runtime_adapt.py:249-268, the sRGB loss branch and trunc_exp at its clamp.  The CPU oracle runs each case in seconds."""
import math

import pytest
import torch

from oracle import acnerf_oracle as O
from test_gpu_parity import _big_scene, close, rel_close, table_grad_close

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _oracle_grads(st):
    for k, v in st.items():
        if "weight" in k or "bias" in k or "hash_table" in k:
            v.requires_grad_(True)


# ----------------------------------------------------------------------------- config 3 at its own size
def test_config3_full_size_8192x96_eight_experts_t20_vs_oracle():
    """BASELINE configs[2] exactly: K=8 (2x4 grid, margin 1.05), T=2^20 (eight 128 MiB tables, per-expert table pointers in
    the kernel-parameter block), 8192 aerial rays x 96 samples, training mode.  fp32 mode: rgb/depth/acc 5e-6, loss 1e-6,
    every gradient; tensor-core mode (the product default): rgb/depth/acc 2e-3 (meta_container.py:275-343)."""
    from nerf_sys_b200 import ops
    from nerf_sys_b200.nerfs.ray_rendering import render_rays
    m, st, cfg, mins, aabb = _big_scene(8, 2, 4, 20, seed=3)
    N, S = 8192, 96
    rays = O.aerial_rays(N, aabb, seed=1000)
    g = torch.Generator().manual_seed(9)
    u, gt = torch.rand(N, S, generator=g), torch.rand(N, 3, generator=g)
    _oracle_grads(st)
    rgb_o, depth_o, w_o, acc_o = O.render_rays(rays, st, mins, cfg, S=S, margin=1.05, u=u, direct_routing=True)
    loss_o = O.mse_loss(rgb_o, gt)
    loss_o.backward()
    m.train()
    with ops.mlp_math("fp32"):
        rgb, depth, w, acc = render_rays(m, rays.to(DEV), ray_samples=S, u=u.to(DEV))
        loss = ops.mse_loss(rgb, gt.to(DEV))
        loss.backward()
    close(rgb, rgb_o, 5e-6, "rgb")
    close(depth, depth_o, 5e-6, "depth")
    close(acc, acc_o, 5e-6, "acc")
    close(w, w_o, 5e-6, "weights")
    close(loss, loss_o, 1e-6, "loss")
    n_checked = 0
    for n, p in m.named_parameters():
        ref = st[n].grad
        assert ref is not None and p.grad is not None, n  # every expert receives samples from aerial rays
        if n.endswith("hash_table"):
            table_grad_close(p.grad, ref, 5e-5, n)
        else:
            rel_close(p.grad, ref, 5e-4, n)
            assert ((p.grad.cpu() - ref).norm() / ref.norm()).item() < 5e-4, n
        n_checked += 1
    assert n_checked == 8 * 15 + 4
    m.zero_grad(set_to_none=True)
    with ops.mlp_math("f16tc"), torch.no_grad():
        rgb_tc, depth_tc, _, acc_tc = render_rays(m, rays.to(DEV), ray_samples=S, u=u.to(DEV))
    close(rgb_tc, rgb_o, 2e-3, "tensor-core rgb")
    close(depth_tc, depth_o, 2e-3, "tensor-core depth")
    close(acc_tc, acc_o, 2e-3, "tensor-core acc")


# ----------------------------------------------------------------------------- config 5: one 1080p row block
def test_1080p_row_block_sample_major_t20_vs_oracle():
    """One 136-row block of a 1920x1080 frame x 64 samples (16.7 M samples, the renderer's block size) through
    render_image (sample-major point ids, speculative routing capacity, in-place compositor) with eight 2^20-row tables,
    against the oracle's render_rays on the same pixels (eval mode: no jitter).  ray_rendering.py:577-627."""
    from nerf_sys_b200 import ops
    from nerf_sys_b200.nerfs.ray_rendering import render_image
    from nerf_sys_b200.nerfs.scene_box import SceneBox
    m, st, cfg, mins, aabb = _big_scene(8, 2, 4, 20, seed=3)
    m.eval()
    H, W, f, S = 1080, 1920, 1480.0, 64
    r0, r1 = 472, 608
    c2w = O.ring_pose(2, 20, radius=0.9)
    kw = dict(H=H, W=W, fx=f, fy=f, cx=W / 2, cy=H / 2, c2w=c2w, scene_box=SceneBox(aabb.to(DEV)), ray_samples=S)
    out = {}
    for mode in ("fp32", "f16tc"):
        with ops.mlp_math(mode):
            out[mode] = [x.cpu() for x in render_image(m, row_range=(r0, r1), **kw)]
    dirs = O.ray_directions(H, W, f, f, W / 2, H / 2, True)
    rays = O.make_rays(dirs, c2w, aabb)[r0:r1].reshape(-1, 8)
    # the oracle in ray chunks (it materialises (P,16,8) int64 index tensors)
    step = 8 * W
    ref = [[], [], []]
    with torch.no_grad():
        for i in range(0, rays.shape[0], step):
            c, d, _, a = O.render_rays(rays[i:i + step], st, mins, cfg, S=S, margin=1.05, u=None, direct_routing=True)
            for lst, v in zip(ref, (c.clamp(0, 1), d, a)):
                lst.append(v)
    ref = [torch.cat(x) for x in ref]
    assert float((ref[2] > 0.01).float().mean()) > 0.5  # the block looks at the scene
    for name, got, want in zip(("rgb", "depth", "acc"), out["fp32"], ref):
        close(got.reshape(want.shape), want, 5e-6, f"fp32 {name}")
    for name, got, want in zip(("rgb", "depth", "acc"), out["f16tc"], ref):
        close(got.reshape(want.shape), want, 2e-3, f"tensor-core {name}")


# ----------------------------------------------------------------------------- routing vs torch.cdist at 2^20 points
@pytest.mark.parametrize("K,c2d,margin", [(8, True, 1.05), (8, False, 1.05), (8, True, 1.0)])
def test_routing_2_20_points_vs_torch_cdist(K, c2d, margin):
    """The kernel's distances are sqrt(sum (x-c)^2); the reference calls torch.cdist (meta_container.py:122), whose result
    can differ in the last bits.  On 2^20 uniform points the integer decisions (membership mask / argmin) must agree
    everywhere except where a distance sits within 4e-6 (relative) of the decision threshold, and such points must be rare."""
    from nerf_sys_b200 import ops
    P = 1 << 20
    g = torch.Generator().manual_seed(100 + K)
    aabb, cent, _ = O.grid_scene(2, 4)
    x = aabb[0] + (aabb[1] - aabb[0]) * torch.rand(P, 3, generator=g)
    x6 = torch.cat([x, torch.zeros(P, 3)], 1).to(DEV)
    w_ref, hard_ref = O.routing(x, cent, margin, c2d, direct=False)  # torch.cdist, as the reference
    w, hard = ops.routing(ops.PointSet(x6=x6), cent.reshape(-1).tolist(), K, margin, c2d)
    axes = [1, 2] if c2d else [0, 1, 2]
    d = torch.cdist(x[:, axes].double(), cent[:, axes].double())
    dmin = d.min(1, keepdim=True).values
    if margin > 1.0:
        mism = ((w.cpu() > 0) != (w_ref > 0))
        near_thr = (d - margin * dmin).abs() <= 4e-6 * d
        assert not (mism & ~near_thr).any(), int((mism & ~near_thr).sum())
        assert mism.any(1).sum().item() <= 64
        agree = ~mism.any(1)
        close(w.cpu()[agree], w_ref[agree], 2e-6, "weights")
    else:
        mism = hard.cpu() != hard_ref
        second = d.topk(2, dim=1, largest=False).values
        tie = (second[:, 1] - second[:, 0]) <= 4e-6 * second[:, 1]
        assert not (mism & ~tie).any() and mism.sum().item() <= 64


# ----------------------------------------------------------------------------- AMP contract
@pytest.mark.parametrize("flow", ["reference", "fused"])
def test_autocast_and_gradscaler_step_matches_the_plain_fp32_step(flow):
    """runtime_adapt.py:249-268: forward under torch.autocast(fp16), scaler.scale(loss).backward(), scaler.unscale_(opt),
    clip_grad_norm_, scaler.step(opt), scaler.update().  The ops are opaque to autocast (fp32 in, fp32 out, fp32
    gradients), so three such steps must land where three plain fp32 steps land.  "reference": the reference's own call
    sequence around FusedAdam; "fused": FusedAdam clips and unscales inside its kernel (GradScaler hands it the scale)."""
    from nerf_sys_b200 import ops
    from nerf_sys_b200.nerfs.losses import compute_mse_loss
    from nerf_sys_b200.optim import build_optimizer

    class HP:
        ray_samples, chunk_points, color_space = 32, 1 << 20, "linear"

    N = 1024
    g = torch.Generator().manual_seed(21)
    batches = []
    _, _, _, _, aabb = _big_scene(2, 1, 2, 14, seed=8)
    for i in range(3):
        batches.append({"rays": O.aerial_rays(N, aabb, seed=90 + i).to(DEV), "rgbs": torch.rand(N, 3, generator=g).to(DEV)})

    def run(amp):
        m, *_ = _big_scene(2, 1, 2, 14, seed=8)
        m.train()
        fused_clip = amp and flow == "fused"
        opt = build_optimizer(m, max_grad_norm=0.05 if (fused_clip or not amp) else None)
        scaler = torch.amp.GradScaler("cuda", init_scale=4096.0, enabled=amp)
        params = [p for gr in opt.param_groups for p in gr["params"]]
        losses = []
        for b in batches:
            torch.manual_seed(1234)  # the stratified jitter (torch.rand on the device) is the same in both runs
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.float16, enabled=amp):
                loss = compute_mse_loss(HP, m, b)
            assert loss.dtype == torch.float32
            scaler.scale(loss).backward()
            if amp and not fused_clip:
                scaler.unscale_(opt)
                torch.nn.utils.clip_grad_norm_(params, 0.05)
            scaler.step(opt)
            scaler.update()
            losses.append(loss.item())
        for p in params:
            assert p.grad is None or p.grad.dtype == torch.float32
        return {n: p.detach().clone() for n, p in m.named_parameters()}, losses

    with ops.mlp_math("fp32"):
        ref_p, ref_l = run(False)
        got_p, got_l = run(True)
    for a, b in zip(got_l, ref_l):
        assert abs(a - b) <= 2e-6 * max(1.0, abs(b)), (got_l, ref_l)
    for n in ref_p:
        close(got_p[n], ref_p[n], 5e-6, n)


# ----------------------------------------------------------------------------- loss colour spaces
@pytest.mark.parametrize("space", ["linear", "srgb", "identity"])
def test_mse_loss_colour_spaces_fwd_bwd_vs_oracle(space):
    """losses.py:10-32 + color_space.py:22-66, every branch of the kernel incl. the sRGB transfer-curve knee."""
    from nerf_sys_b200 import ops
    g = torch.Generator().manual_seed(3)
    N = 5000
    pred = torch.rand(N, 3, generator=g) * 1.2 - 0.1  # includes values the clamp cuts
    gt = torch.rand(N, 3, generator=g)
    pred[:64] = torch.linspace(0.0025, 0.0038, 64)[:, None]  # around the knee of linear_to_srgb (0.0031308)
    gt[:64] = torch.linspace(0.035, 0.045, 64)[:, None]      # around the knee of srgb_to_linear (0.04045)
    p_o = pred.clone().requires_grad_(True)
    loss_o = O.mse_loss(p_o, gt, space)
    loss_o.backward()
    p = pred.clone().to(DEV).requires_grad_(True)
    loss = ops.mse_loss(p, gt.to(DEV), space)
    loss.backward()
    close(loss, loss_o, 1e-6, f"{space} loss")
    rel_close(p.grad, p_o.grad, 2e-5, f"{space} d loss / d pred")


# ----------------------------------------------------------------------------- trunc_exp at the clamp
def test_trunc_exp_at_the_fp32_clamp_in_the_fused_expert_kernel():
    """trunc_exp.py:30-57: sigma = exp(clamp(raw, +-88.7228...)); the backward multiplies by exp(clamped raw) and is NOT
    zeroed outside the clamp.  The raw density is driven through sigma_head.bias with zero head weights, so the fused
    expert kernel sees exactly the boundary values.  The fp32 nearest to the clamp constant (88.72283935...) lies just
    above ln(FLT_MAX), so at and beyond the clamp the reference's sigma is saturated (inf on the CPU): there the checks
    are "saturated, and identical to the value at the clamp"; below it, forward and bias gradient vs the oracle."""
    from nerf_sys_b200 import ops
    lim = O.EXP_CLAMP_F32
    res = {}
    for raw in (-100.0, -lim, -88.0, 80.0, 88.0, 88.7, lim, 100.0):
        m, st, cfg, mins, aabb = _big_scene(1, 1, 1, 12, seed=2)
        with torch.no_grad():
            m.submodules[0].sigma_head.weight.zero_()
            m.submodules[0].sigma_head.bias.fill_(raw)
        st["submodules.0.sigma_head.weight"].zero_()
        st["submodules.0.sigma_head.bias"].fill_(raw)
        g = torch.Generator().manual_seed(4)
        x = aabb[0] + (aabb[1] - aabb[0]) * torch.rand(256, 3, generator=g)
        x6 = torch.cat([x, torch.nn.functional.normalize(torch.randn(256, 3, generator=g), dim=-1)], 1)
        cot = torch.zeros(256, 4)
        cot[:, 3] = (torch.rand(256, generator=g) + 0.5) * (1e-36 if raw > 0 else 1.0)  # keeps g * exp(88) finite
        with ops.mlp_math("fp32"):
            y = m(x6.to(DEV))
            (y * cot.to(DEV)).sum().backward()
        b_o = st["submodules.0.sigma_head.bias"].requires_grad_(True)
        y_o = O.container_forward(x6, st, mins, cfg, margin=1.05, cluster_2d=True)
        (y_o * cot).sum().backward()
        sig, gb = y[:, 3].detach().cpu(), m.submodules[0].sigma_head.bias.grad.detach().cpu()
        res[raw] = (sig, gb)
        if raw >= lim:
            assert (sig >= 3.4e38).all() and (y_o[:, 3] >= 3.4e38).all(), raw      # saturated on both sides
            assert float(gb) > 0 and float(b_o.grad) > 0                           # the gradient is not zeroed beyond the clamp
        elif raw > 0:
            want = math.exp(raw)
            rel_close(sig, torch.full((256,), want), 4e-6, f"sigma at raw={raw}")
            rel_close(sig, y_o[:, 3], 4e-6, f"sigma vs oracle at raw={raw}")
            rel_close(gb, b_o.grad, 2e-5, f"d/d bias at raw={raw}")
        else:  # exp(-88) is at the edge of the denormals: absolute comparison
            close(sig, y_o[:, 3], 1e-6, f"sigma vs oracle at raw={raw}")
            close(gb, b_o.grad, 1e-6, f"d/d bias at raw={raw}")
            assert (sig >= 0).all() and (sig < 1e-37).all()
    assert torch.equal(res[100.0][0], res[lim][0]) and torch.equal(res[-100.0][0], res[-lim][0])  # the clamp is in force
    assert torch.equal(res[100.0][1], res[lim][1])
