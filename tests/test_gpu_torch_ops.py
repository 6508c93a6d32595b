"""torch.ops.acnerf.* on the GPU: torch.library.opcheck (schema, fake-tensor consistency, autograd registration) and equality with
the ops.py wrappers they share kernels with."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_volume_render_and_mse_custom_ops_opcheck_and_match_the_wrappers():
    from nerf_sys_b200 import ops, torch_ops  # noqa: F401
    g = torch.Generator().manual_seed(0)
    N, S = 257, 33
    rs = torch.rand(N, S, 4, generator=g).to(DEV).requires_grad_(True)
    t = (torch.rand(N, S, generator=g).cumsum(1) * 0.05).to(DEV)
    bg = torch.rand(N, 3, generator=g).to(DEV).requires_grad_(True)
    torch.library.opcheck(torch.ops.acnerf.volume_render.default, (rs, t, bg, 1.5), test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
    out = torch.ops.acnerf.volume_render(rs, t, bg, 1.5)
    gt = torch.rand(N, 3, generator=g).to(DEV)
    loss, _ = torch.ops.acnerf.mse_loss(out[0], gt, 0)
    (loss + out[1].sum() * 0.1 + out[3].sum() * 0.01).backward()
    rs2, bg2 = rs.detach().clone().requires_grad_(True), bg.detach().clone().requires_grad_(True)
    ref = ops.composite(rs2, t, bg2, 1.5)
    (ops.mse_loss(ref[0], gt, "linear") + ref[1].sum() * 0.1 + ref[3].sum() * 0.01).backward()
    for a, b in zip(out, ref):
        assert torch.equal(a, b)
    assert torch.equal(rs.grad, rs2.grad) and torch.equal(bg.grad, bg2.grad)
    torch.library.opcheck(torch.ops.acnerf.mse_loss.default, (out[0].detach().requires_grad_(True), gt, 1),
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
    assert torch.equal(torch.ops.acnerf.sh16(bg.detach()), ops.sh16(bg.detach()))


def test_packed_render_custom_op_matches_the_autograd_function():
    from nerf_sys_b200 import ops, torch_ops  # noqa: F401
    g = torch.Generator().manual_seed(1)
    counts = torch.randint(0, 9, (64,), generator=g).int()
    M = int(counts.sum())
    offsets = (torch.cumsum(counts.long(), 0) - counts.long())
    t0 = torch.rand(M, generator=g).to(DEV)
    seg = ops.PackedSegments(64, torch.repeat_interleave(torch.arange(64), counts.long()).int().to(DEV), t0, t0 + 0.05, offsets.to(DEV), counts.to(DEV))
    sigma = (torch.rand(M, generator=g) * 5).to(DEV).requires_grad_(True)
    rgb = torch.rand(M, 3, generator=g).to(DEV).requires_grad_(True)
    a = torch.ops.acnerf.packed_render(sigma, rgb, seg.t0, seg.t1, seg.offsets, seg.counts)
    (a[0].sum() + a[1].sum() + a[3].sum()).backward()
    s2, c2 = sigma.detach().clone().requires_grad_(True), rgb.detach().clone().requires_grad_(True)
    b = ops.packed_render(s2, c2, seg)
    (b[0].sum() + b[1].sum() + b[3].sum()).backward()
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    assert torch.equal(sigma.grad, s2.grad) and torch.equal(rgb.grad, c2.grad)
    torch.library.opcheck(torch.ops.acnerf.packed_render.default, (sigma.detach().requires_grad_(True), rgb.detach(), seg.t0, seg.t1, seg.offsets, seg.counts),
                          test_utils=("test_schema", "test_faketensor", "test_autograd_registration"))
