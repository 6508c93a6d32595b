"""torch.ops.acnerf.* are registered with fake (meta) implementations: shapes propagate under FakeTensorMode without a GPU."""
import torch
from torch._subclasses.fake_tensor import FakeTensorMode


def test_custom_ops_are_registered_and_propagate_shapes():
    from nerf_sys_b200 import torch_ops
    assert all(hasattr(torch.ops.acnerf, n) for n in torch_ops.OPS)
    with FakeTensorMode():
        rs, t, bg = torch.empty(10, 8, 4, device="cuda"), torch.empty(10, 8, device="cuda"), torch.empty(10, 3, device="cuda")
        assert [tuple(o.shape) for o in torch.ops.acnerf.volume_render(rs, t, bg, 1.0)] == [(10, 3), (10,), (10, 8), (10,)]
        loss, d = torch.ops.acnerf.mse_loss(torch.empty(7, 3, device="cuda"), torch.empty(7, 3, device="cuda"), 0)
        assert loss.shape == () and d.shape == (7, 3)
        assert torch.ops.acnerf.sh16(torch.empty(5, 3, device="cuda")).shape == (5, 16)
        assert torch.ops.acnerf.freq_encode(torch.empty(5, 3, device="cuda"), 10, True).shape == (5, 63)
        out = torch.ops.acnerf.packed_render(torch.empty(20, device="cuda"), torch.empty(20, 3, device="cuda"), torch.empty(20, device="cuda"),
                                            torch.empty(20, device="cuda"), torch.empty(6, dtype=torch.int64, device="cuda"),
                                            torch.empty(6, dtype=torch.int32, device="cuda"))
        assert [tuple(o.shape) for o in out] == [(6, 3), (6,), (20,), (6,)]
    s = str(torch.ops.acnerf.volume_render.default._schema)
    assert "Tensor rgb_sigma" in s and "float sigma_scale" in s
