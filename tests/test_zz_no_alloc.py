"""acn_mse_loss_static: the entry point that allocates nothing (scratch in a file-scope `static __device__` array; no
counterpart in the reference, include/acnerf.h).  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from nerf_sys_b200._lib import lib

ACN_EINVAL = 1  # include/acnerf.h
MAX_RAYS = 1 << 24  # ACN_MSE_STATIC_MAX_RAYS


def _p(t):
    return C.c_void_p(t.data_ptr())


def test_static_entry_rejects_sizes_beyond_its_limit_before_touching_the_gpu():
    """Argument checks return before any CUDA call, so this runs without a device; the pointers are never dereferenced."""
    L = lib()
    buf = (C.c_float * 4)()
    fake = C.cast(buf, C.c_void_p)
    rc = L.acn_mse_loss_static(fake, fake, MAX_RAYS + 1, 0, fake, None, None)
    assert rc == ACN_EINVAL
    msg = L.acn_last_error().decode()
    assert "exceeds" in msg and str(MAX_RAYS) in msg and "acn_mse_loss_static" in msg
    assert L.acn_mse_loss_static(None, fake, 16, 0, fake, None, None) == ACN_EINVAL
    assert L.acn_mse_loss_static(fake, fake, 0, 0, fake, None, None) == ACN_EINVAL
    assert L.acn_mse_loss_static(fake, fake, 16, 3, fake, None, None) == ACN_EINVAL


@pytest.mark.gpu
def test_static_entry_matches_the_workspace_entry_on_two_calls_in_a_row():
    """Through the C ABI: loss and d_pred of acn_mse_loss_static == acn_mse_loss (same kernels, same grid, same order of
    summation: bit-equal, and within 1e-6 of the fp64 value), for a large call followed by a small one (the shared array
    still holds the first call's partial sums: the second must not read them), in all three colour spaces; and the
    device's free memory does not move across the static calls (cudaMalloc works in 2 MiB steps)."""
    L = lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator().manual_seed(123)
    cases = []
    for N in (300_001, 777):  # 3516 CTAs' worth (grid capped by the SM count / 2048), then 10 CTAs
        pred = (torch.rand(N, 3, generator=g) * 1.4 - 0.2).to(dev)  # values outside [0,1] exercise the clamps
        gt = torch.rand(N, 3, generator=g).to(dev)
        cases.append((N, pred, gt))
    ws = torch.empty(2048, device=dev)
    for space in (0, 1, 2):
        ref, got = [], []
        for N, pred, gt in cases:
            loss, d = torch.empty(1, device=dev), torch.empty_like(pred)
            assert L.acn_mse_loss(_p(pred), _p(gt), N, space, _p(loss), _p(d), _p(ws), stream) == 0, L.acn_last_error()
            ref.append((loss, d))
        outs = [(torch.empty(1, device=dev), torch.empty_like(pred)) for _, pred, _ in cases]
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for (N, pred, gt), (loss, d) in zip(cases, outs):  # two calls in a row on one stream
            assert L.acn_mse_loss_static(_p(pred), _p(gt), N, space, _p(loss), _p(d), stream) == 0, L.acn_last_error()
            got.append((loss, d))
        torch.cuda.synchronize()
        assert abs(free0 - torch.cuda.mem_get_info()[0]) < (1 << 20)
        for (N, pred, gt), (l_ref, d_ref), (l_got, d_got) in zip(cases, ref, got):
            assert torch.equal(l_got, l_ref) and torch.equal(d_got, d_ref), (space, N, float(l_got), float(l_ref))
            if space == 2:  # identity colour space: the closed form, in fp64
                want = ((pred.double() - gt.double().clamp(0, 1)) ** 2).mean()
                assert abs(float(l_got) - float(want)) <= 1e-6 * max(1.0, float(want))
                torch.testing.assert_close(d_got.double(), 2 * (pred.double() - gt.double().clamp(0, 1)) / (3 * N), rtol=1e-5, atol=1e-12)
    # d_pred is optional, as in acn_mse_loss
    N, pred, gt = cases[1]
    loss = torch.empty(1, device=dev)
    assert L.acn_mse_loss_static(_p(pred), _p(gt), N, 2, _p(loss), None, stream) == 0
    torch.cuda.synchronize()
    assert torch.equal(loss, got[1][0])
