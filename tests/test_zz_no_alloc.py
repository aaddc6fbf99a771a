"""GPU test of l1_loss_forward_static (include/gsplat_b200.h) -- the entry point that allocates nothing: its scratch is a
file-scope `static __device__` array inside the library.  Called through the C ABI (ctypes on libgsplat_b200.so) and checked
against the existing l1_loss_forward (caller workspace) on the same inputs: loss and per-pixel gradient bit-identical (same
kernel, same block count, same summation order; the L1 tests' own tolerance is 1e-6 relative), on two calls in a row with
different inputs -- the second call finds the first call's partial sums and ticket in the shared array.  Collected last
(`zz`)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
vp = ctypes.c_void_p


def _lib():
    from diff_gaussian_rasterization import _C as _gs
    L = _gs.lib()
    L.l1_loss_workspace_bytes.restype = ctypes.c_size_t
    L.l1_loss_workspace_bytes.argtypes = []
    L.l1_loss_forward.restype = ctypes.c_int
    L.l1_loss_forward.argtypes = [vp, vp, ctypes.c_int64, vp, vp, vp, ctypes.c_size_t, vp]
    L.l1_loss_forward_static.restype = ctypes.c_int
    L.l1_loss_forward_static.argtypes = [vp, vp, ctypes.c_int64, vp, vp, vp]
    L.l1_loss_static_max_n.restype = ctypes.c_int64
    L.l1_loss_static_max_n.argtypes = []
    return L


def test_l1_loss_forward_static_matches_workspace_variant(built_lib):
    L = _lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    ws = torch.zeros(L.l1_loss_workspace_bytes(), dtype=torch.uint8, device=dev)
    gen = torch.Generator(device=dev).manual_seed(5)
    for call, shape in enumerate([(3, 512, 512), (3, 333, 517)]):  # two calls in a row; the second with a ragged tail (n % 4 != 0)
        img = torch.rand(*shape, device=dev, generator=gen)
        gt = torch.rand(*shape, device=dev, generator=gen)
        img.view(-1)[::5] = gt.view(-1)[::5]  # exact zeros: sign(0) = 0
        n = img.numel()
        loss_a, loss_b = torch.full((1,), -1.0, device=dev), torch.full((1,), -2.0, device=dev)
        grad_a, grad_b = torch.full_like(img, 7.0), torch.full_like(img, 9.0)
        assert L.l1_loss_forward(img.data_ptr(), gt.data_ptr(), n, loss_a.data_ptr(), grad_a.data_ptr(), ws.data_ptr(), ws.numel(), st) == 0
        assert L.l1_loss_forward_static(img.data_ptr(), gt.data_ptr(), n, loss_b.data_ptr(), grad_b.data_ptr(), st) == 0
        torch.cuda.synchronize(dev)
        ref = (img.double() - gt.double()).abs().mean()
        assert abs(float(loss_b) - float(ref)) <= 1e-6 * float(ref), (call, float(loss_b), float(ref))
        assert float(loss_a) == float(loss_b), (call, float(loss_a), float(loss_b))
        assert torch.equal(grad_a, grad_b), call
        assert torch.equal(grad_b, torch.sign(img - gt) / n), call


def test_l1_loss_forward_static_rejects_sizes_beyond_its_limit(built_lib):
    L = _lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    limit = L.l1_loss_static_max_n()
    assert limit == 1 << 26
    x = torch.zeros(16, device=dev)
    loss = torch.full((1,), 3.0, device=dev)
    # rejected before anything is launched: the (too small) buffers are never touched
    assert L.l1_loss_forward_static(x.data_ptr(), x.data_ptr(), limit + 4, loss.data_ptr(), x.data_ptr(), st) == -5  # GS_ERR_STATIC_LIMIT
    assert L.l1_loss_forward_static(x.data_ptr(), x.data_ptr(), 0, loss.data_ptr(), x.data_ptr(), st) == -1          # GS_ERR_BAD_ARG
    torch.cuda.synchronize(dev)
    assert float(loss) == 3.0 and float(x.abs().sum()) == 0.0
