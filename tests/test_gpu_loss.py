"""GPU test of the fused L1 loss (l1_loss_forward / l1_loss_backward through avatar_pose.loss.l1_loss) against the plain
PyTorch fp32 expression of the reference, utils/loss_utils.py:19-20: value <= 1e-6 relative (different summation order),
gradient bit-identical (sign(diff) * g / n)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(3, 512, 512), (3, 77, 53), (1,), (3, 1080, 1920)])
def test_l1_loss_matches_torch(built_lib, shape):
    from avatar_pose.loss import l1_loss
    dev = torch.device("cuda:0")
    torch.manual_seed(1)
    img = torch.rand(*shape, device=dev)
    gt = torch.rand(*shape, device=dev)
    img.view(-1)[::7] = gt.view(-1)[::7]  # exact zeros: sign(0) = 0
    a = img.clone().requires_grad_(True)
    b = img.clone().requires_grad_(True)
    la = l1_loss(a, gt)
    lb = (b - gt).abs().mean()
    (la * 3.0).backward()
    (lb * 3.0).backward()
    ref64 = (img.double() - gt.double()).abs().mean()
    assert abs(float(la) - float(ref64)) <= 1e-6 * float(ref64) + 1e-12
    assert torch.equal(a.grad, b.grad)
    # deterministic: the same bits on every call
    assert all(float(l1_loss(img, gt)) == float(la) for _ in range(3))
