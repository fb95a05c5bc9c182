"""ResNet-family kernels (BatchNorm+ReLU, residual add, general max-pool, global average pool)
vs PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _g(seed):
    return torch.Generator(device="cuda").manual_seed(seed)


@pytest.mark.parametrize("M,C,relu", [(64 * 49, 64, True), (3000, 256, False), (50, 12, True)])
def test_batchnorm_train_forward_backward(M, C, relu):
    from distributed_b200.ops import nn
    g = _g(M + C)
    x = (torch.randn((M, C), device="cuda", generator=g) * 2 + 1).requires_grad_()
    gamma = (torch.rand(C, device="cuda", generator=g) + 0.5).requires_grad_()
    beta = torch.randn(C, device="cuda", generator=g).requires_grad_()
    eps, mom = 1e-3, 0.9
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    ref = F.batch_norm(x, rm, rv, gamma, beta, training=True, momentum=1 - mom, eps=eps)
    if relu:
        ref = torch.relu(ref)
    stats, saved, red = (torch.zeros(2 * C, device="cuda") for _ in range(3))
    moving = torch.cat([torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")])
    y = torch.empty_like(x)
    nn.bn_fwd(x.detach(), gamma.detach(), beta.detach(), stats, saved, moving, y, eps, mom, relu)
    assert torch.allclose(y, ref, atol=2e-4, rtol=1e-4)
    assert torch.allclose(moving[:C], rm, atol=1e-4) and torch.allclose(moving[C:], rv, atol=1e-3)
    assert stats.abs().max().item() == 0
    dy = torch.randn((M, C), device="cuda", generator=g)
    ref.backward(dy)
    dx = torch.empty_like(x)
    dgamma, dbeta = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    nn.bn_bwd(x.detach(), y, dy, gamma.detach(), saved, red, dx, dgamma, dbeta, relu)
    assert torch.allclose(dx, x.grad, atol=1e-4 * max(1.0, x.grad.abs().max().item()), rtol=1e-3)
    assert torch.allclose(dgamma, gamma.grad, atol=1e-3 * gamma.grad.abs().max().item() + 1e-4)
    assert torch.allclose(dbeta, beta.grad, atol=1e-3 * beta.grad.abs().max().item() + 1e-4)
    assert red.abs().max().item() == 0


@pytest.mark.parametrize("shape,k,s,p", [((2, 112, 112, 64), 3, 2, 1), ((3, 9, 11, 8), 3, 2, 1), ((2, 8, 8, 4), 2, 2, 0)])
def test_general_maxpool(shape, k, s, p):
    from distributed_b200.ops import nn
    g = _g(sum(shape))
    x = torch.randn(shape, device="cuda", generator=g).requires_grad_()
    ref = F.max_pool2d(x.permute(0, 3, 1, 2), k, s, p).permute(0, 2, 3, 1).contiguous()
    y = torch.empty_like(ref)
    code = torch.empty(ref.shape, device="cuda", dtype=torch.uint8)
    nn.maxpool_fwd(x.detach(), y, code, k, s, p)
    assert torch.equal(y, ref)
    dy = torch.randn(ref.shape, device="cuda", generator=g)
    ref.backward(dy)
    dx = torch.empty_like(x)
    nn.maxpool_bwd(dy, code, dx, k, s, p)
    assert torch.allclose(dx, x.grad, atol=1e-6)


def test_add_relu_and_global_average_pool():
    from distributed_b200.ops import nn
    g = _g(3)
    a, b = torch.randn((4, 7, 7, 64), device="cuda", generator=g), torch.randn((4, 7, 7, 64), device="cuda", generator=g)
    out = torch.empty_like(a)
    nn.add_act(a, b, out, True)
    assert torch.equal(out, torch.relu(a + b))
    nn.add_act(a, b, out, False)
    assert torch.equal(out, a + b)
    y = torch.empty((4, 64), device="cuda")
    nn.gap_fwd(a, y)
    assert torch.allclose(y, a.mean((1, 2)), atol=1e-5)
    dx = torch.empty_like(a)
    nn.gap_bwd(y, dx)
    assert torch.allclose(dx, (y / 49)[:, None, None, :].expand_as(a), atol=1e-7)
