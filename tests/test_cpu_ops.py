"""C++ CPU ops vs torch.nn.functional on CPU, incl. backward via autograd."""
import pytest
import torch
import torch.nn.functional as F

from distributed_b200.ops import cpu


def _ref_conv(x, w, b, stride, pad):
    return F.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), b, stride=stride,
                    padding=pad).permute(0, 2, 3, 1).contiguous()


@pytest.mark.parametrize("N,H,W,Ci,Co,K,stride,pad", [
    (4, 28, 28, 1, 32, 3, 1, 0),      # README first layer (README.md:293)
    (2, 13, 13, 8, 16, 3, 1, 0), (2, 16, 16, 3, 8, 7, 2, 3), (3, 9, 9, 5, 6, 1, 2, 0),
    (2, 8, 8, 4, 4, 3, 1, 1)])
def test_conv2d_fwd_dgrad_wgrad(N, H, W, Ci, Co, K, stride, pad):
    g = torch.Generator().manual_seed(N + H + Co)
    x = torch.randn((N, H, W, Ci), generator=g).requires_grad_()
    w = (torch.randn((K, K, Ci, Co), generator=g) * 0.2).requires_grad_()
    b = torch.randn((Co,), generator=g).requires_grad_()
    ref = torch.relu(_ref_conv(x, w, b, stride, pad))
    y = torch.empty_like(ref)
    cpu.conv2d_fwd(x.detach(), w.detach(), b.detach(), y, (stride, stride), (pad, pad), relu=True)
    assert torch.allclose(y, ref, atol=1e-5, rtol=1e-4)
    dy = torch.randn(ref.shape, generator=g)
    ref.backward(dy)
    dpre = torch.empty_like(dy)
    cpu.relu_bwd(dy, y, dpre)
    dx = torch.empty_like(x)
    cpu.conv2d_dgrad(dpre, w.detach(), dx, (stride, stride), (pad, pad))
    assert torch.allclose(dx, x.grad, atol=1e-4, rtol=1e-4)
    dw, db = torch.zeros_like(w), torch.zeros_like(b)
    cpu.conv2d_wgrad(x.detach(), dpre, dw, db, (stride, stride), (pad, pad))
    assert torch.allclose(dw, w.grad, atol=1e-3, rtol=1e-4)
    assert torch.allclose(db, b.grad, atol=1e-3, rtol=1e-4)


@pytest.mark.parametrize("shape", [(4, 26, 26, 32), (2, 7, 9, 5)])
def test_maxpool2(shape):
    g = torch.Generator().manual_seed(1)
    x = torch.randn(shape, generator=g).requires_grad_()
    N, H, W, C = shape
    y = torch.empty((N, H // 2, W // 2, C))
    code = torch.empty(y.shape, dtype=torch.uint8)
    cpu.maxpool2_fwd(x.detach(), y, code)
    ref = F.max_pool2d(x.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert torch.equal(y, ref)
    dy = torch.randn(y.shape, generator=g)
    ref.backward(dy)
    dx = torch.empty(shape)
    cpu.maxpool2_bwd(dy, code, y, dx, False)
    assert torch.equal(dx, x.grad)


def test_gemm_variants_and_xent_and_sgd():
    g = torch.Generator().manual_seed(2)
    x, w, b = torch.randn(64, 100, generator=g), torch.randn(100, 10, generator=g), torch.randn(10, generator=g)
    d = torch.empty(64, 10)
    cpu.gemm(x, w, d, bias=b, relu=True)
    assert torch.allclose(d, torch.relu(x @ w + b), atol=1e-4)
    dw = torch.ones(100, 10)
    cpu.gemm(x, d, dw, trans_a=True, accumulate=True)
    assert torch.allclose(dw, x.t() @ d + 1, atol=1e-3)
    dx = torch.empty(64, 100)
    cpu.gemm(d, w, dx, trans_b=True)
    assert torch.allclose(dx, d @ w.t(), atol=1e-3)

    z = torch.randn(64, 10, generator=g).requires_grad_()
    labels = torch.randint(0, 10, (64,), generator=g, dtype=torch.int32)
    loss = F.cross_entropy(z, labels.long(), reduction="sum")
    (loss / 64).backward()
    dz, m = torch.empty_like(z), torch.zeros(4)
    cpu.softmax_xent(z.detach(), labels, dz, m, 1 / 64)
    assert torch.allclose(dz, z.grad, atol=1e-6)
    assert abs(m[0].item() - loss.item()) < 1e-3
    assert m[1].item() == (z.argmax(1) == labels).sum().item() and m[2].item() == 64

    wv, gv = torch.randn(104, generator=g), torch.randn(104, generator=g)
    w0, g0 = wv.clone(), gv.clone()
    cpu.sgd_apply(wv, gv, 100, 0.01)
    assert torch.allclose(wv[:100], w0[:100] - 0.01 * g0[:100])
    assert torch.allclose(wv[100:], w0[100:] + g0[100:]) and gv.abs().max() == 0
    cs = torch.zeros(10)
    cpu.colsum(d, cs)
    assert torch.allclose(cs, d.sum(0), atol=1e-3)
