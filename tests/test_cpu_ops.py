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


# ---- ResNet-family layers on the CPU (csrc/cpu/resnet_ops.cpp), shapes drawn by hypothesis ----------------
from hypothesis import given, settings, strategies as st_  # noqa: E402


@settings(max_examples=25, deadline=None)
@given(M=st_.integers(2, 200), C=st_.integers(1, 17), relu=st_.booleans(), with_res=st_.booleans(),
       seed=st_.integers(0, 10_000))
def test_batchnorm_cpu_matches_autograd(M, C, relu, with_res, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn((M, C), generator=g) * 2 + 0.5).requires_grad_()
    gamma = (torch.rand(C, generator=g) + 0.5).requires_grad_()
    beta = torch.randn(C, generator=g).requires_grad_()
    res = torch.randn((M, C), generator=g) if with_res else None
    eps, mom, lr = 1e-3, 0.9, 0.05
    rm, rv = torch.randn(C, generator=g), torch.rand(C, generator=g) + 0.5
    mm, mv = rm.clone(), rv.clone()
    pre = F.batch_norm(x, rm, rv, gamma, beta, training=True, momentum=1 - mom, eps=eps)
    ref = pre + res if with_res else pre
    if relu:
        ref = torch.relu(ref)
    y, saved = torch.empty_like(ref), torch.zeros(2 * C)
    gm, gv = torch.zeros(C), torch.zeros(C)
    cpu.bn_fwd(x.detach(), gamma.detach(), beta.detach(), y, saved, (mm, mv), (gm, gv), eps, mom, 1 / lr, relu,
               False, res)
    assert torch.allclose(y, ref, atol=1e-4, rtol=1e-4)
    # the optimizer step on the pseudo-gradient is the moving-average update
    assert torch.allclose(mm - lr * gm, rm, atol=1e-5) and torch.allclose(mv - lr * gv, rv, atol=1e-4)
    dy = torch.randn((M, C), generator=g)
    if with_res:            # fused block tail: the caller masks the gradient, the kernel must not
        if relu:
            dy = dy * (ref > 0)
        pre.backward(dy)
    else:
        ref.backward(dy)
    dx, dgamma, dbeta = torch.empty_like(x), torch.zeros(C), torch.zeros(C)
    cpu.bn_bwd(x.detach(), y, dy, gamma.detach(), saved, dx, dgamma, dbeta, relu and not with_res)
    scale = max(1.0, x.grad.abs().max().item())
    assert torch.allclose(dx, x.grad, atol=2e-4 * scale, rtol=1e-3)
    assert torch.allclose(dgamma, gamma.grad, atol=1e-3 * max(1.0, gamma.grad.abs().max().item()))
    assert torch.allclose(dbeta, beta.grad, atol=1e-3 * max(1.0, beta.grad.abs().max().item()))
    # inference on the moving statistics
    y2 = torch.empty_like(ref)
    cpu.bn_fwd(x.detach(), gamma.detach(), beta.detach(), y2, None, (rm, rv), None, eps, mom, 0.0, relu, True, res)
    ref2 = F.batch_norm(x.detach(), rm, rv, gamma.detach(), beta.detach(), training=False, eps=eps)
    ref2 = ref2 + res if with_res else ref2
    assert torch.allclose(y2, torch.relu(ref2) if relu else ref2, atol=1e-4, rtol=1e-4)


@settings(max_examples=25, deadline=None)
@given(N=st_.integers(1, 3), H=st_.integers(3, 12), W=st_.integers(3, 12), C=st_.integers(1, 6),
       k=st_.sampled_from([2, 3]), s=st_.sampled_from([1, 2]), seed=st_.integers(0, 10_000))
def test_general_maxpool_gap_and_add_cpu(N, H, W, C, k, s, seed):
    g = torch.Generator().manual_seed(seed)
    p = k // 2 if k == 3 else 0
    x = torch.randn((N, H, W, C), generator=g).requires_grad_()
    ref = F.max_pool2d(x.permute(0, 3, 1, 2), k, s, p).permute(0, 2, 3, 1).contiguous()
    y, code = torch.empty_like(ref), torch.empty(ref.shape, dtype=torch.uint8)
    cpu.maxpool_fwd(x.detach(), y, code, k, s, p)
    assert torch.equal(y, ref)
    dy = torch.randn(ref.shape, generator=g)
    ref.backward(dy)
    dx = torch.empty_like(x)
    cpu.maxpool_bwd(dy, code, dx, k, s, p)
    assert torch.allclose(dx, x.grad, atol=1e-6)
    # global average pooling with the consumer's ReLU mask
    m = torch.randn((N, H, W, C), generator=g)
    ya = torch.empty((N, C))
    cpu.gap_fwd(x.detach(), ya)
    assert torch.allclose(ya, x.detach().mean((1, 2)), atol=1e-6)
    da = torch.empty_like(x)
    cpu.gap_bwd(ya, da, m)
    assert torch.allclose(da, (ya / (H * W))[:, None, None, :] * (m > 0), atol=1e-7)
    # add modes
    a, b, out = torch.randn(x.shape, generator=g), torch.randn(x.shape, generator=g), torch.empty_like(x)
    cpu.add_act(a, b, out, 1)
    assert torch.equal(out, torch.relu(a + b))
    cpu.add_act(a, b, out, 2, m)
    assert torch.equal(out, (a + b) * (m > 0))
    cpu.add_act(a, None, out, 2, m)
    assert torch.equal(out, a * (m > 0))


def test_conv_tf_same_padding_with_explicit_output_size():
    """TensorFlow 'same' with stride 2 on an even input pads (0 top-left, 1 bottom-right) for a 3x3 kernel:
    expressed by the explicit output size, not by symmetric padding."""
    g = torch.Generator().manual_seed(5)
    x = torch.randn((2, 8, 8, 3), generator=g).requires_grad_()
    w = (torch.randn((3, 3, 3, 4), generator=g) * 0.3).requires_grad_()
    xp = F.pad(x.permute(0, 3, 1, 2), (0, 1, 0, 1))
    ref = F.conv2d(xp, w.permute(3, 2, 0, 1), None, stride=2).permute(0, 2, 3, 1).contiguous()
    assert ref.shape[1:3] == (4, 4)
    y = torch.empty_like(ref)
    cpu.conv2d_fwd(x.detach(), w.detach(), None, y, (2, 2), (0, 0), relu=False)
    assert torch.allclose(y, ref, atol=1e-5)
    dy = torch.randn(ref.shape, generator=g)
    ref.backward(dy)
    dx, dw = torch.empty_like(x), torch.zeros_like(w)
    cpu.conv2d_dgrad(dy, w.detach(), dx, (2, 2), (0, 0))
    cpu.conv2d_wgrad(x.detach(), dy, dw, None, (2, 2), (0, 0))
    assert torch.allclose(dx, x.grad, atol=1e-5) and torch.allclose(dw, w.grad, atol=1e-4)
