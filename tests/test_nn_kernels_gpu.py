"""Non-GEMM sm_100a kernels vs plain PyTorch fp32 references of the same ops."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _gen(seed):
    return torch.Generator(device="cuda").manual_seed(seed)


def _randn(shape, g, scale=1.0):
    return torch.randn(shape, device="cuda", generator=g) * scale


def _ref_conv(x_nhwc, w_hwio, b, stride=(1, 1), padding=(0, 0)):
    y = F.conv2d(x_nhwc.permute(0, 3, 1, 2), w_hwio.permute(3, 2, 0, 1), b, stride=stride,
                 padding=padding)
    return y.permute(0, 2, 3, 1).contiguous()


@pytest.mark.parametrize("N,H,W,Co,KS", [(64, 28, 28, 32, 3), (5, 28, 28, 64, 3), (300, 28, 28, 32, 3),
                                         (3, 17, 21, 48, 5)])
def test_conv_c1_pool_forward_and_backward(N, H, W, Co, KS):
    from distributed_b200.ops import nn
    g = _gen(N + Co)
    x = torch.rand((N, H, W, 1), device="cuda", generator=g)
    w = _randn((KS, KS, 1, Co), g, 0.3).requires_grad_()
    b = _randn((Co,), g, 0.1).requires_grad_()
    PH, PW = (H - KS + 1) // 2, (W - KS + 1) // 2
    y = torch.empty((N, PH, PW, Co), device="cuda")
    code = torch.empty((N, PH, PW, Co), device="cuda", dtype=torch.uint8)
    nn.conv_c1_pool_fwd(x, w.detach(), b.detach(), y, code)
    conv = torch.relu(_ref_conv(x, w, b))
    ref = F.max_pool2d(conv.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert torch.allclose(y, ref, atol=1e-4, rtol=1e-4)
    dy = _randn(ref.shape, g)
    ref.backward(dy)
    dw = torch.zeros_like(w)
    db = torch.zeros_like(b)
    nn.conv_c1_pool_bwd(x, dy, code, dw, db)
    assert torch.allclose(dw, w.grad, atol=2e-3 * w.grad.abs().max().item(), rtol=1e-3)
    assert torch.allclose(db, b.grad, atol=2e-3 * b.grad.abs().max().item(), rtol=1e-3)


@pytest.mark.parametrize("B,K,C,relu_mask", [(64, 64, 10, True), (1024, 1024, 10, True), (37, 50, 7, False), (5000, 256, 32, True),
                                              (1000, 512, 16, True), (300, 96, 3, False)])   # wide-batch variant, ragged
def test_dense_xent_head_matches_autograd(B, K, C, relu_mask):
    from distributed_b200.ops import nn
    g = _gen(B + K)
    pre = _randn((B, K), g).requires_grad_()
    h = torch.relu(pre) if relu_mask else pre
    w = _randn((K, C), g, 0.2).requires_grad_()
    b = _randn((C,), g, 0.1).requires_grad_()
    labels = torch.randint(0, C, (B,), device="cuda", generator=g, dtype=torch.int32)
    logits_ref = h @ w + b
    loss = F.cross_entropy(logits_ref, labels.long(), reduction="sum")
    (loss / B).backward()
    dh = torch.empty((B, K), device="cuda")
    dw, db = torch.zeros_like(w), torch.zeros_like(b)
    metrics = torch.zeros(4, device="cuda")
    logits = torch.empty((B, C), device="cuda")
    dprev = torch.ones(K, device="cuda")
    nn.dense_xent_head(h.detach().contiguous(), w.detach(), b.detach(), labels, dh, dw, db, metrics,
                       1.0 / B, relu_mask, logits, dh_colsum=dprev)
    assert torch.allclose(dprev, pre.grad.sum(0) + 1, atol=1e-5, rtol=1e-3)
    assert torch.allclose(logits, logits_ref, atol=1e-3, rtol=1e-3)
    assert abs(metrics[0].item() - loss.item()) < 1e-3 * abs(loss.item())
    assert metrics[1].item() == (logits_ref.argmax(1) == labels).sum().item()
    assert metrics[2].item() == B
    assert torch.allclose(dw, w.grad, atol=1e-4, rtol=1e-3)
    assert torch.allclose(db, b.grad, atol=1e-5, rtol=1e-3)
    assert torch.allclose(dh, pre.grad, atol=1e-5, rtol=1e-3)


@pytest.mark.parametrize("B,C", [(64, 10), (100, 1000), (3, 33)])
def test_softmax_xent(B, C):
    from distributed_b200.ops import nn
    g = _gen(B * C)
    z = _randn((B, C), g, 2.0).requires_grad_()
    labels = torch.randint(0, C, (B,), device="cuda", generator=g, dtype=torch.int32)
    loss = F.cross_entropy(z, labels.long(), reduction="sum")
    (loss / B).backward()
    dz = torch.empty_like(z)
    metrics = torch.zeros(4, device="cuda")
    nn.softmax_xent(z.detach(), labels, dz, metrics, 1.0 / B)
    assert torch.allclose(dz, z.grad, atol=1e-6, rtol=1e-3)
    assert abs(metrics[0].item() - loss.item()) < 1e-3 * abs(loss.item())
    assert metrics[1].item() == (z.argmax(1) == labels).sum().item()


@pytest.mark.parametrize("shape", [(4, 26, 26, 32), (3, 11, 11, 128), (2, 7, 9, 5)])
def test_maxpool2_forward_backward(shape):
    from distributed_b200.ops import nn
    g = _gen(sum(shape))
    x = _randn(shape, g).requires_grad_()
    N, H, W, Cc = shape
    y = torch.empty((N, H // 2, W // 2, Cc), device="cuda")
    code = torch.empty(y.shape, device="cuda", dtype=torch.uint8)
    nn.maxpool2_fwd(x.detach(), y, code)
    ref = F.max_pool2d(x.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert torch.equal(y, ref)
    dy = _randn(y.shape, g)
    ref.backward(dy)
    dx = torch.full(shape, float("nan"), device="cuda")
    nn.maxpool2_bwd(dy, code, y, dx, act_relu=False)
    assert torch.equal(dx, x.grad)
    # fused ReLU backward: gradient is dropped where the pooled activation is not positive
    nn.maxpool2_bwd(dy, code, y, dx, act_relu=True)
    xr = x.detach().clone().requires_grad_()
    F.max_pool2d(torch.relu(xr).permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1).backward(dy)
    assert torch.equal(dx, xr.grad)


def test_colsum_and_relu_bwd():
    from distributed_b200.ops import nn
    g = _gen(5)
    dy = _randn((1000, 70), g)
    db = torch.ones(70, device="cuda")
    nn.colsum(dy, db)
    assert torch.allclose(db, dy.sum(0) + 1, atol=1e-3)
    y = _randn((1000, 70), g)
    dx = torch.empty_like(dy)
    nn.relu_bwd(dy, y, dx)
    assert torch.equal(dx, dy * (y > 0))


@pytest.mark.parametrize("N,H,W,Ci,Co,K,stride,pad", [(4, 13, 13, 8, 16, 3, 1, 0), (2, 16, 16, 3, 8, 7, 2, 3),
                                                      (3, 9, 9, 5, 6, 1, 2, 0), (2, 8, 8, 4, 4, 3, 1, 1)])
def test_direct_conv_fallbacks(N, H, W, Ci, Co, K, stride, pad):
    from distributed_b200.ops import nn
    g = _gen(N + H + Ci)
    x = _randn((N, H, W, Ci), g).requires_grad_()
    w = _randn((K, K, Ci, Co), g, 0.2).requires_grad_()
    b = _randn((Co,), g).requires_grad_()
    ref = _ref_conv(x, w, b, (stride, stride), (pad, pad))
    y = torch.empty_like(ref)
    nn.conv2d_fwd_direct(x.detach(), w.detach(), b.detach(), y, (stride, stride), (pad, pad))
    assert torch.allclose(y, ref, atol=1e-4, rtol=1e-4)
    dy = _randn(ref.shape, g)
    ref.backward(dy)
    dx = torch.empty_like(x)
    nn.conv2d_dgrad_direct(dy, w.detach(), dx, (stride, stride), (pad, pad))
    assert torch.allclose(dx, x.grad, atol=1e-4, rtol=1e-4)
    dw = torch.zeros_like(w)
    nn.conv2d_wgrad_direct(x.detach(), dy, dw, (stride, stride), (pad, pad))
    assert torch.allclose(dw, w.grad, atol=1e-3, rtol=1e-3)


def test_sgemm_simt_and_sgd_apply():
    from distributed_b200.ops import nn
    g = _gen(9)
    a, b = _randn((70, 50), g), _randn((50, 10), g)
    bias = _randn((10,), g)
    d = torch.empty((70, 10), device="cuda")
    nn.sgemm_simt(a, b, d, bias=bias, relu=True)
    assert torch.allclose(d, torch.relu(a @ b + bias), atol=1e-4)
    d2 = torch.zeros((50, 10), device="cuda")
    nn.sgemm_simt(a, d, d2, trans_a=True, accumulate=True)
    assert torch.allclose(d2, a.t() @ d, atol=1e-3)
    d3 = torch.empty((70, 50), device="cuda")
    nn.sgemm_simt(d, b, d3, trans_b=True)
    assert torch.allclose(d3, d @ b.t(), atol=1e-3)
    w = _randn((1000 + 4,), g)
    gr = _randn((1000 + 4,), g)
    w0, g0 = w.clone(), gr.clone()
    nn.sgd_apply(w, gr, 1000, 0.1)
    assert torch.allclose(w[:1000], w0[:1000] - 0.1 * g0[:1000])
    assert torch.allclose(w[1000:], w0[1000:] + g0[1000:])
    assert gr.abs().max().item() == 0
