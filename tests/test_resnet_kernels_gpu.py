"""ResNet-family kernels (BatchNorm+ReLU, residual add, general max-pool, global average pool,
casts, bias-gradient column sums) vs PyTorch fp32 references, in fp32 and bf16 storage."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DTYPES = [torch.float32, torch.bfloat16]


def _g(seed):
    return torch.Generator(device="cuda").manual_seed(seed)


def _q(t, dtype):
    """Round to the storage type and come back: the values the kernel actually sees."""
    return t.to(dtype).float()


def _tol(dtype, ref):
    return (1e-4 if dtype == torch.float32 else 1.2e-2) * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("M,C,relu", [(64 * 49, 64, True), (3000, 256, False), (50, 24, True), (1568, 2048, True)])
def test_batchnorm_train_forward_backward(M, C, relu, dtype):
    from distributed_b200.ops import nn
    g = _g(M + C)
    x = _q(torch.randn((M, C), device="cuda", generator=g) * 2 + 1, dtype).requires_grad_()
    gamma = (torch.rand(C, device="cuda", generator=g) + 0.5).requires_grad_()
    beta = torch.randn(C, device="cuda", generator=g).requires_grad_()
    eps, mom = 1e-3, 0.9
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    ref = F.batch_norm(x, rm, rv, gamma, beta, training=True, momentum=1 - mom, eps=eps)
    if relu:
        ref = torch.relu(ref)
    stats, saved, red = (torch.zeros(2 * C, device="cuda") for _ in range(3))
    red.fill_(7.0)                               # the forward pass must clear it for the backward pass
    mm, mv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    xs = x.detach().to(dtype)
    y, dx = torch.empty_like(xs), torch.empty_like(xs)
    dy = _q(torch.randn((M, C), device="cuda", generator=g), dtype)
    dy_buf = torch.empty_like(xs)
    dgamma, dbeta = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    bn = nn.BatchNorm(xs, y, gamma.detach(), beta.detach(), eps=eps, momentum=mom, relu=relu, stats=stats,
                      saved=saved, red=red, moving=(mm, mv), dy2d=dy_buf, dx2d=dx, dgamma=dgamma, dbeta=dbeta)
    bn.forward()
    assert (y.float() - ref).abs().max().item() < 2 * _tol(dtype, ref)
    assert torch.allclose(mm, rm, atol=1e-4) and torch.allclose(mv, rv, atol=1e-3)      # in-place moving update
    assert red.abs().max().item() == 0
    if relu:    # borderline outputs may round to the other side of zero: take them out of the comparison
        dy = dy * ((y.float() > 0) == (ref > 0))
    dy_buf.copy_(dy)
    ref.backward(dy)
    bn.backward()
    assert stats.abs().max().item() == 0         # cleared for the next forward pass
    gtol = 1e-4 if dtype == torch.float32 else 2e-2
    assert (dx.float() - x.grad).abs().max().item() < gtol * max(1.0, x.grad.abs().max().item())
    assert torch.allclose(dgamma, gamma.grad, atol=(1e-3 if dtype == torch.float32 else 2e-2) * gamma.grad.abs().max().item() + 1e-4)
    assert torch.allclose(dbeta, beta.grad, atol=(1e-3 if dtype == torch.float32 else 2e-2) * beta.grad.abs().max().item() + 1e-4)
    # second step: scratch buffers are self-cleaning
    bn.forward()
    assert (y.float() - ref).abs().max().item() < 2 * _tol(dtype, ref)


def test_batchnorm_pseudo_gradient_and_inference():
    """The moving-statistics update expressed as an optimizer step, and inference on those statistics."""
    from distributed_b200.ops import nn
    g = _g(11)
    M, C, mom, lr, eps = 2048, 64, 0.9, 0.05, 1e-3
    x = torch.randn((M, C), device="cuda", generator=g) * 1.5 + 0.3
    gamma, beta = torch.rand(C, device="cuda", generator=g) + 0.5, torch.randn(C, device="cuda", generator=g)
    mm, mv = torch.randn(C, device="cuda", generator=g), torch.rand(C, device="cuda", generator=g) + 0.5
    gm, gv = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    stats, saved, red = (torch.zeros(2 * C, device="cuda") for _ in range(3))
    y = torch.empty_like(x)
    nn.BatchNorm(x, y, gamma, beta, eps=eps, momentum=mom, relu=False, stats=stats, saved=saved, red=red,
                 moving=(mm, mv), moving_grads=(gm, gv), inv_lr=1 / lr).forward()
    rm, rv = mm.clone(), mv.clone()
    F.batch_norm(x, rm, rv, gamma, beta, training=True, momentum=1 - mom, eps=eps)
    assert torch.allclose(mm - lr * gm, rm, atol=1e-5) and torch.allclose(mv - lr * gv, rv, atol=1e-4)
    y2 = torch.empty_like(x)
    nn.BatchNorm(x, y2, gamma, beta, eps=eps, momentum=mom, relu=True, moving=(rm, rv), inference=True).forward()
    ref = torch.relu(F.batch_norm(x, rm, rv, gamma, beta, training=False, eps=eps))
    assert torch.allclose(y2, ref, atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("shape,k,s,p", [((2, 112, 112, 64), 3, 2, 1), ((3, 9, 11, 8), 3, 2, 1), ((2, 8, 8, 8), 2, 2, 0)])
def test_general_maxpool(shape, k, s, p, dtype):
    from distributed_b200.ops import nn
    g = _g(sum(shape))
    x = _q(torch.randn(shape, device="cuda", generator=g), dtype).requires_grad_()
    ref = F.max_pool2d(x.permute(0, 3, 1, 2), k, s, p).permute(0, 2, 3, 1).contiguous()
    y = torch.empty(ref.shape, device="cuda", dtype=dtype)
    code = torch.empty(ref.shape, device="cuda", dtype=torch.uint8)
    nn.maxpool_fwd(x.detach().to(dtype), y, code, k, s, p)
    assert torch.equal(y.float(), ref)
    dy = _q(torch.randn(ref.shape, device="cuda", generator=g), dtype)
    ref.backward(dy)
    dx = torch.empty(shape, device="cuda", dtype=dtype)
    nn.maxpool_bwd(dy.to(dtype), code, dx, k, s, p)
    # ties (equal values in one window) are broken the same way by both: first maximum wins
    assert (dx.float() - x.grad).abs().max().item() < (1e-6 if dtype == torch.float32 else 1e-2 * x.grad.abs().max().item())


@pytest.mark.parametrize("dtype", DTYPES)
def test_add_modes_and_global_average_pool(dtype):
    from distributed_b200.ops import nn
    g = _g(3)
    a = _q(torch.randn((4, 7, 7, 64), device="cuda", generator=g), dtype)
    b = _q(torch.randn((4, 7, 7, 64), device="cuda", generator=g), dtype)
    m = _q(torch.randn((4, 7, 7, 64), device="cuda", generator=g), dtype)
    A, B, Mk = a.to(dtype), b.to(dtype), m.to(dtype)
    out = torch.empty_like(A)
    nn.add_act(A, B, out, 1)
    assert torch.equal(out, torch.relu(a + b).to(dtype))
    nn.add_act(A, B, out, 0)
    assert torch.equal(out, (a + b).to(dtype))
    nn.add_act(A, B, out, 2, Mk)
    assert torch.equal(out, ((a + b) * (m > 0)).to(dtype))
    nn.add_act(A, None, out, 2, Mk)              # stand-alone ReLU backward
    assert torch.equal(out, (a * (m > 0)).to(dtype))
    y = torch.empty((4, 64), device="cuda")
    nn.gap_fwd(A, y)
    assert torch.allclose(y, a.mean((1, 2)), atol=1e-5)
    dx = torch.empty_like(A)
    nn.gap_bwd(y, dx)
    ref = (y / 49)[:, None, None, :].expand_as(a)
    assert torch.allclose(dx.float(), ref, atol=1e-7 if dtype == torch.float32 else 1e-2 * ref.abs().max().item())
    nn.gap_bwd(y, dx, Mk)
    assert torch.allclose(dx.float(), ref * (m > 0), atol=1e-7 if dtype == torch.float32 else 1e-2 * ref.abs().max().item())


def test_cast_and_colsum():
    from distributed_b200.ops import nn
    g = _g(5)
    for n in (8, 1000, 4099, 1 << 20):
        src = torch.randn(n + 8, device="cuda", generator=g)[:n]
        dst = torch.empty(n, device="cuda", dtype=torch.bfloat16)
        nn.cast(src, dst)
        assert torch.equal(dst, src.to(torch.bfloat16)), n
        back = torch.empty(n, device="cuda")
        nn.cast(dst, back)
        assert torch.equal(back, dst.float()), n
    for dtype in DTYPES:
        for M, C in ((5000, 64), (77, 24), (1568, 512)):
            dy = _q(torch.randn((M, C), device="cuda", generator=g), dtype)
            db = torch.ones(C, device="cuda")
            nn.colsum_t(dy.to(dtype), db)
            ref = dy.double().sum(0).float() + 1
            assert torch.allclose(db, ref, atol=1e-3 * max(1.0, ref.abs().max().item())), (dtype, M, C)


def test_im2col_bf16_matches_unfold():
    from distributed_b200.ops import nn
    g = _g(9)
    x = torch.randn((2, 20, 20, 3), device="cuda", generator=g)
    col = torch.full((2 * 10 * 10, 152), float("nan"), device="cuda", dtype=torch.bfloat16)
    nn.im2col(x, col, (7, 7), (2, 2), (2, 2), (10, 10))
    xp = F.pad(x.permute(0, 3, 1, 2), (2, 3, 2, 3))
    ref = F.unfold(xp, 7, stride=2)                                  # [N, C*49, L], channel-major patches
    ref = ref.view(2, 3, 49, 100).permute(0, 3, 2, 1).reshape(200, 147)      # -> (kh, kw, ci) order
    assert torch.equal(col[:, :147], ref.to(torch.bfloat16))
    assert col[:, 147:].abs().max().item() == 0


@pytest.mark.parametrize("dtype", DTYPES)
def test_batchnorm_with_residual_tail(dtype):
    """y = relu(bn(x) + residual) in one kernel; its backward must not mask (dy arrives masked)."""
    from distributed_b200.ops import nn
    g = _g(21)
    M, C, eps, mom = 3136, 256, 1e-3, 0.9
    x = _q(torch.randn((M, C), device="cuda", generator=g) * 2 + 1, dtype).requires_grad_()
    res = _q(torch.randn((M, C), device="cuda", generator=g), dtype)
    gamma = (torch.rand(C, device="cuda", generator=g) + 0.5).requires_grad_()
    beta = torch.randn(C, device="cuda", generator=g).requires_grad_()
    pre = F.batch_norm(x, None, None, gamma, beta, training=True, eps=eps)
    ref = torch.relu(pre + res)
    stats, saved, red = (torch.zeros(2 * C, device="cuda") for _ in range(3))
    xs = x.detach().to(dtype)
    y, dx, dy_buf = torch.empty_like(xs), torch.empty_like(xs), torch.empty_like(xs)
    dgamma, dbeta = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    bn = nn.BatchNorm(xs, y, gamma.detach(), beta.detach(), eps=eps, momentum=mom, relu=True, stats=stats, saved=saved,
                      red=red, residual=res.to(dtype), dy2d=dy_buf, dx2d=dx, dgamma=dgamma, dbeta=dbeta)
    bn.forward()
    assert (y.float() - ref).abs().max().item() < 2 * _tol(dtype, ref)
    dz = _q(torch.randn((M, C), device="cuda", generator=g), dtype)       # gradient of the pre-activation sum
    dy_buf.copy_(dz)
    pre.backward(dz)
    bn.backward()
    gtol = 1e-4 if dtype == torch.float32 else 2e-2
    assert (dx.float() - x.grad).abs().max().item() < gtol * max(1.0, x.grad.abs().max().item())
    assert torch.allclose(dgamma, gamma.grad, atol=(1e-3 if dtype == torch.float32 else 2e-2) * gamma.grad.abs().max().item())
