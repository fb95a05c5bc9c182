"""Implicit-GEMM convolution on tcgen05 (forward, dgrad, wgrad) vs PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

SHAPES = [
    # N, H, W, Ci, Co, K, pad
    (4, 13, 13, 64, 128, 3, 0),      # scaled-CNN conv2 (BASELINE.json config 3)
    (130, 13, 13, 64, 128, 3, 0),
    (3, 14, 14, 32, 64, 3, 1),       # 'same' padding: zero fill by TMA
    (5, 7, 7, 64, 64, 1, 0),         # 1x1, several images per M tile
    (2, 56, 56, 64, 64, 3, 1),       # ResNet-style 56x56
    (2, 9, 20, 96, 36, 5, 2),
]


def _ref(x, w, b, pad):
    return F.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), b, padding=pad).permute(0, 2, 3, 1).contiguous()


@pytest.mark.parametrize("N,H,W,Ci,Co,K,pad", SHAPES)
def test_conv_forward_dgrad_wgrad(N, H, W, Ci, Co, K, pad):
    from distributed_b200.ops.conv import ConvDgradPlan, ConvFwdPlan, ConvWgradPlan, tc_supported
    g = torch.Generator(device="cuda").manual_seed(N * H + Ci)
    x = (torch.randn((N, H, W, Ci), device="cuda", generator=g) * 0.5).requires_grad_()
    w = (torch.randn((K, K, Ci, Co), device="cuda", generator=g) * 0.1).requires_grad_()
    b = torch.randn((Co,), device="cuda", generator=g).requires_grad_()
    assert tc_supported(x.shape, w.shape, (pad, pad), torch.float32)
    ref = torch.relu(_ref(x, w, b, pad))
    y = torch.full(ref.shape, float("nan"), device="cuda")
    ConvFwdPlan(x.detach(), w.detach(), y, (pad, pad), bias=b.detach(), relu=True).run()
    tol = 3e-3 * ref.abs().max().item()
    assert (y - ref).abs().max().item() < tol
    dy = torch.randn(ref.shape, device="cuda", generator=g)
    ref.backward(dy)
    dpre = (dy * (ref > 0)).contiguous()
    if tc_supported(x.shape, w.shape, (pad, pad), torch.float32, dgrad=True):
        dx = torch.full(x.shape, float("nan"), device="cuda")
        ConvDgradPlan(dpre, w.detach(), dx, (pad, pad)).run()
        assert (dx - x.grad).abs().max().item() < 3e-3 * x.grad.abs().max().item()
        mask = torch.randn(x.shape, device="cuda", generator=g)
        ConvDgradPlan(dpre, w.detach(), dx, (pad, pad), mask=mask).run()
        assert (dx - x.grad * (mask > 0)).abs().max().item() < 3e-3 * x.grad.abs().max().item()
    dw = torch.ones_like(w)
    ConvWgradPlan(x.detach(), dpre, dw, (pad, pad)).run()
    assert (dw - 1 - w.grad).abs().max().item() < 3e-3 * w.grad.abs().max().item()


@pytest.mark.parametrize("N,H,W,Ci,Co,K,stride", [(4, 56, 56, 64, 128, 1, 2), (3, 28, 28, 128, 128, 3, 2),
                                                  (2, 14, 14, 256, 64, 3, 2), (2, 15, 15, 32, 64, 3, 2)])
def test_strided_conv_forward_and_wgrad(N, H, W, Ci, Co, K, stride):
    """TensorFlow 'same' padding with stride 2: top/left pad = total // 2, the rest at the end."""
    import math
    from distributed_b200.ops.conv import ConvFwdPlan, ConvWgradPlan, tc_supported
    g = torch.Generator(device="cuda").manual_seed(N + H + Co)
    x = (torch.randn((N, H, W, Ci), device="cuda", generator=g) * 0.5).requires_grad_()
    w = (torch.randn((K, K, Ci, Co), device="cuda", generator=g) * 0.1).requires_grad_()
    OH = math.ceil(H / stride)
    total = max((OH - 1) * stride + K - H, 0)
    p0 = total // 2
    xp = F.pad(x.permute(0, 3, 1, 2), (p0, total - p0, p0, total - p0))
    ref = F.conv2d(xp, w.permute(3, 2, 0, 1), None, stride=stride).permute(0, 2, 3, 1).contiguous()
    assert ref.shape[1] == OH
    assert tc_supported(x.shape, w.shape, (p0, p0), torch.float32, strides=(stride, stride), out_hw=(OH, OH))
    y = torch.full(ref.shape, float("nan"), device="cuda")
    ConvFwdPlan(x.detach(), w.detach(), y, (p0, p0), stride=stride).run()
    assert (y - ref).abs().max().item() < 3e-3 * ref.abs().max().item()
    dy = torch.randn(ref.shape, device="cuda", generator=g)
    ref.backward(dy)
    dw = torch.zeros_like(w)
    ConvWgradPlan(x.detach(), dy, dw, (p0, p0), stride=stride).run()
    assert (dw - w.grad).abs().max().item() < 3e-3 * w.grad.abs().max().item()
    from distributed_b200.ops.conv import ConvDgradStridedPlan, dgrad_strided_supported
    assert dgrad_strided_supported(x.shape, w.shape, stride, torch.float32)
    dx = torch.full(x.shape, float("nan"), device="cuda")
    ConvDgradStridedPlan(dy, w.detach(), dx, (p0, p0), stride).run()
    assert (dx - x.grad).abs().max().item() < 3e-3 * x.grad.abs().max().item()
    mask = torch.randn(x.shape, device="cuda", generator=g)
    ConvDgradStridedPlan(dy, w.detach(), dx, (p0, p0), stride, mask=mask).run()
    assert (dx - x.grad * (mask > 0)).abs().max().item() < 3e-3 * x.grad.abs().max().item()


# ---- bf16 storage (mixed_bfloat16 policy): bf16 operands, fp32 TMEM accumulation -------------------------
def _bf(t):
    return t.to(torch.bfloat16)


BF16_SHAPES = [
    # N, H, W, Ci, Co, K, pad
    (4, 56, 56, 64, 64, 1, 0),        # ResNet-50 stage-1 1x1
    (2, 56, 56, 64, 64, 3, 1),        # ResNet-50 stage-1 3x3
    (3, 14, 14, 256, 1024, 1, 0),
    (5, 7, 7, 512, 512, 3, 1),        # several images per M tile
    (2, 13, 13, 64, 128, 3, 0),
    (32, 56, 56, 64, 256, 1, 0),      # 896 x 2 tiles on <= 296 persistent CTAs
]


@pytest.mark.parametrize("N,H,W,Ci,Co,K,pad", BF16_SHAPES)
def test_conv_bf16_forward_dgrad_wgrad(N, H, W, Ci, Co, K, pad):
    from distributed_b200.ops.conv import ConvDgradPlan, ConvFwdPlan, ConvWgradPlan, tc_supported
    g = torch.Generator(device="cuda").manual_seed(N * H + Ci + 1)
    x = _bf(torch.randn((N, H, W, Ci), device="cuda", generator=g) * 0.5).float().requires_grad_()
    w = _bf(torch.randn((K, K, Ci, Co), device="cuda", generator=g) * 0.1).float().requires_grad_()
    b = torch.randn((Co,), device="cuda", generator=g)
    assert tc_supported(x.shape, w.shape, (pad, pad), torch.bfloat16)
    ref = torch.relu(_ref(x, w, b, pad))
    y = torch.full(ref.shape, float("nan"), device="cuda", dtype=torch.bfloat16)
    ConvFwdPlan(_bf(x.detach()), _bf(w.detach()), y, (pad, pad), bias=b, relu=True).run()
    assert (y.float() - ref).abs().max().item() < 1e-2 * ref.abs().max().item()
    dy = _bf(torch.randn(ref.shape, device="cuda", generator=g)).float()
    dpre = _bf(dy * (ref > 0)).float().contiguous()
    ref.backward(dpre)            # ReLU mask already folded into dpre; zero where ref == 0 anyway
    if tc_supported(x.shape, w.shape, (pad, pad), torch.bfloat16, dgrad=True):
        dx = torch.full(x.shape, float("nan"), device="cuda", dtype=torch.bfloat16)
        ConvDgradPlan(_bf(dpre), _bf(w.detach()), dx, (pad, pad)).run()
        assert (dx.float() - x.grad).abs().max().item() < 1e-2 * x.grad.abs().max().item()
        mask = _bf(torch.randn(x.shape, device="cuda", generator=g))
        ConvDgradPlan(_bf(dpre), _bf(w.detach()), dx, (pad, pad), mask=mask).run()
        assert (dx.float() - x.grad * (mask.float() > 0)).abs().max().item() < 1e-2 * x.grad.abs().max().item()
    dw = torch.ones_like(w)       # fp32 master gradient, accumulated
    ConvWgradPlan(_bf(x.detach()), _bf(dpre), dw, (pad, pad)).run()
    assert (dw - 1 - w.grad).abs().max().item() < 5e-3 * w.grad.abs().max().item()


@pytest.mark.parametrize("N,H,W,Ci,Co,K,stride", [(4, 56, 56, 64, 128, 1, 2), (3, 28, 28, 128, 128, 3, 2),
                                                  (2, 14, 14, 256, 64, 3, 2)])
def test_strided_conv_bf16(N, H, W, Ci, Co, K, stride):
    import math
    from distributed_b200.ops.conv import (ConvDgradStridedPlan, ConvFwdPlan, ConvWgradPlan, dgrad_strided_supported,
                                           tc_supported)
    g = torch.Generator(device="cuda").manual_seed(N + H + Co + 1)
    x = _bf(torch.randn((N, H, W, Ci), device="cuda", generator=g) * 0.5).float().requires_grad_()
    w = _bf(torch.randn((K, K, Ci, Co), device="cuda", generator=g) * 0.1).float().requires_grad_()
    OH = math.ceil(H / stride)
    total = max((OH - 1) * stride + K - H, 0)
    p0 = total // 2
    xp = F.pad(x.permute(0, 3, 1, 2), (p0, total - p0, p0, total - p0))
    ref = F.conv2d(xp, w.permute(3, 2, 0, 1), None, stride=stride).permute(0, 2, 3, 1).contiguous()
    assert tc_supported(x.shape, w.shape, (p0, p0), torch.bfloat16, strides=(stride, stride), out_hw=(OH, OH))
    y = torch.full(ref.shape, float("nan"), device="cuda", dtype=torch.bfloat16)
    ConvFwdPlan(_bf(x.detach()), _bf(w.detach()), y, (p0, p0), stride=stride).run()
    assert (y.float() - ref).abs().max().item() < 1e-2 * ref.abs().max().item()
    dy = _bf(torch.randn(ref.shape, device="cuda", generator=g)).float()
    ref.backward(dy)
    dw = torch.zeros_like(w)
    ConvWgradPlan(_bf(x.detach()), _bf(dy), dw, (p0, p0), stride=stride).run()
    assert (dw - w.grad).abs().max().item() < 5e-3 * w.grad.abs().max().item()
    assert dgrad_strided_supported(x.shape, w.shape, stride, torch.bfloat16)
    dx = torch.full(x.shape, float("nan"), device="cuda", dtype=torch.bfloat16)
    mask = _bf(torch.randn(x.shape, device="cuda", generator=g))
    ConvDgradStridedPlan(_bf(dy), _bf(w.detach()), dx, (p0, p0), stride, mask=mask).run()
    assert (dx.float() - x.grad * (mask.float() > 0)).abs().max().item() < 1e-2 * x.grad.abs().max().item()


def test_stem_conv_bf16_via_im2col():
    """3-channel 7x7/2 stem under the bf16 policy: fp32 image -> bf16 patch matrix -> bf16 GEMMs."""
    from distributed_b200.engine.cuda_backend import CudaBackend
    g = torch.Generator(device="cuda").manual_seed(6)
    x = torch.randn((8, 64, 64, 3), device="cuda", generator=g)
    params = torch.zeros(7 * 7 * 3 * 64 + 4, device="cuda")
    w = params[:7 * 7 * 3 * 64].view(7, 7, 3, 64)
    w.copy_(torch.randn((7, 7, 3, 64), device="cuda", generator=g) * 0.1)
    dy = _bf(torch.randn((8, 32, 32, 64), device="cuda", generator=g))
    be = CudaBackend("cuda:0", torch.bfloat16)
    be.begin_step(params)()
    y = be.alloc((8, 32, 32, 64), torch.bfloat16)
    dw = be.alloc((7, 7, 3, 64))
    be.conv2d_fwd(x, w, None, y, (2, 2), (2, 2), True)()
    be.conv2d_wgrad(x, dy, dw, None, (2, 2), (2, 2))()
    xq, wq = _bf(x).float().requires_grad_(), _bf(w).float().requires_grad_()
    xp = F.pad(xq.permute(0, 3, 1, 2), (2, 3, 2, 3))
    pre = F.conv2d(xp, wq.permute(3, 2, 0, 1), None, stride=2).permute(0, 2, 3, 1)
    ref = torch.relu(pre)
    assert (y.float() - ref).abs().max().item() < 1e-2 * ref.abs().max().item()
    pre.backward(dy.float())
    assert (dw - wq.grad).abs().max().item() < 5e-3 * wq.grad.abs().max().item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_dgrad_epilogue_adds_the_other_branch_gradient(dtype):
    """Residual blocks: dx = (conv_transpose(dy, w) + addend) * (mask > 0) in one kernel."""
    from distributed_b200.ops.conv import ConvDgradPlan
    g = torch.Generator(device="cuda").manual_seed(17)
    N, H, W, Ci, Co = 4, 28, 28, 256, 64
    q = (lambda t: t.to(dtype).float())
    x = torch.zeros((N, H, W, Ci), device="cuda", requires_grad=True)
    w = q(torch.randn((1, 1, Ci, Co), device="cuda", generator=g) * 0.1)
    dy = q(torch.randn((N, H, W, Co), device="cuda", generator=g))
    addend = q(torch.randn((N, H, W, Ci), device="cuda", generator=g))
    mask = q(torch.randn((N, H, W, Ci), device="cuda", generator=g))
    _ref(x, w, None, 0).backward(dy)
    ref = (x.grad + addend) * (mask > 0)
    dx = torch.full(x.shape, float("nan"), device="cuda", dtype=dtype)
    ConvDgradPlan(dy.to(dtype), w.to(dtype), dx, (0, 0), mask=mask.to(dtype), addend=addend.to(dtype)).run()
    tol = (3e-3 if dtype == torch.float32 else 1e-2) * ref.abs().max().item()
    assert (dx.float() - ref).abs().max().item() < tol
    ConvDgradPlan(dy.to(dtype), w.to(dtype), dx, (0, 0), addend=addend.to(dtype)).run()
    assert (dx.float() - (x.grad + addend)).abs().max().item() < tol
