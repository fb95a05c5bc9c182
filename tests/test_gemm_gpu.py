"""tcgen05 GEMM core vs a plain PyTorch fp32 reference (GPU unit tier)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn, bias, relu, mask):
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    d = A.double() @ B.double()
    if bias is not None:
        d = d + bias.double()
    if relu:
        d = d.clamp_min(0)
    if mask is not None:
        d = d * (mask > 0)
    return d.float()


def _mk(shape, dtype, gen):
    return (torch.randn(shape, device="cuda", generator=gen) * 0.5).to(dtype)


CASES = [
    # M, N, K, a_mn, b_mn
    (128, 64, 32, False, False),
    (128, 64, 256, False, False),
    (64, 64, 5408, False, True),     # README dense-64 forward (x[B,in] . W[in,out]) — split-K
    (64, 5408, 64, False, False),    # README dense-64 dgrad   (dy . W^T)
    (5408, 64, 64, True, True),      # README dense-64 wgrad   (x^T . dy)
    (300, 200, 1000, False, True),
    (300, 200, 1000, True, False),
    (1024, 1024, 3200, False, True),
    (260, 36, 72, True, True),
    (8192, 2048, 256, False, True),   # 1024 tiles: persistent CTAs walk several tiles, double-buffered TMEM
    (4096, 640, 8192, False, False),  # 160 tiles, long K loop (deep ring, one CTA per SM)
    # CTA-pair kernel (tcgen05 cta_group::2, 256-row tiles): every operand layout, ragged M / N tails
    (4096, 2048, 2048, False, False),
    (2304 + 40, 2176 + 24, 2048, False, True),
    (2304 + 40, 2176 + 24, 2048, True, False),
    (2560, 2304, 2048, True, True),
    (8192, 1152, 1024, False, False),  # last wave too ragged for pair tiles: stays on the single-CTA kernel
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,N,K,a_mn,b_mn", CASES)
def test_gemm_matches_fp32_reference(M, N, K, a_mn, b_mn, dtype):
    from distributed_b200.ops.gemm import GemmPlan
    if dtype == torch.bfloat16 and (M % 8 or N % 8 or K % 8):
        pytest.skip("bf16 rows must be 16-byte aligned")
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K)
    a = _mk((K, M) if a_mn else (M, K), dtype, g)
    b = _mk((K, N) if b_mn else (N, K), dtype, g)
    d = torch.full((M, N), float("nan"), device="cuda", dtype=torch.float32)
    plan = GemmPlan(a, b, d, a_mn_major=a_mn, b_mn_major=b_mn)
    plan.run()
    plan.run()  # second launch exercises the self-cleaning split-K workspace
    torch.cuda.synchronize()
    ref = _ref(a, b, a_mn, b_mn, None, False, None)
    tol = 2e-3 if dtype == torch.float32 else 1e-2   # tf32 inputs truncate to 10 mantissa bits
    err = (d - ref).abs().max().item() / max(ref.abs().max().item(), 1e-6)
    assert err < tol, f"rel err {err} geometry={[getattr(plan.geometry, f) for f, _ in plan.geometry._fields_]}"


def test_pair_kernel_is_selected_for_long_k_and_epilogue_variants_match():
    """The CTA-pair kernel with bias + ReLU + mask, bf16 output, and accumulation, against fp64."""
    from distributed_b200.ops.gemm import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(11)
    M, N, K = 4096, 2304, 2048
    a, b = _mk((M, K), torch.bfloat16, g), _mk((N, K), torch.bfloat16, g)
    bias, mask = _mk((N,), torch.float32, g), _mk((M, N), torch.bfloat16, g)
    d = torch.empty((M, N), device="cuda", dtype=torch.bfloat16)
    plan = GemmPlan(a, b, d, bias=bias, relu=True, mask=mask)
    assert plan.geometry.pair == 1 and plan.geometry.ctas % 2 == 0
    plan.run()
    ref = _ref(a, b, False, False, bias, True, mask)
    assert (d.float() - ref).abs().max().item() < 0.02 * ref.abs().max().item()
    acc = torch.ones((M, N), device="cuda")
    a32, b32 = _mk((M, K), torch.float32, g), _mk((N, K), torch.float32, g)
    GemmPlan(a32, b32, acc, accumulate=True).run()
    ref = _ref(a32, b32, False, False, None, False, None) + 1
    assert (acc - ref).abs().max().item() < 2e-3 * ref.abs().max().item()


def test_gemm_epilogue_bias_relu_mask_and_bf16_out():
    from distributed_b200.ops.gemm import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(1)
    M, N, K = 200, 128, 512
    a, b = _mk((M, K), torch.float32, g), _mk((K, N), torch.float32, g)
    bias = _mk((N,), torch.float32, g)
    mask = _mk((M, N), torch.float32, g)
    d = torch.empty((M, N), device="cuda", dtype=torch.bfloat16)
    GemmPlan(a, b, d, b_mn_major=True, bias=bias, relu=True, mask=mask).run()
    ref = _ref(a, b, False, True, bias, True, mask)
    assert (d.float() - ref).abs().max().item() < 0.05 * ref.abs().max().item()


def test_gemm_split_k_finalize_with_bias_relu_is_repeatable():
    from distributed_b200.ops.gemm import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(2)
    M, N, K = 64, 64, 5408
    a, b = _mk((M, K), torch.float32, g), _mk((K, N), torch.float32, g)
    bias = _mk((N,), torch.float32, g)
    d = torch.empty((M, N), device="cuda")
    plan = GemmPlan(a, b, d, b_mn_major=True, bias=bias, relu=True)
    assert plan.geometry.splits > 1
    ref = _ref(a, b, False, True, bias, True, None)
    for _ in range(3):
        d.fill_(float("nan"))
        plan.run()
        assert (d - ref).abs().max().item() < 2e-3 * ref.abs().max().item()
    assert plan.workspace.abs().max().item() == 0 and plan.counters.abs().max().item() == 0


def test_gemm_accumulate_adds_into_output():
    from distributed_b200.ops.gemm import GemmPlan
    g = torch.Generator(device="cuda").manual_seed(3)
    M, N, K = 5408, 64, 64
    a, b = _mk((K, M), torch.float32, g), _mk((K, N), torch.float32, g)
    d = torch.ones((M, N), device="cuda")
    GemmPlan(a, b, d, a_mn_major=True, b_mn_major=True, accumulate=True).run()
    ref = _ref(a, b, True, True, None, False, None) + 1
    assert (d - ref).abs().max().item() < 2e-3 * ref.abs().max().item()
