import sys, os, torch
sys.path.insert(0, ".")
from distributed_b200.ops.gemm import GemmPlan
g = torch.Generator(device="cuda").manual_seed(0)
for (M, N, K, bmn) in ((64, 64, 5408, True), (512, 512, 512, False)):
    a = torch.rand((M, K), device="cuda", generator=g)         # positive values: bias shows up
    b = torch.rand((K, N) if bmn else (N, K), device="cuda", generator=g)
    d = torch.empty((M, N), device="cuda")
    GemmPlan(a, b, d, b_mn_major=bmn).run()
    ref = (a.double() @ (b.double() if bmn else b.double().t()))
    rel = ((d.double() - ref) / ref)
    print(f"truncate={os.environ.get('B2D_TF32_TRUNCATE')} M={M} K={K}: mean rel err {rel.mean().item():+.3e}  max |rel| {rel.abs().max().item():.3e}")
