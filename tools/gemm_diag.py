"""Verbose GEMM self-test: keeps going after a failure and prints the error structure, so
one GPU call is enough to tell a descriptor bug from a pipeline bug."""
import sys
import torch
sys.path.insert(0, ".")
from distributed_b200.ops.gemm import GemmPlan


def run(M, N, K, a_mn, b_mn, dtype, splits=0):
    g = torch.Generator(device="cuda").manual_seed(0)
    a = (torch.randn((K, M) if a_mn else (M, K), device="cuda", generator=g)).to(dtype)
    b = (torch.randn((K, N) if b_mn else (N, K), device="cuda", generator=g)).to(dtype)
    d = torch.full((M, N), float("nan"), device="cuda")
    try:
        p = GemmPlan(a, b, d, a_mn_major=a_mn, b_mn_major=b_mn, splits=splits)
        p.run()
        torch.cuda.synchronize()
    except Exception as e:  # noqa
        print(f"M={M} N={N} K={K} a_mn={a_mn} b_mn={b_mn} {dtype}: EXC {e}")
        return
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    ref = (A.double() @ B.double()).float()
    err = (d - ref).abs()
    rel = err.max().item() / ref.abs().max().item()
    geo = {f: getattr(p.geometry, f) for f, _ in p.geometry._fields_}
    status = "ok " if rel < 1e-2 else "BAD"
    print(f"{status} M={M} N={N} K={K} a_mn={int(a_mn)} b_mn={int(b_mn)} {str(dtype)[6:]} rel={rel:.3e} "
          f"nan={torch.isnan(d).sum().item()} geo={geo}")
    if rel >= 1e-2 or torch.isnan(d).any():
        bad = (err > 1e-2 * ref.abs().max()) | torch.isnan(d)
        rows = bad.any(1).nonzero().flatten()[:16].tolist()
        cols = bad.any(0).nonzero().flatten()[:16].tolist()
        print(f"    bad frac={bad.float().mean().item():.3f} first bad rows={rows} cols={cols}")
        print("    d[0,:8]  =", d[0, :8].tolist())
        print("    ref[0,:8]=", ref[0, :8].tolist())


if __name__ == "__main__":
    for dtype in (torch.float32, torch.bfloat16):
        for (a_mn, b_mn) in ((False, False), (False, True), (True, False), (True, True)):
            run(128, 64, 64, a_mn, b_mn, dtype)       # single k-block pair, single tile
            run(128, 128, 512, a_mn, b_mn, dtype)
            run(256, 256, 1024, a_mn, b_mn, dtype, splits=1)
    run(64, 64, 5408, False, True, torch.float32)
    run(4096, 4096, 4096, False, False, torch.bfloat16)
    for dtype in (torch.float32, torch.bfloat16):          # CTA-pair kernel
        for (a_mn, b_mn) in ((False, False), (False, True), (True, False), (True, True)):
            run(2344, 2200, 2048, a_mn, b_mn, dtype)
    # throughput: bf16 / tf32 square GEMMs vs torch.matmul (cuBLAS), L2 flushed between iterations
    import subprocess
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timed(fn, iters=10):
        for _ in range(3):
            fn()
        e = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
        for e0, e1 in e:
            flush.zero_()
            e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts = sorted(a.elapsed_time(b) for a, b in e)
        return ts[len(ts) // 2]

    torch.backends.cuda.matmul.allow_tf32 = True
    for dtype, M, out_dtype in ((torch.bfloat16, 8192, torch.bfloat16), (torch.bfloat16, 8192, torch.float32),
                                (torch.bfloat16, 4096, torch.bfloat16), (torch.float32, 4096, torch.float32),
                                (torch.float32, 8192, torch.float32)):
        a = torch.randn(M, M, device="cuda").to(dtype); b = torch.randn(M, M, device="cuda").to(dtype)
        d = torch.empty(M, M, device="cuda", dtype=out_dtype)
        p = GemmPlan(a, b, d)
        ms = timed(p.run)
        ms_lib = timed(lambda: torch.matmul(a, b.t()))
        clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,clocks_throttle_reasons.active",
                              "--format=csv,noheader", "-i", "0"], capture_output=True, text=True).stdout.strip()
        geo = {f: getattr(p.geometry, f) for f, _ in p.geometry._fields_}
        print(f"perf {str(dtype)[6:]}->{str(out_dtype)[6:]} {M}^3: {ms:.3f} ms {2*M**3/ms/1e9:.1f} TFLOP/s | "
              f"cuBLAS {ms_lib:.3f} ms {2*M**3/ms_lib/1e9:.1f} TFLOP/s | ratio {ms_lib/ms:.3f} | pair={geo['pair']} "
              f"bn={geo['block_n']} stages={geo['stages']} ctas={geo['ctas']} | clocks {clk}")
