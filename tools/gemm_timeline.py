"""Where do the microseconds of a small tcgen05 GEMM go?  Per-CTA SM-clock timeline."""
import ctypes as C
import sys
import torch
sys.path.insert(0, ".")
from distributed_b200.ops.gemm import GemmPlan
from distributed_b200._native import cuda_lib as nl

lib = nl.lib()
lib.b2d_gemm_plan_set_timeline.argtypes = [C.c_void_p, C.c_void_p]
lib.b2d_gemm_plan_set_timeline.restype = None
NAMES = ["start", "prologue", "pdl_wait", "tiles_landed", "mma_issued", "acc_ready", "epi_done", "end", "ld0", "st0", "ld1", "st1"]


def run(name, a, b, d, **kw):
    p = GemmPlan(a, b, d, **kw)
    g = p.geometry
    n_cta = g.ctas          # persistent CTAs (<= tiles); stamps 3-5 describe each CTA's FIRST tile
    buf = torch.zeros(n_cta * 16, dtype=torch.int64, device="cuda")
    for _ in range(3):
        p.run()
    lib.b2d_gemm_plan_set_timeline(p._h, buf.data_ptr())
    p.run()
    torch.cuda.synchronize()
    first = buf.view(n_cta, 16).cpu().double().clone()
    p.run(); p.run(); p.run()          # three back-to-back launches: the buffer keeps the last one
    torch.cuda.synchronize()
    lib.b2d_gemm_plan_set_timeline(p._h, None)
    t = buf.view(n_cta, 16).cpu().double()
    rel = (t - t[:, :1]) / 1e3          # per-CTA, us since its own start (globaltimer, ns)
    print(f"{name}: ctas={n_cta} stages={g.stages} smem={g.smem_bytes}")
    print("   median us since CTA start: " + "  ".join(f"{n}={rel[:, i].median().item():.2f}" for i, n in enumerate(NAMES)))
    t0 = t[:, 0].min()
    print(f"   grid: first CTA start=0  last CTA start={(t[:, 0].max() - t0) / 1e3:.2f}  first end={(t[:, 7].min() - t0) / 1e3:.2f}"
          f"  last end={(t[:, 7].max() - t0) / 1e3:.2f} us;  pdl_wait exit spread {(t[:, 2].min() - t0) / 1e3:.2f}..{(t[:, 2].max() - t0) / 1e3:.2f}")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        p.run()
    e1.record(); torch.cuda.synchronize()
    print(f"   back-to-back: {e0.elapsed_time(e1) / 50 * 1e3:.2f} us/launch")


g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn(64, 5408, device="cuda", generator=g); w1 = torch.randn(5408, 64, device="cuda", generator=g)
b1 = torch.randn(64, device="cuda", generator=g); h = torch.empty(64, 64, device="cuda")
dh = torch.randn(64, 64, device="cuda", generator=g); dw1 = torch.zeros(5408, 64, device="cuda")
dx = torch.empty(64, 5408, device="cuda")
run("dense64 fwd  (64x64x5408 split-K)", x, w1, h, b_mn_major=True, bias=b1, relu=True)
run("dense64 wgrad(5408x64x64)", x, dh, dw1, a_mn_major=True, b_mn_major=True, accumulate=True)
run("dense64 dgrad(64x5408x64)", dh, w1, dx)
