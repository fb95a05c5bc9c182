"""One GEMM configuration, a few launches: the target of `ncu --set full -k regex:gemm_tc` captures.
usage: python tools/gemm_one.py [bf16|tf32] [M] [out: bf16|f32]"""
import sys
import torch
sys.path.insert(0, ".")
from distributed_b200.ops.gemm import GemmPlan

kind = sys.argv[1] if len(sys.argv) > 1 else "bf16"
M = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
out = sys.argv[3] if len(sys.argv) > 3 else ("bf16" if kind == "bf16" else "f32")
dt = torch.bfloat16 if kind == "bf16" else torch.float32
a = torch.randn(M, M, device="cuda").to(dt)
b = torch.randn(M, M, device="cuda").to(dt)
d = torch.empty(M, M, device="cuda", dtype=torch.bfloat16 if out == "bf16" else torch.float32)
p = GemmPlan(a, b, d)
for _ in range(6):
    p.run()
torch.cuda.synchronize()
print("geometry", {f: getattr(p.geometry, f) for f, _ in p.geometry._fields_})
