"""Floor of a dependent chain of N tiny kernels inside a CUDA graph, with and without PDL."""
import os
import sys
import torch
sys.path.insert(0, ".")
from distributed_b200.ops import nn

x = torch.zeros(256, device="cuda")
y = torch.zeros(256, device="cuda")


def chain(n):
    for _ in range(n):
        nn.relu_bwd(x[:0], y[:0], x[:0])      # n = 0 elements: one block, wait/trigger, exit


for n in (1, 7, 14):
    chain(n)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        chain(n)
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 200 * 1e3
    print(f"PDL={os.environ.get('B2D_PDL', '1')} graph of {n:2d} empty kernels: {us:7.2f} us/replay  ({us / n:.2f} us/kernel)")
