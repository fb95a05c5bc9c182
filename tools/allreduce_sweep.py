"""Gradient all-reduce(+SGD) sweep 4 KB - 1 GB vs NCCL (BASELINE.json config 5).

    python -m torch.distributed.run --nproc-per-node N tools/allreduce_sweep.py [--max-mb 1024]

bus GB/s = 2 (N-1)/N * S / t, device-timed (CUDA events), max over ranks.  The fused kernel
does strictly more work than NCCL's all-reduce (it also averages, applies SGD and zeroes the
gradients); the NCCL arm gets a separate fused `w -= lr*g; g = 0` kernel for the like-for-like row.
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    nccl = dist.new_group(backend="nccl")
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.ops import nn
    from distributed_b200.parallel import symm
    from distributed_b200.parallel.strategy import CollectiveCommunication

    class S:      # minimal strategy stand-in
        pass
    s = S()
    s.rank, s.num_workers, s.communication = rank, world, CollectiveCommunication.AUTO
    max_floats = args.max_mb * (1 << 20) // 4
    comm = symm.SymmComm(s, f"sweep-{os.environ.get('MASTER_PORT', '0')}", max_floats + 4, f"cuda:{local}")
    lib = nl.lib()
    w = comm.alloc(comm.n, None)
    g = comm.alloc(comm.n, None)
    gn = torch.zeros(max_floats + 4, device="cuda")
    wn = torch.zeros(max_floats + 4, device="cuda")
    sizes = [1 << k for k in range(10, 31) if (1 << k) // 4 <= max_floats]      # floats*4 bytes: 4 KB ... 1 GB
    sizes = [b for b in sizes if b >= 4096]
    rows = []
    st = torch.cuda.current_stream().cuda_stream

    def timed(fn, iters):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]) * 1e-3

    for nbytes in sizes:
        n = nbytes // 4
        row = {"bytes": nbytes}
        iters = args.iters if nbytes <= (64 << 20) else max(4, args.iters // 5)
        algos = [("oneshot", 0), ("twoshot", 1), ("pull", 3)] + ([("nvls", 2)] if comm.mc_base else [])
        picked = int(lib.b2d_allreduce_pick_algo(n + 4, world, 1 if comm.mc_base else 0))
        row["picked"] = {0: "oneshot", 1: "twoshot", 2: "nvls", 3: "pull"}[picked]
        for name, algo in algos:
            if algo in (0, 3) and nbytes > (16 << 20) and algo != picked:
                continue                      # latency variants: pointless (and slow) far above their range
            d = symm.AllreduceDesc()
            for p in range(world):
                d.w[p], d.g[p] = comm.base[p] + comm.off_w, comm.base[p] + comm.off_g
                d.sig[p], d.epoch[p] = comm.base[p] + comm.off_sig, comm.base[p] + comm.off_epoch
            if comm.mc_base:
                d.w_mc, d.g_mc = comm.mc_base + comm.off_w, comm.mc_base + comm.off_g
            d.n_params, d.n_total, d.lr = n, n + 4, 1e-3
            d.rank, d.world, d.algo, d.emulate, d.plain, d.blocks = rank, world, algo, 0, 0, 0
            if algo == picked:
                # correctness of the dispatched variant at this size: w -= lr/N * sum(g) against NCCL's sum
                gen = torch.Generator(device="cuda").manual_seed(1000 * rank + n % 9973)
                src = torch.randn(min(n, 1 << 22), device="cuda", generator=gen)
                m = src.numel()
                g[:m].copy_(src); g[m:n + 4].zero_(); w[:n + 4].zero_()
                ref = src.clone()
                dist.all_reduce(ref, group=nccl)
                comm.barrier()
                nl.check(lib.b2d_allreduce_sgd(C.byref(d), st), "ar")
                comm.barrier()
                err = float((w[:m] + (1e-3 / world) * ref).abs().max() / ((1e-3 / world) * ref.abs().max()))
                row["picked_rel_err_vs_nccl"] = err
                assert err < 1e-4, (nbytes, name, err)
                assert float(g[:n + 4].abs().max()) == 0.0, "gradient arena not zeroed"
            t = timed(lambda: nl.check(lib.b2d_allreduce_sgd(C.byref(d), st), "ar"), iters)
            row[name + "_us"] = t * 1e6
            row[name + "_busbw"] = 2 * (world - 1) / world * nbytes / t / 1e9
        t = timed(lambda: dist.all_reduce(gn[:n], group=nccl), iters)
        row["nccl_us"], row["nccl_busbw"] = t * 1e6, 2 * (world - 1) / world * nbytes / t / 1e9

        def nccl_sgd():
            dist.all_reduce(gn[:n + 4], group=nccl)
            nn.sgd_apply(wn[:n + 4], gn[:n + 4], n, 1e-3 / world)
        t = timed(nccl_sgd, iters)
        row["nccl_sgd_us"], row["nccl_sgd_busbw"] = t * 1e6, 2 * (world - 1) / world * nbytes / t / 1e9
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0:
        print(json.dumps({"world": world, "multicast": bool(comm.mc_base), "rows": len(rows)}))
    comm.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
