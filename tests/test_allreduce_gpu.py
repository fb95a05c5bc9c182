"""Fused all-reduce + average + SGD kernels (csrc/cuda/allreduce.cu).

Single-GPU tier: all ranks are emulated inside ONE cooperative launch over per-rank buffers on
the same device (kernels of different ranks must never wait on each other across separate
launches on one GPU).  The multi-GPU tier over real NVLink peer memory is test_multigpu.py.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _emulated(world, n_params, n_tail, algo, lr=0.1, rounds=2, blocks=0):
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    lib = nl.lib()
    symm._declare(lib)
    n_total = n_params + n_tail
    g = torch.Generator(device="cuda").manual_seed(world * 1000 + n_total)
    w0 = torch.randn(n_total, device="cuda", generator=g)
    W = [w0.clone() for _ in range(world)]
    G = [torch.randn(n_total, device="cuda", generator=g) for _ in range(world)]
    sig = [torch.zeros(lib.b2d_allreduce_signal_bytes() // 4, dtype=torch.int32, device="cuda") for _ in range(world)]
    ep = [torch.zeros(lib.b2d_allreduce_epoch_bytes() // 4, dtype=torch.int32, device="cuda") for _ in range(world)]
    d = symm.AllreduceDesc()
    for p in range(world):
        d.w[p], d.g[p], d.sig[p], d.epoch[p] = W[p].data_ptr(), G[p].data_ptr(), sig[p].data_ptr(), ep[p].data_ptr()
    d.n_params, d.n_total, d.lr = n_params, n_total, lr
    d.rank, d.world, d.algo, d.emulate, d.plain, d.blocks = 0, world, algo, 1, 0, blocks
    expect = w0.double().clone()
    for r in range(rounds):
        gs = torch.stack(G).double().sum(0)
        expect[:n_params] -= lr / world * gs[:n_params]
        expect[n_params:] += gs[n_params:]
        nl.check(lib.b2d_allreduce_sgd(C.byref(d), torch.cuda.current_stream().cuda_stream), "allreduce")
        torch.cuda.synchronize()
        for p in range(world):
            assert torch.equal(W[p], W[0]), f"round {r}: replica {p} differs from replica 0 (mirrored contract)"
            assert G[p].abs().max().item() == 0, f"round {r}: gradient arena of rank {p} was not zeroed"
        assert torch.allclose(W[0].double(), expect, atol=1e-5, rtol=1e-5), (W[0].double() - expect).abs().max()
        for p in range(world):          # next round's gradients
            G[p].copy_(torch.randn(n_total, device="cuda", generator=g))


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("algo", [0, 1, 3])
@pytest.mark.parametrize("n_params", [64, 347200, 1000 * 4 + 4])
def test_emulated_fused_allreduce_sgd(world, algo, n_params):
    _emulated(world, n_params, 4, algo)


def test_emulated_many_blocks_and_rounds():
    _emulated(8, 1 << 20, 4, algo=1, rounds=4, blocks=16)
    _emulated(4, 1 << 20, 4, algo=0, rounds=3, blocks=16)


def test_algo_dispatch_by_size():
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    lib = nl.lib()
    symm._declare(lib)
    assert lib.b2d_allreduce_pick_algo(1024, 8, 0) == 0            # 4 KB: one-shot
    assert lib.b2d_allreduce_pick_algo(347204, 8, 0) == 3          # README gradient, no multicast: pull
    assert lib.b2d_allreduce_pick_algo(347204, 8, 1) == 2          # README gradient with multicast: NVLS
    assert lib.b2d_allreduce_pick_algo(1 << 24, 8, 1) == 2         # 64 MB with multicast: NVLS
    assert lib.b2d_allreduce_pick_algo(1 << 24, 8, 0) == 1         # 64 MB without multicast: two-shot
    assert lib.b2d_allreduce_pick_algo(1 << 24, 2, 0) == 0         # 2 ranks: one-shot == two-shot traffic
