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


def test_emulated_pull_with_delayed_blocks(monkeypatch):
    """The pull variant synchronises block b of a reader with block b of the owner only, so a reader may pull
    only what that block wrote.  Debug mode pauses pseudo-random (rank, block, epoch) combinations for up to
    20 us at every phase and checks flag monotonicity: with a wrong element mapping, replicas would diverge."""
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    monkeypatch.setenv("B200DIST_AR_DEBUG", "1")
    lib = nl.lib()
    symm._declare(lib)
    for algo in (3, 1, 0):
        _emulated(8, 347200, 4, algo=algo, rounds=6)           # per = 10851 float4: not a multiple of the grid stride
        _emulated(3, 1000 * 4 + 4, 4, algo=algo, rounds=6)


def _ll_emulated(world, segs, n_params, n_total, rounds=4, lr=0.05):
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    lib = nl.lib()
    symm._declare(lib)
    g = torch.Generator(device="cuda").manual_seed(world * 77 + n_total)
    w0 = torch.randn(n_total, device="cuda", generator=g)
    W = [[w0.clone() for _ in range(world)] for _ in range(2)]          # weight ping-pong
    G = [[torch.zeros(n_total, device="cuda") for _ in range(world)] for _ in range(2)]
    inbox = [torch.zeros(lib.b2d_ll_inbox_bytes() // 8, dtype=torch.int64, device="cuda") for _ in range(world)]
    ep = [torch.zeros(lib.b2d_ll_epoch_bytes() // 4, dtype=torch.int32, device="cuda") for _ in range(world)]
    expect = w0.double().clone()
    mask = torch.zeros(n_total, dtype=torch.bool, device="cuda")
    for b, e in segs:
        mask[b:e] = True
    for r in range(rounds):
        s = r & 1
        for p in range(world):
            G[s][p].copy_(torch.randn(n_total, device="cuda", generator=g))
            G[1 - s][p].fill_(3.0)                                       # must come back zeroed inside the segments
            W[1 - s][p].copy_(W[s][p])                                   # untouched outside the segments
        d = symm.LlDesc()
        for p in range(world):
            d.inbox[p], d.epoch[p] = inbox[p].data_ptr(), ep[p].data_ptr()
            d.w[p], d.w_out[p] = W[s][p].data_ptr(), W[1 - s][p].data_ptr()
            d.g[p], d.g_zero[p] = G[s][p].data_ptr(), G[1 - s][p].data_ptr()
        d.nseg = len(segs)
        for k, (b, e) in enumerate(segs):
            d.seg_begin[k], d.seg_end[k] = b, e
        d.n_params, d.lr, d.rank, d.world, d.emulate = n_params, lr, 0, world, 1
        nl.check(lib.b2d_ll_allreduce_sgd(C.byref(d), torch.cuda.current_stream().cuda_stream), "ll")
        torch.cuda.synchronize()
        gs = torch.stack(G[s]).double().sum(0)
        upd = expect.clone()
        upd[:n_params] -= lr / world * gs[:n_params]
        upd[n_params:] += gs[n_params:]
        expect = torch.where(mask, upd, expect)
        for p in range(world):
            assert torch.equal(W[1 - s][p], W[1 - s][0]), f"round {r}: replica {p} differs (mirrored contract)"
            assert G[1 - s][p][mask].abs().max().item() == 0 and (G[1 - s][p][~mask] == 3.0).all()
        assert torch.allclose(W[1 - s][0].double(), expect, atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_emulated_ll_exchange(world):
    """First-layer bucket + metric tail of the README model (320 + 4 values), and the largest launch."""
    _ll_emulated(world, [(0, 320), (347168, 347172)], 347168, 347172)
    _ll_emulated(world, [(64, 64 + 16380), (20000, 20004)], 20000, 20004, rounds=3)
    _ll_emulated(world, [(5, 6)], 8, 8, rounds=3)


def test_ll_exchange_rejects_oversized_launches():
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    lib = nl.lib()
    symm._declare(lib)
    d = symm.LlDesc()
    d.nseg, d.world, d.emulate = 1, 2, 1
    d.seg_begin[0], d.seg_end[0] = 0, lib.b2d_ll_max_floats() + 1
    assert lib.b2d_ll_allreduce_sgd(C.byref(d), None) != 0
    assert b"limit" in lib.b2d_last_error()


def test_nvls_kernel_on_one_gpu():
    """The NVLS (multimem.ld_reduce / multimem.st) variant on a one-device multicast object: the only way to
    execute those instructions without a second GPU.  With one replica the switch-side sum is the identity, so
    the result must equal plain SGD; the 2/4/8-GPU behaviour is tests/test_multigpu.py."""
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    lib = nl.lib()
    symm._declare(lib)
    n_params, n_total, lr = 347168, 347172, 0.1
    arena = (n_total * 4 + 4095) // 4096 * 4096
    sig_b, ep_b = lib.b2d_allreduce_signal_bytes(), lib.b2d_allreduce_epoch_bytes()
    total = 3 * arena + 8192 + 8192
    h = C.c_void_p()
    import os
    nl.check(lib.b2d_symm_create(C.byref(h), f"nvls1-{os.getpid()}".encode(), 0, 1, torch.cuda.current_device(),
                                 total, 30.0), "symm_create")
    try:
        if not lib.b2d_symm_multicast_supported(h):
            pytest.skip("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is 0 on this device")
        if lib.b2d_symm_multicast_phase1(h) != 0 or lib.b2d_symm_multicast_phase2(h) != 0:
            pytest.skip("the driver refused a one-device multicast object: " + lib.b2d_last_error().decode())
        base, mc = int(lib.b2d_symm_ptr(h, 0)), int(lib.b2d_symm_multicast_ptr(h))
        assert mc
        W = torch.as_tensor(symm._Raw(base, n_total), device="cuda")
        W2 = torch.as_tensor(symm._Raw(base + arena, n_total), device="cuda")
        G = torch.as_tensor(symm._Raw(base + 2 * arena, n_total), device="cuda")
        gen = torch.Generator(device="cuda").manual_seed(5)
        W.copy_(torch.randn(n_total, device="cuda", generator=gen))
        G.copy_(torch.randn(n_total, device="cuda", generator=gen))
        w0, g0 = W.clone(), G.clone()
        d = symm.AllreduceDesc()
        d.w[0], d.g[0] = base, base + 2 * arena
        d.sig[0], d.epoch[0] = base + 3 * arena, base + 3 * arena + 8192
        assert sig_b <= 8192 and ep_b <= 8192
        d.w_mc, d.g_mc = mc, mc + 2 * arena
        d.w_out[0], d.w_out_mc = base + arena, mc + arena               # weight ping-pong: write the other buffer
        d.n_params, d.n_total, d.lr = n_params, n_total, lr
        d.rank, d.world, d.algo, d.emulate, d.plain, d.blocks = 0, 1, 2, 0, 0, 0
        nl.check(lib.b2d_allreduce_sgd(C.byref(d), torch.cuda.current_stream().cuda_stream), "allreduce nvls")
        torch.cuda.synchronize()
        expect = w0.clone()
        expect[:n_params] -= lr * g0[:n_params]
        expect[n_params:] += g0[n_params:]
        assert torch.equal(W, w0), "the source weight buffer must stay untouched"
        assert torch.allclose(W2, expect, atol=1e-6, rtol=1e-6)
        assert G.abs().max().item() == 0
    finally:
        lib.b2d_symm_destroy(h)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_gemm_reduce_scatter_epilogue_then_scattered_update(world):
    """GEMM -> reduce-scatter in one kernel, then the update that finishes it (single-GPU emulation: the
    replicas' gradient arenas are separate buffers on this device).  Every replica's weight-gradient GEMM adds
    each 16-byte vector of its result into the arena of the replica owning that vector's shard; afterwards the
    owners hold the complete sums, everything else is untouched, and the algo-4 update reproduces plain SGD on
    the summed gradient."""
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.ops.gemm import GemmPlan
    from distributed_b200.parallel import symm
    lib = nl.lib()
    symm._declare(lib)
    M, N, K, off, lr = 5408, 64, 64, 320, 0.1                            # dense-64 weight gradient of the README model
    n_params = off + M * N + 96
    n_total = n_params + 4
    gen = torch.Generator(device="cuda").manual_seed(world)
    G = [torch.zeros(n_total, device="cuda") for _ in range(world)]
    xs = [torch.randn(K, M, device="cuda", generator=gen) for _ in range(world)]      # x^T . dy, both MN-major
    dys = [torch.randn(K, N, device="cuda", generator=gen) for _ in range(world)]
    per = -(-(M * N // 4) // world) * 4
    plans = []
    for r in range(world):
        dw = G[r][off:off + M * N].view(M, N)
        p = GemmPlan(xs[r], dys[r], dw, a_mn_major=True, b_mn_major=True, accumulate=True)
        assert p.geometry.splits == 1
        p.set_scatter([g.data_ptr() for g in G], off, off, per, r)
        plans.append(p)
    for p in plans:
        p.run()
    torch.cuda.synchronize()
    total = sum(x.double().t() @ dy.double() for x, dy in zip(xs, dys)).reshape(-1)
    for r in range(world):
        got = G[r][off:off + M * N].double()
        lo, hi = min(M * N, per * r), min(M * N, per * (r + 1))
        scale = total.abs().max().item()
        assert (got[lo:hi] - total[lo:hi]).abs().max().item() < 2e-2 * scale      # tf32 products, fp32 sums
        assert got[:lo].abs().max().item() == 0 if lo else True
        assert got[hi:].abs().max().item() == 0 if hi < M * N else True
        assert G[r][:off].abs().max().item() == 0 and G[r][off + M * N:].abs().max().item() == 0
    # the update that finishes the bucket
    w0 = torch.randn(n_total, device="cuda", generator=gen)
    W = [w0.clone() for _ in range(world)]
    sig = [torch.zeros(lib.b2d_allreduce_signal_bytes() // 4, dtype=torch.int32, device="cuda") for _ in range(world)]
    ep = [torch.zeros(lib.b2d_allreduce_epoch_bytes() // 4, dtype=torch.int32, device="cuda") for _ in range(world)]
    summed = torch.zeros(M * N, device="cuda")
    for r in range(world):
        summed += G[r][off:off + M * N]
    d = symm.AllreduceDesc()
    for p in range(world):
        d.w[p], d.g[p], d.sig[p], d.epoch[p] = W[p].data_ptr(), G[p].data_ptr(), sig[p].data_ptr(), ep[p].data_ptr()
    d.n_params, d.n_total, d.lr = n_params, n_total, lr
    d.rank, d.world, d.algo, d.emulate, d.plain, d.blocks = 0, world, 4, 1, 0, 0
    d.nseg, d.seg_begin[0], d.seg_end[0] = 1, off, off + M * N
    nl.check(lib.b2d_allreduce_sgd(C.byref(d), torch.cuda.current_stream().cuda_stream), "allreduce scattered")
    torch.cuda.synchronize()
    expect = w0.clone()
    expect[off:off + M * N] -= lr / world * summed
    for p in range(world):
        assert torch.equal(W[p], W[0])
        assert G[p].abs().max().item() == 0
    assert torch.allclose(W[0], expect, atol=1e-6, rtol=1e-6)


def test_algo_dispatch_by_size():
    from distributed_b200._native import cuda_lib as nl
    from distributed_b200.parallel import symm
    lib = nl.lib()
    symm._declare(lib)
    assert lib.b2d_allreduce_pick_algo(1024, 8, 0) == 0            # 4 KB: one-shot
    assert lib.b2d_allreduce_pick_algo(1024, 8, 1) == 0
    assert lib.b2d_allreduce_pick_algo(16384, 8, 1) == 2           # 64 KB with multicast: NVLS (not pull)
    assert lib.b2d_allreduce_pick_algo(347204, 8, 0) == 1          # README gradient, no multicast: two-shot
    assert lib.b2d_allreduce_pick_algo(347204, 8, 1) == 2          # README gradient with multicast: NVLS
    assert lib.b2d_allreduce_pick_algo(1 << 24, 8, 1) == 2         # 64 MB with multicast: NVLS
    assert lib.b2d_allreduce_pick_algo(1 << 24, 8, 0) == 1         # 64 MB without multicast: two-shot
    assert lib.b2d_allreduce_pick_algo(347204, 2, 1) == 0          # 2 ranks, README gradient: one-shot (one barrier)
    assert lib.b2d_allreduce_pick_algo(1 << 24, 2, 0) == 1         # 2 ranks, 64 MB: two-shot uses both directions
