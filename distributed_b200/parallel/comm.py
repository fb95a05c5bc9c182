"""Replica communicators: who applies the averaged gradient, and how.

Every communicator exposes the same fused operation, ``make_update``: sum the gradient arena
over all replicas, scale by 1/num_workers, apply ``w -= lr * g``, add the metric tail slots
into the running sums, and zero the gradient arena — i.e. the reference's
"batch_all_reduce: 6 all-reduces" plus its four metric all-reduces plus the optimizer step
(README.md:403-407, 302) as ONE operation.

* ``SingleComm``   one replica: just the SGD kernel.
* ``ShmComm``      CPU workers on one host: POSIX shared memory (csrc/cpu/shm_allreduce.cpp).
* ``TorchDistComm`` torch.distributed all_reduce + separate SGD: the gloo fallback for CPU
                   workers on several hosts and the NCCL *baseline* on GPUs.
* ``SymmComm``     GPUs of one NVSwitch box: peer-memory kernels (parallel/symm.py).
"""
from __future__ import annotations

import ctypes as C

import torch


class SingleComm:
    name = "single"
    world, rank = 1, 0
    num_grad_buffers = 1

    def alloc(self, n, be):
        return be.alloc((n,))

    def make_update(self, be, params, grads, n_params, lr):
        return be.sgd_apply(params, grads, n_params, lr)

    def barrier(self):
        pass

    def all_equal(self, tensor) -> bool:
        return True

    def close(self):
        pass

    capturable = True        # the update may be captured in a CUDA graph
    description = "none (single replica): SGD kernel"


class ShmComm:
    """CPU replicas on one host; fused all-reduce + average + SGD through shared memory."""
    name = "shm"
    capturable = False
    num_grad_buffers = 1

    def __init__(self, job_id: str, rank: int, world: int, capacity: int, timeout_s: float = 120.0):
        from .._native import cpu_lib
        self._lib = cpu_lib.lib()
        self.rank, self.world = rank, world
        self.capacity = capacity
        self._h = self._lib.b2c_shm_create(f"/b200dist-{job_id}".encode(), rank, world, capacity, timeout_s)
        if not self._h:
            raise RuntimeError("shared-memory communicator: " + self._lib.b2c_last_error().decode())
        self.description = f"POSIX shared memory, two-phase reduce, {world} workers"

    def alloc(self, n, be):
        return be.alloc((n,))

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError("shared-memory communicator: " + self._lib.b2c_last_error().decode())

    def make_update(self, be, params, grads, n_params, lr):
        n_total = params.numel()
        assert n_total <= self.capacity
        return lambda: self._check(self._lib.b2c_shm_allreduce_sgd(
            self._h, params.data_ptr(), grads.data_ptr(), n_params, n_total, float(lr)))

    def allreduce(self, t):
        self._check(self._lib.b2c_shm_allreduce(self._h, t.data_ptr(), t.numel()))
        return t

    def barrier(self):
        self._check(self._lib.b2c_shm_barrier(self._h))

    def gather(self, t):
        """[world, n] stack of every replica's copy of ``t`` (through the shared segment)."""
        flat = t.detach().reshape(-1).float().contiguous()
        stacked = torch.zeros(self.world * flat.numel())
        if stacked.numel() > self.capacity:
            raise RuntimeError("gather: tensor too large for the shared-memory segment")
        stacked[self.rank * flat.numel():(self.rank + 1) * flat.numel()] = flat
        self.allreduce(stacked)          # x + 0 + ... + 0 is exact
        return stacked.view(self.world, -1)

    def all_equal(self, tensor) -> bool:
        """True when ``tensor`` is bit-identical on every replica (the mirrored contract)."""
        flat = tensor.detach().reshape(-1)
        step = max(1, self.capacity // self.world)
        for off in range(0, flat.numel(), step):
            m = self.gather(flat[off:off + step])
            if not bool((m == m[0:1]).all()):
                return False
        return True

    def close(self):
        if self._h:
            self._lib.b2c_shm_destroy(self._h)
            self._h = None


class TorchDistComm:
    """torch.distributed all_reduce followed by the SGD kernel (gloo on CPU, NCCL on GPU).

    On GPUs this is the *baseline* arrangement (NCCL all-reduce + separate optimizer kernel);
    the product path is SymmComm.
    """
    capturable = False
    num_grad_buffers = 1

    def __init__(self, group=None):
        import torch.distributed as dist
        self._dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.name = dist.get_backend(group)
        self.description = f"torch.distributed {self.name} all_reduce + separate SGD kernel"

    def alloc(self, n, be):
        return be.alloc((n,))

    def make_update(self, be, params, grads, n_params, lr):
        # w -= (lr / N) * sum(g); the tail slots are added unscaled by the SGD kernel
        sgd = be.sgd_apply(params, grads, n_params, lr / self.world)

        def run():
            self._dist.all_reduce(grads, group=self.group)
            sgd()
        return run

    def barrier(self):
        self._dist.barrier(group=self.group)

    def all_equal(self, tensor) -> bool:
        flat = tensor.detach().reshape(-1).float().contiguous()
        mx, mn = flat.clone(), flat.clone()
        self._dist.all_reduce(mx, op=self._dist.ReduceOp.MAX, group=self.group)
        self._dist.all_reduce(mn, op=self._dist.ReduceOp.MIN, group=self.group)
        return bool(torch.equal(mx, mn))

    def close(self):
        pass
