#!/usr/bin/env python
"""Headline benchmark: training throughput (images/s) of the README's MNIST CNN under the
multi-worker mirrored strategy, global batch 64 x N, on N B200s of one node
(BASELINE.json metric / config 2; model README.md:292-302).

    python bench.py --gpus 1 --steps 300 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 300 --warmup 20

Prints ONE JSON line on rank 0.  `value` is the whole-job images/s, timed on the device with
CUDA events around exactly K steps (barrier + synchronize on both sides, max over ranks).
Inputs: every step reads a fresh batch out of a 60000-image synthetic MNIST-shaped data set
(188 MB fp32, larger than the 126 MB L2), resident on the device for `value` and in pinned
host memory for `e2e` (per-step H2D copy of the batch, per-step D2H read of the metric sums,
through the public `model.fit`).

--impl reference : the reference repository is a README with no installable code, so this arm
                   reports {"impl": "reference", "unavailable": ...}.
--impl torch     : the stock-PyTorch baseline (cuDNN/cuBLAS + DDP/NCCL) for the same model.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


# ----------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clocks and throttle reasons while the timed region runs.

    NVML is polled from a thread every few milliseconds (the timed region of the README model is
    only tens of milliseconds long, shorter than `nvidia-smi -lms` takes to start); `nvidia-smi`
    is the fallback when the NVML bindings are missing."""
    REASONS = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
               "hw_power_brake_slowdown": 0x80, "sw_power_cap": 0x4}

    def __init__(self, index: int, period_ms: float = 0.5):
        self.index, self.period = index, period_ms / 1000.0
        self.sm: list[float] = []
        self.mask = 0
        self.max_mhz = None
        self.power_w = 0.0
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _poll(self):
        n = self._nvml
        k = 0
        while not self._stop.is_set():
            try:
                # the clock on every poll (one NVML call, so that a timed region of a few milliseconds still gets
                # several samples); throttle reasons and power on every fourth
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
                if k % 4 == 0:
                    self.mask |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                    self.power_w = max(self.power_w, n.nvmlDeviceGetPowerUsage(self._h) / 1000.0)
            except Exception:
                pass
            k += 1
            self._stop.wait(self.period)
        try:                      # reasons once more at the end, so that a short region cannot miss them
            self.mask |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
        except Exception:
            pass

    def __enter__(self):
        if self._nvml is not None:
            self._thread = threading.Thread(target=self._poll, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)
        elif self._nvml is None:
            self._smi_once()
        return False

    def _smi_once(self):
        try:
            out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits",
                                  "-i", str(self.index)], capture_output=True, text=True, timeout=10).stdout
            a, b = [float(v) for v in out.strip().split(",")[:2]]
            self.sm.append(a)
            self.max_mhz = b
        except Exception:
            pass

    def summary(self) -> dict:
        sm = sorted(self.sm)
        reasons = [k for k, bit in self.REASONS.items() if self.mask & bit]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(sm), "power_w_max": round(self.power_w, 1)}


def _dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    return rank, world, local


def _max_over_ranks(value: float, world: int) -> float:
    if world == 1:
        return value
    import torch
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def _barrier(world: int):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


# ----------------------------------------------------------------------------------------
def run_ours(args) -> dict:
    import numpy as np
    import torch
    import distributed_b200 as tf
    from distributed_b200._native import cuda_lib

    rank, world, local = _dist_env()
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run"
    torch.cuda.set_device(local)
    per_gpu = args.batch_per_gpu
    global_batch = per_gpu * world
    if world > 1:
        strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
        strategy._ensure_group()
    else:
        strategy = tf.distribute.OneDeviceStrategy()
    shape, classes = MODEL_INPUT[args.model]
    tf.keras.mixed_precision.set_global_policy("mixed_bfloat16" if args.dtype == "bf16" else "float32")
    with strategy.scope():
        model = {"mnist_cnn": tf.models.mnist_cnn, "scaled_cnn": tf.models.scaled_cnn,
                 "resnet50": lambda: tf.models.resnet50(learning_rate=0.01)}[args.model]()

    # synthetic data of the model's input shape, fp32, larger than L2; labels int32
    n_images = args.dataset_images - args.dataset_images % global_batch
    row_bytes = 4 * shape[0] * shape[1] * shape[2]
    g = torch.Generator().manual_seed(0)
    x_host = torch.rand((n_images, *shape), generator=g, dtype=torch.float32).pin_memory()
    y_host = torch.randint(0, classes, (n_images,), generator=g, dtype=torch.int32).pin_memory()
    steps_in_src = n_images // global_batch

    # --- warm-up through the public API (builds the trainer, captures the step graphs) ---
    model.fit(x_host, y_host, batch_size=global_batch, epochs=1, steps_per_epoch=max(args.warmup, 3), verbose=0)
    trainer = model._trainer
    x_dev, y_dev = x_host.to(trainer.be.device), y_host.to(trainer.be.device)
    torch.cuda.synchronize()

    # --- device-timed region: exactly K steps, CUDA events on the launching stream --------
    # The processes leave the host barrier milliseconds apart, which is longer than the whole timed region
    # of a short run.  ALIGN untimed steps are therefore enqueued on the same stream right before the start
    # event, with no host synchronisation in between: their fused all-reduce lines the replicas up on the
    # device, so the K timed steps are steady-state steps on every rank.
    align = ALIGN_STEPS if args.model == "mnist_cnn" else 32      # ~10 ms / ~20-180 ms of untimed steady-state steps
    _barrier(world)
    torch.cuda.synchronize()
    with ClockSampler(local) as clocks:
        launches0 = cuda_lib.launch_count()
        ms = trainer.run_steps(x_dev, y_dev, model._global_step, steps_in_src, args.steps, global_batch,
                               timed=True, pre_steps=align)
        torch.cuda.synchronize()
        launches = (cuda_lib.launch_count() - launches0) * args.steps // (args.steps + align)
        _barrier(world)
        model._global_step += args.steps + align
        ms = _max_over_ranks(ms, world)
        value = args.steps * global_batch / (ms / 1e3)

        # --- end to end through model.fit: per-step H2D from pinned memory + per-step D2H -------
        _barrier(world)
        # an untimed fit() of a few steps right before the timed one: its last all-reduce and final sync leave
        # every process within microseconds of the others, instead of the host barrier's release skew (~ms)
        model.fit(x_host, y_host, batch_size=global_batch, epochs=1, steps_per_epoch=8, verbose=0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        hist = model.fit(x_host, y_host, batch_size=global_batch, epochs=1, steps_per_epoch=args.steps, verbose=0)
        torch.cuda.synchronize()
        e2e_s = _max_over_ranks(time.perf_counter() - t0, world)
        e2e = args.steps * global_batch / e2e_s

    out = {
        "metric": "images/sec (device-timed, max over ranks)", "value": value, "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": ("bf16 (bf16 activations / activation gradients / conv operands, fp32 accumulate, fp32 master "
                  "weights and statistics; Dense head tf32)" if args.dtype == "bf16"
                  else "tf32 (fp32 storage, fp32 accumulate; first conv in fp32)"), "data": "synthetic",
        "impl": "distributed_b200",
        "config": {"model": MODEL_NAMES[args.model],
                   "global_batch": global_batch, "batch_per_gpu": per_gpu, "input": "x".join(map(str, shape)),
                   "parallelism": f"dp{world}",
                   "all_reduce": trainer.comm.description + (
                       "; schedule: " + trainer.update_schedule if hasattr(trainer, "update_schedule") else ""),
                   "l2_policy": f"inputs larger than L2: each step reads a fresh batch of a {n_images}-image "
                                f"({n_images * row_bytes / 1e6:.0f} MB) data set",
                   "kernels_per_step": trainer.kernels_per_step, "cuda_graph": trainer.use_graph,
                   "rank_alignment": f"{align} untimed steps enqueued on the same stream before the start event"},
        "clocks": clocks.summary(),
        "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": per_gpu * row_bytes + per_gpu * 4,
                "d2h_bytes_per_step": 16, "ms_per_step": 1e3 * e2e_s / args.steps,
                "api": "model.fit(x_pinned, y_pinned, batch_size, epochs=1, steps_per_epoch=K)"},
        "gpu_launches": int(launches),
        "final_loss": hist.history["loss"][-1],
    }
    # vs_baseline: value / the BASELINE.md section 3 number for this model and GPU count (stock PyTorch,
    # cuDNN/cuBLAS + DDP/NCCL, best recorded arrangement).  The reference itself publishes nothing.
    base = BASELINE_MD.get((args.model, world))
    if base is not None and args.batch_per_gpu == DEFAULT_BATCH[args.model]:
        out["vs_baseline"] = value / base
        out["baseline_md_images_per_s"] = base
    # Same-box baseline: the stock-PyTorch arm measured in this very process right after the framework's arm
    # (same steps / warm-up / data policy / GPU count), so the line carries a ratio taken on this box.
    if args.same_box_baseline:
        model._trainer.close()
        del x_dev, y_dev
        torch.cuda.empty_cache()
        same = {}
        variants = [("eager", False)] + ([("graph", True)] if world == 1 else [])
        # NCCL prints its version banner to stdout when DDP initialises it: this process must print exactly
        # ONE line (the JSON), so stdout is pointed at stderr for the duration of the baseline arm.
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)

        # The framework's own result must not depend on the baseline arm finishing: if it hangs (a NCCL problem on
        # some box, a rank that failed while the others wait in a collective), every rank's watchdog prints / exits.
        def bail():
            out["baseline_same_box"] = {"error": f"the baseline arm did not finish within {args.baseline_timeout} s"}
            os.dup2(saved_stdout, 1)
            if rank == 0:
                os.write(1, (json.dumps(out) + "\n").encode())
            os._exit(0)
        watchdog = threading.Timer(args.baseline_timeout, bail)
        watchdog.daemon = True
        watchdog.start()
        for name, use_graph in variants:
            try:
                a2 = argparse.Namespace(**vars(args))
                a2.torch_graph, a2.channels_last = use_graph, (args.dtype == "bf16")
                a2.steps = min(args.steps, args.baseline_steps)
                r = run_torch(a2)
                if rank == 0:
                    same[name] = {"value": r["value"], "ms_per_step": r["ms_per_step"], "e2e": r["e2e"]["value"],
                                  "steps": a2.steps, "clocks": r["clocks"], "dtype": r["dtype"]}
                    same[f"ratio_vs_{name}"] = value / r["value"]
                    same[f"e2e_ratio_vs_{name}"] = e2e / r["e2e"]["value"]
            except Exception as exc:                                   # the framework's own line must survive
                same[name] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        watchdog.cancel()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
        out["baseline_same_box"] = same
    return out if rank == 0 else {}


# BASELINE.md section 3, stock PyTorch on the same 8 x B200 box (images/s; best recorded arrangement:
# whole-step CUDA graph at 1 GPU, eager DDP/NCCL otherwise)
BASELINE_MD = {("mnist_cnn", 1): 366905.0, ("mnist_cnn", 2): 141688.0, ("mnist_cnn", 4): 271473.0,
               ("mnist_cnn", 8): 506281.0, ("scaled_cnn", 1): 505431.0, ("scaled_cnn", 8): 3619515.0,
               ("resnet50", 1): 3608.0, ("resnet50", 2): None, ("resnet50", 8): 15880.0}
DEFAULT_BATCH = {"mnist_cnn": 64, "scaled_cnn": 1024, "resnet50": 32}

ALIGN_STEPS = 200    # untimed steps enqueued right before the start event: device-side rank alignment, clocks up,
                     # and a region long enough for the clock sampler to see several samples

MODEL_INPUT = {"mnist_cnn": ((28, 28, 1), 10), "scaled_cnn": ((28, 28, 1), 10), "resnet50": ((224, 224, 3), 1000)}
MODEL_NAMES = {"mnist_cnn": "README MNIST CNN (conv32-3x3/maxpool/dense64/dense10, SGD lr=1e-3)",
               "scaled_cnn": "scaled CNN (conv64/conv128/dense1024)",
               "resnet50": "ResNet-50 v1.5, synthetic 224x224x3, 1000 classes"}


def run_reference(args) -> dict:
    return {"impl": "reference",
            "unavailable": "reference repo is a README + LICENSE with no installable code (pip install of "
                           "/root/reference fails: no setup.py/pyproject); it depends on TensorFlow 2.0, "
                           "which is not in the offline wheelhouse"}


def run_torch(args) -> dict:
    from baseline.torch_baseline import run
    return run(args, ClockSampler, _dist_env, _max_over_ranks, _barrier)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch"])
    ap.add_argument("--model", default="mnist_cnn", choices=["mnist_cnn", "scaled_cnn", "resnet50"])
    ap.add_argument("--batch-per-gpu", type=int, default=None)
    ap.add_argument("--dataset-images", type=int, default=None)
    ap.add_argument("--dtype", default=None, choices=["tf32", "bf16"],
                    help="ours: storage/compute type of the convolutional stack (default: bf16 for resnet50 as "
                         "BASELINE.json names it, tf32-on-fp32 for the MNIST CNNs as the reference trains in float32)")
    ap.add_argument("--channels-last", action="store_true", help="torch baseline: channels_last + bf16 autocast")
    ap.add_argument("--torch-graph", action="store_true", help="torch baseline: capture the step in a CUDA graph")
    ap.add_argument("--same-box-baseline", type=int, default=1,
                    help="ours: also run the stock-PyTorch arm in this process and report the ratio (0 = skip)")
    ap.add_argument("--baseline-steps", type=int, default=200, help="steps of the same-box baseline arm")
    ap.add_argument("--baseline-timeout", type=float, default=300.0,
                    help="seconds after which a stuck baseline arm is abandoned (the framework's line is printed anyway)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.dtype is None:
        args.dtype = "bf16" if args.model == "resnet50" else "tf32"
    if args.batch_per_gpu is None:
        args.batch_per_gpu = DEFAULT_BATCH[args.model] if args.model != "scaled_cnn" else 64
    if args.dataset_images is None:
        args.dataset_images = 512 if args.model == "resnet50" else 60000        # both > 126 MB of L2
    if args.impl == "reference":
        out = run_reference(args)
        if int(os.environ.get("RANK", "0")) == 0:
            print(json.dumps(out), flush=True)
        return
    if args.impl == "torch":
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)                     # NCCL's version banner goes to stderr; stdout carries the JSON line only
        try:
            out = run_torch(args)
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    else:
        out = run_ours(args)
    if out:
        print(json.dumps(out), flush=True)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
