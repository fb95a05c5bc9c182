"""The bar to beat: the same model and strategy written
in stock PyTorch — cuDNN convolution, cuBLAS GEMM, ATen elementwise/loss kernels, SGD, and
DistributedDataParallel's bucketed NCCL all-reduce.  No kernel, engine or communicator of
distributed_b200 is on this path.  TF32 is enabled for both cuDNN and cuBLAS so the precision
matches the framework's tf32 dense path.

Same measurement protocol as bench.py: K steps timed with CUDA events between barriers, max
over ranks; a fresh batch of a 60000-image device-resident data set every step; plus an
end-to-end loop with a per-step H2D copy from pinned memory and a per-step D2H loss read.
"""
from __future__ import annotations

import time


def build_model(name: str):
    import torch.nn as nn
    if name == "resnet50":
        import torchvision
        return torchvision.models.resnet50(weights=None)
    if name == "mnist_cnn":
        return nn.Sequential(nn.Conv2d(1, 32, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Flatten(),
                             nn.Linear(5408, 64), nn.ReLU(), nn.Linear(64, 10))
    return nn.Sequential(nn.Conv2d(1, 64, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Conv2d(64, 128, 3), nn.ReLU(),
                         nn.MaxPool2d(2), nn.Flatten(), nn.Linear(3200, 1024), nn.ReLU(), nn.Linear(1024, 10))


def run(args, ClockSampler, dist_env, max_over_ranks, barrier) -> dict:
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F

    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.benchmark = True
    pg = None
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=dev)    # gloo only for the timing reductions
    elif world > 1 and "nccl" not in str(dist.get_backend()):
        pg = dist.new_group(backend="nccl")      # same process as the framework's arm (gloo control plane)
    torch.manual_seed(0)
    model = build_model(args.model).to(dev)
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], process_group=pg)
    opt = torch.optim.SGD(model.parameters(), lr=1e-3)
    per_gpu = args.batch_per_gpu
    global_batch = per_gpu * world
    n_images = args.dataset_images - args.dataset_images % global_batch
    g = torch.Generator().manual_seed(0)
    shape, classes = ((3, 224, 224), 1000) if args.model == "resnet50" else ((1, 28, 28), 10)
    x_host = torch.rand((n_images, *shape), generator=g).pin_memory()              # NCHW, torch's native layout
    y_host = torch.randint(0, classes, (n_images,), generator=g).pin_memory()
    bf16 = bool(getattr(args, "channels_last", False))
    if bf16:
        model = model.to(memory_format=torch.channels_last)
    x_dev, y_dev = x_host.to(dev), y_host.to(dev)
    steps_in_src = n_images // global_batch

    def step(xb, yb):
        opt.zero_grad(set_to_none=True)
        if bf16:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = F.cross_entropy(model(xb.contiguous(memory_format=torch.channels_last)), yb)
        else:
            loss = F.cross_entropy(model(xb), yb)
        loss.backward()
        opt.step()
        return loss

    def batch(src_x, src_y, t):
        lo = (t % steps_in_src) * global_batch + rank * per_gpu
        return src_x[lo:lo + per_gpu], src_y[lo:lo + per_gpu]

    graph = None
    if args.torch_graph:
        # whole-step CUDA graph over static input buffers (the best stock-PyTorch arrangement)
        sx, sy = x_dev[:per_gpu].clone(), y_dev[:per_gpu].clone()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(11):
                step(sx, sy)
        torch.cuda.current_stream().wait_stream(s)
        graph = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(graph):
            static_loss = step(sx, sy)

        def run_step(src_x, src_y, t):
            xb, yb = batch(src_x, src_y, t)
            sx.copy_(xb, non_blocking=True)
            sy.copy_(yb, non_blocking=True)
            graph.replay()
            return static_loss
    else:
        def run_step(src_x, src_y, t):
            xb, yb = batch(src_x, src_y, t)
            if not xb.is_cuda:
                xb, yb = xb.to(dev, non_blocking=True), yb.to(dev, non_blocking=True)
            return step(xb, yb)

    t = 0
    for _ in range(max(args.warmup, 3)):
        run_step(x_dev, y_dev, t)
        t += 1
    torch.cuda.synchronize()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        e0.record()
        for _ in range(args.steps):
            run_step(x_dev, y_dev, t)
            t += 1
        e1.record()
        torch.cuda.synchronize()
    barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    value = args.steps * global_batch / (ms / 1e3)

    barrier(world)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    losses = []
    for _ in range(args.steps):
        losses.append(run_step(x_host, y_host, t).item())       # per-step D2H read of the loss
        t += 1
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)

    out = {
        "metric": "images/sec (device-timed, max over ranks)", "value": value, "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16 autocast, channels_last" if bf16 else "tf32", "data": "synthetic",
        "impl": "torch-baseline" + ("+cudagraph" if graph is not None else ""),
        "config": {"model": args.model, "global_batch": global_batch, "batch_per_gpu": per_gpu,
                   "parallelism": f"dp{world}", "all_reduce": "DDP bucketed NCCL all-reduce" if world > 1 else "none",
                   "l2_policy": "inputs larger than L2 (fresh batch of a 60000-image data set every step)"},
        "clocks": clocks.summary(),
        "e2e": {"value": args.steps * global_batch / e2e_s, "unit": "images/s",
                "h2d_bytes_per_step": per_gpu * 4 * shape[0] * shape[1] * shape[2] + per_gpu * 8, "d2h_bytes_per_step": 4},
        "gpu_launches": 0, "final_loss": losses[-1],
    }
    return out if rank == 0 else {}
