"""Markdown tables of `profiles/r02_*.md` from the raw lines under `profiles/r02_raw/` (or another directory).

    python tools/profile_tables.py sweep   [dir]     # config 5: one table per GPU count from sweep{2,4,8}.jsonl
    python tools/profile_tables.py bench   [dir]     # one row per bench JSON line found in the directory tree

The tables in the profiles were produced with exactly this; keeping it in the tree makes them reproducible from
the tracked raw lines (and from the next lease's `gpurun_out/`)."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def dispatched(world: int, nbytes: int, multicast: bool = True) -> str:
    """Mirror of b2d_allreduce_pick_algo (csrc/cuda/allreduce.cu)."""
    if world <= 2:
        return "twoshot" if nbytes >= (4 << 20) else "oneshot"
    if multicast:
        return "nvls" if nbytes >= (8 << 10) else "oneshot"
    return "twoshot" if nbytes >= (16 << 10) else "oneshot"


def sweep(directory: Path) -> None:
    for n in (8, 4, 2):
        f = directory / f"sweep{n}.jsonl"
        if not f.exists():
            continue
        print(f"\n### {n} GPUs\n")
        print("| bytes | one-shot | two-shot | pull | NVLS | NCCL | NCCL+SGD | dispatched | dispatched / NCCL | "
              "dispatched / (NCCL+SGD) | bus GB/s |")
        print("|---|---|---|---|---|---|---|---|---|---|---|")
        for line in f.read_text().splitlines():
            d = json.loads(line)
            if "bytes" not in d:
                continue
            b = d["bytes"]
            pick = dispatched(n, b)
            cell = lambda k: f"{d[k]:.1f}" if k in d else "—"
            size = f"{b >> 10} KB" if b < (1 << 20) else f"{b >> 20} MB"
            t = d[pick + "_us"]
            print(f"| {size} | {cell('oneshot_us')} | {cell('twoshot_us')} | {cell('pull_us')} | {cell('nvls_us')} | "
                  f"{cell('nccl_us')} | {cell('nccl_sgd_us')} | {pick} | {t / d['nccl_us']:.2f} | "
                  f"{t / d['nccl_sgd_us']:.2f} | {2 * (n - 1) / n * b / t / 1e3:.0f} |")


def bench(directory: Path) -> None:
    print("| file | model | GPUs | steps | step | images/s (device-timed) | images/s e2e | PyTorch eager same box | "
          "PyTorch graph same box | ratio vs eager | ratio vs graph | SM MHz (samples) | reasons |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for f in sorted(directory.rglob("*.json")):
        lines = [l for l in f.read_text().splitlines() if l.startswith("{")]
        if not lines:
            continue
        d = json.loads(lines[-1])
        if "ms_per_step" not in d:
            continue
        sb = d.get("baseline_same_box", {})
        eager, graph = sb.get("eager", {}).get("value"), sb.get("graph", {}).get("value")
        step = d["ms_per_step"]
        fmt = lambda v: f"{v:,.0f}" if v else "—"
        ratio = lambda k: f"{sb[k]:.2f}" if k in sb else "—"
        c = d.get("clocks", {})
        print(f"| {f.relative_to(directory)} | {d['config']['model'][:24]} | {d['n_gpus']} | {d['steps']} | "
              f"{(f'{step * 1e3:.1f} us' if step < 0.2 else f'{step:.3f} ms')} | {d['value']:,.0f} | "
              f"{d['e2e']['value']:,.0f} | {fmt(eager)} | {fmt(graph)} | {ratio('ratio_vs_eager')} | "
              f"{ratio('ratio_vs_graph')} | {c.get('sm_mhz')} ({c.get('samples')}) | {c.get('reasons')} |")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "bench"
    where = Path(sys.argv[2]) if len(sys.argv) > 2 else ROOT / "profiles" / "r02_raw"
    {"sweep": sweep, "bench": bench}[what](where)
