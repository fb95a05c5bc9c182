"""Timeline of the fused all-reduce+SGD kernel inside real training steps (2+ GPUs, torchrun)."""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_b200 as tf

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
with strategy.scope():
    model = tf.models.mnist_cnn()
gb = 64 * world
g = torch.Generator().manual_seed(0)
x = torch.rand((gb * 64, 28, 28, 1), generator=g).pin_memory(); y = torch.randint(0, 10, (gb * 64,), generator=g, dtype=torch.int32).pin_memory()
# attach a timeline buffer to the update descriptors BEFORE the graphs are captured
from distributed_b200.parallel import symm
orig = symm.SymmComm.make_update
bufs = []
def patched(self, *a, **k):
    run = orig(self, *a, **k)
    d = self._descs[-1]
    buf = torch.zeros(144 * 8, dtype=torch.int64, device=self.device)
    bufs.append(buf); d.timeline = buf.data_ptr()
    return run
symm.SymmComm.make_update = patched
model.fit(x, y, batch_size=gb, epochs=1, steps_per_epoch=40, verbose=0)
torch.cuda.synchronize()
for k, b in enumerate(bufs):
    t = b.view(144, 8).cpu().double()
    t = t[t[:, 0] > 0]
    n = t[:, 0:1]
    avg = (t[:, 1:5] / n / 1e3).mean(0)
    print(f"rank {rank} slot {k}: blocks={len(t)} launches/block={int(n[0].item())}  early_launch_wait={avg[0]:.2f}  release={avg[1]:.2f}  "
          f"wait_for_peers={avg[2]:.2f}  data={avg[3]:.2f} us (avg per step)", flush=True)
model._trainer.close()
