"""Where the communication of one training step sits in time (2+ GPUs, torchrun).

Every all-reduce launch of the step (overlapped buckets on the communication branch, closing LL exchange on
the main branch) stamps %globaltimer at its start, once its inputs are ready / its barrier is passed, and at
its end.  Printed per rank, relative to the end of the previous step's closing exchange: the step period, when
each bucket ran, and how long the closing exchange waited for the slowest replica."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_b200 as tf
from distributed_b200.parallel import symm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
model_name = os.environ.get("MODEL", "mnist_cnn")
per_gpu = int(os.environ.get("BATCH", "64"))
strategy = tf.distribute.experimental.MultiWorkerMirroredStrategy()
with strategy.scope():
    model = getattr(tf.models, model_name)()
gb = per_gpu * world
g = torch.Generator().manual_seed(0)
x = torch.rand((gb * 16, 28, 28, 1), generator=g).pin_memory()
y = torch.randint(0, 10, (gb * 16,), generator=g, dtype=torch.int32).pin_memory()
ar_bufs, ll_bufs = [], []
orig_update, orig_step = symm.SymmComm.make_update, symm.SymmComm.make_step_updates

def make_update(self, *a, **k):
    run = orig_update(self, *a, **k)
    buf = torch.zeros(144 * 8, dtype=torch.int64, device=self.device)
    ar_bufs.append(buf); self._descs[-1].timeline = buf.data_ptr()
    return run

def make_step_updates(self, *a, **k):
    n0 = len(ar_bufs)
    out = orig_step(self, *a, **k)
    d = self._descs[-1]
    if isinstance(d, symm.LlDesc):
        buf = torch.zeros(64 * 4, dtype=torch.int64, device=self.device)
        ll_bufs.append((buf, len(ar_bufs) - n0)); d.timeline = buf.data_ptr()
    return out

symm.SymmComm.make_update, symm.SymmComm.make_step_updates = make_update, make_step_updates
steps = 400                               # even: slot 1 holds the last step, slot 0 the one before
model.fit(x, y, batch_size=gb, epochs=1, steps_per_epoch=4, verbose=0)          # builds the trainer / graphs
t = model._trainer
xd, yd = x.to(t.be.device), y.to(t.be.device)                                   # device-resident data, as in bench.py
t.run_steps(xd, yd, 4, 16, steps - 4, gb)
torch.cuda.synchronize()
if len(ll_bufs) == 2:
    ll = [b.view(64, 4).cpu()[0].double() / 1e3 for b, _ in ll_bufs]            # block 0 of each slot, us
    n_early = ll_bufs[0][1]
    t_ref = ll[0][2].item()                                                      # end of step K-2
    msg = [f"rank {rank}: step period {ll[1][2].item() - t_ref:7.2f} us"]
    for k in range(n_early):
        t = ar_bufs[n_early + k].view(144, 8).cpu().double()
        t = t[t[:, 0] > 0] / 1e3
        msg.append(f"  bucket {k}: blocks {len(t)}  start {t[:, 5].min().item() - t_ref:7.2f}  past barrier "
                   f"{t[:, 6].max().item() - t_ref:7.2f}  end {t[:, 7].max().item() - t_ref:7.2f}   "
                   f"(avg/step: upstream wait {t[0, 1].item() / t[0, 0].item():5.2f}, peers "
                   f"{t[0, 3].item() / t[0, 0].item():5.2f}, data {t[0, 4].item() / t[0, 0].item():5.2f})")
    msg.append(f"  closing LL exchange: launched {ll[1][0].item() - t_ref:7.2f}  inputs ready {ll[1][1].item() - t_ref:7.2f}  "
               f"done {ll[1][2].item() - t_ref:7.2f}")
    print("\n".join(msg), flush=True)
else:
    for k, b in enumerate(ar_bufs):
        t = b.view(144, 8).cpu().double()
        t = t[t[:, 0] > 0]
        n = t[:, 0:1]
        avg = (t[:, 1:5] / n / 1e3).mean(0)
        print(f"rank {rank} slot {k}: blocks={len(t)} upstream_wait={avg[0]:.2f} release={avg[1]:.2f} "
              f"wait_for_peers={avg[2]:.2f} data={avg[3]:.2f} us (avg per step)", flush=True)
model._trainer.close()
