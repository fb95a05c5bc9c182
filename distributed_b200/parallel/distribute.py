"""``tf.distribute`` facade: ``distribute.experimental.MultiWorkerMirroredStrategy()`` (README.md:364)."""
from .strategy import (CollectiveCommunication, MultiWorkerMirroredStrategy, OneDeviceStrategy,  # noqa: F401
                       get_strategy, set_seed)
from .strategy import _Experimental as experimental  # noqa: F401
from .cluster import ClusterSpec  # noqa: F401



class MirroredStrategy(OneDeviceStrategy):
    """``tf.distribute.MirroredStrategy`` mirrors variables over the GPUs of ONE process; this framework runs one
    process per GPU instead (the reference's multi-worker layout, README.md:398: one device per worker), so there
    is nothing to mirror inside a process.  With no ``devices`` or a single device this is the one-replica
    strategy; asking for several devices raises and points at the multi-worker path."""

    def __init__(self, devices=None, backend: str | None = None):
        if devices is not None and len(list(devices)) > 1:
            raise NotImplementedError(
                "in-process mirroring over several devices is not provided: run one process per GPU with "
                "MultiWorkerMirroredStrategy (TF_CONFIG, or launch.barrier_apply / torchrun)")
        super().__init__(backend)
