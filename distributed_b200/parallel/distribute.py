"""``tf.distribute`` facade: ``distribute.experimental.MultiWorkerMirroredStrategy()`` (README.md:364)."""
from .strategy import (CollectiveCommunication, MultiWorkerMirroredStrategy, OneDeviceStrategy,  # noqa: F401
                       get_strategy, set_seed)
from .strategy import _Experimental as experimental  # noqa: F401
from .cluster import ClusterSpec  # noqa: F401

MirroredStrategy = OneDeviceStrategy   # in-host multi-device mirroring is not part of the reference
