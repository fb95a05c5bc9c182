import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    import torch
    # References must be true fp32: cuDNN / cuBLAS default to TF32 for convolutions.
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    config.addinivalue_line("markers", "gpu: test needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "multigpu: test needs >= 2 GPUs")


def pytest_collection_modifyitems(config, items):
    import torch
    has_gpu = torch.cuda.is_available()
    ngpu = torch.cuda.device_count() if has_gpu else 0
    skip_gpu = pytest.mark.skip(reason="no GPU in this container")
    skip_multi = pytest.mark.skip(reason="needs >= 2 GPUs")
    for it in items:
        if "gpu" in it.keywords and not has_gpu:
            it.add_marker(skip_gpu)
        if "multigpu" in it.keywords and ngpu < 2:
            it.add_marker(skip_multi)
