"""Environment checks the README performs before training (README.md:35-40, 259-266, 185)."""
from __future__ import annotations


def tf_version() -> str:
    """R's ``tf_version()`` prints '2.0' (README.md:38-40); this framework tracks that API level."""
    return "2.0"


def install_check() -> dict:
    """What ``install_tensorflow()`` / the version print validate on each worker
    (README.md:185 installs lazily on executors; here nothing is downloaded — the native
    libraries are built in-tree on first use)."""
    import torch
    from .._native import build
    info = {"api_level": tf_version(), "torch": torch.__version__,
            "cuda_available": torch.cuda.is_available(),
            "gpus": torch.cuda.device_count() if torch.cuda.is_available() else 0}
    info["cpu_lib"] = str(build.build_cpu())
    lib = build.LIB_DIR / "libb200dist_cuda.so"
    info["cuda_lib"] = str(lib) if lib.exists() else None
    return info
