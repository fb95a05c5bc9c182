"""The C ABI contract: every exported ``b2d_*`` / ``b2c_*`` symbol of the two native libraries is declared in
``include/b200dist.h`` and every declared function is exported (no drift between the header that users
compile against and the kernels that exist)."""
import re
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _exported(lib: Path, prefix: str) -> set[str]:
    out = subprocess.run(["nm", "-D", "--defined-only", str(lib)], capture_output=True, text=True, check=True).stdout
    return {ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-1].startswith(prefix)}


@pytest.mark.skipif(shutil.which("nm") is None, reason="binutils not installed")
def test_header_declares_exactly_the_exported_entry_points():
    from distributed_b200._native import build
    libs = build.build_all(verbose=False)
    header = (ROOT / "include" / "b200dist.h").read_text()
    code = re.sub(r"/\*.*?\*/", "", header, flags=re.S)                 # prose in comments is not a declaration
    declared = set(re.findall(r"\b(b2[dc]_[a-z0-9_]+)\s*\(", code))
    exported = _exported(libs["cuda"], "b2d_") | _exported(libs["cpu"], "b2c_")
    assert exported - declared == set(), f"exported but not declared in b200dist.h: {sorted(exported - declared)}"
    assert declared - exported == set(), f"declared but not exported: {sorted(declared - exported)}"


def test_every_cuda_source_is_built_for_sm_100a_only():
    from distributed_b200._native import build
    flags = " ".join(build.NVCC_FLAGS)
    assert "arch=compute_100a,code=sm_100a" in flags and "-lineinfo" in flags
    assert flags.count("-gencode") == 1
