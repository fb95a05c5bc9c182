#!/usr/bin/env bash
# Race / memory checking of the single-GPU kernels.
# ONE tool per gpurun call (see B200_PROFILING.md): ./tools/sanitize.sh memcheck|racecheck|synccheck
# Runs the smallest test that exercises every kernel family.  The cross-GPU kernels are covered by
# the emulated (single cooperative launch) tests; the C++ CPU backend can be built with
# CXXFLAGS="-fsanitize=thread,address" via `CXX_EXTRA=... python -m distributed_b200._native.build`.
# NOTE (round 1): on the shared B200 pool `compute-sanitizer` is refused ("closed on this pool"), so no
# sanitizer log exists for this round.  Until it is available again the kernels are guarded by their own
# argument checks (B2D_REQUIRE), small-shape tests and comparison with the C++ CPU reference.
set -euo pipefail
tool="${1:-memcheck}"
cd "$(dirname "$0")/.."
exec compute-sanitizer --tool "$tool" --error-exitcode 1 \
  python -m pytest tests/test_nn_kernels_gpu.py tests/test_gemm_gpu.py -m gpu -x -q -k "64 or head or colsum"
