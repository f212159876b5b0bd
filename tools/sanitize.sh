#!/usr/bin/env bash
# NOTE: compute-sanitizer is CLOSED on the B200 pool of this project (profiles/sanitizer/README.md); tools/bounds_check.py is the
# guard-band stand-in that runs there.  On a box where the tool is open:
# compute-sanitizer targets for the kernels.  ONE tool per invocation (and per gpurun call): running several
# tools back to back in one call has left B200 boxes unusable (B200_PROFILING.md), so pick one:
#
#   tools/sanitize.sh memcheck   [pytest -k expression]
#   tools/sanitize.sh racecheck  [pytest -k expression]
#   tools/sanitize.sh synccheck  [pytest -k expression]
#   tools/sanitize.sh initcheck  [pytest -k expression]
#
# The default selection is the small-shape kernel tests (a few seconds each without the tool, minutes with it); the plain
# run comes first and the tool only runs if it passed.  Output: gpurun_out/sanitize_<tool>.log
set -euo pipefail
tool="${1:-memcheck}"
sel="${2:-fedavg or median or xent or adam or gemm_small or im2col or batchnorm or implicit_gemm}"
mkdir -p gpurun_out
python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "$sel" > gpurun_out/sanitize_plain.log 2>&1 &&
compute-sanitizer --tool "$tool" --error-exitcode 1 --log-file "gpurun_out/sanitize_${tool}.log" \
    python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "$sel" > "gpurun_out/sanitize_${tool}_pytest.log" 2>&1
status=$?
tail -n 5 "gpurun_out/sanitize_${tool}.log" || true
exit $status
