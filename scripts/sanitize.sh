#!/bin/bash
# compute-sanitizer over the small-size kernel targets (tests/test_gpu_sanitize.py touches every kernel family once and checks each
# result against the oracle).  ONE tool per invocation and per gpurun call (B200_PROFILING.md: running several tools in one call has
# left a GPU unusable): usage  scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expression]
# Run the plain command first (`python -m pytest tests/test_gpu_sanitize.py -m gpu -q`) and only sanitize what exits 0.
# The mbarrier-parity bug of round 1 (wrong result one run in eight at 2^24) is the kind of thing racecheck / synccheck are for.
set -e
TOOL=${1:-memcheck}
KEXPR=${2:-"not tensor_core"}        # the log-22 tensor-core case is minutes under racecheck: ask for it explicitly with -k tensor_core
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
exec timeout ${SANITIZE_TIMEOUT:-900} compute-sanitizer --tool "$TOOL" --error-exitcode 86 --print-limit 20 \
  --log-file gpurun_out/sanitize_$TOOL.log \
  python -m pytest tests/test_gpu_sanitize.py -m gpu -q -x -k "$KEXPR"
