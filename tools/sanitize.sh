#!/usr/bin/env bash
# compute-sanitizer targets for the native kernels.  ONE tool per GPU-box call (see
# B200_PROFILING.md), on the smallest case that exercises every kernel: a 2-step single-GPU run.
#   tools/sanitize.sh memcheck | racecheck | synccheck | initcheck
# Run the same command once without the sanitizer first; only then wrap it.
set -euo pipefail
tool="${1:-memcheck}"
here="$(cd "$(dirname "$0")/.." && pwd)"
cmd=(python "$here/bench.py" --steps 2 --warmup 3 --no-graphs --no-flush)
"${cmd[@]}" > /dev/null
exec compute-sanitizer --tool "$tool" --error-exitcode 1 --print-limit 20 "${cmd[@]}"
