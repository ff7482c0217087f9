#!/usr/bin/env bash
# usage: ./run.sh <num_ps> <num_workers> [extra role flags, e.g. --synthetic --max-steps 20]
# Same two knobs as the reference's run.sh; ranks [0,P) are parameter servers, [P,P+W) workers.
here="$(cd "$(dirname "$0")" && pwd)"
export PYTHONPATH="$here/..${PYTHONPATH:+:$PYTHONPATH}"
exec python3 -m ddl_b200.parallel.launch --variant-dir "$here" "$1" "$2" "${@:3}"
