#!/bin/bash
# usage: scripts/variant_probe.sh <rays> "<nvcc -D flags A>[@ENV=VAL ...]" ...   per-kernel device times (no PDL overlap) of
# a bench step for each build variant; the default build is restored at the end.
rays=$1; shift
for spec in "$@"; do
  f="${spec%%@*}"; e=""
  [[ "$spec" == *@* ]] && e="${spec#*@}"
  echo "=== variant [$f] env [$e]"
  TRC_NVCC_EXTRA="$f" python tracer_b200/csrc/build.py > /dev/null || { echo BUILD FAILED; continue; }
  env $e TRC_NO_PDL=1 TRC_PROBE_LIST=1 python scripts/timeline_probe.py $rays 2>&1 | grep -v -i warn | grep -A60 "^launches" | grep -E "walk|optics|order|partition|flux" | head -${TRC_PROBE_LINES:-12}
done
python tracer_b200/csrc/build.py > /dev/null
