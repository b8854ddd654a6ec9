#!/bin/bash
# usage: scripts/variant_probe.sh <rays> "<nvcc -D flags A>" ...   per-kernel device times (no PDL overlap) of a
# bench step for each build variant; the default build is restored at the end.
rays=$1; shift
for f in "$@"; do
  echo "=== variant [$f]"
  TRC_NVCC_EXTRA="$f" python tracer_b200/csrc/build.py > /dev/null || { echo BUILD FAILED; continue; }
  TRC_NO_PDL=1 TRC_PROBE_LIST=1 python scripts/timeline_probe.py $rays 2>&1 | grep -v -i warn | grep -B0 -A60 "^span" | grep -v "largest\|after"
done
python tracer_b200/csrc/build.py > /dev/null
