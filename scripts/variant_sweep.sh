#!/bin/bash
# usage: scripts/variant_sweep.sh "<nvcc -D flags A>" "<flags B>" ...
# rebuilds libtracer_b200.so with each flag set and prints the per-call timing of scripts/timing_probe.py;
# the default build is restored at the end.
for f in "$@"; do
  echo "=== variant: $f"
  TRC_NVCC_EXTRA="$f" python tracer_b200/csrc/build.py --force > /dev/null || { echo BUILD FAILED; continue; }
  python scripts/timing_probe.py 1e8 2>&1 | grep -v Warning | head -8
done
python tracer_b200/csrc/build.py --force > /dev/null
