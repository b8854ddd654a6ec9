#!/bin/bash
# usage: scripts/variant_sweep.sh "<nvcc -D flags A>[@ENV=VAL ...]" ...
# rebuilds libtracer_b200.so with each flag set (text before '@'), exports the env assignments after '@',
# and prints the per-call timing of scripts/timing_probe.py; the default build is restored at the end.
for spec in "$@"; do
  f="${spec%%@*}"; e=""
  [[ "$spec" == *@* ]] && e="${spec#*@}"
  echo "=== variant: flags [$f] env [$e]"
  TRC_NVCC_EXTRA="$f" python tracer_b200/csrc/build.py --force > /dev/null || { echo BUILD FAILED; continue; }
  env $e python scripts/timing_probe.py 1e8 2>&1 | grep -v Warning | head -4
done
python tracer_b200/csrc/build.py --force > /dev/null
