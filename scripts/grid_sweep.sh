#!/bin/bash
# usage: scripts/grid_sweep.sh "<nvcc -D flags A>" "<flags B>" ...   rebuilds with each flag set and prints the bench's
# step and first-bounce nearest-hit time; the default build is restored at the end.
for f in "$@"; do
  TRC_NVCC_EXTRA="$f" python tracer_b200/csrc/build.py --force > /dev/null || { echo "[$f] BUILD FAILED"; continue; }
  python bench.py --no-e2e --no-cpu --no-configs --steps 10 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.load(sys.stdin); print('[$f]', 'value %.4g  ms/step %.3f  kernel ms %.3f  frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac']))"
done
python tracer_b200/csrc/build.py --force > /dev/null
