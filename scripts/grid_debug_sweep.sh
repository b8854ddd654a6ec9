#!/bin/bash
# usage: scripts/grid_debug_sweep.sh "<nvcc -D flags>" ...   first-bounce kernel time of the bench per build, as it is
# and with TRC_WALK_DEBUG=4 (the shadow-grid kernel only streams the rays): separates the cost of the stream from the rest
for f in "$@"; do
  TRC_NVCC_EXTRA="$f" python tracer_b200/csrc/build.py --force > /dev/null || { echo "[$f] BUILD FAILED"; continue; }
  for d in 0 4; do
  TRC_WALK_DEBUG=$d python bench.py --no-e2e --no-cpu --no-configs --steps 10 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.load(sys.stdin); print('[$f] debug $d', 'ms/step %.3f  kernel ms %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms']))"
  done
done
python tracer_b200/csrc/build.py --force > /dev/null
