#!/bin/bash
# usage: scripts/tiles_sweep.sh "<nvcc -D flags>" ...  step time at 1e8 and at 1.25e7 rays (the 8-GPU shard) per build
for f in "$@"; do
  TRC_NVCC_EXTRA="$f" python tracer_b200/csrc/build.py --force > /dev/null || { echo "[$f] BUILD FAILED"; continue; }
  for r in 1e8 1.25e7; do
  python bench.py --rays $r --no-e2e --no-cpu --no-configs --steps 30 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.load(sys.stdin); print('[$f] rays $r', 'value %.4g  ms/step %.3f  kernel ms %.3f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms']))"
  done
done
python tracer_b200/csrc/build.py --force > /dev/null
