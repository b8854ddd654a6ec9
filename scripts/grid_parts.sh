#!/bin/bash
# first-bounce kernel time of the bench with parts of the shadow-grid kernel switched off (TRC_WALK_DEBUG bits:
# 4 stream only, 8 no record emission, 16 no float64 stage, 32 no cell fetch); results are wrong by design
for d in 0 8 16 24 32 4; do
  TRC_WALK_DEBUG=$d python bench.py --no-e2e --no-cpu --no-configs --steps 10 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.load(sys.stdin); print('debug $d', 'ms/step %.3f  kernel ms %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms']))"
done
