#!/bin/bash
# usage: scripts/sweep.sh "<bench args A>" "<bench args B>" ...   prints value / ms_per_step / kernel_ms per arm
for a in "$@"; do
  python bench.py $a --no-e2e --no-cpu 2>&1 | tail -1 > /tmp/line.json
  python - "$a" <<'PY'
import json, sys
try:
    d = json.load(open('/tmp/line.json'))
    print(sys.argv[1], '| value %.4g | ms/step %.3f | hit kernel ms %.3f | frac %.3f' % (
        d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e, open('/tmp/line.json').read()[-500:])
PY
done
