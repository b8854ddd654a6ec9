#!/usr/bin/env python
"""profiles/walk_ncu.json from an ncu --set full capture of the tree-walk kernel:

    python scripts/ncu_to_json.py gpurun_out/X.ncu-rep <rays of the captured launch> [launch index]

The record carries the hash of the kernel's sources (bench_gpu.walk_source_hash): bench.py only quotes
`roofline.traffic` and the issue-rate roofline while the sources it runs are the ones that were profiled."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench_gpu import WALK_NCU_JSON, walk_source_hash  # noqa: E402

rep, rays = sys.argv[1], float(sys.argv[2])
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
# the first-bounce nearest-hit kernel of the bench: the shadow-grid kernel when the capture has it, else the tree walk
walks = [r for r in rows[2:] if 'nearest_hit_grid' in r[hdr.index('Kernel Name')]] or \
        [r for r in rows[2:] if 'nearest_hit_walk' in r[hdr.index('Kernel Name')]]
r = walks[which]


def val(name):
    return float(r[hdr.index(name)].replace(',', ''))


def in_bytes(name):
    unit = rows[1][hdr.index(name)].lower()
    scale = {'byte': 1., 'kbyte': 1e3, 'mbyte': 1e6, 'gbyte': 1e9}[unit]
    return val(name) * scale


rec = {
    'capture': os.path.basename(rep), 'rays': rays, 'kernel': r[hdr.index('Kernel Name')].split('(')[0],
    'source_sha1': walk_source_hash(),
    'dram_bytes_per_ray': (in_bytes('dram__bytes_read.sum') + in_bytes('dram__bytes_write.sum')) / rays,
    'warp_inst_per_ray': val('smsp__inst_executed.sum') / rays,
    'duration_us_under_ncu': val('gpu__time_duration.sum') / (1e3 if rows[1][hdr.index('gpu__time_duration.sum')] in ('ns', 'nsecond') else 1.),
    'issue_active_pct': val('smsp__issue_active.avg.pct_of_peak_sustained_active'),
    'warps_active_pct': val('sm__warps_active.avg.pct_of_peak_sustained_active'),
    'registers_per_thread': val('launch__registers_per_thread'),
}
json.dump(rec, open(WALK_NCU_JSON, 'w'), indent=1)
print(json.dumps(rec, indent=1))
