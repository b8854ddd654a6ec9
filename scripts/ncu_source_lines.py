"""Top source lines of a kernel by executed instructions and by stall samples.
    ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > x.csv ; python scripts/ncu_source_lines.py x.csv [n]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
cur, out = None, []
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path':
        cur = r[1].split('/')[-1]
    elif len(r) > 10 and r[0].isdigit():
        try:
            out.append((float(r[7] or 0), float(r[4] or 0), cur, int(r[0]), r[1].strip()[:100], float(r[8] or 0)))
        except ValueError:
            pass
ti, ts = sum(o[0] for o in out), sum(o[1] for o in out)
print('warp instructions %.4g, stall samples %.4g' % (ti, ts))
print('--- by instructions executed (share, share of samples, average active lanes)')
for o in sorted(out, reverse=True)[:top]:
    print('%5.1f%% inst %5.1f%% smp lanes %4.1f  %s:%d  %s' % (100 * o[0] / ti, 100 * o[1] / ts, o[5] / max(o[0], 1), o[2], o[3], o[4]))
print('--- by stall samples')
for o in sorted(out, key=lambda o: -o[1])[:top]:
    print('%5.1f%% inst %5.1f%% smp lanes %4.1f  %s:%d  %s' % (100 * o[0] / ti, 100 * o[1] / ts, o[5] / max(o[0], 1), o[2], o[3], o[4]))
