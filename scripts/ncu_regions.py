"""Stall samples of one kernel by code region.  Input: `ncu -i X.ncu-rep --page source --csv` output.
Regions are cut at SASS landmarks given on the command line as name=substring pairs (first match of the
substring starts the region); prints samples, instructions executed and the main stall reasons per region."""
import csv
import sys


def f(x):
    try:
        return float(x)
    except ValueError:
        return 0.


rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
cuts = []
for arg in sys.argv[2:]:
    name, sub = arg.split('=', 1)
    if sub.startswith('@'):               # explicit row number
        cuts.append((int(sub[1:]), name))
        continue
    nth = 1
    if '#' in sub:
        sub, nth = sub.rsplit('#', 1)
        nth = int(nth)
    seen = 0
    for i, r in enumerate(data):
        if sub in r[ix['Source']]:
            seen += 1
            if seen == nth:
                cuts.append((i, name))
                break
cuts.sort()
cuts = [(0, 'prologue')] + cuts + [(len(data), 'end')]
stalls = [h for h in hdr if h.startswith('stall_')]
tot = sum(f(r[ix['# Samples']]) for r in data)
tot_inst = sum(f(r[ix['Instructions Executed']]) for r in data)
print('total samples %d, warp instructions %d' % (tot, tot_inst))
for (a, name), (b, _) in zip(cuts[:-1], cuts[1:]):
    seg = data[a:b]
    smp = sum(f(r[ix['# Samples']]) for r in seg)
    inst = sum(f(r[ix['Instructions Executed']]) for r in seg)
    thr = sum(f(r[ix['Thread Instructions Executed']]) for r in seg)
    st = sorted(((sum(f(r[ix[s]]) for r in seg), s) for s in stalls), reverse=True)[:4]
    print('%-14s rows %4d-%4d  samples %5.1f%%  inst %5.1f%%  lanes %4.1f  %s' % (
        name, a, b, 100 * smp / tot, 100 * inst / max(tot_inst, 1), thr / max(inst, 1),
        ' '.join('%s %.1f%%' % (s[6:], 100 * v / tot) for v, s in st)))
