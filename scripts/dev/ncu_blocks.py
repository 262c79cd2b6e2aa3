"""Developer helper: per-block instruction accounting of one kernel from `ncu --page source --csv`."""
import csv, collections, sys
src, unit = sys.argv[1], float(sys.argv[2])
rows = list(csv.reader(open(src)))
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr) and r[hdr.index('Instructions Executed')].isdigit()]
iI = hdr.index("Instructions Executed"); iS = hdr.index("Warp Stall Sampling (All Samples)")
tot = sum(int(r[iI]) for r in data); tots = sum(int(r[iS]) for r in data)
print("total warp-inst", tot, "per unit %.1f" % (tot / unit), "samples", tots)
def op(r):
    t = r[1].split()
    return (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
blocks = []; cur = None
for k, r in enumerate(data):
    n = int(r[iI])
    if cur and abs(n - cur['n']) <= 0.02 * max(n, cur['n'], 1):
        cur['len'] += 1; cur['end'] = k; cur['smp'] += int(r[iS]); cur['ops'][op(r)] += 1
    else:
        cur = {'n': n, 'len': 1, 'start': k, 'end': k, 'smp': int(r[iS]), 'ops': collections.Counter()}; blocks.append(cur); cur['ops'][op(r)] += 1
for b in blocks:
    w = b['n'] * b['len']
    if w / tot > 0.004:
        fp = sum(v for o, v in b['ops'].items() if o in ('DFMA', 'DMUL', 'DADD', 'DSETP'))
        print("%5d-%5d len %4d exec/unit %6.2f inst %5.1f%% smp %5.1f%% fp64 %3d %s" % (b['start'], b['end'], b['len'], b['n'] / unit, 100 * w / tot, 100 * b['smp'] / tots, fp, dict(b['ops'].most_common(6))))
