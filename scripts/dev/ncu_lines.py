"""Developer helper: per-source-line instruction / stall-sample accounting of one kernel from
`ncu -i REP --page source --csv --kernel-name regex:K --print-source cuda,sass > file.csv`.
usage: ncu_lines.py file.csv [units] [min_pct]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
unit = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.7
cur_file = None; hdr = None; seen = set()
acc = collections.OrderedDict()
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if len(r) > 10 and r[0] == 'Line No': hdr = r; iI = hdr.index('Instructions Executed'); iS = hdr.index('Warp Stall Sampling (All Samples)'); continue
    if hdr is None or len(r) != len(hdr): continue
    if r[0] != '' and r[2] == '-':      # a source line: its totals
        key = (cur_file, int(r[0]))
        if key in seen: continue
        seen.add(key)
        acc[key] = (int(r[iI]), int(r[iS]), r[1].strip())
tot = sum(v[0] for v in acc.values()); tots = sum(v[1] for v in acc.values()) or 1
print("total warp-inst %d (%.1f per unit), samples %d" % (tot, tot / unit, tots))
for (f, ln), (n, s, src) in acc.items():
    if 100.0 * n / tot >= minp or 100.0 * s / tots >= minp:
        print("%-16s %4d inst %5.1f%% (%7.1f/unit) smp %5.1f%%  %s" % (f[:16], ln, 100.0 * n / tot, n / unit, 100.0 * s / tots, src[:110]))
