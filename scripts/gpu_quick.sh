#!/bin/bash
# quick developer loop under gpurun: GPU tests, bench, launch list of ~50 steps
TAG=${1:-q}
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --steps 1000 --warmup 200 > gpurun_out/${TAG}_bench.json 2>gpurun_out/${TAG}_bench.err; tail -3 gpurun_out/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
r=d["roofline"]
print("value %.0f chain-steps/s  ms/step %.4f  e2e %.0f  eval %.1f us (share %.2f) frac %.3f  cpu %.1f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], r["avg_launch_us"], r["kernel_share_of_step"], r["frac"], d["cpu_baseline"]["value"]))
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 200 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 40 --warmup 110 > /dev/null 2>&1
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/${TAG}_launches.csv")) if len(r)>10 and r[0].isdigit()]
t=collections.defaultdict(list)
for r in rows: t[r[4]].append(float(r[-1])/1e3)
for k,v in t.items(): print("%-40s n=%3d avg %.1f us" % (k[:40], len(v), sum(v)/len(v)))
PY
