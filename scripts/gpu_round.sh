#!/bin/bash
# Run under gpurun: GPU tests, bench, launch list, one ncu --set full capture of the fused kernel.
set -x
mkdir -p gpurun_out
TAG=${1:-run}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python bench.py --steps 1000 --warmup 200 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 300 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 40 --warmup 110 > gpurun_out/${TAG}_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tg_k_eval -s 30 -c 1 -o gpurun_out/${TAG}_eval_full -f python bench.py --steps 40 --warmup 110 > gpurun_out/${TAG}_ncu_full.log 2>&1
ls -la gpurun_out
