#!/bin/bash
# Run under gpurun: GPU tests, bench, launch list, one ncu --set full capture of each kernel of the step.
set -x
mkdir -p gpurun_out
TAG=${1:-run}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
python bench.py --impl reference --steps 8 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm exit $?"
cat gpurun_out/${TAG}_bench_reference.json
# launch list of the same command (after it exited 0 without ncu)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 300 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 40 --warmup 110 > gpurun_out/${TAG}_ncu_launch.log 2>&1
for K in eval propose finish; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:tg_k_${K} -s 130 -c 1 -o gpurun_out/${TAG}_${K}_full -f python bench.py --steps 40 --warmup 110 > gpurun_out/${TAG}_ncu_${K}.log 2>&1
done
ls -la gpurun_out | tail -12
