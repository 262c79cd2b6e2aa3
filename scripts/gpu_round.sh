#!/bin/bash
# Run under gpurun: GPU tests, bench (driver protocol + long run), reference arm, launch list, one ncu --set full capture of the
# kernels of one step.  Usage: gpu_round.sh TAG
mkdir -p gpurun_out
TAG=${1:-run}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_gputests.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_gputests.log
tail -3 gpurun_out/${TAG}_gputests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm exit $?"
python bench.py --no-cfg5 > gpurun_out/${TAG}_bench_2000steps.json 2> gpurun_out/${TAG}_bench_2000steps.err; echo "long bench exit $?"
# launch list of the bench command (after it exited 0 without ncu)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 40 --warmup 40 --no-cfg5 > gpurun_out/${TAG}_ncu_launch.log 2>&1
# one step of cfg3 without graphs: every kernel of the step under --set full
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:tg_k_(ff|eval|finish|propose)" -s 100 -c 5 -o gpurun_out/${TAG}_step_full -f python scripts/dev/steps_for_ncu.py 10 1 > gpurun_out/${TAG}_ncu_step.log 2>&1
# the fused kernel and the pre-pass at 8 stars x 10 chains (cfg5)
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:^tg_k_(ff|eval)$" -s 90 -c 2 -o gpurun_out/${TAG}_cfg5_full -f python scripts/dev/steps_for_ncu.py 10 8 > gpurun_out/${TAG}_ncu_cfg5.log 2>&1
ls -la gpurun_out | tail -14
