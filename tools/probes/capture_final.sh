#!/bin/bash
# Final-build evidence in one gpurun call: test log, bench lines, launch list, one ncu --set full capture per kernel.
# Each ncu run follows a plain run of the same command line that exited 0 (B200_PROFILING.md).
set -u
T=${1:-r02p}
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/${T}_gpu_tests.log 2>&1; tail -2 $O/${T}_gpu_tests.log
python bench.py > $O/${T}_bench_default_1gpu.json 2> $O/${T}_bench.err
python bench.py --steps 20 --warmup 3 > $O/${T}_bench_steps20_1gpu.json 2>> $O/${T}_bench.err
S="--steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2"
python bench.py $S --no-extra-workloads > $O/plain_c3.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches_bench_c3.csv python bench.py $S --no-extra-workloads > /dev/null 2>&1
python bench.py $S --no-extra-workloads > $O/plain_c3.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:env_step_pair -s 5 -c 1 -f -o $O/${T}_c3_pair python bench.py $S --no-extra-workloads > /dev/null 2>&1
python bench.py $S --workload k1 > $O/plain_k1.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:env_step_pair -s 5 -c 1 -f -o $O/${T}_k1_pair python bench.py $S --workload k1 > /dev/null 2>&1
python bench.py $S --workload c5 > $O/plain_c5.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:step_guided -s 5 -c 1 -f -o $O/${T}_c5_guided python bench.py $S --workload c5 > /dev/null 2>&1
python bench.py $S --workload c4 > $O/plain_c4.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:policy_forward_kernel -s 3 -c 1 -f -o $O/${T}_c4_policy python bench.py $S --workload c4 > /dev/null 2>&1
python bench.py $S --workload c4 > $O/plain_c4.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:policy_forward_lean -s 3 -c 1 -f -o $O/${T}_c4_policy_lean python bench.py $S --workload c4 > /dev/null 2>&1
ls -la $O/${T}*
