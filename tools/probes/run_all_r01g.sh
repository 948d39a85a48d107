python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r01g_gpu_tests.log; cat gpurun_out/r01g_gpu_tests.log
python bench.py > gpurun_out/r01g_bench_c3.json 2> gpurun_out/r01g_bench_c3.err; python -c "import json; d=json.load(open('gpurun_out/r01g_bench_c3.json')); print('c3', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'])"
for w in k1 c4 c5; do python bench.py --workload $w > gpurun_out/r01g_bench_$w.json 2> gpurun_out/r01g_bench_$w.err; python -c "import json; d=json.load(open('gpurun_out/r01g_bench_$w.json')); print('$w', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'])"; done
ACB200_ENV_PAIR=0 python bench.py --no-cpu-baseline --steps 300 > gpurun_out/r01g_bench_c3_scalar.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r01g_bench_c3_scalar.json')); print('c3 scalar', d['value'], d['ms_per_step'], d['roofline']['frac'])"
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01g_launches_bench_c3.csv $CMD > gpurun_out/ncu_launches.log 2>&1
./tools/probes/ffma2_probe > gpurun_out/r01g_ffma2_probe.json
