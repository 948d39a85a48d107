# r01g measurement set (one B200): GPU tests, the four bench workloads, the scalar-kernel A/B, launch list, one full ncu capture
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r01g_gpu_tests.log; cat gpurun_out/r01g_gpu_tests.log
python bench.py > gpurun_out/r01g_bench_c3.json 2> gpurun_out/r01g_bench_c3.err
for w in k1 c4 c5; do python bench.py --workload $w > gpurun_out/r01g_bench_$w.json 2> gpurun_out/r01g_bench_$w.err; done
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r01g_bench_c3_reference.json 2>/dev/null
ACB200_ENV_PAIR=0 python bench.py --no-cpu-baseline --steps 300 > gpurun_out/r01g_bench_c3_scalar.json 2>/dev/null
python bench.py --precision 64 --envs 262144 --steps 100 --no-cpu-baseline > gpurun_out/r01g_bench_c3_f64.json 2>/dev/null
for f in c3 k1 c4 c5 c3_scalar c3_f64; do python -c "import json; d=json.load(open('gpurun_out/r01g_bench_$f.json')); print('$f', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['roofline']['kernel'])"; done
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 2"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01g_launches_bench_c3.csv $CMD > gpurun_out/ncu_launches.log 2>&1
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 2"
ncu --set full --clock-control none --import-source on -k regex:env_step_pair -s 4 -c 1 -o gpurun_out/r01g_pair_final $CMD > gpurun_out/ncu.log 2>&1; tail -1 gpurun_out/ncu.log
