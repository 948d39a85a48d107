# GPU suite + the four workloads, pair kernels (default) against the one-aircraft kernels (ACB200_ENV_PAIR=0)
run() { python bench.py --workload $1 --steps 300 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$2', d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['kernel'])"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run c3 "c3 pair"; ACB200_ENV_PAIR=0 run c3 "c3 one"; run k1 "k1 pair"; ACB200_ENV_PAIR=0 run k1 "k1 one"; run c5 "c5 pair"; ACB200_ENV_PAIR=0 run c5 "c5 one"; run c4 "c4"
