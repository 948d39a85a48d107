run() { python bench.py --workload $1 --steps 300 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$2', d['value'], d['ms_per_step'], d['roofline']['frac'])"; }
python -m pytest tests/test_gpu_pair_kernel.py -x -q 2>&1 | tail -3
for v in 0 1 2; do ACB200_PAIR_VARIANT=$v run c3 "c3 rolled variant=$v"; done
for v in 0 1; do ACB200_LIB=$PWD/hybrid-classical-and-reinforcement-learning-aircraft-controllers_b200/libacb200_unroll.so ACB200_PAIR_VARIANT=$v run c3 "c3 unrolled variant=$v"; done
ACB200_PAIR_VARIANT=0 run k1 "k1 variant=0"
ACB200_PAIR_VARIANT=1 run k1 "k1 variant=1"
