# A/B of the pair kernel's block shapes (ACB200_PAIR_VARIANT), C3 workload
run() { python bench.py --workload $1 --steps 300 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$2', d['value'], d['ms_per_step'], d['roofline']['frac'])"; }
python -m pytest tests/test_gpu_pair_kernel.py tests/test_gpu_parity_f32.py -x -q 2>&1 | tail -3
for v in ${VARIANTS:-0 1 2 3 4 5}; do ACB200_PAIR_VARIANT=$v run c3 "c3 variant=$v"; done
for v in ${VARIANTS_K1:-1 3}; do ACB200_PAIR_VARIANT=$v run k1 "k1 variant=$v"; done
run k1 "k1 default"
