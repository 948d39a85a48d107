# A/B of the pair kernel's block shapes (ACB200_PAIR_VARIANT), C3 workload
run() { python bench.py --workload $1 --steps 300 --warmup 5 --no-cpu-baseline --e2e-steps 2 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$2', d['value'], d['ms_per_step'], d['roofline']['frac'])"; }
for v in ${VARIANTS:-0 1 2 5 6 7}; do ACB200_PAIR_VARIANT=$v run c3 "c3 variant=$v"; done
