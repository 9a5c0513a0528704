# A/B of the k = 50 stage layout: 2 x 25 rows, 4-deep ring (default) against 1 x 50 rows, 2-deep ring
run() { python bench.py --workload $1 --steps $2 --warmup 3 --no-cpu-baseline --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3 $1 ms=%.4f k2_stage=%.4f'%(d['ms_per_step'], d['roofline']['stages']['ms']['kernel2']))"; }
for kc in 25 50 25 50; do F2P_GATED_KC=$kc run batch 5 "kc=$kc"; done
for kc in 25 50; do F2P_GATED_KC=$kc run pair 100 "kc=$kc"; done
F2P_GATED_KC=50 python -m pytest tests -m gpu -x -q -k "golden or cfg5 or full_batch" 2>&1 | tail -2
