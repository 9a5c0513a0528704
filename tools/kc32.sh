# A/B of the k = 128 stage layout: 8 x 16 rows, 4-deep ring (default) against 4 x 32 rows, 2-deep ring (same shared memory)
run() { python bench.py --workload large --steps 5 --warmup 3 --no-cpu-baseline --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1 ms=%.4f k2_stage=%.4f'%(d['ms_per_step'], d['roofline']['stages']['ms']['kernel2']))"; }
for kc in 16 32 16 32; do F2P_GATED_KC=$kc run "kc=$kc"; done
F2P_GATED_KC=32 python -m pytest tests/test_gpu_full_size.py -m gpu -x -q -k "cfg4 or stage_length" 2>&1 | tail -2
