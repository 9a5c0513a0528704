# arg-min row-split count on the single 5k x 2k x 50 pair: step time and kernel-2 time per forced split
for ns in 0 8 10 13 14 16 18 20 27 40; do
  F2P_NSPLIT=$ns python bench.py --workload pair --steps 200 --warmup 20 --no-cpu-baseline --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ns=$ns ms=%.4f nn_ms=%.4f k2_stage=%.4f'%(d['ms_per_step'], d['roofline']['ms_per_launch'], d['roofline']['stages']['ms']['kernel2']))"
done
