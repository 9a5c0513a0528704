for ns in 0 5 7 9 10 14 18 20; do
  F2P_NSPLIT=$ns python bench.py --workload pair --steps 100 --warmup 10 --no-cpu-baseline --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ns=$ns ms=%.4f nn_ms=%.4f'%(d['ms_per_step'], d['roofline']['ms_per_launch']))"
done
