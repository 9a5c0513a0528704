# A/B of kernel-2 stage length (KC) and top-5 row-split granularity on the k=50 workloads
mkdir -p gpurun_out
run() { python bench.py --workload $1 --steps $2 --warmup 3 --no-cpu-baseline --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3 $1 ms=%.4f pairs/s=%.1f nn_ms=%.4f'%(d['ms_per_step'], d['pairs_per_s'], d['roofline']['ms_per_launch']))"; }
for kc in 16 17; do for mt in 16 40; do
  F2P_GATED_KC=$kc F2P_TOP5_MIN_TILES=$mt run batch 5 "kc=$kc mt=$mt"
done; done
for kc in 16 17; do F2P_GATED_KC=$kc run pair 50 "kc=$kc"; done
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
F2P_TOP5_MIN_TILES=40 python -m pytest tests -m gpu -x -q -k "spatial or top5" 2>&1 | tail -1
