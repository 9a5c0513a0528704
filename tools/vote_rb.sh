# A/B of the vote kernel's rows-per-block variants: kernel-3 parity tests + batch bench per variant
mkdir -p gpurun_out
for rb in 1 2 4; do
  F2P_VOTE_RB=$rb python -m pytest tests -m gpu -x -q -k "kernel3 or spatial" 2>&1 | tail -1
  F2P_VOTE_RB=$rb python bench.py --workload batch --steps 5 --warmup 3 --no-cpu-baseline --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('rb=$rb ms=%.3f pairs/s=%.1f'%(d['ms_per_step'], d['pairs_per_s']))"
done
