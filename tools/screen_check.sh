# usage: bash tools/screen_check.sh <tag>  (GPU box): spatial parity tests, then the batch workload with and without the screened vote
T=${1:-s}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "spatial or kernel3 or stacked or stage or host or cfg5" 2>&1 | tail -4
for S in 1 0; do
F2P_VOTE_SCREEN=$S python bench.py --workload batch --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_batch_s$S.json 2> gpurun_out/${T}_batch_s$S.err || tail -20 gpurun_out/${T}_batch_s$S.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_batch_s$S.json").read().strip().splitlines()[-1]); r=d["roofline"]
print("screen=$S ms/step=%.3f pairs/s=%.0f e2e=%.1f"%(d["ms_per_step"], d["pairs_per_s"], d["e2e"]["pairs_per_s"]))
print("   stages", {k: round(v,4) for k,v in r["stages"]["ms"].items()}, "amb", r["stages"]["ambiguous_columns_per_step"])
PY
done
