# usage: bash tools/stage_check.sh <tag>  (GPU box): the new tests, then the three bench workloads with the stage breakdown
T=${1:-v11}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "stage_timing or teaser" 2>&1 | tail -5
python bench.py --workload batch --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_batch.json 2> gpurun_out/${T}_batch.err || tail -20 gpurun_out/${T}_batch.err
python bench.py --workload pair --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/${T}_pair.json 2> gpurun_out/${T}_pair.err || tail -20 gpurun_out/${T}_pair.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/${T}_large.json 2> gpurun_out/${T}_large.err || tail -20 gpurun_out/${T}_large.err
python - <<PY
import json
for w in ("batch","pair","large"):
    d=json.loads(open("gpurun_out/${T}_%s.json"%w).read().strip().splitlines()[-1]); r=d["roofline"]
    print(w, "ms/step=%.3f pairs/s=%.0f e2e=%.1f k2: %.2f TF frac=%.3f ms=%.4f"%(d["ms_per_step"], d["pairs_per_s"], d["e2e"]["pairs_per_s"], r["achieved"], r["frac"], r["ms_per_launch"]))
    print("   stages", {k: round(v,4) for k,v in r["stages"]["ms"].items()}, "evals", r["stages"]["filter_evals_per_step"], "amb", r["stages"]["ambiguous_columns_per_step"])
    if "filter" in r: print("   filter", {k:(round(v,4) if isinstance(v,float) else v) for k,v in r["filter"].items() if k!="note"})
PY
