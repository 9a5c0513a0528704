# the driver's round-end sequence: smoke, reference arm, own arm (default flags), plus the two side workloads
T=${1:-v9}
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
( time python bench.py --impl reference > gpurun_out/${T}_ref.json 2> gpurun_out/${T}_ref.err ) 2>&1 | grep real
( time python bench.py > gpurun_out/${T}_default.json 2> gpurun_out/${T}_default.err ) 2>&1 | grep real
python bench.py --workload pair --steps 50 --warmup 5 > gpurun_out/${T}_pair.json 2> gpurun_out/${T}_pair.err
python bench.py --workload batch --steps 5 --warmup 3 > gpurun_out/${T}_batch.json 2> gpurun_out/${T}_batch.err
tail -c 600 gpurun_out/${T}_ref.json; echo
python - <<PY
import json
for w in ("default","pair","batch"):
    d=json.loads(open("gpurun_out/${T}_%s.json"%w).read().strip().splitlines()[-1])
    r=d["roofline"]; print(w, "value=%.1f e2e=%.1f ms=%.3f pairs/s=%.1f frac=%.3f cpu=%s launches=%s clocks=%s"%(d["value"],d["e2e"]["value"],d["ms_per_step"],d.get("pairs_per_s",0),r["frac"],d.get("cpu_baseline",{}).get("value"),d.get("gpu_launches"),d.get("clocks")))
    if w=="default": print(json.dumps(d.get("other_workloads"))[:600])
PY
