# usage: bash tools/gpu_check.sh <tag>   (on the GPU box): parity tests + the three bench workloads
T=${1:-x}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_large.json 2> gpurun_out/${T}_bench_large.err
python bench.py --workload pair --steps 50 --warmup 5 > gpurun_out/${T}_bench_pair.json 2> gpurun_out/${T}_bench_pair.err
python bench.py --workload batch --steps 5 --warmup 3 > gpurun_out/${T}_bench_batch.json 2> gpurun_out/${T}_bench_batch.err
python - <<PY
import json
for w in ("large","pair","batch"):
    try:
        d=json.loads(open("gpurun_out/${T}_bench_%s.json"%w).read().strip().splitlines()[-1])
        r=d["roofline"]
        print(w, "value=%.1f e2e=%.1f ms=%.3f pairs/s=%.1f nn=%.2fTF frac=%.3f nn_ms=%.4f proj=%.0fGB/s"%(d["value"],d["e2e"]["value"],d["ms_per_step"],d["pairs_per_s"],r["achieved"],r["frac"],r["ms_per_launch"],r["projection"]["achieved"]))
    except Exception as e:
        print(w,"FAILED",e); print(open("gpurun_out/${T}_bench_%s.err"%w).read()[-1500:])
PY
