# cfg3 at its full size: 4096 pairs in one call on one GPU
python bench.py --workload batch --pairs-per-gpu 4096 --steps 3 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/${1:-x}_bench_batch4096.json 2> gpurun_out/${1:-x}_bench_batch4096.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${1:-x}_bench_batch4096.json").read().strip().splitlines()[-1])
print("ms", d["ms_per_step"], "pairs/s", d["pairs_per_s"], "e2e pairs/s", d["e2e"]["pairs_per_s"], d["roofline"]["stages"]["ms"])
PY
