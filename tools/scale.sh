# multi-GPU bench lines (torchrun, one rank per GPU): usage bash tools/scale.sh N tag
N=${1:-2}; T=${2:-v9}
mkdir -p gpurun_out
for w in large batch; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-extra > gpurun_out/${T}_n${N}_$w.json 2> gpurun_out/${T}_n${N}_$w.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${T}_n${N}_$w.json").read().strip().splitlines()[-1])
    print("$w n=$N value=%.1f e2e=%.1f ms=%.3f pairs/s=%.1f"%(d["value"],d["e2e"]["value"],d["ms_per_step"],d.get("pairs_per_s",0)))
except Exception as e:
    print("$w FAILED", e); print(open("gpurun_out/${T}_n${N}_$w.err").read()[-1500:])
PY
done
