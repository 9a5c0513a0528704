# launch lists of the three workloads (per-launch durations; the SHARE of each kernel is what is compared)
B="python bench.py --no-cpu-baseline --no-extra"
mkdir -p gpurun_out
for w in large batch pair; do
  $B --workload $w --steps 2 --warmup 1 > gpurun_out/p6_${w}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches6_$w.csv $B --workload $w --steps 2 --warmup 1 > gpurun_out/ncu6_$w.log 2>&1
done
ls -la gpurun_out | tail -4
