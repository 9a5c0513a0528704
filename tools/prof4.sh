B="python bench.py --no-cpu-baseline"
mkdir -p gpurun_out
$B --workload batch --steps 2 --warmup 1 > gpurun_out/p4_b_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 700 --csv --log-file gpurun_out/launches4_batch.csv $B --workload batch --steps 2 --warmup 1 > gpurun_out/ncu4_b.log 2>&1
python tools/launch_summary.py gpurun_out/launches4_batch.csv
