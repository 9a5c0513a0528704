B="python bench.py --no-cpu-baseline"
mkdir -p gpurun_out
$B --steps 1 --warmup 1 > gpurun_out/p2_large_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches2_large.csv $B --steps 1 --warmup 1 > gpurun_out/ncu2_l.log 2>&1
$B --workload batch --pairs-per-gpu 32 --steps 1 --warmup 1 > gpurun_out/p2_b_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches2_batch32.csv $B --workload batch --pairs-per-gpu 32 --steps 1 --warmup 1 > gpurun_out/ncu2_b.log 2>&1
python tools/launch_summary.py gpurun_out/launches2_large.csv
python tools/launch_summary.py gpurun_out/launches2_batch32.csv
