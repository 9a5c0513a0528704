set -x
B="python bench.py --no-cpu-baseline"
$B --steps 2 --warmup 1 > gpurun_out/p_large_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_large.csv $B --steps 2 --warmup 1 > gpurun_out/ncu_l.log 2>&1
$B --workload batch --steps 2 --warmup 1 > gpurun_out/p_batch_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_batch.csv $B --workload batch --steps 2 --warmup 1 > gpurun_out/ncu_b.log 2>&1
$B --workload pair --steps 2 --warmup 1 > gpurun_out/p_pair_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_pair.csv $B --workload pair --steps 2 --warmup 1 > gpurun_out/ncu_p.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:nn_kernel -c 2 -o gpurun_out/prof_nn_large -f $B --steps 2 --warmup 1 > gpurun_out/ncu_f1.log 2>&1
$B --workload batch --pairs-per-gpu 32 --steps 2 --warmup 1 > gpurun_out/p_batch32_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"vote_kernel|nn_kernel|combine_kernel" -c 7 -o gpurun_out/prof_batch32 -f $B --workload batch --pairs-per-gpu 32 --steps 2 --warmup 1 > gpurun_out/ncu_f2.log 2>&1
ls -la gpurun_out
