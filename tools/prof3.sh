B="python bench.py --no-cpu-baseline"
mkdir -p gpurun_out
$B --workload batch --pairs-per-gpu 32 --steps 1 --warmup 1 > gpurun_out/p3_b_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"vote_kernel|nn_gated" -c 4 -o gpurun_out/prof3_batch32 -f $B --workload batch --pairs-per-gpu 32 --steps 1 --warmup 1 > gpurun_out/ncu3_b.log 2>&1
$B --steps 1 --warmup 1 > gpurun_out/p3_l_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"nn_gated" -c 1 -o gpurun_out/prof3_large -f $B --steps 1 --warmup 1 > gpurun_out/ncu3_l.log 2>&1
ls -la gpurun_out | tail -5
