# ncu --set full of the gated top-5 kernel in a 256-pair spatial batch (source-level samples)
B="python bench.py --no-cpu-baseline --no-extra --workload batch --pairs-per-gpu 256 --steps 1 --warmup 1"
mkdir -p gpurun_out
$B > gpurun_out/p10_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"nn_gated" -c 1 -o gpurun_out/prof10_top5 -f $B > gpurun_out/ncu10.log 2>&1
tail -3 gpurun_out/ncu10.log
