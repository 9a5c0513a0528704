# ncu --set full of the gated arg-min kernel on the 100k x 100k x 128 pair (source-level samples)
B="python bench.py --no-cpu-baseline --no-extra --workload large --steps 1 --warmup 1"
mkdir -p gpurun_out
$B > gpurun_out/p12_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"nn_gated" -c 1 -o gpurun_out/prof12_large -f $B > gpurun_out/ncu12.log 2>&1
tail -3 gpurun_out/ncu12.log
