B="python bench.py --no-cpu-baseline --no-extra"
mkdir -p gpurun_out
$B --workload pair --steps 2 --warmup 1 > gpurun_out/p7_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"nn_gated" -c 2 -o gpurun_out/prof7_pair -f $B --workload pair --steps 2 --warmup 1 > gpurun_out/ncu7.log 2>&1
ls -la gpurun_out | tail -3
