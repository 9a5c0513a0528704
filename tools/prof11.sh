# ncu --set full of the first (full-size) vote kernel launch in a 64-pair spatial batch (source-level samples)
B="python bench.py --no-cpu-baseline --no-extra --workload batch --pairs-per-gpu 64 --steps 1 --warmup 1"
mkdir -p gpurun_out
$B > gpurun_out/p11_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"vote_kernel" -c 1 -o gpurun_out/prof11_vote -f $B > gpurun_out/ncu11.log 2>&1
tail -3 gpurun_out/ncu11.log
