# ncu --set full of the prep kernels (projection + transpose) in a 256-pair batch
B="python bench.py --no-cpu-baseline --no-extra --workload batch --pairs-per-gpu 256 --steps 1 --warmup 1"
mkdir -p gpurun_out
$B > gpurun_out/p9_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"project_kernel|transpose_kernel" -c 2 -o gpurun_out/prof9_prep -f $B > gpurun_out/ncu9.log 2>&1
tail -3 gpurun_out/ncu9.log
ls -la gpurun_out | grep prof9
