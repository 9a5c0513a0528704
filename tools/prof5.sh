# batch32 capture of the vote and top-5 kernels + side timings of the ground-truth kernels
B="python bench.py --no-cpu-baseline"
mkdir -p gpurun_out
python tools/time_gt.py > gpurun_out/time_gt.json 2> gpurun_out/time_gt.err
$B --workload batch --pairs-per-gpu 32 --steps 1 --warmup 1 > gpurun_out/p5_b_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"vote_kernel|nn_gated" -c 4 -o gpurun_out/prof5_batch32 -f $B --workload batch --pairs-per-gpu 32 --steps 1 --warmup 1 > gpurun_out/ncu5_b.log 2>&1
ls -la gpurun_out | tail -5
cat gpurun_out/time_gt.json
