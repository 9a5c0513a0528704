# ncu --set full of the screened vote kernel (first round of a 32-pair batch) + the verify kernel
B="python bench.py --no-cpu-baseline --workload batch --pairs-per-gpu 32 --steps 1 --warmup 1"
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "band_is_what" 2>&1 | tail -3
$B > gpurun_out/p8_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"vote_kernel|verify_kernel|classify" -c 6 -o gpurun_out/prof8_batch32 -f $B > gpurun_out/ncu8.log 2>&1
tail -3 gpurun_out/ncu8.log
ls -la gpurun_out | grep prof8
