# one gpurun call: ncu --set full of the kernels of the first two trips (search trip + re-use trip) of one update
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_search|k_fused|k_solve" -s 0 -c 8 -o gpurun_out/trip_kernels $CMD > gpurun_out/ncu_full.log 2>&1; tail -2 gpurun_out/ncu_full.log | cut -c1-160
