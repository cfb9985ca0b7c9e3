# ncu --set full + source of the first k_search launch of the timed workload
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-frontend --no-map-add"
$CMD > gpurun_out/plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"k_search" -s 2 -c 1 -f -o gpurun_out/k_search $CMD > gpurun_out/ncu_search.log 2>&1; tail -2 gpurun_out/ncu_search.log | cut -c1-160
