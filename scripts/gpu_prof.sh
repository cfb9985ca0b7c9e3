# usage: bash scripts/gpu_prof.sh <name> <kernel-regex> <skip> <count>   (one ncu --set full capture; run the plain command first)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$2" -s $3 -c $4 -o gpurun_out/$1 $CMD > gpurun_out/ncu_$1.log 2>&1
tail -2 gpurun_out/ncu_$1.log | cut -c1-200
