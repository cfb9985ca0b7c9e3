timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
CMD="python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
timeout 400 $CMD > gpurun_out/b.json 2> gpurun_out/b.err; echo rc=$?
python - <<EOP
import json
d=json.loads(open("gpurun_out/b.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"], d["roofline"], d["search_stats"])
EOP
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:"flivo|cub::|k_" -c 300 --csv --log-file gpurun_out/launches_r1.csv $CMD > gpurun_out/ncu1.log 2>&1; tail -2 gpurun_out/ncu1.log | cut -c1-200
