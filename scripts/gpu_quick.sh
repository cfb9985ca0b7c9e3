# quick bench (no tests): $1 = extra env assignments (optional, quoted)
env $1 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/q.json 2> gpurun_out/q.err; echo rc=$?
grep "solve phases" gpurun_out/q.err
python - <<EOP
import json
d=json.loads(open("gpurun_out/q.json").read().strip().splitlines()[-1])
print("ms/step %.4f"%d["ms_per_step"], "value %.4g"%d["value"], "e2e %.4g"%d["e2e"]["value"], "launches", d["gpu_launches"])
print(d["roofline"]); print(d["search_stats"])
EOP
