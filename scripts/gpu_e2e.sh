for t in 1 4 8 16; do
  FLIVO_HOST_THREADS=$t timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-frontend > gpurun_out/e2e.json 2> gpurun_out/e2e.err
  python - "$t" <<EOP
import json,sys
d=json.loads(open("gpurun_out/e2e.json").read().strip().splitlines()[-1])
print("host threads", sys.argv[1], "ms/step %.4f"%d["ms_per_step"], "e2e %.4g pts/s"%d["e2e"]["value"], "e2e ms %.4f"%(4*260000/d["e2e"]["value"]*1e3))
EOP
done
nproc
