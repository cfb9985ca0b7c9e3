# A/B of environment switches on one box.  usage: bash scripts/gpu_env_ab.sh <workload> "<ENV=val>"...
W=$1; shift
for rep in 1 2; do
for e in "$@"; do
  env $e timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-frontend --no-ivox --no-map-add --workload $W > gpurun_out/ab.json 2> gpurun_out/ab.err
  python - "$e" "$W" <<EOP
import json,sys
d=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("%-18s %-24s"%(sys.argv[2], sys.argv[1]), "ms/step %.4f"%d["ms_per_step"], "search-trip %.4f"%r["avg_launch_ms"], "reuse %.4f"%r["reuse_trip"]["avg_launch_ms"], "e2e %.4g"%d["e2e"]["value"])
EOP
done; done
