# A/B on one box.  usage: bash scripts/gpu_ab.sh <workload> <lib>...   with <lib> a file under fast-livo-noted_b200/variants/ ("-" = the in-tree build)
W=$1; shift
for a in "$@"; do
  if [ "$a" = "-" ]; then unset FLIVO_LIB; else export FLIVO_LIB=variants/$a; fi
  timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-frontend --workload $W > gpurun_out/ab.json 2> gpurun_out/ab.err
  python - "$a" "$W" <<EOP
import json,sys
try:
    d=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
    r=d["roofline"]
    print("%-14s %-18s"%(sys.argv[2], sys.argv[1]), "ms/step %.4f"%d["ms_per_step"], "search-trip %.4f"%r["avg_launch_ms"], "reuse %.4f"%r["reuse_trip"]["avg_launch_ms"], "solve %.4f"%r["solve_avg_ms"], "e2e %.3g"%d["e2e"]["value"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
EOP
done
