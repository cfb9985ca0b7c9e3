timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for w in ouster260k_map20M avia24k_map1M; do
for e in FLIVO_NO_SEED=1 FLIVO_X=1 FLIVO_NO_SEED=1 FLIVO_X=1; do
  env $e timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-frontend --no-ivox --no-map-add --workload $w > gpurun_out/ab.json 2> gpurun_out/ab.err
  python - "$e" "$w" <<EOP
import json,sys
d=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("%-18s %-16s"%(sys.argv[2], sys.argv[1]), "ms/step %.4f"%d["ms_per_step"], "search-trip %.4f"%r["avg_launch_ms"], "reuse %.4f"%r["reuse_trip"]["avg_launch_ms"], "solve %.4f"%r["solve_avg_ms"], d["search_stats"])
EOP
done; done
