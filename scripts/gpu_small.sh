for w in avia24k_map1M ouster260k_map20M; do
timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-frontend --workload $w > gpurun_out/small.json 2> gpurun_out/small.err
grep "search CTAs\|far queries" gpurun_out/small.err
python - "$w" <<EOP
import json,sys
d=json.loads(open("gpurun_out/small.json").read().strip().splitlines()[-1])
r=d["roofline"]
print(sys.argv[1], "ms/step %.4f"%d["ms_per_step"], "search-trip %.4f reuse %.4f solve %.4f"%(r["avg_launch_ms"], r["reuse_trip"]["avg_launch_ms"], r["solve_avg_ms"]), d["search_stats"])
EOP
done
