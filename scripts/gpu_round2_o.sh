#!/bin/bash
# k_fused as a programmatic dependent of its trip's k_search (early on re-use trips): suite with it on, then A/B FLIVO_PDL=1 (all) / 3 (k_search only)
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/o_pytest.log
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/o_$w.$tag.json 2> gpurun_out/o_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/o_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(8), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'p99 %.4f'%d['latency_ms_per_scan_stats']['p99'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  for rep in a b; do
    run pdl1$rep $w FLIVO_PDL=1
    run pdl3$rep $w FLIVO_PDL=3
  done
done
for m in 0 1; do
  timeout 600 python scripts/replay_bench.py --sequences 8 --frames 60 --concurrent 1 --mode $m > gpurun_out/o_replay_m$m.json 2> gpurun_out/o_replay_m$m.err; echo "replay mode $m rc=$?"; cut -c1-330 gpurun_out/o_replay_m$m.json
done
