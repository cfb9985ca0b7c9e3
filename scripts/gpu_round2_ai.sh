#!/bin/bash
# parallel probes on the first (cold-L2) search trip only: A/B
mkdir -p gpurun_out
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/ai_$w.$tag.json 2> gpurun_out/ai_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/ai_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(8), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'value %.4g'%d['value'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia100k_map5M; do
  for rep in a b; do
    run parf1$rep $w FLIVO_PAR_FIRST=1
    run parf0$rep $w FLIVO_PAR_FIRST=0
  done
done
