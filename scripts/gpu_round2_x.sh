#!/bin/bash
# dear tiles first (second search trip deals its tiles by the first one's measured cost): suite, then A/B FLIVO_DEAR_FIRST=1/0
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/x_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/x_pytest.log
run() {  # tag, workload, extra args, env...
  tag=$1; w=$2; extra=$3; shift 3
  env "$@" timeout 600 python bench.py --workload $w $extra --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/x_$w.$tag.json 2> gpurun_out/x_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/x_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    pc=(d.get('parity_check') or {})
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(10), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'value %.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'x diff', pc.get('x_max_abs_diff'))
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia100k_map5M avia24k_map1M; do
  for rep in a b; do
    run dear1$rep $w "" FLIVO_DEAR_FIRST=1
    run dear0$rep $w "" FLIVO_DEAR_FIRST=0
  done
done
run force_dear1 ouster260k_map20M "--force-search 1" FLIVO_DEAR_FIRST=1
run force_dear0 ouster260k_map20M "--force-search 1" FLIVO_DEAR_FIRST=0
run mega_dear1 mega4M_map100M "--scans 2 --steps 5" FLIVO_DEAR_FIRST=1
run mega_dear0 mega4M_map100M "--scans 2 --steps 5" FLIVO_DEAR_FIRST=0
