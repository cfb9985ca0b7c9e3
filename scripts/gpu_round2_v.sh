#!/bin/bash
# tile size of k_search: 128-point tiles (5 CTAs/SM) against 64-point tiles (10 or 8 CTAs/SM), same box, interleaved
mkdir -p gpurun_out
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/v_$w.$tag.json 2> gpurun_out/v_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/v_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(8), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'value %.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'parity ids', d['parity_check']['first_trip_all_points']['ids_mismatch_rows'] if d.get('parity_check') and d['parity_check'].get('first_trip_all_points') else None)
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
  grep "search CTAs" gpurun_out/v_$w.$tag.err | tail -1 | cut -c1-260
}
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  for rep in a b; do
    run base$rep $w FLIVO_LIB=libflivo_b200.so
    run t64c10$rep $w FLIVO_LIB=variants/t64c10.so
    run t64c8$rep $w FLIVO_LIB=variants/t64c8.so
  done
done
