#!/bin/bash
# heavy queries (more than N candidates in the 27-cell shell) handed to the warp-cooperative search: variants against the base
mkdir -p gpurun_out
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/w_$w.$tag.json 2> gpurun_out/w_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/w_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    pc=(d.get('parity_check') or {}); fa=pc.get('first_trip_all_points') or {}
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(9), 'ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'far', d['search_stats']['far_queries_last_step'], 'ids mismatch', fa.get('ids_mismatch_rows'), 'x diff', pc.get('x_max_abs_diff'))
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
  grep "search CTAs" gpurun_out/w_$w.$tag.err | tail -1 | cut -c60-250
}
for w in avia24k_map1M ouster260k_map20M avia100k_map5M; do
  for rep in a b; do
    run base$rep $w FLIVO_LIB=libflivo_b200.so
    run heavy48$rep $w FLIVO_LIB=variants/heavy48.so
    run heavy96$rep $w FLIVO_LIB=variants/heavy96.so
    run heavy160$rep $w FLIVO_LIB=variants/heavy160.so
  done
done
