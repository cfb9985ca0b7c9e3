#!/bin/bash
# the scan loop of thread_knn5_shell without the per-slot candidate counter / without the found counter: variants against the base
mkdir -p gpurun_out
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/aa_$w.$tag.json 2> gpurun_out/aa_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/aa_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    pc=(d.get('parity_check') or {}); fa=pc.get('first_trip_all_points') or {}
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(9), 'ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], 'frac %.4f'%d['roofline']['frac'], 'cand', d['search_stats']['candidates_last_step'], 'ids mismatch', fa.get('ids_mismatch_rows'), 'x diff', pc.get('x_max_abs_diff'))
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  for rep in a b; do
    run base$rep $w FLIVO_LIB=libflivo_b200.so
    run ncbucket$rep $w FLIVO_LIB=variants/ncbucket.so
    run ncnone$rep $w FLIVO_LIB=variants/ncnone.so
    run nofound$rep $w FLIVO_LIB=variants/nofound.so
  done
done
