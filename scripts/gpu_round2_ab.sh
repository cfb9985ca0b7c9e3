#!/bin/bash
# staging sort without the one-CTA scans (every CTA sums the histogram columns itself): its test (all flavours and sizes), then
# A/B own (R = 4 and R = 2) against the library sort on three workloads
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stage_sort.py tests/test_gpu_upload_paths.py -x -q > gpurun_out/ab_pytest_stage.log 2>&1; echo "stage sort tests rc=$?"; tail -4 gpurun_out/ab_pytest_stage.log
FLIVO_SORT=own FLIVO_STAGE_R=2 timeout 600 python tests/test_gpu_stage_sort.py > gpurun_out/ab_stage_r2.log 2>&1; echo "R=2 flavour rc=$?"; tail -1 gpurun_out/ab_stage_r2.log
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/ab_$w.$tag.json 2> gpurun_out/ab_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/ab_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(8), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'launches', d['gpu_launches'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia100k_map5M avia24k_map1M; do
  for rep in a b; do
    run own4$rep $w FLIVO_SORT=own FLIVO_STAGE_R=4
    run own2$rep $w FLIVO_SORT=own FLIVO_STAGE_R=2
    run cub$rep $w FLIVO_SORT=cub
  done
done
run own_mega mega4M_map100M FLIVO_SORT=own
run cub_mega mega4M_map100M FLIVO_SORT=cub
K='regex:k_stage'
FLIVO_SORT=own timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 30 -c 9 --csv --log-file gpurun_out/ab_stage_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/ab_ncu.log 2>&1; echo "ncu list rc=$?"
python - <<EOP
import csv
rows=[r for r in csv.reader(open('gpurun_out/ab_stage_launches.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
for r in rows[1:]:
    print('%-70s %8.2f us'%(r[ki][:68], float(r[vi].replace(',',''))/1e3))
EOP
