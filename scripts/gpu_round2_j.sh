#!/bin/bash
# PDL (programmatic dependent launch of the next trip's k_search): correctness with it on (default), then A/B against FLIVO_PDL=0
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/j_pytest.log 2>&1; echo "pytest (PDL on) rc=$?"; tail -5 gpurun_out/j_pytest.log
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/j_$w.$tag.json 2> gpurun_out/j_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/j_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(8), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'p99 %.4f'%d['latency_ms_per_scan_stats']['p99'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  run pdl1 $w FLIVO_PDL=1
  run pdl0 $w FLIVO_PDL=0
  run pdl1b $w FLIVO_PDL=1
  run pdl0b $w FLIVO_PDL=0
done
K='regex:k_search|k_fused|k_ctl_reset|k_scan_prepare|k_gather4'
FLIVO_PDL=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 60 -c 36 --csv --log-file gpurun_out/j_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/j_ncu.log 2>&1; echo "ncu list rc=$?"
python - <<EOP
import csv
rows=[r for r in csv.reader(open('gpurun_out/j_launches.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
for r in rows[1:40]:
    print('%-50s %s us'%(r[ki][:48], float(r[vi].replace(',',''))/1e3))
EOP
