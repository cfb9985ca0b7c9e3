#!/bin/bash
# PDL for the first trip (k_search(0) behind k_ctl_reset): correctness with it on (default), A/B FLIVO_PDL=1 (all trips) / 2 (from
# the second trip on) / 0 (off), then the bench lines of the final build for profiles/
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/k_pytest.log 2>&1; echo "pytest (PDL on) rc=$?"; tail -5 gpurun_out/k_pytest.log
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/k_$w.$tag.json 2> gpurun_out/k_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/k_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(8), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'p99 %.4f'%d['latency_ms_per_scan_stats']['p99'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  for rep in a b; do
    run pdl1$rep $w FLIVO_PDL=1
    run pdl2$rep $w FLIVO_PDL=2
    run pdl0$rep $w FLIVO_PDL=0
  done
done
# final lines
timeout 900 python bench.py > gpurun_out/k_bench_default.json 2> gpurun_out/k_bench_default.err; echo "bench default rc=$?"; tail -c 600 gpurun_out/k_bench_default.json | head -c 600; echo
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/k_bench_reference.json 2> gpurun_out/k_bench_reference.err; echo "reference rc=$?"
for w in avia24k_map1M avia100k_map5M; do
  timeout 600 python bench.py --workload $w --no-replay > gpurun_out/k_bench_$w.json 2> gpurun_out/k_bench_$w.err; echo "$w rc=$?"
done
timeout 600 python bench.py --force-search 1 --no-replay --no-cpu-baseline > gpurun_out/k_bench_force.json 2> gpurun_out/k_bench_force.err; echo "force rc=$?"
python - <<EOP
import json,glob
for f in sorted(glob.glob('gpurun_out/k_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1].ljust(34), 'ms/step', d.get('ms_per_step'), 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'frac', (d.get('roofline') or {}).get('frac'))
    except Exception as e:
        print('FAILED', f, e)
EOP
