#!/bin/bash
# noalloc test alone in a fresh process, the full suite, then the bench on three workloads + force-search
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/c_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/c_pytest.log
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/c_$w.json 2> gpurun_out/c_$w.err; echo "bench $w rc=$?"
  python - "$w" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/c_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'ms/step', round(d['ms_per_step'],4), 'median', round(d['latency_ms_per_scan_stats']['median'],4), 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'serial %.4g'%d['e2e']['serial_value'], 'search trip ms', round(d['roofline']['avg_launch_ms'],4), 'reuse', round(d['roofline']['reuse_trip']['avg_launch_ms'],4), 'trips', d["run"]["trips_per_step"])
except Exception as e:
    print('FAILED', e)
EOP
  grep "solve inside" gpurun_out/c_$w.err | head -4
  grep "search CTAs\|far queries, first" gpurun_out/c_$w.err | cut -c1-330
done
timeout 600 python bench.py --workload ouster260k_map20M --force-search 1 --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/c_force.json 2> gpurun_out/c_force.err; echo "force rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/c_force.json').read().strip().splitlines()[-1]); print('force_search ms/step', d['ms_per_step'], 'value %.4g'%d['value'], 'trips', d['run']['trips_per_step'])"
