#!/bin/bash
# the -m gpu suite, then A/B of the solve variants (same box) on two workloads
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -30 gpurun_out/b_pytest.log
for w in ouster260k_map20M avia24k_map1M; do
 for v in - solve_inline.so fused6.so -; do
  if [ "$v" = "-" ]; then unset FLIVO_LIB; else export FLIVO_LIB=variants/$v; fi
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/b_$w.$v.json 2> gpurun_out/b_$w.$v.err; echo "bench $w $v rc=$?"
  python - "$w" "$v" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/b_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1], sys.argv[2], 'ms/step', round(d['ms_per_step'],4), 'median', round(d['latency_ms_per_scan_stats']['median'],4), 'e2e', '%.4g'%d['e2e']['value'], 'search trip ms', d['roofline']['avg_launch_ms'], 'reuse', d['roofline']['reuse_trip']['avg_launch_ms'], 'trips', d["run"]["trips_per_step"], 'launches', d['gpu_launches'])
except Exception as e:
    print('FAILED', e)
EOP
  grep "solve inside" gpurun_out/b_$w.$v.err | head -5
 done
done
