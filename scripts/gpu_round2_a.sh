#!/bin/bash
# GPU pass of round 2: the whole -m gpu suite, then a short bench of two single-GPU workloads
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_smi.txt
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -25 gpurun_out/a_pytest.log
for w in ouster260k_map20M avia24k_map1M; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/a_bench_$w.json 2> gpurun_out/a_bench_$w.err; echo "bench $w rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/a_bench_$w.json').read().strip().splitlines()[-1])
print('$w', 'ms/step', round(d['ms_per_step'],4), 'value', '%.3g'%d['value'], 'e2e', '%.3g'%d['e2e']['value'], 'search trip ms', d['roofline']['avg_launch_ms'], 'reuse', d['roofline']['reuse_trip']['avg_launch_ms'], 'frac', d['roofline']['frac'])
"
  grep "solve phases\|search CTAs" gpurun_out/a_bench_$w.err | head -8
done
