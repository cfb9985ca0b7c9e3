#!/bin/bash
# the default bench line and the reference arm, final build
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/ah_bench_default.json 2> gpurun_out/ah_bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/ah_bench_reference.json 2> gpurun_out/ah_bench_reference.err; echo "reference rc=$?"
python - <<EOP
import json
d=json.loads(open('gpurun_out/ah_bench_default.json').read().strip().splitlines()[-1])
print('ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'frac %.4f'%d['roofline']['frac'], 'launches', d['gpu_launches'])
print('replay', d['replay']); print('cpu', d['cpu_baseline']['value'])
r=json.loads(open('gpurun_out/ah_bench_reference.json').read().strip().splitlines()[-1]); print('reference', r['value'], r['config']==d['config'])
EOP
