#!/bin/bash
# last check of the round: the whole -m gpu suite and smoke() with the final build, plus the default bench line
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/ae_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/ae_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/ae_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/ae_smoke.log
timeout 900 python bench.py > gpurun_out/ae_bench_default.json 2> gpurun_out/ae_bench_default.err; echo "bench rc=$?"
python - <<EOP
import json
d=json.loads(open('gpurun_out/ae_bench_default.json').read().strip().splitlines()[-1])
print('ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'frac %.4f'%d['roofline']['frac'], 'launches', d['gpu_launches'], 'clocks', d['clocks'])
print('replay', d['replay']['frames_per_sec'], 'map_add', d['map_add']['ms_per_scan_median'])
EOP
