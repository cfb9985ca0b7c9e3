#!/bin/bash
# split upload (head repacked by the host, tail sent raw): parity test, then pipelined e2e at N = 1 with 4 host threads
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_upload_paths.py tests/test_gpu_stage_sort.py tests/test_frontend_gpu.py -x -q > gpurun_out/ad_pytest.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/ad_pytest.log
for sp in 0 30 50 70 100; do
  FLIVO_RAW_DMA=2 FLIVO_RAW_SPLIT=$sp FLIVO_HOST_THREADS=4 timeout 600 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/ad_e2e_s$sp.json 2> gpurun_out/ad_e2e_s$sp.err
  python - $sp <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/ad_e2e_s%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print('split %s %% repacked, 4 host threads:'%sys.argv[1], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'h2d', d['e2e']['h2d_bytes_per_step'])
except Exception as e:
    print('FAILED', sys.argv[1], e)
EOP
done
