#!/bin/bash
# replay with one context per driving thread (reuse) against one per sequence
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_replay.py -x -q > gpurun_out/af_pytest.log 2>&1; echo "replay tests rc=$?"; tail -4 gpurun_out/af_pytest.log
for rep in a b; do for ru in 1 0; do for c in 1 8; do
  timeout 600 python scripts/replay_bench.py --sequences 32 --frames 200 --concurrent $c --mode 1 --reuse $ru > gpurun_out/af_replay_c${c}_r$ru$rep.json 2> gpurun_out/af_replay_c${c}_r$ru$rep.err
  python - $c $ru $rep <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/af_replay_c%s_r%s%s.json'%tuple(sys.argv[1:4])).read().strip().splitlines()[-1])
    print('in flight %s reuse %s:'%(sys.argv[1],sys.argv[2]), 'frames/s %.0f'%d['frames_per_sec'], 'pts/s %.4g'%d['value'], 'gpu_busy %.2f'%d['gpu_busy_updates_rank0'])
except Exception as e:
    print('FAILED', sys.argv[1:], e)
EOP
done; done; done
