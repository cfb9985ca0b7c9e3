#!/bin/bash
# blocking waits inside flivo_frame: replay on one GPU with FLIVO_BLOCKING_SYNC=0/1 (8 and 16 in flight), replay tests with it on
mkdir -p gpurun_out
FLIVO_BLOCKING_SYNC=1 timeout 900 python -m pytest tests/test_gpu_replay.py tests/test_gpu_trajectory.py -x -q > gpurun_out/ag_pytest.log 2>&1; echo "tests (blocking) rc=$?"; tail -3 gpurun_out/ag_pytest.log
for b in 0 1; do for c in 8 16; do
  FLIVO_BLOCKING_SYNC=$b timeout 600 python scripts/replay_bench.py --sequences 32 --frames 200 --concurrent $c --mode 1 > gpurun_out/ag_replay_c${c}_b$b.json 2> gpurun_out/ag_replay_c${c}_b$b.err
  python - $c $b <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/ag_replay_c%s_b%s.json'%tuple(sys.argv[1:3])).read().strip().splitlines()[-1])
    print('in flight %s blocking %s:'%(sys.argv[1],sys.argv[2]), 'frames/s %.0f'%d['frames_per_sec'], 'pts/s %.4g'%d['value'], 'gpu_busy %.2f'%d['gpu_busy_updates_rank0'])
except Exception as e:
    print('FAILED', sys.argv[1:], e)
EOP
done; done
