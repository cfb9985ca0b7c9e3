#!/bin/bash
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/test_gpu_replay.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/i_pytest.log
for c in 1 4 8; do timeout 900 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent $c 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('replay conc', d['concurrent_per_rank'], 'pts/s %.4g'%d['value'], 'frames/s %.0f'%d['frames_per_sec'], 'ms/frame %.3f'%d['ms_per_frame_wall'], 'gpu_busy(updates) %.2f'%d['gpu_busy_updates_rank0'], 'map', d['map_points_seq0'])"; done
timeout 1500 python bench.py --steps 20 --warmup 3 > gpurun_out/i_default.json 2> gpurun_out/i_default.err; echo "default bench rc=$?"; 
python - <<EOP
import json
d=json.loads(open('gpurun_out/i_default.json').read().strip().splitlines()[-1])
print('default: value %.4g e2e %.4g ms %.4f frac %.4f'%(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac']))
print('parity', json.dumps(d['parity_check']))
print('replay', json.dumps(d['replay']))
print('clocks', d['clocks'])
EOP
