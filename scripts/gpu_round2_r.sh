#!/bin/bash
# super-block occupancy table for wide search cubes: suite, replay profile (V1), replay throughput, headline sanity
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r_pytest.log
timeout 600 python scripts/replay_profile.py --mode 1 > gpurun_out/r_replay_profile_v1.txt 2> gpurun_out/r_replay_profile_v1.err; echo "profile v1 rc=$?"; grep -A3 "^frame [1234]0:" gpurun_out/r_replay_profile_v1.txt | cut -c1-400
python - <<EOP
import re
ms=[float(m.group(1)) for m in re.finditer(r'device ms ([0-9.]+)', open('gpurun_out/r_replay_profile_v1.txt').read())]
print('graph-path frames: n=%d mean %.3f median %.3f max %.3f ms'%(len(ms), sum(ms)/len(ms), sorted(ms)[len(ms)//2], max(ms)))
EOP
rp() {
  tag=$1; shift
  env FLIVO_FRAME_TRACE=1 "$@" > gpurun_out/r_replay_$tag.json 2> gpurun_out/r_replay_$tag.err; echo "replay $tag rc=$?"
  python - "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/r_replay_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print('   ', sys.argv[1].ljust(16), 'frames/s %.0f'%d['frames_per_sec'], 'pts/s %.4g'%d['value'], 'ms/frame %.3f'%d['ms_per_frame_wall'], 'gpu_busy %.2f'%d['gpu_busy_updates_rank0'])
except Exception as e:
    print('FAILED', sys.argv[1], e)
EOP
  grep "frame trace" gpurun_out/r_replay_$tag.err | head -1
}
rp v1_c1 timeout 600 python scripts/replay_bench.py --sequences 8 --frames 100 --concurrent 1 --mode 1
rp v1_c8 timeout 600 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent 8 --mode 1
rp v3_c1 timeout 600 python scripts/replay_bench.py --sequences 4 --frames 60 --concurrent 1 --mode 0
for w in ouster260k_map20M avia24k_map1M; do
  timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/r_$w.json 2> gpurun_out/r_$w.err
  python - $w <<EOP
import json,sys
d=json.loads(open('gpurun_out/r_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'parity', d.get('parity_check',{}).get('first_trip_all_points'))
EOP
done
