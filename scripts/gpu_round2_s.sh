#!/bin/bash
# replay with voxel-filtered scans (feats_down_body): profile, throughput at 1 / 4 / 8 sequences in flight, V3, and the default bench line
mkdir -p gpurun_out
timeout 600 python scripts/replay_profile.py --mode 1 > gpurun_out/s_replay_profile_v1.txt 2> gpurun_out/s_replay_profile_v1.err; echo "profile v1 rc=$?"; grep -A3 "^frame [1234]0:" gpurun_out/s_replay_profile_v1.txt | cut -c1-400
python - <<EOP
import re
ms=[float(m.group(1)) for m in re.finditer(r'device ms ([0-9.]+)', open('gpurun_out/s_replay_profile_v1.txt').read())]
print('graph-path frames: n=%d mean %.3f median %.3f max %.3f ms'%(len(ms), sum(ms)/len(ms), sorted(ms)[len(ms)//2], max(ms)))
EOP
rp() {
  tag=$1; shift
  env FLIVO_FRAME_TRACE=1 "$@" > gpurun_out/s_replay_$tag.json 2> gpurun_out/s_replay_$tag.err; echo "replay $tag rc=$?"
  python - "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/s_replay_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print('   ', sys.argv[1].ljust(16), 'frames/s %.0f'%d['frames_per_sec'], 'pts/s %.4g'%d['value'], 'ms/frame %.3f'%d['ms_per_frame_wall'], 'gpu_busy %.2f'%d['gpu_busy_updates_rank0'])
except Exception as e:
    print('FAILED', sys.argv[1], e)
EOP
  grep "frame trace" gpurun_out/s_replay_$tag.err | head -1
}
rp v1_c1 timeout 600 python scripts/replay_bench.py --sequences 8 --frames 200 --concurrent 1 --mode 1
rp v1_c4 timeout 600 python scripts/replay_bench.py --sequences 16 --frames 200 --concurrent 4 --mode 1
rp v1_c8 timeout 600 python scripts/replay_bench.py --sequences 32 --frames 200 --concurrent 8 --mode 1
rp v3_c1 timeout 600 python scripts/replay_bench.py --sequences 4 --frames 100 --concurrent 1 --mode 0
timeout 900 python bench.py > gpurun_out/s_bench_default.json 2> gpurun_out/s_bench_default.err; echo "bench rc=$?"
python - <<EOP
import json
d=json.loads(open('gpurun_out/s_bench_default.json').read().strip().splitlines()[-1])
print('ms/step %.4f'%d['ms_per_step'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'frac %.4f'%d['roofline']['frac'])
print('replay', d.get('replay')); print('map_add', d.get('map_add')); print('parity', d.get('parity_check'))
EOP
