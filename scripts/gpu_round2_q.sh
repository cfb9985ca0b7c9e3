#!/bin/bash
# what a replay frame's update is made of + a cleaner A/B of the captured Add_Points sequence (order swapped, longer sequences)
mkdir -p gpurun_out
timeout 600 python scripts/replay_profile.py --mode 1 > gpurun_out/q_replay_profile_v1.txt 2> gpurun_out/q_replay_profile_v1.err; echo "profile v1 rc=$?"; cat gpurun_out/q_replay_profile_v1.txt | cut -c1-420
timeout 600 python scripts/replay_profile.py --mode 0 > gpurun_out/q_replay_profile_v3.txt 2> gpurun_out/q_replay_profile_v3.err; echo "profile v3 rc=$?"; grep -v "^   " gpurun_out/q_replay_profile_v3.txt | tail -12 | cut -c1-300
rp() {
  tag=$1; shift
  env FLIVO_FRAME_TRACE=1 "$@" > gpurun_out/q_replay_$tag.json 2> gpurun_out/q_replay_$tag.err; echo "replay $tag rc=$?"
  python - "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/q_replay_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print('   ', sys.argv[1].ljust(16), 'frames/s %.0f'%d['frames_per_sec'], 'pts/s %.4g'%d['value'], 'ms/frame %.3f'%d['ms_per_frame_wall'], 'gpu_busy %.2f'%d['gpu_busy_updates_rank0'])
except Exception as e:
    print('FAILED', sys.argv[1], e)
EOP
  grep "frame trace" gpurun_out/q_replay_$tag.err | head -1
}
for rep in a b; do
  rp plain_c1$rep FLIVO_ADD_GRAPH=0 timeout 600 python scripts/replay_bench.py --sequences 4 --frames 300 --concurrent 1 --mode 1
  rp graph_c1$rep FLIVO_ADD_GRAPH=1 timeout 600 python scripts/replay_bench.py --sequences 4 --frames 300 --concurrent 1 --mode 1
done
