#!/bin/bash
# Add_Points as one captured graph + k_ds_decide through the block records: suite, config-2 trajectory (map set vs the reference
# tree), replay with the frame trace (graph on / off, V1 and V3, 1 and 8 sequences in flight), map_add line of the bench
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/p_pytest.log
timeout 1200 python scripts/config2_trajectory.py > gpurun_out/p_config2.json 2> gpurun_out/p_config2.err; echo "config2 rc=$?"; tail -c 700 gpurun_out/p_config2.json; echo
rp() {  # tag, args..., env via leading VAR=...
  tag=$1; shift
  env FLIVO_FRAME_TRACE=1 "$@" > gpurun_out/p_replay_$tag.json 2> gpurun_out/p_replay_$tag.err; echo "replay $tag rc=$?"
  python - "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/p_replay_%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
    print('   ', sys.argv[1].ljust(16), 'frames/s %.0f'%d['frames_per_sec'], 'pts/s %.4g'%d['value'], 'ms/frame %.3f'%d['ms_per_frame_wall'], 'gpu_busy %.2f'%d['gpu_busy_updates_rank0'])
except Exception as e:
    print('FAILED', sys.argv[1], e)
EOP
  grep "frame trace" gpurun_out/p_replay_$tag.err | head -2
}
rp v1_c1_graph   FLIVO_ADD_GRAPH=1 timeout 600 python scripts/replay_bench.py --sequences 8 --frames 100 --concurrent 1 --mode 1
rp v1_c1_plain   FLIVO_ADD_GRAPH=0 timeout 600 python scripts/replay_bench.py --sequences 8 --frames 100 --concurrent 1 --mode 1
rp v1_c8_graph   FLIVO_ADD_GRAPH=1 timeout 600 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent 8 --mode 1
rp v1_c8_plain   FLIVO_ADD_GRAPH=0 timeout 600 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent 8 --mode 1
rp v3_c1_graph   FLIVO_ADD_GRAPH=1 timeout 600 python scripts/replay_bench.py --sequences 4 --frames 60 --concurrent 1 --mode 0
for g in 1 0; do
  FLIVO_ADD_GRAPH=$g timeout 600 python bench.py --workload avia100k_map5M --steps 10 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-replay > gpurun_out/p_bench100k_g$g.json 2> gpurun_out/p_bench100k_g$g.err
  python - $g <<EOP
import json,sys
d=json.loads(open('gpurun_out/p_bench100k_g%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print('map_add graph=%s'%sys.argv[1], d.get('map_add'))
EOP
done
