#!/bin/bash
# two GPUs: the default bench exactly as the driver launches it (replica + sharded 260 k + sharded mega + replay), and replay alone
mkdir -p gpurun_out
S=$(date +%s)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/n2b_bench.json 2> gpurun_out/n2b_bench.err; echo "bench N=2 rc=$? wall $(( $(date +%s) - S )) s"
tail -3 gpurun_out/n2b_bench.err | cut -c1-300
python - <<EOP
import json
d=json.loads(open('gpurun_out/n2b_bench.json').read().strip().splitlines()[-1])
print('N=2 value %.4g e2e %.4g ms/step %.4f'%(d['value'], d['e2e']['value'], d['ms_per_step']))
print('sharded', json.dumps(d.get('sharded'))[:2500])
print('replay', json.dumps(d.get('replay')))
EOP
for c in 4 8; do timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601 scripts/replay_bench.py --sequences 32 --frames 100 --concurrent $c 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('replay N=2 conc', d['concurrent_per_rank'], 'pts/s %.4g'%d['value'], 'frames/s %.0f'%d['frames_per_sec'], 'ms/frame/rank %.3f'%d['ms_per_frame_wall'])"; done
for c in 1 4 8; do CUDA_VISIBLE_DEVICES=0 timeout 900 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent $c 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('replay N=1 conc', d['concurrent_per_rank'], 'pts/s %.4g'%d['value'], 'frames/s %.0f'%d['frames_per_sec'], 'ms/frame %.3f'%d['ms_per_frame_wall'])"; done
