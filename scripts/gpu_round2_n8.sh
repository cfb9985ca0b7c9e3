#!/bin/bash
# the default bench line exactly as the driver launches it, N = 8 (then N = 2 if time allows)
mkdir -p gpurun_out
N=${1:-8}
t0=$(date +%s)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/n${N}_bench_default.json 2> gpurun_out/n${N}_bench_default.err; echo "bench N=$N rc=$? wall $(( $(date +%s) - t0 )) s"
python - $N <<EOP
import json,sys
N=sys.argv[1]
try:
    d=json.loads(open('gpurun_out/n%s_bench_default.json'%N).read().strip().splitlines()[-1])
    print('value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ms/step %.4f'%d['ms_per_step'], 'stats', d.get('latency_ms_per_scan_stats'))
    print('sharded', json.dumps(d.get('sharded'))[:1500])
    print('replay', d.get('replay'))
except Exception as e:
    print('FAILED', e)
EOP
tail -5 gpurun_out/n${N}_bench_default.err | cut -c1-300
