#!/bin/bash
mkdir -p gpurun_out
for tag in "--no-sharded" "--no-mega" "--no-cpu-baseline" "--no-mega --no-cpu-baseline"; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 10 --warmup 3 $tag > gpurun_out/n2c.json 2> gpurun_out/n2c.err
  python - "$tag" <<EOP
import json,sys
d=json.loads(open('gpurun_out/n2c.json').read().strip().splitlines()[-1])
r=d.get('replay') or {}
print(sys.argv[1].ljust(32), 'replay frames/s %.0f ms/frame/rank %.3f upd_dev_ms %.3f'%(r.get('frames_per_sec',0), r.get('ms_per_frame_per_rank',0), r.get('update_device_ms_mean',0)))
EOP
done
