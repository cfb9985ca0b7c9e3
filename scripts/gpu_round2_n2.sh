#!/bin/bash
# two GPUs: the two multi-GPU tests, then the default bench at N=2 exactly as the driver launches it (replica + sharded block)
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q > gpurun_out/n2_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/n2_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err; echo "bench N=2 rc=$?"
tail -5 gpurun_out/n2_bench.err | cut -c1-300
python - <<EOP
import json
d=json.loads(open('gpurun_out/n2_bench.json').read().strip().splitlines()[-1])
print('N=2 value %.4g e2e %.4g ms/step %.4f'%(d['value'], d['e2e']['value'], d['ms_per_step']))
print('sharded', json.dumps(d.get('sharded'))[:1500])
EOP
