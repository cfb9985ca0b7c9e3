# usage: bash scripts/gpu_multi.sh N   (run under gpurun --gpus N)
N=$1
timeout 600 python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q 2>&1 | tail -5
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/multi_replica_$N.json 2> gpurun_out/multi_replica_$N.err; echo replica rc=$?
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --shard-scan > gpurun_out/multi_shard_$N.json 2> gpurun_out/multi_shard_$N.err; echo shard rc=$?
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --shard-scan --exchange nccl > gpurun_out/multi_shardnccl_$N.json 2> gpurun_out/multi_shardnccl_$N.err; echo shard-nccl rc=$?
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --shard-scan --exchange nccl --workload mega4M_map100M > gpurun_out/multi_meganccl_$N.json 2> gpurun_out/multi_meganccl_$N.err; echo mega-nccl rc=$?
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --shard-scan --workload mega4M_map100M > gpurun_out/multi_mega_$N.json 2> gpurun_out/multi_mega_$N.err; echo mega rc=$?
python - <<EOP
import json
for k in ("replica","shard","shardnccl","mega","meganccl"):
    try:
        d=json.loads(open("gpurun_out/multi_%s_$N.json"%k).read().strip().splitlines()[-1])
        print(k, "value %.4g"%d["value"], "ms/step %.4f"%d["ms_per_step"], "e2e %.4g"%d["e2e"]["value"], d.get("shard_check"))
    except Exception as e:
        print(k, "FAILED", e)
EOP
tail -3 gpurun_out/multi_mega_$N.err
