# replica-mode bench at N ranks, the way the driver launches it
N=$1
if [ "$N" = "1" ]; then timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err; fi
echo rc=$?; nproc
python - <<EOP
import json
d=json.loads(open("gpurun_out/scale_$N.json").read().strip().splitlines()[-1])
print("N", d["n_gpus"], "value %.4g"%d["value"], "ms/step %.4f"%d["ms_per_step"], "e2e %.4g"%d["e2e"]["value"], d["clocks"])
EOP
