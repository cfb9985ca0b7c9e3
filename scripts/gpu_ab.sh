# A/B of launch-structure switches on the bench workload (same box, same clocks)
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python - "$name" <<EOP
import json,sys
try:
    d=json.loads(open("gpurun_out/ab_%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "ms/step %.4f"%d["ms_per_step"], "e2e %.3g"%d["e2e"]["value"], "launches", d["gpu_launches"], "xdiff", d.get("parity_check"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
EOP
}
run base A=1
run nopdl FLIVO_NO_PDL=1
run nocond FLIVO_NO_COND_GRAPH=1
run nocond_nopdl FLIVO_NO_COND_GRAPH=1 FLIVO_NO_PDL=1
run far1 FLIVO_FAR_LEVELS=1
run far1_nocond FLIVO_FAR_LEVELS=1 FLIVO_NO_COND_GRAPH=1
run far1_nocond_nopdl FLIVO_FAR_LEVELS=1 FLIVO_NO_COND_GRAPH=1 FLIVO_NO_PDL=1
run base2 A=1
