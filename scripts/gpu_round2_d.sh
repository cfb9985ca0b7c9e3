#!/bin/bash
# full suite on the LDG.256 build, then same-box A/B: in-tree, PAR probes on/off, 4 / 6 search CTAs per SM
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/d_pytest.log
run() {  # tag, env..., workload
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/d_$w.$tag.json 2> gpurun_out/d_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/d_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(10), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'serial %.4g'%d['e2e']['serial_value'], 'search-trip %.4f'%d['roofline']['avg_launch_ms'], 'reuse %.4f'%d['roofline']['reuse_trip']['avg_launch_ms'], 'trips', d["run"]["trips_per_step"])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
  grep "search CTAs" gpurun_out/d_$w.$tag.err | cut -c1-260
}
for w in ouster260k_map20M avia24k_map1M; do
  run base $w A=1
  run par1 $w FLIVO_PAR_PROBES=1
  run par0 $w FLIVO_PAR_PROBES=0
  run cta4 $w FLIVO_LIB=variants/search4.so
  run cta6 $w FLIVO_LIB=variants/search6.so
  run base2 $w A=1
done
