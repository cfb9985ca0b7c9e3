#!/bin/bash
# same-box A/B of the barrier-free chunk search and the shorter Gauss-Jordan chain; then the default bench exactly as the
# driver runs it (N=1, cpu baseline, all blocks) and the reference arm
mkdir -p gpurun_out
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/e_$w.$tag.json 2> gpurun_out/e_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/e_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(10), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'search-trip %.4f'%d['roofline']['avg_launch_ms'], 'reuse %.4f'%d['roofline']['reuse_trip']['avg_launch_ms'], 'solve_us', d['roofline']['solve_inside_k_fused_us'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
  grep "search CTAs" gpurun_out/e_$w.$tag.err | cut -c1-260
  grep "solve inside" gpurun_out/e_$w.$tag.err | sed -n 2p | cut -c1-200
}
for w in ouster260k_map20M avia24k_map1M; do
  run base $w A=1
  run chunks $w FLIVO_LIB=variants/chunks.so
  run gj2 $w FLIVO_LIB=variants/gj2.so
  run base2 $w A=1
  run chunks2 $w FLIVO_LIB=variants/chunks.so
done
FLIVO_LIB=variants/chunks.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_map.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -3
FLIVO_LIB=variants/gj2.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_ivox.py -m gpu -x -q 2>&1 | tail -3
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/e_default.json 2> gpurun_out/e_default.err; echo "default bench rc=$?"
timeout 1200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/e_reference.json 2> gpurun_out/e_reference.err; echo "reference arm rc=$?"
python - <<EOP
import json
d=json.loads(open('gpurun_out/e_default.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/e_reference.json').read().strip().splitlines()[-1])
print('default: value %.4g e2e %.4g ms %.4f frac %.4f'%(d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac']))
print('cpu_baseline', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], d['cpu_baseline']['sample'])
print('parity', json.dumps(d['parity_check']))
print('reference arm value %.4g'%r['value'], 'ratio e2e', d['e2e']['value']/r['value'], 'same_config', d['config']==r['config'])
print('clocks', d['clocks'])
EOP
