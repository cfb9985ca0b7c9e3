#!/bin/bash
# A/B: barrier-free chunk search (fixed counter), fold32 + batched final fold, both; then ncu launch list and config-4 replay
mkdir -p gpurun_out
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/f_$w.$tag.json 2> gpurun_out/f_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/f_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(11), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'search-pair %.4f'%d['roofline']['avg_launch_ms'], 'reuse-pair %.4f'%d['roofline']['reuse_trip']['avg_launch_ms'], 'frac %.4f'%d['roofline']['frac'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
  grep "search CTAs" gpurun_out/f_$w.$tag.err | cut -c1-260
}
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  run base $w A=1
  run chunks $w FLIVO_LIB=variants/chunks.so
  run fold32 $w FLIVO_LIB=variants/fold32.so
  run both $w FLIVO_LIB=variants/chunksfold.so
  run base2 $w A=1
done
FLIVO_LIB=variants/chunksfold.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_map.py tests/test_gpu_fullsize.py tests/test_gpu_robustness.py tests/test_multigpu_gpu.py -m gpu -x -q 2>&1 | tail -3
# launch list (per-kernel durations, cold + serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/f_ncu.log 2>&1; echo "ncu rc=$?"
python - <<EOP
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/f_launches.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(',',''))
    except: continue
    agg.setdefault(r[ki][:60],[]).append(v/1e3)
for k,v in agg.items(): print('%-62s n=%3d  mean %.1f us  min %.1f  max %.1f'%(k,len(v),sum(v)/len(v),min(v),max(v)))
EOP
# config 4 in miniature on one GPU: 16 sequences x 100 frames, 1 / 2 / 4 / 8 concurrent
for c in 1 2 4 8; do timeout 900 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent $c 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('replay conc', d['concurrent_per_rank'], 'pts/s %.4g'%d['value'], 'frames/s %.0f'%d['frames_per_sec'], 'ms/frame %.3f'%d['ms_per_frame_wall'], 'gpu_busy(updates) %.2f'%d['gpu_busy_updates_rank0'])"; done
