#!/bin/bash
# full suite (fold32 + V1 range bound + repeatability test), replay (config 4 in miniature), launch list and ncu --set full of
# a SEARCH trip's k_search / k_fused at 260 k and 4 M points
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/h_pytest.log
for c in 1 4 8; do timeout 900 python scripts/replay_bench.py --sequences 16 --frames 100 --concurrent $c 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('replay conc', d['concurrent_per_rank'], 'pts/s %.4g'%d['value'], 'frames/s %.0f'%d['frames_per_sec'], 'ms/frame %.3f'%d['ms_per_frame_wall'], 'gpu_busy(updates) %.2f'%d['gpu_busy_updates_rank0'])"; done
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err
python -c "
import json
d=json.loads(open('gpurun_out/h_bench.json').read().strip().splitlines()[-1]); print('bench 260k ms/step %.4f value %.4g e2e %.4g frac %.4f'%(d['ms_per_step'],d['value'],d['e2e']['value'],d['roofline']['frac']))"
K='regex:k_search|k_fused|k_ctl_reset|k_scan_prepare|k_gather4|DeviceRadixSort'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 60 -c 60 --csv --log-file gpurun_out/h_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/h_ncu.log 2>&1; echo "ncu list rc=$?"
python - <<EOP
import csv
rows=[r for r in csv.reader(open('gpurun_out/h_launches.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
for r in rows[1:40]:
    print('%-70s %s us'%(r[ki][:68], float(r[vi].replace(',',''))/1e3))
EOP
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_search -s 15 -c 1 -o gpurun_out/h_ksearch_260k -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/h_ncu260k.log 2>&1; echo "ncu 260k rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 15 -c 2 -o gpurun_out/h_kfused_260k -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/h_ncufused.log 2>&1; echo "ncu fused rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_search -s 15 -c 1 -o gpurun_out/h_ksearch_4M -f python bench.py --workload mega4M_map100M --scans 2 --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/h_ncu4M.log 2>&1; echo "ncu 4M rc=$?"
