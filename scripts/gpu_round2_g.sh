#!/bin/bash
# long evidence job: config-2 trajectory at size, 1-GPU mega line (config 5 curve), ncu --set full of k_search at 260 k and at
# 4 M points, sanitizer record
mkdir -p gpurun_out
timeout 1500 python scripts/config2_trajectory.py --scans 100 > gpurun_out/g_config2.json 2> gpurun_out/g_config2.err; echo "config2 rc=$?"; cut -c1-1500 gpurun_out/g_config2.json
timeout 1200 python bench.py --workload mega4M_map100M --scans 2 --steps 5 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/g_mega_n1.json 2> gpurun_out/g_mega_n1.err; echo "mega N=1 rc=$?"
python - <<EOP
import json
try:
    d=json.loads(open('gpurun_out/g_mega_n1.json').read().strip().splitlines()[-1])
    print('mega N=1: ms/step %.4f value %.4g e2e %.4g search-pair ms %.4f frac %.4f map_build_s %.1f'%(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['avg_launch_ms'], d['roofline']['frac'], d['map_build_s']))
except Exception as e: print('mega FAILED', e)
EOP
tail -3 gpurun_out/g_mega_n1.err | cut -c1-300
# ncu --set full of the search kernel: first search trip of one step (kernel regex, launch skip past warm-up)
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_search -s 8 -c 1 -o gpurun_out/g_ksearch_260k -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/g_ncu260k.log 2>&1; echo "ncu 260k rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_search -s 8 -c 1 -o gpurun_out/g_ksearch_4M -f python bench.py --workload mega4M_map100M --scans 2 --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/g_ncu4M.log 2>&1; echo "ncu 4M rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 8 -c 2 -o gpurun_out/g_kfused_260k -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/g_ncufused.log 2>&1; echo "ncu fused rc=$?"
ls -la gpurun_out/*.ncu-rep
bash scripts/gpu_sanitize.sh > gpurun_out/g_sanitize_stdout.log 2>&1; echo "sanitize rc=$?"; tail -60 gpurun_out/sanitizer.txt
