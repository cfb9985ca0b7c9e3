#!/bin/bash
# final build, one GPU: the -m gpu suite, the bench lines kept under profiles/, the launch list and the ncu --set full captures
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/f_pytest.log
timeout 900 python bench.py > gpurun_out/f_bench_default.json 2> gpurun_out/f_bench_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f_bench_reference.json 2> gpurun_out/f_bench_reference.err; echo "reference rc=$?"
for w in avia24k_map1M avia100k_map5M; do
  timeout 600 python bench.py --workload $w --no-replay > gpurun_out/f_bench_$w.json 2> gpurun_out/f_bench_$w.err; echo "$w rc=$?"
done
timeout 600 python bench.py --force-search 1 --no-replay --no-cpu-baseline > gpurun_out/f_bench_force.json 2> gpurun_out/f_bench_force.err; echo "force rc=$?"
timeout 900 python bench.py --workload mega4M_map100M --scans 2 --steps 5 --no-replay --no-cpu-baseline --no-ivox --no-frontend --no-map-add > gpurun_out/f_bench_mega.json 2> gpurun_out/f_bench_mega.err; echo "mega rc=$?"
python - <<EOP
import json,glob
for f in sorted(glob.glob('gpurun_out/f_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1].ljust(34), 'ms/step', d.get('ms_per_step'), 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'frac', (d.get('roofline') or {}).get('frac'))
    except Exception as e:
        print('FAILED', f, e)
EOP
K='regex:^k_|^Device'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 200 -c 60 --csv --log-file gpurun_out/f_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/f_ncu.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_search -s 15 -c 1 -o gpurun_out/f_ksearch_260k -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/f_ncu260k.log 2>&1; echo "ncu 260k rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 15 -c 2 -o gpurun_out/f_kfused_260k -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/f_ncufused.log 2>&1; echo "ncu fused rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_search -s 15 -c 1 -o gpurun_out/f_ksearch_4M -f python bench.py --workload mega4M_map100M --scans 2 --steps 2 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/f_ncu4M.log 2>&1; echo "ncu 4M rc=$?"
