#!/bin/bash
# own staging sort (stage.cu): its test, the whole -m gpu suite, A/B against the library sort (FLIVO_SORT=cub), launch lists
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stage_sort.py -x -q > gpurun_out/n_pytest_stage.log 2>&1; echo "stage sort test rc=$?"; tail -4 gpurun_out/n_pytest_stage.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/n_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/n_pytest.log
run() {  # tag, workload, env...
  tag=$1; w=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $w --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/n_$w.$tag.json 2> gpurun_out/n_$w.$tag.err
  python - "$w" "$tag" <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/n_%s.%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print(sys.argv[1].ljust(18), sys.argv[2].ljust(8), 'ms/step %.4f'%d['ms_per_step'], 'median %.4f'%d['latency_ms_per_scan_stats']['median'], 'p99 %.4f'%d['latency_ms_per_scan_stats']['p99'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'launches', d['gpu_launches'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
}
for w in ouster260k_map20M avia24k_map1M avia100k_map5M; do
  for rep in a b; do
    run own$rep $w FLIVO_SORT=own
    run cub$rep $w FLIVO_SORT=cub
  done
done
K='regex:^k_|^Device'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 100 -c 40 --csv --log-file gpurun_out/n_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay --no-sharded > gpurun_out/n_ncu.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 1500 -c 160 --csv --log-file gpurun_out/n_replay_launches.csv python scripts/replay_bench.py --sequences 1 --frames 30 --concurrent 1 > gpurun_out/n_ncu_replay.log 2>&1; echo "ncu replay rc=$?"
python - <<EOP
import csv
for f in ('gpurun_out/n_launches.csv','gpurun_out/n_replay_launches.csv'):
    print('==', f)
    try:
        rows=[r for r in csv.reader(open(f)) if len(r)>10]
        hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
        for r in rows[1:]:
            print('%-60s %8.2f us'%(r[ki][:58], float(r[vi].replace(',',''))/1e3))
    except Exception as e:
        print('FAILED', e)
EOP
