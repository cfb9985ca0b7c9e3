#!/bin/bash
# the new upload-flavour test + the ncu launch list of the update steps of the default bench command (final build)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_upload_paths.py -x -q > gpurun_out/y_pytest.log 2>&1; echo "upload test rc=$?"; tail -5 gpurun_out/y_pytest.log
K='regex:k_search|k_fused|k_ctl_reset|k_scan_prepare|k_gather4|k_stage|DeviceRadixSort'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 60 -c 48 --csv --log-file gpurun_out/y_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/y_ncu.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -s 60 -c 48 --csv --log-file gpurun_out/y_launches_24k.csv python bench.py --workload avia24k_map1M --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/y_ncu24.log 2>&1; echo "ncu list 24k rc=$?"
python - <<EOP
import csv
for f in ('gpurun_out/y_launches.csv','gpurun_out/y_launches_24k.csv'):
    print('==', f)
    rows=[r for r in csv.reader(open(f)) if len(r)>10]
    hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
    for r in rows[1:]:
        print('%-70s %8.2f us'%(r[ki][:68], float(r[vi].replace(',',''))/1e3))
EOP
