#!/bin/bash
# ncu launch list (gpu__time_duration) of the default bench command's update steps, 260 k and 24 k points; upload flavours from
# cudaHostAlloc memory
mkdir -p gpurun_out
K='regex:k_search|k_fused|k_ctl_reset|k_scan_prepare|k_gather4|k_stage|DeviceRadixSort'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 1500 --csv --log-file gpurun_out/z_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/z_ncu.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 1500 --csv --log-file gpurun_out/z_launches_24k.csv python bench.py --workload avia24k_map1M --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/z_ncu24.log 2>&1; echo "ncu list 24k rc=$?"
python - <<EOP
import csv
for f in ('gpurun_out/z_launches.csv','gpurun_out/z_launches_24k.csv'):
    rows=[r for r in csv.reader(open(f)) if len(r)>10]
    hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
    names=[r[ki] for r in rows[1:]]
    resets=[i for i,n in enumerate(names) if 'k_ctl_reset' in n]
    print('==', f, 'kernels', len(names), 'updates', len(resets))
    if len(resets) >= 4:
        a=resets[2]-8 if resets[2] >= 8 else 0
        for r in rows[1+a:1+resets[4] if len(resets) > 4 else None]:
            print('%-70s %8.2f us'%(r[ki][:68], float(r[vi].replace(',',''))/1e3))
EOP
for hm in registered pinned; do
  for m in 0 1; do
    FLIVO_RAW_DMA=$m timeout 600 python bench.py --host-mem $hm --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/z_e2e_${hm}_m$m.json 2> gpurun_out/z_e2e_${hm}_m$m.err
    python - $hm $m <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/z_e2e_%s_m%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print('host-mem=%s RAW_DMA=%s'%(sys.argv[1],sys.argv[2]), 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'serial %.4g'%d['e2e']['serial_value'], 'h2d', d['e2e']['h2d_bytes_per_step'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
  done
done
