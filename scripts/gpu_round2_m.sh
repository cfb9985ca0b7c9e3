#!/bin/bash
# launch list (ncu gpu__time_duration) of a few replay frames: which kernels make up map_incremental's ~0.5 ms at 24 k points
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:^k_|^Device" -s 1500 -c 160 --csv --log-file gpurun_out/m_replay_launches.csv \
  python scripts/replay_bench.py --sequences 1 --frames 30 --concurrent 1 > gpurun_out/m_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'EOP'
import csv
rows=[r for r in csv.reader(open('gpurun_out/m_replay_launches.csv')) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
for r in rows[1:]:
    print('%-70s %8.2f us'%(r[ki][:68], float(r[vi].replace(',',''))/1e3))
EOP
