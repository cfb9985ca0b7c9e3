#!/bin/bash
# full suite with the slot-keyed insert sort, then the e2e upload flavours at N = 1 (repack / raw 48-byte DMA / 2-D gather DMA)
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/u_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/u_pytest.log
for m in 0 1 2; do
  for ht in 16 4; do
    FLIVO_RAW_DMA=$m FLIVO_HOST_THREADS=$ht timeout 600 python bench.py --steps 30 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/u_e2e_m${m}_t$ht.json 2> gpurun_out/u_e2e_m${m}_t$ht.err
    python - $m $ht <<EOP
import json,sys
try:
    d=json.loads(open('gpurun_out/u_e2e_m%s_t%s.json'%(sys.argv[1],sys.argv[2])).read().strip().splitlines()[-1])
    print('RAW_DMA=%s host_threads=%s'%(sys.argv[1],sys.argv[2]), 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'serial %.4g'%d['e2e']['serial_value'], 'h2d', d['e2e']['h2d_bytes_per_step'])
except Exception as e:
    print('FAILED', sys.argv[1], sys.argv[2], e)
EOP
  done
done
for g in 1 0; do
  FLIVO_SLOT_KEYS=$g timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-replay > gpurun_out/u_mapadd_slot$g.json 2> gpurun_out/u_mapadd_slot$g.err
  python - $g <<EOP
import json,sys
d=json.loads(open('gpurun_out/u_mapadd_slot%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print('slot keys=%s map_add'%sys.argv[1], d.get('map_add',{}).get('ms_per_scan_median'), d.get('map_add',{}).get('ms_per_scan'))
EOP
done
for g in 1 0; do
  FLIVO_SLOT_KEYS=$g timeout 600 python scripts/replay_bench.py --sequences 8 --frames 200 --concurrent 1 --mode 1 > gpurun_out/u_replay_slot$g.json 2> gpurun_out/u_replay_slot$g.err
  FLIVO_SLOT_KEYS=$g FLIVO_FRAME_TRACE=1 timeout 600 python scripts/replay_bench.py --sequences 2 --frames 200 --concurrent 1 --mode 1 2>&1 >/dev/null | grep "frame trace" | tail -1
  cut -c1-330 gpurun_out/u_replay_slot$g.json
done
