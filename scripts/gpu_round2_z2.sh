#!/bin/bash
# ncu launch list (gpu__time_duration) of the default bench command's update steps, 260 k and 24 k points (final build)
mkdir -p gpurun_out
K='regex:k_search|k_fused|k_ctl_reset|k_scan_prepare|k_gather4|k_stage|DeviceRadixSort'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 1500 --csv --log-file gpurun_out/z_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/z_ncu.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 1500 --csv --log-file gpurun_out/z_launches_24k.csv python bench.py --workload avia24k_map1M --steps 3 --warmup 3 --no-cpu-baseline --no-ivox --no-frontend --no-map-add --no-replay > gpurun_out/z_ncu24.log 2>&1; echo "ncu list 24k rc=$?"
