#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_superblocks.py -x -q > gpurun_out/ac_pytest.log 2>&1; echo "superblock test rc=$?"; tail -15 gpurun_out/ac_pytest.log
