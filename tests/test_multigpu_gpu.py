"""Two-GPU runs of the sharded path (SURVEY §8e, config 5 style) through bench.py under torchrun.

Skipped on boxes with fewer than two GPUs (the round-end `-m gpu` run uses one); run with `gpurun --gpus 2`.
The sharded update (scan split over the ranks; the 96-double normal-equation record of every rank travels per trip
either through NVLink peer memory from inside the solve kernel or through ncclAllGather; sum in rank order, replicated
solve) must give the single-GPU posterior of the whole scan.
"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_sharded_update_equals_single_gpu(exchange):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "2", "--warmup", "1",
           "--workload", "avia24k_map1M", "--shard-scan", "--exchange", exchange, "--no-cpu-baseline", "--no-frontend"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    chk = line["shard_check"]
    assert chk["trips_sharded"] == chk["trips_single"]
    assert chk["n_eff_sharded"] == chk["n_eff_single"]
    assert chk["x_max_abs_diff_vs_single_gpu"] < 1e-9      # only the summation order of the partial Grams differs
    assert chk["P_max_abs_diff_vs_single_gpu"] < 1e-12
    assert line["n_gpus"] == 2 and line["scaling"] == "strong"
