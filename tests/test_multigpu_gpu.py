"""The sharded update: the scan split over ranks, every rank holding the whole map; per trip the
96-double normal-equation record of every rank travels either through peer memory from inside k_fused's solving CTA
(stores into the peers' exchange buffers + release/acquire sequence flags) or through ncclAllGather; sum in rank order,
replicated solve.

One-GPU cases (run on the round-end 1-GPU lease): all ranks on cuda:0 —
  * in one process (contexts exchanging by device pointer: flivo_peer_attach_local), checked against the ORACLE's
    update of the whole scan; 2 and 4 ranks, both modes;
  * a rank that shows up late: the time-out leaves state and P untouched, refuses further updates, flivo_peer_reset re-arms;
  * in two processes through cudaIpc handles (bench.py --shard-scan under torchrun, both ranks on device 0).
Two-GPU cases (skipped on a 1-GPU lease; run with `gpurun --gpus 2`): bench.py under torchrun over NVLink, peer and NCCL.
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import flivo_parity as par

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("nranks,mode", [(2, 0), (4, 0), (2, 1), (3, 1)])
def test_sharded_update_in_process_matches_oracle(pkg, small_world, small_scan, nranks, mode):
    par.sharded_update_in_process(pkg, small_world, small_scan, nranks=nranks, mode=mode, verbose=True)


def test_peer_timeout_then_reset(pkg, small_world, small_scan):
    """One rank arrives 1.5 s late with a 0.5 s time-out: the punctual rank gives up (FLIVO_E_STATE, nothing written), the
    late one times out in turn; every later update is refused until both ranks reset; then the update is the oracle's."""
    r = par.sharded_update_in_process(pkg, small_world, small_scan, nranks=2, delay_rank=1, delay_s=1.5, timeout_s=0.5)
    ctxs = r["ctxs"]
    try:
        assert all(e is not None and e.code == pkg.E_STATE for e in r["errors"]), r["errors"]
        assert "timed out" in str(r["errors"][0])
        r2 = par.sharded_update_in_process(pkg, small_world, small_scan, nranks=2, ctxs=ctxs)
        assert all(e is not None and e.code == pkg.E_STATE for e in r2["errors"])   # refused on the host, no 2 s wait
        for F in ctxs:
            F.peer_reset()
        r3 = par.sharded_update_in_process(pkg, small_world, small_scan, nranks=2, ctxs=ctxs, timeout_s=2.0, verbose=True)
        assert not any(r3["errors"]) and r3["x_err"] <= 1e-6
    finally:
        for F in ctxs:
            F.close()


def _torchrun_bench(extra, port, env_extra=None, nproc=2, timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", str(nproc), "--steps", "2", "--warmup", "1",
           "--workload", "avia24k_map1M", "--no-cpu-baseline", "--no-frontend", "--no-ivox", "--no-map-add"] + extra
    env = dict(os.environ)
    env.update(env_extra or {})
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT, env=env)
    assert out.returncode == 0, out.stderr[-3000:]
    return json.loads(out.stdout.strip().splitlines()[-1])


def _check_sharded_line(line):
    chk = line["shard_check"]
    assert chk["trips_sharded"] == chk["trips_single"]
    assert chk["n_eff_sharded"] == chk["n_eff_single"]
    assert chk["x_max_abs_diff_vs_single_gpu"] < 1e-9      # only the summation order of the partial Grams differs
    assert chk["P_max_abs_diff_vs_single_gpu"] < 1e-12
    assert line["scaling"] == "strong"


def test_sharded_update_two_processes_one_gpu_ipc():
    """Both ranks on cuda:0 (FLIVO_BENCH_ONE_DEVICE=1): the exchange buffers cross the process boundary as cudaIpc handles,
    as they do between GPUs.  The two contexts time-slice one GPU, so an exchange can cost a scheduling quantum: only
    correctness is asserted.  gloo carries the rendezvous (NCCL refuses two ranks on one device)."""
    line = _torchrun_bench(["--shard-scan", "--exchange", "peer"], 29543, {"FLIVO_BENCH_ONE_DEVICE": "1"})
    _check_sharded_line(line)


@pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_sharded_update_equals_single_gpu(exchange):
    line = _torchrun_bench(["--shard-scan", "--exchange", exchange], 29541)
    _check_sharded_line(line)
    assert line["n_gpus"] == 2
