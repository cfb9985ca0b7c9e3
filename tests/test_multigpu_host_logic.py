"""CPU, world_size 2 over gloo: the host-side logic of the sharded multi-GPU mode (fast-livo-noted_b200/sharding.py).
Each rank reduces its block of the scan with the oracle, the 96-double records are all-gathered and summed in rank
order; every rank must hold bit-identical normal equations, equal (to rounding) to the single-process result, and
the replicated solve must therefore give the same state increment everywhere."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world_size, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    import __graft_entry__ as ge
    import importlib
    ge.load_package()
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    import flivo_parity as par
    import flivo_scenes as scenes
    from oracle import pyoracle as orc
    world = scenes.make_world(seed=7, n_points=20000, ds=0.5)
    scan = scenes.make_scan(world, seed=8, n_points=1501)      # odd size: uneven shards
    lo, hi = sh.shard_range(len(scan["body"]), rank, world_size)
    cache = {}
    o = par.oracle_trip(world["points"], scan["body"][lo:hi], scan["x_prior"], True, cache, orc.MODE_V3)
    mine = torch.from_numpy(sh.pack_block(o["HTH"], o["HTz"], o["res_sum"], o["n_eff"]))
    gathered = [torch.zeros(sh.PSTRIDE, dtype=torch.float64) for _ in range(world_size)]
    dist.all_gather(gathered, mine)
    total = sh.combine_rank_blocks(np.stack([g.numpy() for g in gathered]))
    HTH, HTz, res_sum, n_eff = sh.unpack_block(total)
    x = orc.State.from_array(scan["x_prior"])
    r = orc.solve_v3(x, x.copy(), np.eye(23) * 1e-4, HTH, HTz, 0.001, -1, 4, np.full(23, 0.001), 0)
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), total=total, dx=r["dx"], lo=lo, hi=hi, n_eff=n_eff)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_normal_equations_world_size_2(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(os.path.join(tmp_path, "rank0.npz")), np.load(os.path.join(tmp_path, "rank1.npz"))
    assert (int(r0["lo"]), int(r0["hi"]), int(r1["lo"]), int(r1["hi"])) == (0, 750, 750, 1501)
    assert np.array_equal(r0["total"], r1["total"])            # bit-identical on every rank
    assert np.array_equal(r0["dx"], r1["dx"])
    # against the single-process trip over the whole scan
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import flivo_parity as par
    import flivo_scenes as scenes
    from oracle import pyoracle as orc
    world = scenes.make_world(seed=7, n_points=20000, ds=0.5)
    scan = scenes.make_scan(world, seed=8, n_points=1501)
    o = par.oracle_trip(world["points"], scan["body"], scan["x_prior"], True, {}, orc.MODE_V3)
    import __graft_entry__ as ge
    import importlib
    ge.load_package()
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    HTH, HTz, res_sum, n_eff = sh.unpack_block(r0["total"])
    assert n_eff == o["n_eff"]
    assert par.rel_err_block(HTH, o["HTH"]) < 1e-12 and par.rel_err_block(HTz, o["HTz"]) < 1e-10


def test_partition_helpers(pkg):
    import importlib
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    for n in (0, 1, 7, 1000, 4_000_000):
        for g in (1, 2, 4, 8):
            r = [sh.shard_range(n, k, g) for k in range(g)]
            assert r[0][0] == 0 and r[-1][1] == n and all(r[i][1] == r[i + 1][0] for i in range(g - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
    seqs = [sh.sequences_of_rank(64, k, 8) for k in range(8)]
    assert sorted(sum(seqs, [])) == list(range(64)) and all(len(s) == 8 for s in seqs)
    H = np.arange(144.0).reshape(12, 12); H = H + H.T
    b = sh.pack_block(H, np.arange(12.0), 3.5, 17)
    H2, z2, rs, ne = sh.unpack_block(b)
    assert np.array_equal(H, H2) and np.array_equal(z2, np.arange(12.0)) and rs == 3.5 and ne == 17


class _FakeReport:
    trips = 4


class _FakeState:
    def __init__(self, a):
        self.a = np.asarray(a, dtype=np.float64)

    @staticmethod
    def from_array(a):
        return _FakeState(a)

    def as_array(self):
        return self.a


class _FakeFlivo:
    """Records the calls of sharding.replay_sequences (no device): the call order per sequence is the contract."""
    log = []

    def __init__(self, **kw):
        self.kw, self.n, self.calls = kw, 0, []
        _FakeFlivo.log.append(self.calls)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.calls.append("close")

    def scan_upload(self, body):
        self.calls.append(("upload", len(body)))

    def h_share_model(self, st, mode, do_search, want):
        self.calls.append("trip")
        return {"world": np.zeros((3, 3), np.float32)}

    def Build(self, w):
        self.calls.append("build")
        self.n += len(w)

    def scan_upload_async(self, body):
        self.calls.append(("stage", len(body)))

    def scan_commit(self):
        self.calls.append("commit")

    def frame(self, st, P0, cfg, next_body=None):
        self.calls.append(("frame", None if next_body is None else len(next_body)))
        self.n += 2
        return _FakeState(st.as_array() + 1.0), P0, _FakeReport(), 2

    def iekf_update_begin(self, st, P0, cfg):
        self.calls.append("begin")
        self._pending = (st, P0)

    def iekf_update_end(self):
        self.calls.append("end")
        st, P0 = self._pending
        return _FakeState(st.as_array() + 1.0), P0, _FakeReport()

    def map_add_scan(self, x):
        self.calls.append("add")
        self.n += 2
        return 2

    def size(self):
        return self.n


def test_replay_sequences_control_flow(pkg):
    """config 4: sequence s -> rank s mod G, own context per sequence, first scan builds, later scans update + add."""
    import importlib
    import types
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    fake = types.SimpleNamespace(Flivo=_FakeFlivo, State=_FakeState)
    cfg = types.SimpleNamespace(mode=1)
    seqs = [[{"body": np.zeros((10 + s, 3), np.float32), "x_prior": np.full(24, float(s))} for _ in range(3 + s % 2)] for s in range(5)]
    seen = {}
    for rank in range(2):
        _FakeFlivo.log = []
        r = sh.replay_sequences(fake, seqs, rank, 2, np.eye(18), cfg, ds=0.5)
        assert sorted(r) == list(range(rank, 5, 2))
        assert len(_FakeFlivo.log) == len(r)          # one context per sequence
        for calls, s in zip(_FakeFlivo.log, sorted(r)):
            frames = len(seqs[s])
            assert calls[:3] == [("upload", 10 + s), "trip", "build"]
            # pipelined frame loop: the next scan is staged between begin and end of the running update
            want = [("stage", 10 + s)] + [("frame", 10 + s if k + 1 < frames else None) for k in range(1, frames)]
            assert calls[3:] == want + ["close"]
            assert r[s]["updates"] == frames - 1 and r[s]["trips"] == 4 * (frames - 1)
            assert r[s]["work"] == 4 * (frames - 1) * (10 + s) and r[s]["added"] == [2] * (frames - 1)
            assert r[s]["map_points"] == 3 + 2 * (frames - 1)
            assert np.array_equal(r[s]["x"], np.full(24, float(s) + 1.0))
        seen.update(r)
    assert sorted(seen) == list(range(5))
    # concurrent sequences of one rank (one host thread and context each): same results, one context per sequence
    _FakeFlivo.log = []
    r = sh.replay_sequences(fake, seqs, 0, 1, np.eye(18), cfg, ds=0.5, concurrent=3)
    assert sorted(r) == list(range(5)) and len(_FakeFlivo.log) == 5
    for s in range(5):
        assert r[s]["updates"] == seen[s]["updates"] and r[s]["work"] == seen[s]["work"] and np.array_equal(r[s]["x"], seen[s]["x"])
    for calls in _FakeFlivo.log:
        assert calls[1:3] == ["trip", "build"] and calls[-1] == "close"      # no interleaving inside a context
    # reuse_contexts: one context per driving thread, sequence s of the rank's list -> thread s mod concurrent, Build per sequence
    for conc, n_ctx in ((1, 1), (2, 2), (8, 5)):
        _FakeFlivo.log = []
        r = sh.replay_sequences(fake, seqs, 0, 1, np.eye(18), cfg, ds=0.5, concurrent=conc, reuse_contexts=True)
        assert sorted(r) == list(range(5)) and len(_FakeFlivo.log) == n_ctx
        assert sum(calls.count("build") for calls in _FakeFlivo.log) == 5 and all(calls[-1] == "close" for calls in _FakeFlivo.log)
        for s in range(5):
            assert r[s]["updates"] == seen[s]["updates"] and r[s]["work"] == seen[s]["work"] and np.array_equal(r[s]["x"], seen[s]["x"])
