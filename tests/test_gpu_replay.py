"""-m gpu: batched replay of independent sequences (BASELINE config 4) through sharding.replay_sequences.
Every sequence owns a context; the partition over ranks must not change any sequence's result (bitwise), and one
sequence replayed alone must follow the oracle frame by frame (first scan builds the map, later scans update + add)."""
import importlib

import numpy as np
import pytest

import flivo_parity as par
import flivo_scenes as scenes
from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu


def _sequences(n_seq, n_frames):
    world = scenes.make_world(seed=33, n_points=30000, ds=0.5)
    poses = scenes.street_positions(world, n_seq + n_frames)
    return world, [[scenes.make_scan(world, seed=900 + 17 * s + k, n_points=2000 + 100 * s, pos=poses[s + k]) for k in range(n_frames)]
                   for s in range(n_seq)]


@pytest.mark.parametrize("mode", [0, 1])
def test_partition_does_not_change_a_sequence(pkg, mode):
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    world, seqs = _sequences(4, 4)
    npd = 18 if mode == 1 else 23
    P0, cfg = np.eye(npd) * 1e-4, pkg.make_cfg(mode=mode, max_iter=4)
    whole = sh.replay_sequences(pkg, seqs, 0, 1, P0, cfg, ds=world["ds"])
    assert sorted(whole) == [0, 1, 2, 3]
    parts = {}
    for rank in range(2):
        r = sh.replay_sequences(pkg, seqs, rank, 2, P0, cfg, ds=world["ds"])
        assert sorted(r) == list(range(rank, 4, 2))
        parts.update(r)
    for s in range(4):
        assert whole[s]["updates"] == 3 and whole[s]["map_points"] > 0
        assert np.array_equal(whole[s]["x"], parts[s]["x"])
        assert (whole[s]["map_points"], whole[s]["trips"], whole[s]["added"]) == (parts[s]["map_points"], parts[s]["trips"], parts[s]["added"])


@pytest.mark.parametrize("mode", [0, 1])
def test_reused_contexts_give_the_same_sequences(pkg, mode):
    """reuse_contexts: a driving thread keeps one context for all its sequences (Build replaces the map).  Every sequence
    must come out bitwise as from a context of its own — whatever the context held before — with 1, 2 and 3 threads."""
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    world, seqs = _sequences(5, 4)
    npd = 18 if mode == 1 else 23
    P0, cfg = np.eye(npd) * 1e-4, pkg.make_cfg(mode=mode, max_iter=4)
    fresh = sh.replay_sequences(pkg, seqs, 0, 1, P0, cfg, ds=world["ds"])
    for conc in (1, 2, 3):
        got = sh.replay_sequences(pkg, seqs, 0, 1, P0, cfg, ds=world["ds"], concurrent=conc, reuse_contexts=True)
        assert sorted(got) == sorted(fresh)
        for s in fresh:
            assert np.array_equal(got[s]["x"], fresh[s]["x"]), (conc, s)
            assert (got[s]["map_points"], got[s]["trips"], got[s]["added"]) == (fresh[s]["map_points"], fresh[s]["trips"], fresh[s]["added"]), (conc, s)


def test_one_sequence_follows_the_oracle(pkg):
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    world, seqs = _sequences(1, 4)
    mode = 0
    P0, cfg = np.eye(23) * 1e-4, pkg.make_cfg(mode=mode, max_iter=4)
    got = sh.replay_sequences(pkg, seqs, 0, 1, P0, cfg, ds=world["ds"])[0]
    # oracle: Build from the first scan's world points (no down-sampling), then update + Add_Points(downsample) per frame
    s0 = seqs[0][0]
    M = orc.OracleMap(world["ds"])
    M.build(orc.body_to_world(s0["body"], orc.State.from_array(s0["x_prior"])))
    added, x = [], None
    for scan in seqs[0][1:]:
        cur, _ = M.live()
        o = par.oracle_update(cur, scan["body"], scan["x_prior"], P0, mode, 4, v1_bound=pkg.v1_search_bound_sq(world["ds"]))
        x = o["x"]
        added.append(M.add_points(orc.body_to_world(scan["body"], orc.State.from_array(o["x"])), True))
    assert np.abs(got["x"] - x).max() < 1e-6
    # the maps were grown with each side's own posterior: one ulp of state may move a point across a voxel boundary
    assert all(abs(a - b) <= max(2, b // 200) for a, b in zip(got["added"], added)), (got["added"], added)
    assert abs(got["map_points"] - len(M.live()[0])) <= max(4, len(M.live()[0]) // 200)
