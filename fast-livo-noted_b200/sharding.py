"""Host-side logic of the multi-GPU modes.  No data-path code lives here: the kernels and the
NCCL allgather are inside libflivo_b200.so (flivo_comm_init / flivo_iekf_update).

  replay mode (BASELINE config 4): independent scan sequences are dealt round-robin to ranks; no collective.
  sharded mode (config 5): every rank holds the whole map and a contiguous block of the (Morton-ordered) scan;
      per trip each rank reduces its block to 96 doubles (78 upper H^T H terms, 12 H^T z, res_sum, n_eff, pad),
      the blocks are all-gathered and summed IN RANK ORDER on every rank, so all replicas of the on-device
      solve see bit-identical normal equations (no floating-point reduction order depends on the fabric).
"""
import numpy as np

NTERMS, PSTRIDE = 92, 96


def shard_range(n, rank, world_size):
    """Contiguous block [lo, hi) of a scan of n points owned by `rank` (same formula as bench.py / the C++ caller)."""
    return rank * n // world_size, (rank + 1) * n // world_size


def sequences_of_rank(n_sequences, rank, world_size):
    """Round-robin partition of independent scan sequences (config 4): sequence s -> rank s mod G."""
    return list(range(rank, n_sequences, world_size))


def pack_block(HTH, HTz, res_sum, n_eff):
    """12x12 H^T H, 12 H^T z, residual sum and effective-row count -> the 96-double record the kernels exchange."""
    b = np.zeros(PSTRIDE)
    H = np.asarray(HTH, float).reshape(12, 12)
    b[:78] = H[np.triu_indices(12)]
    b[78:90] = HTz
    b[90], b[91] = res_sum, n_eff
    return b


def unpack_block(b):
    H = np.zeros((12, 12))
    iu = np.triu_indices(12)
    H[iu] = b[:78]
    H = H + np.triu(H, 1).T
    return H, b[78:90].copy(), float(b[90]), int(b[91])


def combine_rank_blocks(blocks):
    """Sum of the gathered per-rank records in rank order (SOLVE_PHASE_FROM_RANKS in csrc/solve.cu)."""
    blocks = np.asarray(blocks, float).reshape(-1, PSTRIDE)
    s = np.zeros(PSTRIDE)
    for r in range(blocks.shape[0]):
        s = s + blocks[r]
    return s


def replay_sequences(pkg, sequences, rank, world_size, P0, cfg, ds, device=0, initial_points=0, cell_ratio=2, sort_scan=True,
                     concurrent=1, reuse_contexts=False):
    """Batched replay of independent scan sequences (BASELINE config 4): sequence s belongs to rank
    s mod G, every sequence gets its own context (its own map in HBM) and no rank ever talks to another.

    One sequence = the frame loop of LaserMapping::Run (src/laser_mapping.cpp:129-251): the first scan only builds the
    map from its world-transformed points (:134-142), every later scan is one iterated update against the map as it
    stands, followed by map_incremental with the posterior (:251, :329-389).

    sequences[s] is an iterable of scans, each a dict with "body" ((n, 3) float32, lidar frame) and "x_prior" (the
    flat state handed to the update).  Returns {s: dict(x=posterior of the last scan, map_points, updates, trips,
    work=sum of trips x scan points, added=[points added per frame])} for this rank's sequences.

    concurrent > 1 drives that many of the rank's sequences at a time, one host thread and one context (one stream) each:
    contexts share nothing, so the results are those of concurrent = 1; the GPU then overlaps the single-CTA solve and
    the search tails of one sequence with the kernels of another.

    reuse_contexts: every driving thread keeps ONE context for all the sequences it is dealt (sequence s of the rank's list
    goes to thread s mod concurrent) instead of creating and destroying one per sequence — Build replaces the map, the
    insertion ids restart at 0, so a sequence's results do not depend on what the context held before
    (tests/test_gpu_replay.py); what is saved is ~17 ms of cudaMalloc / cudaFree per sequence, several times that when
    eight threads contend for the allocator, and the occasional second-long stall of a context teardown.
    """
    mine = sequences_of_rank(len(sequences), rank, world_size)
    if reuse_contexts and mine:
        workers = max(1, min(int(concurrent), len(mine)))
        chunks = [mine[w::workers] for w in range(workers)]

        def work(chunk):
            out = {}
            with pkg.Flivo(device=device, downsample_size=ds, cell_ratio=cell_ratio, initial_points=initial_points, sort_scan=sort_scan) as F:
                for s in chunk:
                    out[s] = _replay_seq(pkg, F, sequences[s], s, P0, cfg)
            return out
        if workers > 1:
            from concurrent.futures import ThreadPoolExecutor
            with ThreadPoolExecutor(max_workers=workers) as ex:
                parts = list(ex.map(work, chunks))
        else:
            parts = [work(chunks[0])]
        merged = {}
        for part in parts:
            merged.update(part)
        return {s: merged[s] for s in mine}
    if concurrent > 1 and len(mine) > 1:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=concurrent) as ex:
            parts = list(ex.map(lambda s: _replay_one(pkg, sequences, s, P0, cfg, ds, device, initial_points, cell_ratio, sort_scan), mine))
        return dict(zip(mine, parts))
    return {s: _replay_one(pkg, sequences, s, P0, cfg, ds, device, initial_points, cell_ratio, sort_scan) for s in mine}


def _replay_seq(pkg, F, frames, s, P0, cfg):
    """The frames of one sequence through context F (see _replay_one for the loop's form)."""
    res = dict(x=None, map_points=0, updates=0, trips=0, work=0, added=[], device_ms=0.0)
    bodies = [np.ascontiguousarray(fr["body"], dtype=np.float32) for fr in frames]
    F.scan_upload(bodies[0])
    w = F.h_share_model(pkg.State.from_array(frames[0]["x_prior"]), mode=cfg.mode, do_search=True, want=("world",))["world"]
    F.Build(w)
    if len(frames) > 1:
        F.scan_upload_async(bodies[1])
    for k in range(1, len(frames)):
        # flivo_frame: commit, update, stage the next scan while the update runs, map_incremental — one call
        x, _, rep, n_add = F.frame(pkg.State.from_array(frames[k]["x_prior"]), P0, cfg, bodies[k + 1] if k + 1 < len(frames) else None)
        res["added"].append(n_add)
        res["updates"] += 1
        res["trips"] += rep.trips
        res["work"] += rep.trips * len(bodies[k])
        res["device_ms"] += getattr(rep, "gpu_ms", 0.0)
        res["x"] = x.as_array()
    res["map_points"] = F.size()
    return res


def _replay_one(pkg, sequences, s, P0, cfg, ds, device, initial_points, cell_ratio, sort_scan):
    """One sequence through its own context (see replay_sequences).

    The frame loop is the pipelined form of LaserMapping::Run's step 5: while the update of frame k runs on the GPU, frame
    k+1 is staged (flivo_scan_upload_async on the context's copy stream); then map_incremental with the posterior.
    Stated deviation from src/laser_mapping.cpp:134-142: the reference starts at the identity pose and builds its first map
    from feats_down_body as it is; a replayed synthetic sequence starts at a known pose of a larger world, so the first
    scan is placed with that pose (LaserMapping::Run of host/laser_mapping.h follows the reference)."""
    import os
    import sys
    import time
    trace = os.environ.get("FLIVO_REPLAY_TRACE")     # per-sequence wall-clock phases on stderr
    t0 = time.perf_counter()
    with pkg.Flivo(device=device, downsample_size=ds, cell_ratio=cell_ratio, initial_points=initial_points, sort_scan=sort_scan) as F:
        t1 = time.perf_counter()
        res = _replay_seq(pkg, F, sequences[s], s, P0, cfg)
        t2 = time.perf_counter()
    if trace:
        t3 = time.perf_counter()
        print("[replay trace] seq %d: start %.3f s, create %.1f ms, %d frames (first scan + Build included) %.1f ms, close %.1f ms"
              % (s, t0, 1e3 * (t1 - t0), len(sequences[s]), 1e3 * (t2 - t1), 1e3 * (t3 - t2)), file=sys.stderr)
    return res
