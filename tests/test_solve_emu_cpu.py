"""CPU-only: the LOGIC of the on-device solve (csrc/solve_dev.cuh: solve_trip / solve_pre, single-warp Gauss-Jordan, stop
policy, next-trip preparation) compiled for the host and run by 128 threads (tests/cpp/solve_emu.cpp), against the oracle's
restated solve (esekfom.hpp:1619-1928 / src/laser_mapping.cpp:178-238) trip by trip on the oracle's own normal equations."""
import os
import struct
import subprocess

import numpy as np
import pytest

import flivo_parity as par

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "fast-livo-noted_b200", "csrc")
SRC = os.path.join(ROOT, "tests", "cpp", "solve_emu.cpp")
EXE = os.path.join(ROOT, "fast-livo-noted_b200", "build", "solve_emu")


def _build():
    deps = [SRC] + [os.path.join(CSRC, f) for f in ("solve_dev.cuh", "fused.cuh", "flivo_dev.cuh", "solve.cuh")]
    if os.path.exists(EXE) and os.path.getmtime(EXE) >= max(os.path.getmtime(d) for d in deps):
        return EXE
    inc = "/usr/local/cuda/include"
    if not os.path.exists(os.path.join(inc, "cuda_runtime.h")):
        pytest.skip("CUDA headers not present")
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++20", "-O1", "-pthread", "-ffp-contract=off", "-w", "-I" + inc, "-I" + CSRC, SRC, "-o", EXE])
    return EXE


def _pack_sums(trip):
    s = np.zeros(96)
    s[:78] = trip["HTH"][np.triu_indices(12)]
    s[78:90] = trip["HTz"]
    s[91] = trip["n_eff"]
    return s


@pytest.mark.parametrize("mode", [0, 1])
def test_emulated_solve_follows_the_oracle_trip_by_trip(small_world, small_scan, mode, tmp_path):
    exe = _build()
    npd = 18 if mode == 1 else 23
    P0 = np.eye(npd) * 1e-4
    P0[0, 3] = P0[3, 0] = 2e-5          # not diagonal: the manifold Jacobians must hit off-diagonal blocks too
    o = par.oracle_update(small_world["points"], small_scan["body"], small_scan["x_prior"], P0, mode, 4)
    trips = o["trip_out"]
    Ppad = np.zeros(529); Ppad[:npd * npd] = P0.reshape(-1)   # n x n row-major prefix, as the C ABI passes it
    inp, outp = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(inp, "wb") as f:
        f.write(struct.pack("4i", mode, 4, len(trips) + 1, 0))
        f.write(np.float64(0.001).tobytes()); f.write(np.full(23, 0.001).tobytes())
        f.write(np.asarray(small_scan["x_prior"], np.float64).tobytes()); f.write(Ppad.tobytes())
        for t in trips + [trips[-1]]:     # one surplus trip: must not run (stopped)
            f.write(_pack_sums(t).tobytes())
    r = subprocess.run([exe, str(inp), str(outp)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert ("trips %d stopped 1" % len(trips)) in r.stdout, r.stdout
    raw = open(outp, "rb").read()
    rec = 36 * 8 + 23 * 8 + 8 * 4
    for k, t in enumerate(trips):
        b = raw[k * rec:(k + 1) * rec]
        x = np.frombuffer(b[:288], np.float64); dx = np.frombuffer(b[288:288 + 184], np.float64); c = np.frombuffer(b[472:], np.int32)
        assert np.abs(x - t["x_after"]).max() <= 1e-9, "trip %d: state differs by %g" % (k, np.abs(x - t["x_after"]).max())
        assert np.abs(dx[:npd] - o["dx"][k][:npd]).max() <= 1e-9
        assert c[5] == k + 1 and c[7] == t["n_eff"]
        assert c[1] == (1 if k == len(trips) - 1 else 0)                       # stopped exactly where the oracle stops
        if k + 1 < len(trips):
            assert bool(c[0]) == bool(trips[k + 1]["searched"])                 # the re-match decision for the next trip
    Pout = np.frombuffer(raw[len(trips) * rec:], np.float64)[:npd * npd].reshape(npd, npd)
    assert np.abs(Pout - o["P"]).max() <= 1e-9 * np.abs(o["P"]).max() + 1e-15
