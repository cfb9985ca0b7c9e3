"""-m gpu: behaviour at the edges of the C ABI — call order, empty inputs, repeated use, device-memory hygiene."""
import numpy as np
import pytest
import torch

import flivo_frontend as fe
import flivo_scenes as scenes

pytestmark = pytest.mark.gpu


def test_call_order_errors_are_reported_not_fatal(pkg, small_world, small_scan):
    with pkg.Flivo(downsample_size=0.5) as F:
        st = pkg.State.from_array(small_scan["x_prior"])
        with pytest.raises(pkg.FlivoError):                       # update before any scan
            F.iekf_update(st, np.eye(23) * 1e-4, mode=0, max_iter=4)
        with pytest.raises(pkg.FlivoError):                       # down-sample before a raw scan
            F.raw_downsample(0.5)
        assert F.scan_download().shape == (0, 4)
        F.scan_upload(small_scan["body"])
        # empty map: every point is masked, the update leaves the state where it was and says so
        x, P, rep = F.iekf_update(st, np.eye(23) * 1e-4, mode=0, max_iter=4)
        assert rep.n_eff == 0 and np.abs(x.as_array() - small_scan["x_prior"]).max() < 1e-12
        with pytest.raises(pkg.FlivoError):
            F.iekf_update(st, np.eye(23) * 1e-4, mode=0, max_iter=99)   # max_iter out of range
        F.Build(small_world["points"])                           # the context is still usable afterwards
        x, P, rep = F.iekf_update(st, np.eye(23) * 1e-4, mode=0, max_iter=4)
        assert rep.n_eff > 0


def test_results_do_not_depend_on_what_ran_before(pkg, small_world, small_scan):
    """The captured graph, the plane cache and the work counters are per-update state: interleaving scans of other
    sizes, the other flavour and front-end calls must not change a result."""
    st = pkg.State.from_array(small_scan["x_prior"])
    P0 = np.eye(23) * 1e-4
    other = scenes.make_scan(small_world, seed=91, n_points=777)
    with pkg.Flivo(downsample_size=small_world["ds"]) as F:
        F.Build(small_world["points"])
        F.scan_upload(small_scan["body"])
        x0, Pa, r0 = F.iekf_update(st, P0, mode=0, max_iter=4)
        for k in range(3):
            F.scan_upload(other["body"][: 200 + 100 * k])
            F.iekf_update(pkg.State.from_array(other["x_prior"]), np.eye(18) * 1e-4, mode=1, max_iter=3)
            xyz, t, inten = fe.make_raw(k, 5000)
            F.raw_upload(fe.raw_records(xyz, t, inten), time_col=9, intensity_col=8)
            F.raw_downsample(0.5)
            F.h_share_model(pkg.State.from_array(other["x_prior"]), 0, True)
            F.scan_upload(small_scan["body"])
            x1, Pb, r1 = F.iekf_update(st, P0, mode=0, max_iter=4)
            assert np.array_equal(x0.as_array(), x1.as_array()) and np.array_equal(Pa, Pb)
            assert (r0.trips, r0.searches, r0.n_eff) == (r1.trips, r1.searches, r1.n_eff)


def test_contexts_release_their_device_memory(pkg, small_world, small_scan):
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        with pkg.Flivo(downsample_size=small_world["ds"], initial_points=len(small_world["points"])) as F:
            F.Build(small_world["points"])
            F.scan_upload(small_scan["body"])
            F.iekf_update(pkg.State.from_array(small_scan["x_prior"]), np.eye(23) * 1e-4, mode=0, max_iter=4)
            xyz, t, inten = fe.make_raw(3, 20000)
            F.raw_upload(fe.raw_records(xyz, t, inten), time_col=9, intensity_col=8)
            F.raw_downsample(0.5)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free0 - free1 < 64 << 20          # nothing but allocator slack stays behind


def test_two_contexts_are_independent(pkg, small_world, small_scan):
    st = pkg.State.from_array(small_scan["x_prior"])
    P0 = np.eye(23) * 1e-4
    with pkg.Flivo(downsample_size=small_world["ds"]) as A, pkg.Flivo(downsample_size=small_world["ds"]) as B:
        A.Build(small_world["points"])
        B.Build(small_world["points"][::2])
        A.scan_upload(small_scan["body"]); B.scan_upload(small_scan["body"])
        xa, _, ra = A.iekf_update(st, P0, mode=0, max_iter=4)
        xb, _, rb = B.iekf_update(st, P0, mode=0, max_iter=4)
        xa2, _, ra2 = A.iekf_update(st, P0, mode=0, max_iter=4)
        assert np.array_equal(xa.as_array(), xa2.as_array()) and ra.n_eff == ra2.n_eff
        assert ra.n_eff != rb.n_eff                      # B's map is thinner


def test_staging_graph_across_scan_sizes(pkg, small_world):
    """The scan staging (keys, sort on n padded to 8192, gather) is a captured graph keyed by the padded size: sizes
    inside one bucket, on its edges and in other buckets, uploads that are overwritten before any update, and both
    upload paths (host AoS, device pointer) must give the posterior of the unsorted path."""
    big = scenes.make_scan(small_world, seed=77, n_points=17000)
    P0 = np.eye(23) * 1e-4
    st = pkg.State.from_array(big["x_prior"])
    with pkg.Flivo(downsample_size=small_world["ds"], sort_scan=True) as F, \
            pkg.Flivo(downsample_size=small_world["ds"], sort_scan=False) as G:
        F.Build(small_world["points"]); G.Build(small_world["points"])
        for n in (3000, 8191, 8192, 8193, 3000, 17000, 16385, 1):
            body = np.ascontiguousarray(big["body"][:n])
            F.scan_upload(np.ascontiguousarray(big["body"][: max(1, n // 2)]))   # overwritten at once: never used
            if n % 2:
                F.scan_upload(body)
                keep = None
            else:
                keep = torch.zeros((n, 4), dtype=torch.float32, device="cuda")
                keep[:, :3] = torch.from_numpy(body).cuda()
                torch.cuda.synchronize()
                F.scan_set_device(keep.data_ptr(), n)
            G.scan_upload(body)
            xf, Pf, rf = F.iekf_update(st, P0, mode=0, max_iter=4)
            xg, Pg, rg = G.iekf_update(st, P0, mode=0, max_iter=4)
            assert (rf.trips, rf.searches, rf.n_eff) == (rg.trips, rg.searches, rg.n_eff), n
            assert np.abs(xf.as_array() - xg.as_array()).max() < 1e-9, n
            hf = F.h_share_model(st, 0, True, want=("nbr_id", "selected"))
            hg = G.h_share_model(st, 0, True, want=("nbr_id", "selected"))
            assert np.array_equal(hf["nbr_id"], hg["nbr_id"]) and np.array_equal(hf["selected"], hg["selected"]), n
