"""GPU parity of the scan path (K1 / K2) through the C-ABI, against the tier-B
oracle, the committed golden vectors of the reference (tier A), and
size-independent properties at BASELINE.json's full sizes.

Tolerances (BASELINE.json north_star): per-ray range within 1e-5 relative of the
oracle; hit-segment index identical except at certified ties (two candidates
whose ranges at that ray agree within 1e-5 relative; the lowest index wins in
the oracle).
"""
import ctypes as C
import glob
import os

import numpy as np
import pytest

import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import _lib, synth

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
SCAN_GOLDEN = sorted(g for g in glob.glob(os.path.join(HERE, "golden", "*.npz"))
                     if "primitives" not in g)
REL_TOL = 1e-5


def _load(path):
    d = np.load(path)
    return {k: d[k] for k in d.files}


def check_against_tierb(oracle, segs, obj, fan, range_max, n, r_gpu, i_gpu, skip=(0, 0)):
    """Returns (#rays bit-identical, #certified ties); asserts the contract."""
    r_o, i_o = oracle.scan_tierb(segs, obj, fan[:2], range_max, fan[2], fan[3], n, skip=skip)
    rel = np.abs(r_gpu.astype(np.float64) - r_o) / np.maximum(np.abs(r_o.astype(np.float64)), 1e-9)
    assert rel.max() <= REL_TOL, "range off by %.3g rel at ray %d" % (rel.max(), rel.argmax())
    ties = 0
    allseg = np.concatenate([np.asarray(segs, np.float32).reshape(-1, 4),
                             np.asarray(obj if obj is not None else np.zeros((0, 4)), np.float32).reshape(-1, 4)])
    for k in np.nonzero(i_gpu != i_o)[0]:
        assert i_gpu[k] >= 0 and i_o[k] >= 0, "hit/no-hit disagreement at ray %d" % k
        ra = oracle.candidate_range(allseg[i_gpu[k]], fan[:2], fan[2], fan[3], n, int(k))
        assert np.isfinite(ra) and abs(ra - r_o[k]) <= REL_TOL * abs(r_o[k]), \
            "idx mismatch at ray %d is not a tie: %r vs %r" % (k, ra, r_o[k])
        ties += 1
    return int((r_gpu.view(np.uint32) == r_o.view(np.uint32)).sum()), ties


@pytest.mark.parametrize("path", SCAN_GOLDEN, ids=[os.path.basename(p) for p in SCAN_GOLDEN])
def test_predicted_scan_matches_golden(oracle, path):
    """Drop-in umrs_predicted_scan (VectorMap::GetPredictedScan's argument list) on
    the committed fixtures: tier B within tolerance + exact idx; every ray that
    disagrees with the reference's own output (tier A) is one where tier B
    disagrees with it too (SURVEY.md 8c exception classes D1-D6)."""
    g = _load(path)
    n, rmax = int(g["num_rays"]), float(g["range_max"])
    vmap = umrs.VectorMap(lines=g["raw"], cleaned=False)
    assert np.array_equal(vmap.lines.view(np.uint32), g["segs"].view(np.uint32))
    vmap.object_lines = g["obj"]
    exact = total = 0
    for f, ra, rb, ib in zip(g["fans"], g["tier_a"], g["tier_b"], g["idx_b"]):
        r, i = vmap.GetPredictedScan(f[:2], 0.4, rmax, f[2], f[3], n, return_index=True)
        e, _ = check_against_tierb(oracle, g["segs"], g["obj"], f, rmax, n, r, i)
        exact += e
        total += n
        rel_b = np.abs(r - rb) / np.maximum(np.abs(rb), 1e-9)
        assert rel_b.max() <= REL_TOL
        rel_a = np.abs(r.astype(np.float64) - ra) / np.maximum(np.abs(ra), 1e-9)
        rel_ab = np.abs(rb.astype(np.float64) - ra) / np.maximum(np.abs(ra), 1e-9)
        # GPU-vs-reference disagreements are a subset of tierB-vs-reference ones
        # (2e-5 on the left so a ray sitting at the threshold cannot flip sides)
        assert not np.any((rel_a > 2e-5) & (rel_ab <= 1e-5))
    assert exact / total > 0.999, "only %.4f of rays bit-identical to the oracle" % (exact / total)


def test_scan_batch_equals_single_scans(oracle):
    g = _load(os.path.join(HERE, "golden", "floorplan_256.npz"))
    n, rmax = int(g["num_rays"]), float(g["range_max"])
    vmap = umrs.VectorMap(lines=g["segs"])
    rb, ib = vmap.ScanBatch(g["fans"], rmax, n, return_index=True)
    for k, f in enumerate(g["fans"]):
        r, i = vmap.GetPredictedScan(f[:2], 0.4, rmax, f[2], f[3], n, return_index=True)
        assert np.array_equal(r.view(np.uint32), rb[k].view(np.uint32))
        assert np.array_equal(i, ib[k])


def test_empty_map_and_edge_sizes(oracle):
    """Empty candidate set -> every ray keeps the fill value range_max
    (vector_map.cpp:433); num_rays == 1; zero fans; hits beyond range_max inside
    the AABB are returned unclamped (class D7)."""
    vmap = umrs.VectorMap(lines=np.zeros((0, 4), np.float32))
    r, i = vmap.GetPredictedScan((0, 0), 0.4, 10.0, -1.0, 1.0, 7, return_index=True)
    assert np.all(r == np.float32(10.0)) and np.all(i == -1)
    assert vmap.ScanBatch(np.zeros((0, 4), np.float32), 10.0, 5).shape == (0, 5)
    # one wall at x = 9.9, seen from the origin with range_max 10: the corner
    # rays hit it beyond 10 m and are NOT clamped
    wall = np.array([[9.9, -9.9, 9.9, 9.9]], np.float32)
    vmap = umrs.VectorMap(lines=wall)
    r, i = vmap.GetPredictedScan((0, 0), 0.4, 10.0, -0.7, 0.7, 64, return_index=True)
    check_against_tierb(oracle, wall, None, np.array([0, 0, -0.7, 0.7], np.float32), 10.0, 64, r, i)
    assert r.max() > 10.0 and np.all(i == 0)
    r1 = vmap.GetPredictedScan((0, 0), 0.4, 10.0, 0.0, 0.1, 1)
    assert r1.shape == (1,) and r1[0] == np.float32(9.9)
    # AABB cull: a wall just outside [loc +- range_max]^2 is not a candidate
    far = np.array([[10.5, -1, 10.5, 1]], np.float32)
    r = umrs.VectorMap(lines=far).GetPredictedScan((0, 0), 0.4, 10.0, -0.5, 0.5, 9)
    assert np.all(r == np.float32(10.0))


def test_wraparound_and_full_circle_fans(oracle):
    """Segments straddling the -x axis (a0 > a1 after the swap) and fans whose
    world-frame limits exceed +-pi (angle_min + theta), incl. the 360 deg jackal."""
    rng = np.random.default_rng(2)
    segs = umrs.map_cleanup(synth.floorplan_segments(128, extent=60.0, seed=8))
    vmap = umrs.VectorMap(lines=segs)
    poses = synth.sample_free_poses(segs, 12, 60.0, seed=3)
    for p in poses:
        for amin, amax, n in ((-2.356, 2.356, 1081), (-np.pi, np.pi, 3601), (-0.3, 0.2, 33)):
            th = rng.uniform(-7, 7)  # un-normalised heading
            f = np.array([p[0], p[1], np.float32(amin) + np.float32(th), np.float32(amax) + np.float32(th)], np.float32)
            r, i = vmap.GetPredictedScan(f[:2], 0.4, 100.0, f[2], f[3], n, return_index=True)
            check_against_tierb(oracle, segs, None, f, 100.0, n, r, i)


def test_ties_resolve_to_lowest_index(oracle):
    """Two identical walls: the oracle keeps the first (strict <); so must the GPU."""
    segs = np.array([[5, -3, 5, 3], [5, -3, 5, 3], [7, -3, 7, 3]], np.float32)
    vmap = umrs.VectorMap(lines=segs)
    r, i = vmap.GetPredictedScan((0, 0), 0.4, 50.0, -0.4, 0.4, 101, return_index=True)
    assert np.all(i == 0)
    check_against_tierb(oracle, segs, None, np.array([0, 0, -0.4, 0.4], np.float32), 50.0, 101, r, i)


def test_object_lines_occlude(oracle):
    g = _load(os.path.join(HERE, "golden", "clutter_objects.npz"))
    n, rmax = int(g["num_rays"]), float(g["range_max"])
    vmap = umrs.VectorMap(lines=g["segs"])
    f = g["fans"][0]
    r0, i0 = vmap.GetPredictedScan(f[:2], 0.4, rmax, f[2], f[3], n, return_index=True)
    vmap.object_lines = g["obj"]
    r1, i1 = vmap.GetPredictedScan(f[:2], 0.4, rmax, f[2], f[3], n, return_index=True)
    assert np.all(r1 <= r0)
    assert np.all(i0 < len(g["segs"]))
    hits_obj = i1 >= len(g["segs"])
    assert np.array_equal(r1[~hits_obj].view(np.uint32), r0[~hits_obj].view(np.uint32))


def test_invalid_arguments_return_status():
    lib = _lib.load()
    vmap = umrs.VectorMap(lines=np.array([[0, 0, 1, 0]], np.float32))
    out = np.zeros(4, np.float32)
    fp = out.ctypes.data_as(C.POINTER(C.c_float))
    assert lib.umrs_predicted_scan(None, None, 0, 0, 0, 0.4, 10, -1, 1, 4, fp, None) == _lib.ERR_INVALID
    assert lib.umrs_predicted_scan(vmap._h, None, 0, 0, 0, 0.4, 10, -1, 1, 0, fp, None) == _lib.ERR_INVALID
    assert lib.umrs_predicted_scan(vmap._h, None, 0, 0, 0, 0.4, 10, -1, 1, 4, None, None) == _lib.ERR_INVALID
    assert lib.umrs_predicted_scan(vmap._h, None, 3, 0, 0, 0.4, 10, -1, 1, 4, fp, None) == _lib.ERR_INVALID
    assert lib.umrs_map_create(None, 5, C.byref(C.c_void_p())) == _lib.ERR_INVALID


def test_map_load_file_roundtrip(tmp_path, oracle):
    """VectorMap::Load: text format + Cleanup (vector_map.cpp:397-411)."""
    raw = synth.floorplan_segments(64, extent=50.0, seed=4)
    path = tmp_path / "m.vectormap.txt"
    synth.write_map(str(path), raw)
    vmap = umrs.VectorMap(file=str(path))
    assert np.array_equal(vmap.lines.view(np.uint32), oracle.map_cleanup(raw).view(np.uint32))


@pytest.mark.parametrize("kernel", [0, 1], ids=["k1_smem", "k2_grid"])
def test_forced_kernels_agree(oracle, kernel):
    """The shared-memory kernel and the uniform-grid walk are the same function."""
    lib = _lib.load()
    g = _load(os.path.join(HERE, "golden", "floorplan_1024.npz"))
    n, rmax = int(g["num_rays"]), float(g["range_max"])
    vmap = umrs.VectorMap(lines=g["segs"])
    _lib.check(lib.umrs_map_set_scan_kernel(vmap._h, kernel), "set_scan_kernel")
    r, i = vmap.ScanBatch(g["fans"], rmax, n, return_index=True)
    for k, f in enumerate(g["fans"]):
        check_against_tierb(oracle, g["segs"], None, f, rmax, n, r[k], i[k])


def test_large_map_grid_walk(oracle):
    """BASELINE config 4 (scaled for the checker): 100k-segment city-block map, the
    candidate set does not fit shared memory -> uniform-grid kernel.  Full tier-B
    check on 8 fans; properties on 925 fans x 1081 rays (~1M rays)."""
    segs = synth.cityblock_segments(100000, extent=2000.0, seed=99)
    vmap = umrs.VectorMap(lines=segs)  # blocks never touch: already "clean"
    rng = np.random.default_rng(4)
    n_fans, n = 925, 1081
    poses = np.stack([rng.uniform(100, 1900, n_fans), rng.uniform(100, 1900, n_fans),
                      rng.uniform(-np.pi, np.pi, n_fans)], 1).astype(np.float32)
    d135 = np.float32(np.pi * 135.0 / 180.0)
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - d135, poses[:, 2] + d135], 1).astype(np.float32)
    r, i = vmap.ScanBatch(fans, 100.0, n, return_index=True)
    assert r.shape == (n_fans, n)
    for k in range(8):
        check_against_tierb(oracle, segs, None, fans[k], 100.0, n, r[k], i[k])
    # properties over the full 1M rays: finite, positive, bounded by the AABB corner;
    # no-hit rays carry exactly range_max; idx in range
    assert np.isfinite(r).all() and (r > 0).all() and r.max() <= 100.0 * np.sqrt(2) * (1 + 1e-5)
    assert np.all(r[i < 0] == np.float32(100.0)) and i.max() < len(segs)
    # determinism
    r2 = vmap.ScanBatch(fans, 100.0, n)
    assert np.array_equal(r.view(np.uint32), r2.view(np.uint32))


@pytest.mark.parametrize("kernel", [0, 1], ids=["k1_smem", "k2_grid"])
def test_multipass_when_candidates_exceed_the_slots(oracle, kernel):
    """The "cap" tuning hook bounds the shared-memory candidate slots per pass: with 96 slots a
    1024-segment fan takes ~11 passes; results (incl. ties across passes and the
    object lines, which live in the last source row) must not change."""
    lib = _lib.load()
    g = _load(os.path.join(HERE, "golden", "clutter_objects.npz"))
    n, rmax = int(g["num_rays"]), float(g["range_max"])
    vmap = umrs.VectorMap(lines=g["segs"])
    vmap.object_lines = g["obj"]
    _lib.check(lib.umrs_map_set_scan_kernel(vmap._h, kernel), "set_scan_kernel")
    with _lib.tuning(cap=96):
        r, i = vmap.ScanBatch(g["fans"], rmax, n, return_index=True)
    for k, f in enumerate(g["fans"]):
        check_against_tierb(oracle, g["segs"], g["obj"], f, rmax, n, r[k], i[k])
    r1, i1 = vmap.ScanBatch(g["fans"], rmax, n, return_index=True)
    assert np.array_equal(r.view(np.uint32), r1.view(np.uint32)) and np.array_equal(i, i1)
    # duplicates split across passes still resolve to the lowest index
    with _lib.tuning(cap=32):
        segs = np.array([[5, -3, 5, 3]] * 70 + [[7, -3, 7, 3]], np.float32)
        vm2 = umrs.VectorMap(lines=segs)
        _lib.check(lib.umrs_map_set_scan_kernel(vm2._h, kernel), "set_scan_kernel")
        r2, i2 = vm2.GetPredictedScan((0, 0), 0.4, 50.0, -0.4, 0.4, 101, return_index=True)
    assert np.all(i2 == 0)
    # the hooks validate their arguments and never read the environment
    assert lib.umrs_tuning_set(b"cap", 7) == _lib.ERR_INVALID
    assert lib.umrs_tuning_set(b"cta_shape", 9) == _lib.ERR_INVALID
    assert lib.umrs_tuning_set(b"no_such_key", 1) == _lib.ERR_INVALID
    assert lib.umrs_tuning_set(b"split_step", 2) == _lib.ERR_INVALID
    v = C.c_int32(7)
    assert lib.umrs_tuning_get(b"split_step", C.byref(v)) == 0 and v.value == -1


def test_grid_kernel_with_long_segments_and_far_poses(oracle):
    """K2 corner cases: segments longer than two cells (always-visited tail), fans
    outside the map's bounding box, fans that see nothing."""
    lib = _lib.load()
    rng = np.random.default_rng(12)
    blocks = synth.cityblock_segments(6000, extent=400.0, seed=5)
    longs = np.array([[0, 0, 400, 0], [400, 0, 400, 400], [400, 400, 0, 400], [0, 400, 0, 0],
                      [10, 200, 390, 203]], np.float32)
    segs = np.concatenate([longs, blocks]).astype(np.float32)
    vmap = umrs.VectorMap(lines=segs)
    _lib.check(lib.umrs_map_set_scan_kernel(vmap._h, 1), "set_scan_kernel")
    fans = []
    for _ in range(10):
        x, y, th = rng.uniform(5, 395), rng.uniform(5, 395), rng.uniform(-3, 3)
        fans.append([x, y, th - 2.356, th + 2.356])
    fans += [[-50, -50, 0.0, 1.5], [450, 200, 2.0, 4.0], [200, -30, 0.5, 2.5], [5000, 5000, 0, 1]]
    fans = np.array(fans, np.float32)
    r, i = vmap.ScanBatch(fans, 60.0, 541, return_index=True)
    for k, f in enumerate(fans):
        check_against_tierb(oracle, segs, None, f, 60.0, 541, r[k], i[k])
    assert np.all(r[-1] == np.float32(60.0)) and np.all(i[-1] == -1)
    assert (i[:10] < 5).any() and (i[:10] >= 5).any()   # both long walls and blocks are hit


@pytest.mark.parametrize("n_walls", [100, 400], ids=["spill_list", "sweep_every_slot"])
def test_bucket_overflow_paths(oracle, n_walls):
    """Many candidates behind one another inside a few 32-ray chunks: more than a
    bucket holds (they spill to the span-checked list; with 400 also more than that
    list holds, so the ray loop sweeps every slot).  Same answers either way, incl.
    the duplicated tests of a candidate that sits in a bucket AND the spill list."""
    k = np.arange(n_walls, dtype=np.float32)
    # short walls at x = 3 + 0.05 k, each spanning y in [-0.3, 0.3] (plus a few long ones)
    segs = np.stack([3 + 0.05 * k, -0.3 + 0 * k, 3 + 0.05 * k, 0.3 + 0 * k], 1).astype(np.float32)
    segs = np.concatenate([segs, np.array([[-5, -9, 30, -9], [30, -9, 30, 9], [30, 9, -5, 9], [-5, 9, -5, -9]],
                                          np.float32)])
    vmap = umrs.VectorMap(lines=segs)
    for loc, th in (((0.0, 0.0), 0.0), ((1.0, 0.1), 0.05), ((2.9, 0.0), 3.0), ((10.0, 0.05), 3.1)):
        f = np.array([loc[0], loc[1], np.float32(-2.356) + np.float32(th), np.float32(2.356) + np.float32(th)],
                     np.float32)
        r, i = vmap.GetPredictedScan(f[:2], 0.4, 100.0, f[2], f[3], 1081, return_index=True)
        check_against_tierb(oracle, segs, None, f, 100.0, 1081, r, i)


def test_back_wedge_cull_keeps_every_covering_segment(oracle):
    """The gather drops segments that lie entirely in the wedge of directions behind a
    fan wider than half a turn.  Segments that straddle either fan edge, touch the
    wedge margin, pass through the laser origin's neighbourhood or end exactly on
    the edge rays must survive: tier B on hand-placed segments around both edges,
    for several headings and fan widths (just above pi .. almost a full turn)."""
    rng = np.random.default_rng(77)
    for span_deg in (181.0, 200.0, 270.0, 340.0, 359.0):
        span = np.deg2rad(span_deg)
        for th in (0.0, 1.234, -2.9, 3.1, 6.0):
            amin, amax = np.float32(th - span / 2), np.float32(th + span / 2)
            n = 1081
            da = (float(amax) - float(amin)) / n
            edges = [float(amin), float(amin) + (n - 1) * da]
            segs = []
            for e in edges:
                for d_ang in (-3e-3, -1.1e-3, -1e-3, -9e-4, -1e-6, 0.0, 1e-6, 9e-4, 1e-3, 1.1e-3, 3e-3):
                    for r0, r1 in ((0.5, 3.0), (2.0, 2.5), (1e-3, 5.0)):
                        a, b = e + d_ang, e + d_ang + rng.uniform(-0.4, 0.4)
                        segs.append([r0 * np.cos(a), r0 * np.sin(a), r1 * np.cos(b), r1 * np.sin(b)])
            # segments wholly behind the fan, and ones crossing the whole wedge
            back = th + np.pi
            half = (2 * np.pi - span) / 2
            for f0, f1 in ((-0.9, 0.9), (-0.5, 0.2), (-1.2, 1.2), (-1.0, 1.0), (0.99, 1.01)):
                a, b = back + f0 * half, back + f1 * half
                segs.append([4 * np.cos(a), 4 * np.sin(a), 4 * np.cos(b), 4 * np.sin(b)])
            segs = np.asarray(segs, np.float32) + np.array([3.0, -2.0, 3.0, -2.0], np.float32)
            vmap = umrs.VectorMap(lines=segs)
            fan = np.array([3.0, -2.0, amin, amax], np.float32)
            r, i = vmap.GetPredictedScan(fan[:2], 0.4, 50.0, fan[2], fan[3], n, return_index=True)
            check_against_tierb(oracle, segs, None, fan, 50.0, n, r, i)


def _nested_and_grazing_segments():
    segs = []
    for s in (1.0, 1.0005, 1.001, 2.0, 5.0, 20.0):       # nested boxes, two almost coincident
        segs += [[-s, -s, s, -s], [s, -s, s, s], [s, s, -s, s], [-s, s, -s, -s]]
    segs = np.asarray(segs, np.float32)
    # open a 1-ray-wide gap in the inner boxes' +x wall: replace wall 1 (x = s) pieces
    gap = 2e-3
    extra = []
    for s in (1.0, 1.0005, 1.001):
        extra += [[s, -s, s, -gap], [s, gap, s, s]]
    keep = np.ones(len(segs), bool)
    keep[[1, 5, 9]] = False
    segs = np.concatenate([segs[keep], np.asarray(extra, np.float32)])
    # grazing walls and walls through the origin
    graz = np.array([[0.3, 1e-4, 30.0, 2e-4], [-25.0, 0.0, 25.0, 0.0], [0.2, 0.2, 15.0, 15.0001],
                     [1e-3, -3.0, 1.2e-3, 3.0], [0.5, 0.0, 0.9, 0.0]], np.float32)
    return np.concatenate([segs, graz])


def test_occlusion_cull_against_grazing_and_nested_walls(oracle):
    """Per-chunk occlusion cull: candidates farther than a wall that certainly covers
    a whole 32-ray chunk are dropped from that chunk.  Nested boxes (everything
    behind the inner box is culled), walls seen at grazing angles (ill-conditioned:
    must neither occlude nor be culled), walls through the laser origin, segments
    ending inside a chunk, and a far wall visible only through a one-ray gap."""
    segs = _nested_and_grazing_segments()
    vmap = umrs.VectorMap(lines=segs)
    for loc in ((0.0, 0.0), (0.013, -0.4), (0.9, 0.9), (-0.999, 0.5), (1.5, 0.0), (0.0, 1.00025)):
        for amin, amax, n in ((-2.356, 2.356, 1081), (-np.pi, np.pi, 3601), (0.5, 7.0, 2048), (-0.05, 0.05, 64)):
            fan = np.array([loc[0], loc[1], amin, amax], np.float32)
            r, i = vmap.GetPredictedScan(fan[:2], 0.4, 30.0, fan[2], fan[3], n, return_index=True)
            check_against_tierb(oracle, segs, None, fan, 30.0, n, r, i)


def test_coarse_precull_keeps_every_winner(oracle):
    """The coarse occlusion pre-cull (S1b) runs once a fan gathers more candidates than the CTA
    has threads.  The nested / grazing / through-the-origin walls above, buried in 700 far
    segments so that it does run: short far segments whose end-point angles differ by less than
    the coarse atan error (the arc-order guard), radial segments (nearest point = an end point),
    tangential ones (nearest point interior), long far walls spanning many chunks, and segments
    placed at 1.0005x / 1.002x the distance of a wall that covers their chunk (either side of
    the cull margin).  Tier-B contract on every ray, hit index included."""
    rng = np.random.default_rng(2024)
    base = _nested_and_grazing_segments()
    far = []
    for k in range(300):                                   # short, any orientation, ring 21..29 m
        a, r = rng.uniform(-np.pi, np.pi), rng.uniform(21.0, 29.0)
        c = np.array([r * np.cos(a), r * np.sin(a)])
        d = rng.normal(0, 1.0, 2)
        d *= rng.choice([0.03, 0.06, 0.5, 2.0]) / np.linalg.norm(d)
        far.append(np.concatenate([c - d, c + d]))
    for k in range(150):                                   # radial: seen end-on
        a, r0, ln = rng.uniform(-np.pi, np.pi), rng.uniform(5.5, 25.0), rng.uniform(0.1, 4.0)
        off = rng.choice([0.0, 1e-3, 0.05, 0.3])           # line passes this close to the origin
        u, v = np.array([np.cos(a), np.sin(a)]), np.array([-np.sin(a), np.cos(a)])
        far.append(np.concatenate([r0 * u + off * v, (r0 + ln) * u + off * v]))
    for k in range(150):                                   # tangential, some long
        a, r = rng.uniform(-np.pi, np.pi), rng.uniform(5.5, 28.0)
        h = rng.choice([0.2, 2.0, 12.0])
        u, v = np.array([np.cos(a), np.sin(a)]), np.array([-np.sin(a), np.cos(a)])
        far.append(np.concatenate([r * u - h * v, r * u + h * v]))
    for f in (1.0002, 1.0005, 1.001, 1.002, 1.01):        # just behind the s = 5 box's corners
        for sx, sy in ((1, 1), (-1, 1), (-1, -1), (1, -1)):
            cx, cy = 5.0 * f * sx, 5.0 * f * sy
            far.append([cx - 0.2 * sy, cy + 0.2 * sx, cx + 0.2 * sy, cy - 0.2 * sx])
        far.append([5.0 * f, -1.0, 5.0 * f, 1.0])         # and parallel to its +x wall
    segs = np.concatenate([base, np.asarray(far, np.float32)]).astype(np.float32)
    vmap = umrs.VectorMap(lines=segs)
    for loc in ((0.0, 0.0), (0.013, -0.4), (0.9, 0.9), (-0.999, 0.5), (1.5, 0.0), (0.0, 1.00025),
                (3.0, 3.0), (-7.0, 2.0), (12.0, -12.0)):
        for amin, amax, n in ((-2.356, 2.356, 1081), (-np.pi, np.pi, 3601), (0.5, 7.0, 2048),
                              (-0.05, 0.05, 64), (2.0, 4.5, 37)):
            fan = np.array([loc[0], loc[1], amin, amax], np.float32)
            r, i = vmap.GetPredictedScan(fan[:2], 0.4, 30.0, fan[2], fan[3], n, return_index=True)
            check_against_tierb(oracle, segs, None, fan, 30.0, n, r, i)


@pytest.mark.parametrize("style", ["floorplan", "clutter_scales", "star", "rings"])
def test_precull_fuzz_map_styles(oracle, style):
    """Fuzz of the culls on maps that gather several hundred candidates per fan (the coarse
    pre-cull is on): axis-aligned floor plans (collinear and abutting walls: ties), clutter at
    three length scales, a star of radial walls around the pose region (all ill-conditioned from
    the centre), concentric broken rings (occluders that cover whole chunks, gaps one or two rays
    wide).  32 poses each, 270 and 360 degree fans; tier-B contract on every ray."""
    rng = np.random.default_rng({"floorplan": 1, "clutter_scales": 2, "star": 3, "rings": 4}[style])
    if style == "floorplan":
        segs = umrs.map_cleanup(synth.floorplan_segments(1500, extent=120.0, seed=77))
        lo, hi, rmax = 5.0, 115.0, 60.0
    elif style == "clutter_scales":
        parts = []
        for n, ln in ((500, 0.05), (500, 1.0), (300, 8.0)):
            c = rng.uniform(-40, 40, (n, 2))
            d = rng.normal(0, 1, (n, 2))
            d *= ln / np.linalg.norm(d, axis=1, keepdims=True)
            parts.append(np.concatenate([c - d, c + d], 1))
        segs = np.concatenate(parts).astype(np.float32)
        lo, hi, rmax = -35.0, 35.0, 50.0
    elif style == "star":
        a = rng.uniform(-np.pi, np.pi, 900)
        r0, r1 = rng.uniform(2.0, 30.0, 900), rng.uniform(0.2, 10.0, 900)
        segs = np.stack([r0 * np.cos(a), r0 * np.sin(a), (r0 + r1) * np.cos(a), (r0 + r1) * np.sin(a)], 1)
        segs = segs.astype(np.float32)
        lo, hi, rmax = -1.5, 1.5, 45.0
    else:
        parts = []
        for r in (3.0, 3.002, 6.0, 9.0, 14.0, 20.0, 27.0):
            k = 96
            a = np.linspace(-np.pi, np.pi, k + 1)
            gap = rng.choice([0.0, 2e-3, 6e-3], k)           # rad, some arcs leave a slit
            parts.append(np.stack([r * np.cos(a[:-1] + gap), r * np.sin(a[:-1] + gap),
                                   r * np.cos(a[1:]), r * np.sin(a[1:])], 1))
        segs = np.concatenate(parts).astype(np.float32)
        lo, hi, rmax = -2.5, 2.5, 40.0
    vmap = umrs.VectorMap(lines=segs)
    m = 32
    loc = rng.uniform(lo, hi, (m, 2))
    if style in ("star", "rings"):
        loc[0] = 0.0                                         # exactly the centre
    th = rng.uniform(-np.pi, np.pi, m)
    for half, n in ((np.float32(np.pi * 0.75), 1081), (np.float32(np.pi), 3601)):
        fans = np.stack([loc[:, 0], loc[:, 1], th - half, th + half], 1).astype(np.float32)
        r, i = vmap.ScanBatch(fans, rmax, n, return_index=True)
        for k, f in enumerate(fans):
            r_o, _ = oracle.scan_tierb(segs, None, f[:2], rmax, f[2], f[3], n)
            fin = np.isfinite(r_o)
            assert np.array_equal(np.isfinite(r[k]), fin), (style, k)
            if fin.all():
                check_against_tierb(oracle, segs, None, f, rmax, n, r[k], i[k])
            else:
                rel = np.abs(r[k][fin] - r_o[fin]) / np.maximum(np.abs(r_o[fin]), 1e-9)
                assert rel.max() <= REL_TOL, (style, k)


def test_occlusion_cull_random_clutter_many_fans(oracle):
    """Dense random clutter (short segments at all depths) from many poses: the
    cull must never remove a winner; idx compared exactly (ties certified)."""
    rng = np.random.default_rng(5)
    c = rng.uniform(-30, 30, (1500, 2))
    d = rng.normal(0, 1.5, (1500, 2))
    segs = np.concatenate([c - d, c + d], 1).astype(np.float32)
    vmap = umrs.VectorMap(lines=segs)
    fans = np.stack([rng.uniform(-25, 25, 24), rng.uniform(-25, 25, 24), rng.uniform(-3, 3, 24),
                     np.zeros(24)], 1).astype(np.float32)
    fans[:, 3] = fans[:, 2] + np.float32(4.712389)
    rb, ib = vmap.ScanBatch(fans, 40.0, 1081, return_index=True)
    for k, f in enumerate(fans):
        check_against_tierb(oracle, segs, None, f, 40.0, 1081, rb[k], ib[k])


def test_random_raw_maps_with_degenerate_geometry(oracle):
    """The default scan on 150 small RAW lattice maps (crossings, shared end points, collinear
    overlaps, duplicates, zero-length lines) from poses that are often exactly on a line or an end
    point: the tier-B contract holds on every ray (non-finite oracle ranges must be non-finite
    here too)."""
    rng = np.random.default_rng(78)
    for trial in range(150):
        n = int(rng.integers(1, 50))
        step = float(rng.choice([0.5, 1.0, 2.0]))
        raw = (rng.integers(-10, 11, (n, 4)) * step).astype(np.float32)
        if trial % 4 == 0:
            raw[rng.integers(0, n)] = np.repeat(raw[rng.integers(0, n), :2], 2)[[0, 2, 1, 3]]
        vmap = umrs.VectorMap(lines=raw)
        m = 8
        locs = rng.integers(-10, 11, (m, 2)) * step * rng.choice([1.0, 0.5], (m, 1))
        if trial % 2 == 0:
            locs = locs + rng.normal(0, 0.3, (m, 2))
        th = rng.uniform(-np.pi, np.pi, m)
        d135 = np.float32(np.pi * 0.75)
        fans = np.stack([locs[:, 0], locs[:, 1], th - d135, th + d135], 1).astype(np.float32)
        r, i = vmap.ScanBatch(fans, 25.0, 181, return_index=True)
        for k, f in enumerate(fans):
            r_o, i_o = oracle.scan_tierb(raw, None, f[:2], 25.0, f[2], f[3], 181)
            fin = np.isfinite(r_o)
            assert np.array_equal(np.isfinite(r[k]), fin), (trial, k)
            if fin.all():
                check_against_tierb(oracle, raw, None, f, 25.0, 181, r[k], i[k])
            else:
                rel = np.abs(r[k][fin] - r_o[fin]) / np.maximum(np.abs(r_o[fin]), 1e-9)
                assert rel.max() <= REL_TOL, (trial, k)
