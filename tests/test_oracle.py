"""CPU tests of the checkers themselves (no GPU, no product compute).

The oracle (oracle/umrs_oracle.c, plain-C restatement) is pinned here against
  * the committed golden fixtures tests/golden/*.npz, which were generated from
    the UNMODIFIED reference scan (oracle/_ref, vector_map.cpp compiled in place;
    tests/golden/make_golden.py), and
  * the reference itself, live, wherever oracle/_ref/libumrs_ref.so exists.
"""
import ctypes as C
import glob
import os

import numpy as np
import pytest

from tests import oracle_lib

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))
SCAN_GOLDEN = [g for g in GOLDEN if "primitives" not in g and "kinematics" not in g]


def _load(path):
    d = np.load(path)
    return {k: d[k] for k in d.files}


def test_golden_fixtures_present():
    names = {os.path.basename(g) for g in GOLDEN}
    for n in ("config1_campus.npz", "floorplan_64.npz", "floorplan_256.npz", "floorplan_1024.npz",
              "jackal_360.npz", "clutter_objects.npz", "primitives.npz"):
        assert n in names


@pytest.mark.parametrize("path", SCAN_GOLDEN, ids=[os.path.basename(p) for p in SCAN_GOLDEN])
def test_tierb_oracle_reproduces_golden(oracle, path):
    """Bit-exact: tier-B restatement today == tier-B when the fixtures were made."""
    g = _load(path)
    n = int(g["num_rays"])
    for f, rb, ib in zip(g["fans"], g["tier_b"], g["idx_b"]):
        r, i = oracle.scan_tierb(g["segs"], g["obj"], f[:2], float(g["range_max"]), f[2], f[3], n)
        assert np.array_equal(r.view(np.uint32), rb.view(np.uint32))
        assert np.array_equal(i, ib)


@pytest.mark.parametrize("path", SCAN_GOLDEN, ids=[os.path.basename(p) for p in SCAN_GOLDEN])
def test_reference_reproduces_golden_tier_a(ref, path):
    """The reference's GetPredictedScan run here == the committed tier-A vectors."""
    g = _load(path)
    n = int(g["num_rays"])
    rmap = ref.map_from_segments(g["raw"], True)
    assert np.array_equal(rmap.lines.view(np.uint32), g["segs"].view(np.uint32))
    rmap.set_object_lines(g["obj"])
    for f, ra in zip(g["fans"], g["tier_a"]):
        r = rmap.predicted_scan(f[:2], 0.4, float(g["range_max"]), f[2], f[3], n)
        assert np.array_equal(r.view(np.uint32), ra.view(np.uint32))


@pytest.mark.parametrize("path", SCAN_GOLDEN, ids=[os.path.basename(p) for p in SCAN_GOLDEN])
def test_tier_a_vs_tier_b_contract(path):
    """the reference's analytic render (tier A) and the per-ray
    minimum (tier B) are the same visible surface except for the documented
    classes D1-D6; the disagreement rate is small and agreeing rays differ by
    at most ~1e-4 relative (class D6)."""
    g = _load(path)
    a, b = g["tier_a"].astype(np.float64), g["tier_b"].astype(np.float64)
    rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-9)
    disagree = rel > 1e-5
    limit = 0.03 if "clutter" in path else 0.005
    assert disagree.mean() < limit, disagree.mean()
    # the bulk of the disagreements are "reference farther" (near piece dropped by
    # TrimOcclusion / eps rules: classes D1, D3)
    if disagree.sum() >= 20:
        assert (a[disagree] > b[disagree]).mean() > 0.5


def test_cleanup_restatements_match_reference_cleanup(oracle):
    """VectorMap::Cleanup (vector_map.cpp:367-395): oracle port and the product's
    host-side restatement both reproduce the reference's output bit for bit
    (golden `segs` came out of the reference's own Cleanup)."""
    import ut_multirobot_sim_b200 as umrs
    for path in SCAN_GOLDEN:
        g = _load(path)
        o = oracle.map_cleanup(g["raw"])
        assert np.array_equal(o.view(np.uint32), g["segs"].view(np.uint32)), path
        p = umrs.map_cleanup(g["raw"])
        assert np.array_equal(p.view(np.uint32), g["segs"].view(np.uint32)), path


def test_cleanup_splits_at_crossings(oracle):
    # two crossing segments -> 3 pieces (the second one split), all shrunk by 1e-4
    raw = np.array([[0, 0, 2, 0], [1, -1, 1, 1]], np.float32)
    out = oracle.map_cleanup(raw)
    assert len(out) == 3
    assert np.allclose(out[0], [1e-4, 0, 2 - 1e-4, 0], atol=1e-7)
    # short lines (< 5 cm) are dropped
    assert len(oracle.map_cleanup(np.array([[0, 0, 0.04, 0]], np.float32))) == 0
    assert len(oracle.map_cleanup(np.zeros((0, 4), np.float32))) == 0


def test_primitives_known_answers(oracle):
    """line2d_tests.cc:32-87 inputs through Line2f::Intersection (the predicate
    Cleanup uses), reference outputs committed in primitives.npz."""
    g = _load(os.path.join(HERE, "golden", "primitives.npz"))
    out = np.zeros(2, np.float32)
    for a, b, hit, pt in zip(g["a"], g["b"], g["intersection_hit"], g["intersection_pt"]):
        h = oracle.lib.orc_line_intersection(oracle_lib._fp(np.ascontiguousarray(a)),
                                             oracle_lib._fp(np.ascontiguousarray(b)),
                                             oracle_lib._fp(out))
        assert h == hit, (a, b)
        if hit:
            assert np.array_equal(out.view(np.uint32), pt.view(np.uint32)), (a, b)
    # the gtest truth table for Intersects (expected values from the reference's test)
    assert g["intersects_expected"].shape == g["intersects_ref"][: len(g["intersects_expected"])].shape
    assert np.array_equal(g["intersects_expected"], g["intersects_ref"][: len(g["intersects_expected"])])
    # AngleMod<float> / ray angles
    for ang, want in zip(g["angles"], g["angle_mod_f"]):
        got = oracle.lib.orc_ray_angle(float(ang), float(ang), 1, 0)
        assert np.float32(got).view(np.uint32) == want.view(np.uint32)


def test_primitives_live_against_reference(oracle, ref):
    rng = np.random.default_rng(3)
    segs = rng.uniform(-5, 5, (2000, 8)).astype(np.float32)
    segs[:200] = np.round(segs[:200])  # many exact touches / collinear cases
    out_o = np.zeros(2, np.float32)
    out_r = np.zeros(2, np.float32)
    for s in segs:
        a, b = np.ascontiguousarray(s[:4]), np.ascontiguousarray(s[4:])
        ho = oracle.lib.orc_line_intersection(oracle_lib._fp(a), oracle_lib._fp(b), oracle_lib._fp(out_o))
        hr = ref.lib.umrs_ref_line_intersection(oracle_lib._fp(a), oracle_lib._fp(b), oracle_lib._fp(out_r))
        assert bool(ho) == bool(hr)
        if hr:
            assert np.array_equal(out_o.view(np.uint32), out_r.view(np.uint32))
    for ang in rng.uniform(-50, 50, 2000).astype(np.float32):
        assert np.float32(oracle.lib.orc_ray_angle(float(ang), float(ang), 1, 0)).view(np.uint32) == \
            np.float32(ref.lib.umrs_ref_angle_mod_f(float(ang))).view(np.uint32)


def test_square_template_matches_reference_short_term_object(oracle, ref):
    """ShortTermObject (short_term_object.cpp:39-67) compiled from the reference:
    template and pose lines == oracle restatement."""
    tmpl = np.zeros((4, 4), np.float32)
    lines = np.zeros((4, 4), np.float32)
    pose = np.zeros(3, np.float32)
    n = ref.lib.umrs_ref_short_term_object(oracle_lib._fp(tmpl), oracle_lib._fp(lines), oracle_lib._fp(pose))
    assert n == 4
    assert np.array_equal(oracle.square_template().view(np.uint32), tmpl.view(np.uint32))
    assert np.array_equal(oracle.transform_lines(tmpl, pose).view(np.uint32), lines.view(np.uint32))


def test_square_template_golden(oracle):
    g = _load(os.path.join(HERE, "golden", "primitives.npz"))
    assert np.array_equal(oracle.square_template().view(np.uint32), g["square_template"].view(np.uint32))
    assert np.array_equal(oracle.transform_lines(g["square_template"], g["square_pose"]).view(np.uint32),
                          g["square_lines"].view(np.uint32))


def test_human_template_shape(oracle):
    """human_object.cpp:95-112: 20 chords of radius 0.1, 1 mm end gaps except the
    closing chord."""
    import ut_multirobot_sim_b200 as umrs
    hp = umrs.default_human()
    t = oracle.human_template(hp)
    assert t.shape == (20, 4)
    r0 = np.hypot(t[:, 0], t[:, 1])
    r1 = np.hypot(t[:, 2], t[:, 3])
    assert np.all(np.abs(r0 - 0.1) < 2e-4) and np.all(np.abs(r1 - 0.1) < 2e-4)  # chord ends pulled in 1 mm
    gaps = np.hypot(t[1:19, 0] - t[:18, 2], t[1:19, 1] - t[:18, 3])
    assert np.allclose(gaps, 0.002 * np.cos(np.pi / 20), atol=2e-6)  # two 1 mm pull-backs meeting at 18 deg
    # closing chord ends exactly at (r, 0) and starts exactly at the last vertex
    assert t[19, 2] == np.float32(0.1) and t[19, 3] == 0.0


def _mk_state(oracle, pose):
    s = oracle_lib.RobotState()
    oracle.lib.orc_robot_init(C.byref(s), oracle_lib._fp(np.asarray(pose, np.float32)))
    return s


def test_ackermann_step_known_answers(oracle):
    """ackermann_model.cpp:52-96 with ut_automata_config.lua limits."""
    import ut_multirobot_sim_b200 as umrs
    car = umrs.default_robot("ut_automata")
    dt = np.float32(0.025)
    s = _mk_state(oracle, [1.0, 2.0, 0.0])
    cmd = np.array([1.0, 0.0, 0.0], np.float32)
    oracle.lib.orc_robot_command(C.byref(s), C.byref(car), oracle_lib._fp(cmd))
    oracle.lib.orc_robot_step(C.byref(s), C.byref(car), dt, 0.0, 0.0)
    # accel limited: dv = dt * 3.0 = 0.075; straight line
    assert s.vel[0] == np.float32(np.float64(dt) * 3.0)
    assert s.pose[0] == np.float32(1.0) + np.float32(np.float64(s.vel[0]) * np.float64(dt))
    assert s.pose[1] == 2.0 and s.pose[2] == 0.0
    # speed clamp: after many steps v == ak_max_speed
    for _ in range(40):
        oracle.lib.orc_robot_command(C.byref(s), C.byref(car), oracle_lib._fp(np.array([5.0, 0, 0], np.float32)))
        oracle.lib.orc_robot_step(C.byref(s), C.byref(car), dt, 0.0, 0.0)
    assert abs(s.vel[0] - 1.2) < 1e-6
    # arc: curvature clamp at 1/0.98, heading advances by dist * kappa
    th0, v = s.pose[2], s.vel[0]
    oracle.lib.orc_robot_command(C.byref(s), C.byref(car), oracle_lib._fp(np.array([5.0, 9.0, 0], np.float32)))
    oracle.lib.orc_robot_step(C.byref(s), C.byref(car), dt, 0.0, 0.0)
    kappa = np.float32(1.0 / np.float64(np.float32(0.98)))
    assert abs((s.pose[2] - th0) - v * dt * kappa) < 1e-6
    # watchdog (sim time): no command for > 0.1 s -> velocity command zeroed, car decelerates
    for _ in range(6):
        oracle.lib.orc_robot_step(C.byref(s), C.byref(car), dt, 0.0, 0.0)
    assert s.vel[0] < 1.2 - 0.07
    # non-finite commands are ignored (ackermann_model.cpp:42-47)
    age = s.cmd_age
    oracle.lib.orc_robot_command(C.byref(s), C.byref(car), oracle_lib._fp(np.array([np.nan, 0, 0], np.float32)))
    assert s.cmd_age == age


def test_diff_drive_and_omni_quirks(oracle):
    import ut_multirobot_sim_b200 as umrs
    dt = np.float32(0.025)
    jk = umrs.default_robot("ut_jackal")
    s = _mk_state(oracle, [0, 0, 0])
    oracle.lib.orc_robot_command(C.byref(s), C.byref(jk), oracle_lib._fp(np.array([9.0, 0.5, 0], np.float32)))
    assert s.cmd[0] == 2.0  # max_linear_vel clamp in the callback (diff_drive_model.cpp:181-186)
    oracle.lib.orc_robot_step(C.byref(s), C.byref(jk), dt, 0.0, 0.0)
    # per-TICK accel limit quirk: |target - v| = 2.0 < limit 3.0 -> jumps to the target at once
    assert s.linear_vel == 2.0 and s.angular_vel == 0.5
    # trapezoid: dx = v_old*dt + 0.5*(v_new - v_old)*dt = 0.025
    assert abs(s.pose[0] - 0.025) < 1e-7
    co = umrs.default_robot("cobot")
    s = _mk_state(oracle, [0, 0, 0])
    oracle.lib.orc_robot_command(C.byref(s), C.byref(co), oracle_lib._fp(np.array([3.0, 4.0, 0.01], np.float32)))
    oracle.lib.orc_robot_step(C.byref(s), C.byref(co), dt, 0.0, 0.0)
    # speed cap by norm then accel cap by norm: |v| = co_max_accel * dt along (3,4)/5
    assert abs(np.hypot(s.vel[0], s.vel[1]) - 3.0 * 0.025) < 1e-6
    assert abs(s.vel[0] / s.vel[1] - 0.75) < 1e-5
    # omnidirectional_model.cpp:115 quirk: |desired| (0.01) is below the accel cap, so the
    # whole delta is applied in one tick
    assert abs(s.vel[2] - 0.01) < 1e-8


def test_human_step_goal_swap(oracle):
    import ut_multirobot_sim_b200 as umrs
    hp = umrs.default_human()
    s = oracle_lib.HumanState()
    oracle.lib.orc_human_init(C.byref(s), C.byref(hp))
    assert list(s.pose) == [hp.start_x, hp.start_y, hp.start_theta]
    d0 = np.hypot(hp.goal_x - hp.start_x, hp.goal_y - hp.start_y)
    n = 0
    while list(s.goal) == [hp.goal_x, hp.goal_y, hp.goal_theta]:
        oracle.lib.orc_human_step(C.byref(s), C.byref(hp), 0.025)
        n += 1
        assert n < 1000
    # walks at avg_speed 1 m/s until within 0.3 m of the goal, then swaps (Repeat mode)
    assert abs(n * 0.025 - (d0 - 0.3)) < 0.05
    assert list(s.start) == [hp.goal_x, hp.goal_y, hp.goal_theta]


def test_laser_postprocess_rule(oracle):
    """simulator.cpp:412-418."""
    r = np.array([0.5, 99.89, 99.91, 100.0, 141.0, 0.001], np.float32)
    out = oracle.laser_postprocess(r, 100.0, 0.01, normals=np.array([1, 1, 1, 1, 1, -5], np.float32))
    assert out[0] == np.float32(0.5) + np.float32(0.01)
    assert out[1] == np.float32(99.89) + np.float32(0.01)
    assert out[2] == 0 and out[3] == 0 and out[4] == 0
    assert out[5] == 0  # max(0, r + noise)


PHILOX_KAT = [  # Random123 kat_vectors, philox4x32-10
    ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
     [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
]


def test_philox_known_answers(oracle):
    for ctr, key, want in PHILOX_KAT:
        assert oracle.philox(ctr, key) == want


def test_philox_matches_numpy(oracle):
    # numpy's Philox is 4x64; only check our 4x32 against the KATs and the avalanche property
    a = oracle.philox([1, 2, 3, 4], [5, 6])
    b = oracle.philox([1, 2, 3, 5], [5, 6])
    diff = sum(bin(x ^ y).count("1") for x, y in zip(a, b))
    assert 40 < diff < 90


def test_box_muller_distribution(oracle):
    rng = np.random.default_rng(0)
    u = rng.integers(0, 2 ** 32, (20000, 2), dtype=np.uint64)
    g = np.array([oracle.box_muller(int(a), int(b)) for a, b in u]).ravel()
    assert abs(g.mean()) < 0.02 and abs(g.std() - 1.0) < 0.02
    from scipy import stats
    assert stats.kstest(g, "norm").pvalue > 1e-3


def _crossing_maps():
    """Raw maps that make Cleanup work: random clutter with thousands of crossings,
    an axis-aligned lattice (T-junctions, shared end points, collinear overlaps,
    duplicates), long diagonals across everything, and mixed coordinate scales."""
    rng = np.random.default_rng(17)
    c = rng.uniform(-40, 40, (1500, 2))
    d = rng.normal(0, 3.0, (1500, 2))
    clutter = np.concatenate([c - d, c + d], 1).astype(np.float32)
    lat = []
    for k in range(25):
        lat += [[0, k, 24, k], [k, 0, k, 24]]            # crossings at integer points
    lat += [[2, 3, 9, 3], [2, 3, 9, 3], [5, 3, 14, 3]]    # duplicate + collinear overlap
    lat += [[-5, -5, 30, 30], [-5, 30, 30, -5], [0.5, 0.5, 0.5, 0.52]]   # diagonals, a < 5 cm stub
    lattice = np.asarray(lat, np.float32)
    # (coordinates stay below ~500 m: beyond that the 1e-4 shrink is smaller than a float
    # ulp and the reference's own loop re-splits the same piece forever)
    mixed = np.concatenate([clutter[:400] * 0.01, clutter[400:800] * 5.0, lattice * 3.0 + 100.0])
    return {"clutter": clutter, "lattice": lattice, "mixed": mixed.astype(np.float32)}


def test_grid_cleanup_equals_quadratic_and_reference(ref):
    """the grid-accelerated Cleanup returns the reference's segments in the
    reference's order, bit for bit -- against the product's quadratic restatement and
    against the compiled reference's own Cleanup (oracle/_ref) on crossing-heavy maps."""
    import ut_multirobot_sim_b200 as umrs
    for name, raw in _crossing_maps().items():
        q = umrs.map_cleanup(raw, algorithm=1)
        g = umrs.map_cleanup(raw, algorithm=2)
        a = umrs.map_cleanup(raw, algorithm=0)
        assert len(q) > len(raw) or name == "mixed", name      # splits happened
        assert np.array_equal(q.view(np.uint32), g.view(np.uint32)), name
        assert np.array_equal(q.view(np.uint32), a.view(np.uint32)), name
        if ref is not None:
            r = ref.map_from_segments(raw, True).lines
            assert np.array_equal(np.asarray(r, np.float32).view(np.uint32), g.view(np.uint32)), name


def test_grid_cleanup_degenerate_inputs():
    import ut_multirobot_sim_b200 as umrs
    assert len(umrs.map_cleanup(np.zeros((0, 4), np.float32), algorithm=2)) == 0
    one = np.array([[1, 1, 2, 2]], np.float32)
    assert np.array_equal(umrs.map_cleanup(one, 2).view(np.uint32), umrs.map_cleanup(one, 1).view(np.uint32))
    # all segments identical points / zero extent: dropped by the 5 cm rule
    assert len(umrs.map_cleanup(np.ones((10, 4), np.float32), 2)) == 0
    # non-finite coordinates: the grid is not used (same output as the quadratic loop);
    # no crossing with the finite segment so the reference's loop terminates
    bad = np.array([[0, 0, 1, 0], [np.inf, 5, np.inf, 6]], np.float32)
    with np.errstate(invalid="ignore"):
        assert np.array_equal(umrs.map_cleanup(bad, 2).view(np.uint32), umrs.map_cleanup(bad, 1).view(np.uint32))


def test_grid_cleanup_large_map_fast_and_identical_on_a_sample():
    """100k-segment city-block map + 500 long streets crossing it (config 4's segment count,
    coordinates within +-450 m so the reference's own loop terminates): the grid finishes in
    seconds; on a crop of a few thousand segments the quadratic loop gives the same bits.
    Config 4's own 2 km map has no crossings, so it goes through either search unchanged."""
    import time
    import ut_multirobot_sim_b200 as umrs
    from ut_multirobot_sim_b200 import synth
    E = 900.0
    blocks = synth.cityblock_segments(100000, extent=E, seed=99) - np.float32(E / 2)
    # streets cross the blocks' axis-aligned walls at 20..70 degrees: at grazing angles the
    # 1e-4 m cut leaves the piece within one float ulp of the wall it was cut at and the
    # reference's loop (and this one, until its queue cap) re-splits it forever
    rng = np.random.default_rng(3)
    c = rng.uniform(-E / 2, E / 2, (500, 2))
    ang = rng.uniform(0.35, 1.2, 500) * rng.choice([-1.0, 1.0], 500)
    half = rng.uniform(50, 200, 500)[:, None] * np.stack([np.cos(ang), np.sin(ang)], 1)
    streets = np.concatenate([c - half, c + half], 1).astype(np.float32)
    raw = np.concatenate([blocks, streets])
    t0 = time.perf_counter()
    g = umrs.map_cleanup(raw, algorithm=2)
    dt = time.perf_counter() - t0
    assert len(g) > len(raw) and dt < 60.0, (len(g), dt)
    crop = raw[(np.abs(raw[:, 0]) < 60) & (np.abs(raw[:, 1]) < 60)]
    q = umrs.map_cleanup(crop, 1)
    assert len(q) > len(crop)
    assert np.array_equal(q.view(np.uint32), umrs.map_cleanup(crop, 2).view(np.uint32))
    # config 4's map (2 km, nothing crosses): identical through both searches on a sample
    c4 = synth.cityblock_segments(4000, extent=2000.0, seed=99)
    assert np.array_equal(umrs.map_cleanup(c4, 1).view(np.uint32), umrs.map_cleanup(c4, 2).view(np.uint32))


def test_cleanup_reports_inputs_the_reference_never_finishes():
    """Beyond ~1 km the 1e-4 m cut is below the float spacing: the piece pushed back equals
    the line that was cut and the reference's loop (vector_map.cpp:371-392) spins forever.
    Both searches here detect the cycle and return an error instead."""
    import ut_multirobot_sim_b200 as umrs
    from ut_multirobot_sim_b200 import _lib
    far = np.array([[5000, 4990, 5000, 5010], [4990, 5000, 5010, 5000]], np.float32)
    for alg in (1, 2):
        with pytest.raises(_lib.UmrsError):
            umrs.map_cleanup(far, algorithm=alg)


def test_grid_cleanup_property_random_lattice_maps(ref):
    """Property check of SURVEY 8f-3 on 300 random maps whose end points sit on a coarse lattice
    (so T-junctions, shared end points, collinear overlaps and duplicates are the norm, not the
    exception): grid search == quadratic search == the compiled reference's Cleanup, bit for bit,
    whenever the reference's loop terminates (the queue cap reports the others identically)."""
    import ut_multirobot_sim_b200 as umrs
    from ut_multirobot_sim_b200 import _lib
    rng = np.random.default_rng(2024)
    done = failed = 0
    for trial in range(300):
        n = int(rng.integers(1, 60))
        step = float(rng.choice([0.25, 0.5, 1.0, 2.5]))
        raw = (rng.integers(-12, 13, (n, 4)) * step).astype(np.float32)
        if trial % 3 == 0:       # off-lattice perturbation of some end points
            raw += (rng.random((n, 4)) < 0.3) * rng.normal(0, 0.01, (n, 4)).astype(np.float32)
        out = []
        for alg in (1, 2):
            try:
                out.append(umrs.map_cleanup(raw, algorithm=alg))
            except _lib.UmrsError:
                out.append(None)
        assert (out[0] is None) == (out[1] is None), trial
        if out[0] is None:
            failed += 1
            continue
        assert np.array_equal(out[0].view(np.uint32), out[1].view(np.uint32)), trial
        if ref is not None:
            r = np.asarray(ref.map_from_segments(raw, True).lines, np.float32)
            assert np.array_equal(r.view(np.uint32), out[1].view(np.uint32)), trial
        done += 1
    assert done >= 250, (done, failed)
