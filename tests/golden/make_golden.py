"""Generates tests/golden/*.npz from the UNMODIFIED reference scan
(oracle/_ref/libumrs_ref.so, built in place from /root/reference by oracle/Makefile)
plus the tier-B restatement.  Run in the build container (it needs
/root/reference); the resulting fixtures travel with the repo.

  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from tests import oracle_lib  # noqa: E402
from ut_multirobot_sim_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
F32 = np.float32


def fan_for(pose, amin, amax):
    return np.array([pose[0], pose[1], F32(amin) + pose[2], F32(amax) + pose[2]], F32)


def make_case(name, raw, extent, n_poses, seed, amin, amax, num_rays, range_max, obj=None):
    ref = oracle_lib.load_ref()
    orc = oracle_lib.load_oracle()
    assert ref is not None, "needs oracle/_ref (build container only)"
    rmap = ref.map_from_segments(raw, True)           # the reference's own Cleanup
    segs = rmap.lines
    poses = synth.sample_free_poses(segs, n_poses, extent, seed)
    fans = np.stack([fan_for(p, amin, amax) for p in poses])
    obj = np.zeros((0, 4), F32) if obj is None else np.asarray(obj, F32)
    rmap.set_object_lines(obj)
    tier_a = np.stack([rmap.predicted_scan(f[:2], 0.4, range_max, f[2], f[3], num_rays) for f in fans])
    tb = [orc.scan_tierb(segs, obj, f[:2], range_max, f[2], f[3], num_rays) for f in fans]
    tier_b = np.stack([t[0] for t in tb])
    idx_b = np.stack([t[1] for t in tb])
    rel = np.abs(tier_a - tier_b) / np.maximum(np.abs(tier_b), 1e-9)
    print("%-22s segs=%4d obj=%3d fans=%3d rays=%d  tierA!=tierB(>1e-5): %.4f%%  max rel among agreeing %.2e"
          % (name, len(segs), len(obj), len(fans), num_rays, 100.0 * (rel > 1e-5).mean(),
             rel[rel <= 1e-5].max()))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), raw=raw, segs=segs, obj=obj, fans=fans,
                        num_rays=num_rays, range_max=F32(range_max), tier_a=tier_a,
                        tier_b=tier_b, idx_b=idx_b)


def clutter(n, extent, seed):
    """Random crossing segments: exercises Cleanup's split-at-crossing path."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(0, extent, (n, 2))
    a = rng.uniform(0, np.pi, n)
    l = rng.uniform(0.5, 6.0, n)
    d = np.stack([np.cos(a), np.sin(a)], 1) * l[:, None] * 0.5
    return np.concatenate([c - d, c + d], 1).astype(F32)


# line2d_tests.cc:32-87 -- (segment a, segment b, expected a.Intersects(b))
INTERSECTS_TABLE = [
    ((1, 4, 4, 1), (6, 6, 2.5, 2.5), True), ((1, 1, 2, 2), (3, 3, 4, 4), False),
    ((1, 1, 3, 3), (2, 2, 4, 4), True), ((1, 1, 3, 3), (3, 3, 4, 4), True),
    ((1.1, 0.1, 1.9, -0.1), (1, 0, 2, 0), True), ((1.1, 0.1, 1.9, 0), (1, 0, 2, 0), True),
    ((1.1, 0.1, 1.9, 0.001), (1, 0, 2, 0), False), ((0, 0.1, 1.1, -0.1), (1, 0, 2, 0), False),
    ((1, 1, 2, 2), (3, 3, 6, 6), False), ((0, 0, 0, 5), (0, 5, 6, 6), True),
    ((0, 0, 0, 5), (0, 4, 0, 6), True), ((0, 0, 0, 5), (0, 5, 0, 6), True),
    ((0, 0, 2, 3), (4, 6, 6, 9), False), ((1, 1, 2, 2), (1, 1, 2, 2), True),
    ((1, 1, 2, 2), (2, 1, 3, 2), False), ((5, 0, 5, 10), (0, 5, 10, 5), True),
    ((5, 0, 5, 10), (50, 50, 60, 60), False), ((-1, -1, 1, 1), (-2, -2, 3, 3), True),
    ((0, 0, 0, 5), (1, 1, 5, 5), False), ((0, 0, 0, 5), (0, 5, 0, 10), True),
    ((10, 10, 10, 10), (30, 4, 30, 10), False), ((0, 200, 300, 200), (100, 300, 100, 200), True),
    ((0, 0, 0, 0), (0, 0, 0, 0), True), ((0, 0, 0, 0), (0, 3, 0, 3), False),
    ((0, 0, 0, 0.4), (0, 3, 0, 3), False), ((0.4, 0, 0, 0), (3, 0, 3, 0), False),
]


def make_primitives():
    """Reference outputs of the primitives the path uses, on the reference's own
    gtest inputs plus seeded random ones: Line2f::Intersects / Intersection
    (line2d.h:123-160), AngleMod<float> (math_util.h:55-59), ShortTermObject
    template + pose lines (short_term_object.cpp:39-67)."""
    import ctypes as C
    ref = oracle_lib.load_ref()
    assert ref is not None
    fp = oracle_lib._fp
    rng = np.random.default_rng(17)
    a = np.array([t[0] for t in INTERSECTS_TABLE], F32)
    b = np.array([t[1] for t in INTERSECTS_TABLE], F32)
    expected = np.array([t[2] for t in INTERSECTS_TABLE], np.int32)
    extra = rng.uniform(-5, 5, (400, 8)).astype(F32)
    extra[:100] = np.round(extra[:100])
    a = np.concatenate([a, extra[:, :4]])
    b = np.concatenate([b, extra[:, 4:]])
    hit = np.zeros(len(a), np.int32)
    pts = np.zeros((len(a), 2), F32)
    inter = np.zeros(len(a), np.int32)
    for k in range(len(a)):
        aa, bb = np.ascontiguousarray(a[k]), np.ascontiguousarray(b[k])
        inter[k] = ref.lib.umrs_ref_line_intersects(fp(aa), fp(bb))
        out = np.zeros(2, F32)
        hit[k] = ref.lib.umrs_ref_line_intersection(fp(aa), fp(bb), fp(out))
        pts[k] = out
    assert np.array_equal(inter[: len(expected)], expected), "reference fails its own gtest table?"
    angles = np.concatenate([rng.uniform(-60, 60, 500), [0, np.pi, -np.pi, 3 * np.pi, 1e-8]]).astype(F32)
    am = np.array([ref.lib.umrs_ref_angle_mod_f(float(x)) for x in angles], F32)
    tmpl = np.zeros((4, 4), F32)
    lines = np.zeros((4, 4), F32)
    pose = np.zeros(3, F32)
    ref.lib.umrs_ref_short_term_object(fp(tmpl), fp(lines), fp(pose))
    np.savez_compressed(os.path.join(HERE, "primitives.npz"), a=a, b=b, intersects_expected=expected,
                        intersects_ref=inter, intersection_hit=hit, intersection_pt=pts,
                        angles=angles, angle_mod_f=am, square_template=tmpl, square_pose=pose,
                        square_lines=lines)
    print("primitives: %d segment pairs (%d from the gtest table), %d angles" % (len(a), len(expected), len(angles)))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "primitives":
        make_primitives()
        return
    make_primitives()
    d135 = F32(np.pi * 135.0 / 180.0)
    # config 1: single car, 1081-ray 270 deg scan, "campus" floor plan
    make_case("config1_campus", synth.floorplan_segments(1500, 200.0, 1234), 200.0, 64, 1234,
              -d135, d135, 1081, 100.0)
    for m in (64, 256, 1024):
        make_case("floorplan_%d" % m, synth.floorplan_segments(m, 200.0, 1234), 200.0, 16, 7,
                  -d135, d135, 1081, 100.0)
    # jackal laser: 360 deg, 3601 rays, 150 m
    make_case("jackal_360", synth.floorplan_segments(256, 120.0, 3), 120.0, 8, 5,
              -F32(np.pi), F32(np.pi), 3601, 150.0)
    # crossing clutter (Cleanup splits) + object lines (a square and a 20-gon human)
    orc = oracle_lib.load_oracle()
    from ut_multirobot_sim_b200 import default_human
    hum = orc.transform_lines(orc.human_template(default_human()), np.array([20.0, 21.0, 0.3], F32))
    sq = orc.transform_lines(orc.square_template(), np.array([24.0, 18.0, 0.0], F32))
    make_case("clutter_objects", clutter(120, 40.0, 9), 40.0, 16, 21, -d135, d135, 1081, 100.0,
              obj=np.concatenate([sq, hum]))


if __name__ == "__main__":
    main()
