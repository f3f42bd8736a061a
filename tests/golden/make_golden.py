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


def main():
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
