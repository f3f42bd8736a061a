"""TEST INFRASTRUCTURE (CPU): independent checks of the scan oracles and the per-ray classifier
of reference disagreements (classes D1-D6).

  brute_force_f64   an INDEPENDENT nearest-hit ray caster: float64, cross-product sidedness
                    (no atan2, no angular intervals, no unit normals -- none of the reference's
                    or the oracle port's formulae).  Pins tier B ("the exact nearest hit").
  Classifier        runs the host build of the product's restated SceneRender/TrimOcclusion
                    (csrc/umrs_render.h) with provenance (tests/cpp/render_classify.cpp) and
                    names, for every ray on which the reference's GetPredictedScan (tier A)
                    differs from tier B by more than 1e-5 relative, the documented defect
                    of the reference that produced the difference.

Classes (file:line in the reference tree):
  D1  a visible piece <= 1 cm was dropped      vector_map.cpp:222,238,256 (eps on squared length)
  D2  angular span < 1e-4 rad dropped          vector_map.cpp:451 (on a trimmed piece; whole
                                               segments are dropped identically by both tiers)
  D3  TrimOcclusion trimmed away (or invented) a part that is visible: wrong-case / over-trim
      defects, the class the reference documents itself ("closed shape would have wrong
      occlusion", human_object.cpp:106)        vector_map.cpp:55-189
  D4  the 2000-candidate cap                   vector_map.cpp:223,229,248
  D5  first rendered piece that covers the ray is not the nearest one   vector_map.cpp:474
  D6  same segment, range evaluated on a trimmed sub-segment (end-point rounding moves the
      unit normal)                             vector_map.cpp:471-473 on pieces of :150-180
  E   end-point rounding: the ray lies within 3e-6 rad of the end of the winning segment or
      piece, where float(atan2) of a trimmed end point and of the original one order differently
"""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

from tests import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_FP = C.POINTER(C.c_float)
_IP = C.POINTER(C.c_int)
CLASSES = ("D1", "D2", "D3", "D4", "D5", "D6", "E")
ANG_TOL = 3e-6          # rad: float(atan2) + ray-lattice rounding (csrc/umrs_device.cuh: 3.6e-6 budget)


def ray_angles(oracle, fan, n):
    """The fan's ray angles as the reference computes them (vector_map.cpp:465-467), float32."""
    return np.array([oracle.lib.orc_ray_angle(float(fan[2]), float(fan[3]), n, i) for i in range(n)], np.float32)


def aabb_candidates(segs, obj, loc, range_max, skip=(0, 0)):
    """GetSceneLines (vector_map.cpp:192-215): indices into map||objects that pass the AABB cull, in order."""
    m = np.asarray(segs, np.float32).reshape(-1, 4)
    o = np.zeros((0, 4), np.float32) if obj is None else np.asarray(obj, np.float32).reshape(-1, 4)
    al = np.concatenate([m, o])
    lx, ly, r = np.float32(loc[0]), np.float32(loc[1]), np.float32(range_max)
    x0, x1, y0, y1 = lx - r, lx + r, ly - r, ly + r
    drop = ((al[:, 0] < x0) & (al[:, 2] < x0)) | ((al[:, 1] < y0) & (al[:, 3] < y0)) | \
           ((al[:, 0] > x1) & (al[:, 2] > x1)) | ((al[:, 1] > y1) & (al[:, 3] > y1))
    keep = ~drop
    k = np.arange(len(al))
    keep &= ~((k >= len(m) + skip[0]) & (k < len(m) + skip[1]))
    idx = np.nonzero(keep)[0]
    return idx, al[idx]


def _wrap(x):
    return (x + np.pi) % (2 * np.pi) - np.pi


def spans64(cands, loc):
    """Angle each candidate subtends from loc (float64, in [0, pi])."""
    p = cands.astype(np.float64)
    a0 = np.arctan2(p[:, 1] - loc[1], p[:, 0] - loc[0])
    a1 = np.arctan2(p[:, 3] - loc[1], p[:, 2] - loc[0])
    return np.abs(_wrap(a1 - a0))


def brute_force_f64(cands, loc, angles, min_span=1e-4):
    """Nearest hit per ray over `cands` ([n][4] float32), float64, sidedness test.
    Segments that subtend less than min_span rad are left out: the reference drops them before the
    ray fill (vector_map.cpp:451, class D2 -- in BOTH tiers, it is part of the scan's definition).
    Returns (range [inf = none], candidate index [-1 = none], second-nearest range)."""
    p = cands.astype(np.float64)
    p0 = p[:, 0:2] - np.asarray(loc, np.float64)
    p1 = p[:, 2:4] - np.asarray(loc, np.float64)
    e = p1 - p0
    seen = spans64(cands, loc) >= min_span
    a = angles.astype(np.float64)
    d = np.stack([np.cos(a), np.sin(a)], 1)                       # [R][2]
    # ray o + t d meets p0 + u e:  t = cross(p0, e) / cross(d, e),  u = cross(p0, d) / cross(d, e)
    den = d[:, None, 0] * e[None, :, 1] - d[:, None, 1] * e[None, :, 0]      # [R][S]
    ce = (p0[:, 0] * e[:, 1] - p0[:, 1] * e[:, 0])[None, :]
    cd = p0[None, :, 0] * d[:, None, 1] - p0[None, :, 1] * d[:, None, 0]
    with np.errstate(divide="ignore", invalid="ignore"):
        t = ce / den
        u = cd / den
    hit = (t > 0) & (u >= 0) & (u <= 1) & np.isfinite(t) & seen[None, :]
    t = np.where(hit, t, np.inf)
    order = np.argsort(t, axis=1)[:, :2]
    best = np.take_along_axis(t, order[:, :1], 1)[:, 0]
    second = np.take_along_axis(t, order[:, 1:2], 1)[:, 0] if t.shape[1] > 1 else np.full(len(a), np.inf)
    idx = np.where(np.isfinite(best), order[:, 0], -1)
    return best, idx, second


def _end_angles(seg, loc):
    s = np.asarray(seg, np.float64)
    return (np.arctan2(s[1] - loc[1], s[0] - loc[0]), np.arctan2(s[3] - loc[1], s[2] - loc[0]))



def covers64(seg, loc, a, tol=ANG_TOL):
    """Does direction `a` fall inside the angular interval the segment subtends from loc (float64,
    widened by tol at both ends)?  Returns (covers, distance in rad to the nearer end, span)."""
    a0, a1 = _end_angles(seg, loc)
    span = _wrap(a1 - a0)
    if span < 0:
        a0, span = a1, -span
    off = _wrap(a - a0)
    if off < -tol:
        off += 2 * np.pi
    inside = -tol <= off <= span + tol
    return inside, min(abs(off), abs(off - span)), span


def range64(seg, loc, a):
    """Distance along direction a to the segment's supporting line (float64)."""
    s = np.asarray(seg, np.float64)
    p0 = s[0:2] - np.asarray(loc, np.float64)
    e = s[2:4] - s[0:2]
    d = np.array([np.cos(a), np.sin(a)])
    den = d[0] * e[1] - d[1] * e[0]
    return (p0[0] * e[1] - p0[1] * e[0]) / den if den != 0 else np.inf


class Classifier:
    def __init__(self):
        so = os.path.join(tempfile.mkdtemp(prefix="umrs_classify_"), "librender_classify.so")
        orc = oracle_lib.load_oracle()        # builds oracle/liboracle.so if needed
        subprocess.run(["g++", "-std=c++14", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-Wall", "-Werror",
                        os.path.join(ROOT, "tests", "cpp", "render_classify.cpp"),
                        "-L" + oracle_lib.ORACLE_DIR, "-l:liboracle.so", "-Wl,-rpath," + oracle_lib.ORACLE_DIR,
                        "-o", so], check=True)
        self.oracle = orc
        self.lib = C.CDLL(so)
        self.lib.umrs_test_render_provenance.restype = None
        self.lib.umrs_test_render_provenance.argtypes = [
            _FP, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, _FP, _IP,
            _FP, _IP, C.c_int, _FP, _IP, C.c_int, _IP]

    def render(self, cands, fan, range_max, n):
        """Tier A restated with provenance: ranges, per-ray piece, pieces + parents, dead pieces + parents, cap flag."""
        c = np.ascontiguousarray(cands, np.float32)
        ranges = np.empty(n, np.float32)
        ray_piece = np.empty(n, np.int32)
        cap_p, cap_d = 2000, 8000
        pieces = np.empty((cap_p, 4), np.float32)
        ppar = np.empty(cap_p, np.int32)
        dead = np.empty((cap_d, 4), np.float32)
        dpar = np.empty(cap_d, np.int32)
        counts = np.zeros(4, np.int32)
        fp = lambda a: a.ctypes.data_as(_FP)
        ip = lambda a: a.ctypes.data_as(_IP)
        self.lib.umrs_test_render_provenance(fp(c), len(c), float(fan[0]), float(fan[1]), float(range_max),
                                             float(fan[2]), float(fan[3]), n, fp(ranges), ip(ray_piece),
                                             fp(pieces), ip(ppar), cap_p, fp(dead), ip(dpar), cap_d, ip(counts))
        npc, nd = int(counts[0]), min(int(counts[1]), cap_d)
        return {"ranges": ranges, "ray_piece": ray_piece, "pieces": pieces[:npc], "piece_parent": ppar[:npc],
                "dead": dead[:nd], "dead_parent": dpar[:nd], "cap": bool(counts[2]), "sentinels": int(counts[3])}

    def classify_fan(self, segs, obj, fan, range_max, n, r_b, i_b, r_a=None, skip=(0, 0)):
        """Classifies every ray with |r_a - r_b| > 1e-5 |r_b|.  r_b/i_b: tier B (or the GPU's output,
        which is tier B bit for bit); r_a: the reference's own output (checked equal to the restated
        render when given).  Returns (counts per class + 'unclassified' + 'ref_nearer_unexplained',
        details list)."""
        loc = (float(fan[0]), float(fan[1]))
        cand_idx, cands = aabb_candidates(segs, obj, loc, range_max, skip)
        rd = self.render(cands, fan, range_max, n)
        if r_a is not None:
            assert np.array_equal(rd["ranges"].view(np.uint32), np.asarray(r_a, np.float32).view(np.uint32)), \
                "restated render != reference output"
        r_a = rd["ranges"]
        ang = ray_angles(self.oracle, fan, n).astype(np.float64)
        den = np.maximum(np.abs(r_b.astype(np.float64)), 1e-9)
        rel = np.abs(r_a.astype(np.float64) - r_b) / den
        rays = np.nonzero(rel > 1e-5)[0]
        out = {k: 0 for k in CLASSES + ("unclassified", "ref_nearer", "ref_farther")}
        details = []
        if len(rays) == 0:
            return out, details
        inv = {int(g): k for k, g in enumerate(cand_idx)}          # map||obj index -> candidate number
        r_bf, i_bf, _ = brute_force_f64(cands, loc, ang[rays].astype(np.float32))
        n_map_c = int((cand_idx < len(np.asarray(segs).reshape(-1, 4))).sum())
        for ri, i in enumerate(rays):
            a = ang[i]
            W = inv.get(int(i_b[i]), -1) if i_b[i] >= 0 else -1
            pk = int(rd["ray_piece"][i])
            A = int(rd["piece_parent"][pk]) if pk >= 0 else -1
            farther = (pk < 0) or (float(r_a[i]) > float(r_b[i]))
            out["ref_farther" if farther else "ref_nearer"] += 1
            # the independent float64 ray cast must back tier B on this ray (else only an end-point ray is excused)
            if W >= 0:
                confirmed = abs(r_bf[ri] - float(r_b[i])) <= 2e-5 * max(abs(float(r_b[i])), 1e-9)
            else:
                confirmed = not np.isfinite(r_bf[ri])
            cls = None
            if rd["cap"]:
                cls = "D4"
            elif A == W and W >= 0:
                cls = "D6"
            elif not confirmed:
                # excused only at an end-point ray of either winner, or when the ray cast's winner sits on the
                # :451 threshold (its float angles decide whether both tiers drop it)
                both = [k for k in (W, int(i_bf[ri])) if k >= 0]
                ends = [covers64(cands[k], loc, a)[1] for k in both]
                on_451 = [k for k in both if abs(covers64(cands[k], loc, a)[2] - 1e-4) <= 4e-6]
                cls = "E" if (ends and min(ends) <= ANG_TOL) else ("D2" if on_451 else None)
            elif farther:
                mine = [k for k in range(len(rd["pieces"])) if rd["piece_parent"][k] == W]
                live_cov = [k for k in mine if covers64(rd["pieces"][k], loc, a)[0]]
                if live_cov:
                    k = live_cov[0]
                    rr = self.oracle.candidate_range(rd["pieces"][k], loc, fan[2], fan[3], n, int(i))
                    if rr == rr:
                        cls = "D5"          # a piece of the true winner is rendered and covers the ray; an earlier piece matched first
                    else:
                        span = covers64(rd["pieces"][k], loc, a)[2]
                        cls = "D2" if span < 1.05e-4 else "E"
                else:
                    dead_cov = [k for k in range(len(rd["dead"])) if rd["dead_parent"][k] == W and
                                covers64(rd["dead"][k], loc, a, tol=2e-5)[0]]
                    cls = "D1" if dead_cov else "D3"
            else:
                # the reference returns a range NEARER than the exact nearest hit, and the independent ray cast
                # sides with tier B: the piece the reference used is not geometry in this direction
                dist = covers64(cands[A], loc, a)[1]
                cls = "E" if dist <= 5e-5 else "D3"
            if cls is None:
                out["unclassified"] += 1
            else:
                out[cls] += 1
                if cls == "D3" and (W >= n_map_c or A >= n_map_c):
                    out["D3_closed_shape"] = out.get("D3_closed_shape", 0) + 1
            details.append((int(i), cls, float(r_a[i]), float(r_b[i]), W, A, float(rel[i])))
        return out, details
