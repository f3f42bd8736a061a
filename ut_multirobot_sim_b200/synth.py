"""Seeded synthetic inputs of the shapes BASELINE.json names.

The reference's default map (UT_Campus.vectormap.txt, default_init_config.lua:1)
lives in the external amrl_maps package and is not available, so every config
uses a seeded floor plan written in the reference's text format
("x1,y1,x2,y2" per line, vector_map.cpp:404-407).  Pure host-side numpy; nothing
here is on the hot path.
"""
import numpy as np


def floorplan_segments(n_target, extent=200.0, seed=1234, door=1.0):
    """Rooms-and-corridors floor plan: axis-aligned wall pieces on a g x g grid
    over [0, extent]^2.  Walls never cross; they meet at grid nodes (shared
    end points, like real vector maps), which VectorMap::Cleanup turns into
    2e-4 m gaps.  Returns RAW segments, float32 [n, 4], n >= n_target pieces
    trimmed to exactly n_target (outer boundary kept first)."""
    rng = np.random.default_rng(seed)
    g = 2
    while True:
        # expected pieces: boundary 4g + interior 2g(g-1) * 0.6 * 1.5
        if 4 * g + 2 * g * (g - 1) * 0.9 >= 1.25 * n_target + 8:
            break
        g += 1
    cell = extent / g
    boundary, interior = [], []
    for i in range(g):
        a, b = i * cell, (i + 1) * cell
        boundary += [(a, 0.0, b, 0.0), (a, extent, b, extent),
                     (0.0, a, 0.0, b), (extent, a, extent, b)]

    def wall(x0, y0, x1, y1):
        if rng.random() >= 0.6:
            return
        if rng.random() < 0.5 and cell > 3 * door:
            t = rng.uniform(0.15, 0.85)
            h = 0.5 * door / cell
            ax, ay = x0 + (x1 - x0) * (t - h), y0 + (y1 - y0) * (t - h)
            bx, by = x0 + (x1 - x0) * (t + h), y0 + (y1 - y0) * (t + h)
            interior.append((x0, y0, ax, ay))
            interior.append((bx, by, x1, y1))
        else:
            interior.append((x0, y0, x1, y1))

    for i in range(1, g):
        for j in range(g):
            wall(i * cell, j * cell, i * cell, (j + 1) * cell)   # vertical piece
            wall(j * cell, i * cell, (j + 1) * cell, i * cell)   # horizontal piece
    interior = np.array(interior, np.float64).reshape(-1, 4)
    boundary = np.array(boundary, np.float64)
    if len(boundary) >= n_target:
        segs = boundary[:n_target]
    else:
        need = n_target - len(boundary)
        if need > len(interior):
            raise ValueError("floor plan too small for n_target=%d" % n_target)
        pick = np.sort(rng.permutation(len(interior))[:need])
        segs = np.concatenate([boundary, interior[pick]])
    return segs.astype(np.float32)


def cityblock_segments(n_target, extent=2000.0, seed=99):
    """Large-map generator (config 4): rectangular blocks (4 walls each, 1 mm
    corner gaps so nothing touches) scattered on a jittered grid."""
    rng = np.random.default_rng(seed)
    n_blocks = (n_target + 3) // 4
    g = int(np.ceil(np.sqrt(n_blocks)))
    cell = extent / g
    segs = []
    eps = 0.001
    for k in range(n_blocks):
        i, j = divmod(k, g)
        cx = (i + 0.5) * cell + rng.uniform(-0.1, 0.1) * cell
        cy = (j + 0.5) * cell + rng.uniform(-0.1, 0.1) * cell
        hw = rng.uniform(0.15, 0.35) * cell
        hh = rng.uniform(0.15, 0.35) * cell
        segs += [(cx + hw - eps, cy + hh, cx - hw + eps, cy + hh),
                 (cx - hw, cy + hh - eps, cx - hw, cy - hh + eps),
                 (cx - hw + eps, cy - hh, cx + hw - eps, cy - hh),
                 (cx + hw, cy - hh + eps, cx + hw, cy + hh - eps)]
    return np.array(segs[:n_target], np.float32)


def write_map(path, segs):
    """The reference's text format, one 'x1,y1,x2,y2' per line."""
    with open(path, "w") as f:
        for s in np.asarray(segs, np.float32):
            f.write("%.9g,%.9g,%.9g,%.9g\n" % (s[0], s[1], s[2], s[3]))


def _dist_to_segments(p, segs):
    a = segs[:, 0:2].astype(np.float64)
    d = segs[:, 2:4].astype(np.float64) - a
    l2 = np.maximum((d * d).sum(1), 1e-12)
    t = np.clip(((p[None, :] - a) * d).sum(1) / l2, 0.0, 1.0)
    q = a + t[:, None] * d
    return np.sqrt(((q - p[None, :]) ** 2).sum(1)).min()


def sample_free_poses(segs, n, extent, seed, clearance=0.4, margin=1.0):
    """n poses {x, y, theta} uniform in free space (>= clearance from any wall)."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, 3), np.float32)
    k = 0
    segs = np.asarray(segs, np.float32)
    while k < n:
        p = rng.uniform(margin, extent - margin, 2)
        if len(segs) and _dist_to_segments(p, segs) < clearance:
            continue
        out[k] = (p[0], p[1], rng.uniform(-np.pi, np.pi))
        k += 1
    return out


def sample_free_poses_fast(segs, n, extent, seed, clearance=0.4, margin=1.0):
    """Vectorised variant for tens of thousands of poses (same contract)."""
    rng = np.random.default_rng(seed)
    segs = np.asarray(segs, np.float64)
    a = segs[:, 0:2]
    d = segs[:, 2:4] - a
    l2 = np.maximum((d * d).sum(1), 1e-12)
    out = np.empty((0, 3), np.float32)
    while len(out) < n:
        m = max(256, int(1.3 * (n - len(out))))
        p = rng.uniform(margin, extent - margin, (m, 2))
        th = rng.uniform(-np.pi, np.pi, m)
        keep = np.ones(m, bool)
        for s in range(0, len(segs), 256):
            aa, dd, ll = a[s:s + 256], d[s:s + 256], l2[s:s + 256]
            t = np.clip(((p[:, None, :] - aa[None]) * dd[None]).sum(2) / ll[None], 0, 1)
            q = aa[None] + t[..., None] * dd[None]
            dist = np.sqrt(((q - p[:, None, :]) ** 2).sum(2)).min(1)
            keep &= dist >= clearance
        new = np.concatenate([p[keep], th[keep, None]], 1).astype(np.float32)
        out = np.concatenate([out, new])
    return out[:n]


def ackermann_commands(n_worlds, n_robots, step, seed=42):
    """Config 3/5 commands: v ~ U[-1.2, 1.2], curvature ~ U[-1.02, 1.02] per world
    per step.  Host-side numpy Philox keyed by (seed, step)."""
    bg = np.random.Philox(key=seed, counter=[step, 0, 0, 0])
    rng = np.random.Generator(bg)
    c = np.zeros((n_worlds, n_robots, 3), np.float32)
    c[..., 0] = rng.uniform(-1.2, 1.2, (n_worlds, n_robots))
    c[..., 1] = rng.uniform(-1.02, 1.02, (n_worlds, n_robots))
    return c
