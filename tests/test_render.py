"""The opt-in reference render (SceneRender + TrimOcclusion, vector_map.cpp:55-258).
CPU: the restatement in csrc/umrs_render.h, built for the host, returns the compiled
reference's rendered line list bit for bit.  GPU: umrs_scan_batch_reference_render returns the
reference's GetPredictedScan ranges bit for bit (tier A), through the C-ABI."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from ut_multirobot_sim_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
D135 = np.float32(np.pi * 135.0 / 180.0)


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("render") / "librender_check.so")
    subprocess.run(["g++", "-std=c++14", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-Wall", "-Werror",
                    os.path.join(ROOT, "tests", "cpp", "render_check.cpp"), "-o", so], check=True)
    lib = C.CDLL(so)
    fp = C.POINTER(C.c_float)
    lib.umrs_test_scene_render.argtypes = [fp, C.c_int, C.c_float, C.c_float, fp, C.c_int]
    return lib


def _maps():
    import ut_multirobot_sim_b200 as umrs
    rng = np.random.default_rng(5)
    c = rng.uniform(5, 55, (220, 2))
    d = rng.normal(0, 2.5, (220, 2))
    clutter = umrs.map_cleanup(np.concatenate([c - d, c + d], 1).astype(np.float32))
    return {"floorplan_300": (umrs.map_cleanup(synth.floorplan_segments(300, extent=60.0, seed=1234)), 60.0),
            "floorplan_1024": (umrs.map_cleanup(synth.floorplan_segments(1024, extent=200.0, seed=1234)), 200.0),
            "clutter": (clutter, 60.0)}


def test_host_build_of_the_render_equals_the_reference_scene_render(harness, ref):
    fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))
    total = 0
    for name, (segs, extent) in _maps().items():
        rm = ref.map_from_segments(segs, False)
        # humans (closed 20-gons) + a square: the shapes whose occlusion the reference gets wrong
        th = np.linspace(0, 2 * np.pi, 21)
        ring = np.stack([np.cos(th[:-1]), np.sin(th[:-1]), np.cos(th[1:]), np.sin(th[1:])], 1) * 0.3
        obj = np.concatenate([ring + np.array([x, y, x, y]) for x, y in
                              [(extent * 0.4, extent * 0.5), (extent * 0.55, extent * 0.45)]]).astype(np.float32)
        rm.set_object_lines(obj)
        poses = synth.sample_free_poses_fast(segs, 24, extent, seed=3)
        for p in poses:
            scene_in = rm.scene_lines(p[:2], 100.0)
            out = np.empty((4000, 4), np.float32)
            n = harness.umrs_test_scene_render(fp(scene_in), len(scene_in), p[0], p[1], fp(out), 4000)
            want = np.empty((4000, 4), np.float32)
            m = ref.lib.umrs_ref_scene_render(rm.h, p[0], p[1], 100.0, p[2] - D135, p[2] + D135, fp(want), 4000)
            assert n == m, (name, n, m)
            assert np.array_equal(out[:n].view(np.uint32), want[:m].view(np.uint32)), name
            total += n
    assert total > 1000


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["floorplan_300", "floorplan_1024", "clutter"])
def test_gpu_reference_render_equals_the_reference_scan_bit_for_bit(ref, name):
    """umrs_scan_batch_reference_render vs the UNMODIFIED reference GetPredictedScan on the same
    map, object lines (closed human shapes) and fans: every range identical to the bit."""
    import ut_multirobot_sim_b200 as umrs
    segs, extent = _maps()[name]
    vmap = umrs.VectorMap(lines=segs)
    rm = ref.map_from_segments(segs, False)
    th = np.linspace(0, 2 * np.pi, 21)
    ring = np.stack([np.cos(th[:-1]), np.sin(th[:-1]), np.cos(th[1:]), np.sin(th[1:])], 1) * 0.3
    poses = synth.sample_free_poses_fast(segs, 96, extent, seed=7)
    # people standing around the first poses, so that scans see closed shapes up close
    obj = np.concatenate([ring + np.array([x, y, x, y]) for x, y in poses[:6, :2] + 0.9]).astype(np.float32)
    vmap.object_lines = obj
    n = 1081
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - D135, poses[:, 2] + D135], 1).astype(np.float32)
    got = vmap.ScanBatchReferenceRender(fans, 100.0, n)
    # the bridge takes object lines per fan: the same set for every fan
    want = rm.scan_batch(fans, 0.4, 100.0, n, threads=4, obj=np.tile(obj, (len(fans), 1)),
                         obj_offset=np.arange(len(fans) + 1, dtype=np.int32) * len(obj))
    same = got.view(np.uint32) == want.view(np.uint32)
    assert same.all(), "%d of %d rays differ, first fan %d" % ((~same).sum(), same.size, np.argwhere(~same)[0][0])
    # the default scan (exact nearest hit) agrees with it on most rays, not all
    exact = vmap.ScanBatch(fans, 100.0, n)
    agree = (np.abs(exact - got) <= 1e-5 * np.abs(got)).mean()
    assert agree > 0.9
    print("%s: exact nearest hit within 1e-5 of the reference render on %.4f %% of rays" % (name, 100 * agree))


@pytest.mark.gpu
def test_gpu_reference_render_full_circle_jackal_and_the_2000_line_cap(ref):
    """360-degree / 3601-ray fans (ut_jackal laser) and a clutter map whose render runs into the
    reference's 2000-candidate cap (vector_map.cpp:221-223): still the reference's bits."""
    import ut_multirobot_sim_b200 as umrs
    rng = np.random.default_rng(9)
    c = rng.uniform(0, 80, (2600, 2))
    d = rng.normal(0, 0.6, (2600, 2))
    raw = np.concatenate([c - d, c + d], 1).astype(np.float32)
    segs = umrs.map_cleanup(raw)
    assert len(segs) > 2300
    vmap = umrs.VectorMap(lines=segs)
    rm = ref.map_from_segments(segs, False)
    poses = synth.sample_free_poses_fast(segs, 12, 80.0, seed=2, clearance=0.2)
    lz = umrs.default_laser(jackal=True)
    n = umrs.laser_num_rays(lz)
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] + lz.angle_min, poses[:, 2] + lz.angle_max], 1).astype(np.float32)
    got = vmap.ScanBatchReferenceRender(fans, 150.0, n)
    want = rm.scan_batch(fans, 0.4, 150.0, n, threads=4)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_gpu_reference_render_argument_validation_and_empty_map():
    import ut_multirobot_sim_b200 as umrs
    vmap = umrs.VectorMap(lines=np.zeros((0, 4), np.float32))
    r = vmap.ScanBatchReferenceRender(np.array([[0, 0, -1, 1]], np.float32), 30.0, 16)
    assert np.array_equal(r, np.full((1, 16), 30.0, np.float32))
    assert vmap.ScanBatchReferenceRender(np.zeros((0, 4), np.float32), 30.0, 16).shape == (0, 16)
    with pytest.raises(umrs._lib.UmrsError):
        vmap.ScanBatchReferenceRender(np.array([[0, 0, -1, 1]], np.float32), 30.0, 0)


@pytest.mark.gpu
def test_world_step_in_reference_render_mode_equals_the_reference_scan(ref):
    """umrs_world_desc.reference_render: the fused tick (entity steps, footprints, laser origin)
    followed by the reference's own render per robot.  Crowded worlds (8 robots of three kinds,
    16 humans, 1 object): every scan of every step equals the unmodified reference
    GetPredictedScan on the step's own fan and object lines (minus the robot's own body) to
    the bit."""
    import ut_multirobot_sim_b200 as umrs
    extent = 60.0
    segs = umrs.map_cleanup(synth.floorplan_segments(300, extent=extent, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    rps = [umrs.default_robot("ut_jackal") for _ in range(3)] + [umrs.default_robot("bwibot") for _ in range(3)] + \
          [umrs.default_robot("cobot") for _ in range(2)]
    hps = [umrs.default_human() for _ in range(16)]
    W, R = 3, 8
    world = umrs.World(vmap, rps, W, humans=hps, static_objects=1, seed=9, reference_render=True)
    free = synth.sample_free_poses_fast(segs, W * R, extent, seed=21, clearance=0.6).reshape(W, R, 3)
    hs = synth.sample_free_poses_fast(segs, W * 32, extent, seed=22).reshape(W, 32, 3)
    st = synth.sample_free_poses_fast(segs, W, extent, seed=23).reshape(W, 1, 3)
    for arr in (free, hs, st):
        c = free[:, :1, :2]
        arr[..., :2] = c + (arr[..., :2] - c) * 0.2
    world.SetPose(free)
    world.SetHumans(hs[:, :16], hs[:, 16:])
    world.SetStaticObjects(st)
    rm = ref.map_from_segments(segs, False)
    rng = np.random.default_rng(0)
    bodies = [k for k, q in enumerate(rps) if q.footprint_enabled]
    checked = 0
    for step in range(3):
        ranges = world.Step(rng.uniform(-1, 1, (W, R, 3)).astype(np.float32))
        lines, fp = world.GetLines(), world.fan_params()
        for w in range(W):
            for r, rp in enumerate(rps):
                hdr = world.scan_header(r)
                o = lines[w]
                if rp.footprint_enabled:
                    b0 = 4 + 20 * 16 + 4 * bodies.index(r)
                    o = np.concatenate([o[:b0], o[b0 + 4:]])
                rm.set_object_lines(o)
                f = fp[w, r]
                want = rm.predicted_scan(f[:2], 0.4, rp.laser.range_max, f[2], f[3], hdr.num_rays)
                got = ranges[w, hdr.ranges_offset:hdr.ranges_offset + hdr.num_rays]
                assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (step, w, r)
                checked += hdr.num_rays
    assert checked == 3 * W * world.rays_per_world
    with pytest.raises(umrs._lib.UmrsError):      # the reference has no hit index
        world.Step(np.zeros((W, R, 3), np.float32), want_idx=True)


@pytest.mark.gpu
def test_reference_render_worlds_draw_the_same_noise_as_default_worlds():
    """Laser noise in reference-render worlds is the step kernel's Philox stream: wherever the
    clean ranges of the two modes agree, the noisy outputs agree to the bit as well."""
    import ut_multirobot_sim_b200 as umrs
    segs = umrs.map_cleanup(synth.floorplan_segments(256, extent=100.0, seed=5))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    W = 64
    poses = synth.sample_free_poses_fast(segs, W, 100.0, seed=11).reshape(W, 1, 3)
    cmds = synth.ackermann_commands(W, 1, 0)
    out = {}
    for render in (False, True):
        for noise in (False, True):
            world = umrs.World(vmap, [car], W, seed=77, laser_noise=noise, reference_render=render,
                               write_hit_idx=False)
            world.SetPose(poses)
            world.Step(cmds)
            out[(render, noise)] = world.Step(cmds).copy()
    same_clean = out[(False, False)].view(np.uint32) == out[(True, False)].view(np.uint32)
    assert same_clean.mean() > 0.9      # bitwise: the trimmed-piece drift (class D6) counts as different
    a, b = out[(False, True)], out[(True, True)]
    assert np.array_equal(a.view(np.uint32)[same_clean], b.view(np.uint32)[same_clean])
    assert (a != out[(False, False)]).mean() > 0.9      # the noise is on


@pytest.mark.gpu
def test_gpu_reference_render_picket_fence_overflows_the_speculative_staging(ref):
    """A long wall behind 70 posts: the wall is split 70 times (every post pushes a piece onto the list being
    iterated, vector_map.cpp:174), far more than the 32 pieces a speculative candidate may stage, so the
    CTA-per-fan form has to give the round up and redo the wall in order; walls in front of and behind it
    exercise the commit order.  Both forms must return the reference's bits."""
    import ut_multirobot_sim_b200 as umrs
    posts = [[x, 5.0, x + 0.12, 5.0] for x in np.linspace(-14, 14, 70)]
    raw = np.array([[-3, 2.5, 3, 2.6]] + posts + [[-20, 9.0, 20, 9.3], [-25, 12.0, 25, 12.0], [-6, 7.0, -2, 7.2]], np.float32)
    segs = umrs.map_cleanup(raw)
    vmap = umrs.VectorMap(lines=segs)
    rm = ref.map_from_segments(segs, False)
    rng = np.random.default_rng(3)
    poses = np.stack([rng.uniform(-4, 4, 1200), rng.uniform(-3, 2, 1200), rng.uniform(0.6, 2.5, 1200)], 1).astype(np.float32)
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - D135, poses[:, 2] + D135], 1).astype(np.float32)
    want = rm.scan_batch(fans, 0.4, 100.0, 1081, threads=4)
    big = vmap.ScanBatchReferenceRender(fans, 100.0, 1081)              # 1200 fans: warp per fan
    small = np.concatenate([vmap.ScanBatchReferenceRender(fans[k:k + 400], 100.0, 1081) for k in range(0, 1200, 400)])
    assert np.array_equal(big.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(small.view(np.uint32), want.view(np.uint32))
    assert (want < 8.0).mean() > 0.1 and (want > 8.5).mean() > 0.1    # rays end on posts and, between them, on the wall


@pytest.mark.gpu
def test_gpu_reference_render_warp_per_fan_and_cta_per_fan_forms_agree(ref):
    """Big batches take the warp-per-fan kernel, small ones the CTA-per-fan kernel: same bits,
    and the reference's on a sample."""
    import ut_multirobot_sim_b200 as umrs
    segs, extent = _maps()["floorplan_300"]
    vmap = umrs.VectorMap(lines=segs)
    rm = ref.map_from_segments(segs, False)
    poses = synth.sample_free_poses_fast(segs, 2000, extent, seed=13)
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - D135, poses[:, 2] + D135], 1).astype(np.float32)
    big = vmap.ScanBatchReferenceRender(fans, 100.0, 1081)                        # 2000 fans: warp per fan
    small = np.concatenate([vmap.ScanBatchReferenceRender(fans[k:k + 500], 100.0, 1081)
                            for k in range(0, 2000, 500)])                        # 500 fans: CTA per fan
    assert np.array_equal(big.view(np.uint32), small.view(np.uint32))
    want = rm.scan_batch(fans[::50], 0.4, 100.0, 1081, threads=4)
    assert np.array_equal(big[::50].view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_gpu_reference_render_random_raw_maps_with_degenerate_geometry(ref):
    """Stress of the trimming predicates: 150 small RAW maps (no Cleanup) on a coarse lattice --
    crossing segments, shared end points, collinear overlaps, duplicates, zero-length lines, the
    laser exactly on a line or on an end point -- every range equals the reference's, bit for bit
    (NaN and infinite ranges included, compared as bit patterns)."""
    import ut_multirobot_sim_b200 as umrs
    rng = np.random.default_rng(77)
    rays = 0
    for trial in range(150):
        n = int(rng.integers(1, 50))
        step = float(rng.choice([0.5, 1.0, 2.0]))
        raw = (rng.integers(-10, 11, (n, 4)) * step).astype(np.float32)
        if trial % 4 == 0:
            raw[rng.integers(0, n)] = np.repeat(raw[rng.integers(0, n), :2], 2)[[0, 2, 1, 3]]   # zero-length line
        vmap = umrs.VectorMap(lines=raw)          # taken as is: the render has to cope, like the reference
        rm = ref.map_from_segments(raw, False)
        m = 12
        locs = rng.integers(-10, 11, (m, 2)) * step * rng.choice([1.0, 0.5], (m, 1))   # often ON the lattice
        if trial % 2 == 0:
            locs = locs + rng.normal(0, 0.3, (m, 2))
        th = rng.uniform(-np.pi, np.pi, m)
        fans = np.stack([locs[:, 0], locs[:, 1], th - D135, th + D135], 1).astype(np.float32)
        nr = 181
        got = vmap.ScanBatchReferenceRender(fans, 25.0, nr)
        with np.errstate(all="ignore"):
            want = rm.scan_batch(fans, 0.4, 25.0, nr, threads=1)
        same = got.view(np.uint32) == want.view(np.uint32)
        both_nan = np.isnan(got) & np.isnan(want)      # NaN payload/sign may differ between x86 and the GPU
        assert (same | both_nan).all(), (trial, np.argwhere(~(same | both_nan))[:3], got[~(same | both_nan)][:3],
                                         want[~(same | both_nan)][:3])
        rays += got.size
    assert rays == 150 * 12 * 181
