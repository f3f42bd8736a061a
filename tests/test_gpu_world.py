"""GPU parity of the fused world step (Simulator::update + publishLaser batched
over worlds) through the C-ABI, against the oracle's restated Step()s, footprint
generation and tier-B scan; noise distribution; full-size properties.

Tolerances: scans -- as tests/test_gpu_scan.py (1e-5 relative, exact idx but for
certified ties), always on the fan the GPU itself used (umrs_world_get_fan_params)
so pose rounding cannot leak into the scan comparison.  Poses -- the kinematic
steps use single-precision sinf/cosf, which CUDA and glibc may round differently
in the last place: |pose_gpu - pose_oracle| <= 2e-6 * (1 + |pose|) per step when
both start the step from the same state, 1e-4 after 200 free-running steps.
"""
import ctypes as C

import numpy as np
import pytest

import ut_multirobot_sim_b200 as umrs
from tests import oracle_lib
from tests.test_gpu_scan import check_against_tierb
from ut_multirobot_sim_b200 import _lib, synth

pytestmark = pytest.mark.gpu
DT = np.float32(0.025)


def _robot_states(oracle, poses):
    out = []
    for p in poses:
        s = oracle_lib.RobotState()
        oracle.lib.orc_robot_init(C.byref(s), oracle_lib._fp(np.ascontiguousarray(p, np.float32)))
        out.append(s)
    return out


def _oracle_step(oracle, s, rp, cmd, n1=0.0, n2=0.0):
    if cmd is not None:
        oracle.lib.orc_robot_command(C.byref(s), C.byref(rp), oracle_lib._fp(np.ascontiguousarray(cmd, np.float32)))
    oracle.lib.orc_robot_step(C.byref(s), C.byref(rp), DT, n1, n2)


@pytest.mark.parametrize("kind", ["ut_automata", "ut_jackal", "bwibot", "cobot"])
def test_kinematic_step_parity(oracle, kind):
    """200 free-running steps of every model family with random commands, command
    gaps (watchdog) and non-finite commands, vs the restated Step()s."""
    rp = umrs.default_robot(kind)
    rp.laser.angle_increment = np.float32(0.1)  # few rays: this test is about poses
    W = 64
    vmap = umrs.VectorMap(lines=np.array([[-50, -50, 50, -50]], np.float32))
    world = umrs.World(vmap, [rp], W, dt=float(DT), seed=3)
    rng = np.random.default_rng(11)
    poses = np.stack([rng.uniform(-5, 5, W), rng.uniform(-5, 5, W), rng.uniform(-3.1, 3.1, W)], 1).astype(np.float32)
    world.SetPose(poses.reshape(W, 1, 3))
    states = _robot_states(oracle, poses)
    worst_step = 0.0
    for step in range(200):
        cmds = np.zeros((W, 1, 3), np.float32)
        cmds[..., 0] = rng.uniform(-2.5, 2.5, (W, 1))
        cmds[..., 1] = rng.uniform(-1.5, 1.5, (W, 1))
        cmds[..., 2] = rng.uniform(-4, 4, (W, 1))
        if step % 17 == 3:
            cmds[::5, 0, 0] = np.nan
        if step % 23 == 5:
            cmds[::3, 0, 1] = 0.0  # exactly straight (Ackermann linear branch)
        mask = (rng.random((W, 1)) < 0.6).astype(np.uint8)
        if 80 <= step < 90:
            mask[:] = 0  # > 0.1 s without commands: watchdog zeroes them
        world.Step(cmds, cmd_mask=mask)
        gp, gv = world.GetPose()
        for w in range(W):
            _oracle_step(oracle, states[w], rp, cmds[w, 0] if mask[w, 0] else None)
            op = np.array(list(states[w].pose), np.float64)
            ov = np.array(list(states[w].vel), np.float64)
            # the omni model stores non-finite commands (omnidirectional_model.cpp:49-58
            # only prints): NaN then propagates identically on both sides
            assert np.array_equal(np.isnan(gp[w, 0]), np.isnan(op)), (kind, step, w, gp[w, 0], op)
            assert np.array_equal(np.isnan(gv[w, 0]), np.isnan(ov)), (kind, step, w, gv[w, 0], ov)
            fin = ~np.isnan(op)
            tol = 1e-4 * (1 + np.abs(op))
            assert np.all(np.abs(gp[w, 0] - op)[fin] <= tol[fin]), (kind, step, w, gp[w, 0], op)
            fv = ~np.isnan(ov)
            assert np.all(np.abs(gv[w, 0] - ov)[fv] <= (1e-5 * (1 + np.abs(ov)))[fv]), (kind, step, w, gv[w, 0], ov)
            if fin.all():
                worst_step = max(worst_step, float(np.max(np.abs(gp[w, 0] - op))))
    assert worst_step < 1e-4


@pytest.mark.parametrize("kind", ["ut_automata", "ut_jackal", "bwibot", "cobot"])
def test_kinematic_step_vs_compiled_reference_models(kind):
    """The fused step kernel against the UNMODIFIED reference models themselves
    (ackermann/diff_drive/omnidirectional_model.cpp compiled into oracle/_ref, see
    tests/test_ref_models.py), not the restatement: 120 free-running steps of 32 robots with
    a drive message before every step (the reference ages commands on the wall clock, the
    batched world in ticks: the watchdog is pinned on the CPU, tests/test_ref_models.py::
    test_command_watchdog).  Same tolerances as above; most poses must be bit-identical.

    Why not all of them, and why the tolerance is 1e-4 * (1 + |pose|) here: the ONLY arithmetic that
    differs between the kernel and the compiled model is sinf/cosf of the heading -- CUDA's and
    glibc's single-precision implementations are both faithful but round differently on about one
    argument in ten (1 ulp).  A 1-ulp difference in sin(theta) moves x or y by ~1e-9 m per step; the
    runs are FREE-RUNNING (each side integrates its own state), so once a pose differs in the last
    bit it stays different and the gap performs a random walk over the 120 steps.  From identical
    state a single step agrees to 2e-6 * (1 + |pose|) with > 90 % of poses bit-identical: see
    test_single_step_vs_compiled_reference_models below."""
    ref = oracle_lib.load_ref()
    if ref is None or not getattr(ref, "has_models", False):
        pytest.skip("oracle/_ref/libumrs_ref.so without the model bridge")
    rp = umrs.default_robot(kind)
    # the reference Ackermann model always draws its angular error from its own engine; the
    # draws are pinned in tests/test_ref_models.py::test_ackermann_noise_draws, here they are off
    rp.ak_angular_error_rate = 0.0
    rp.laser.angle_increment = np.float32(0.1)
    W = 32
    vmap = umrs.VectorMap(lines=np.array([[-50, -50, 50, -50]], np.float32))
    world = umrs.World(vmap, [rp], W, dt=float(DT), seed=3)
    rng = np.random.default_rng(21)
    poses = np.stack([rng.uniform(-5, 5, W), rng.uniform(-5, 5, W), rng.uniform(-3.1, 3.1, W)], 1).astype(np.float32)
    world.SetPose(poses.reshape(W, 1, 3))
    models = []
    for w in range(W):
        m = ref.robot_model(rp)
        m.SetPose(poses[w])
        models.append(m)
    exact = total = 0
    for step in range(120):
        cmds = np.zeros((W, 1, 3), np.float32)
        cmds[..., 0] = rng.uniform(-2.5, 2.5, (W, 1))
        cmds[..., 1] = rng.uniform(-1.5, 1.5, (W, 1))
        cmds[..., 2] = rng.uniform(-4, 4, (W, 1))
        if step % 19 == 7:
            cmds[::4, 0, 1] = 0.0   # exactly straight
        world.Step(cmds)
        gp, gv = world.GetPose()
        for w, m in enumerate(models):
            m.DriveCallback(cmds[w, 0])
            m.Step(float(DT))
            rpose, rvel = m.GetPose().astype(np.float64), m.GetVel().astype(np.float64)
            assert np.all(np.abs(gp[w, 0] - rpose) <= 1e-4 * (1 + np.abs(rpose))), (kind, step, w, gp[w, 0], rpose)
            assert np.all(np.abs(gv[w, 0] - rvel) <= 1e-5 * (1 + np.abs(rvel))), (kind, step, w, gv[w, 0], rvel)
            exact += int(np.array_equal(gp[w, 0].astype(np.float32).view(np.uint32),
                                        m.GetPose().view(np.uint32)))
            total += 1
    assert exact >= 0.5 * total, "only %d/%d poses bit-identical to the reference model" % (exact, total)


@pytest.mark.parametrize("kind", ["ut_automata", "ut_jackal", "cobot"])
def test_single_step_vs_compiled_reference_models(kind):
    """ONE step from bit-identical state against the UNMODIFIED compiled model (oracle/_ref):
    |pose difference| <= 2e-6 * (1 + |pose|) -- a few ulp of the 100 m coordinates, from CUDA vs glibc
    sinf/cosf -- velocities bit-identical, and > 90 % of the poses bit-identical."""
    ref = oracle_lib.load_ref()
    if ref is None or not getattr(ref, "has_models", False):
        pytest.skip("oracle/_ref/libumrs_ref.so without the model bridge")
    rp = umrs.default_robot(kind)
    rp.ak_angular_error_rate = 0.0      # the reference model draws from its own engine (pinned on the CPU)
    rp.laser.angle_increment = np.float32(0.1)
    W = 256
    vmap = umrs.VectorMap(lines=np.array([[-50, -50, 50, -50]], np.float32))
    world = umrs.World(vmap, [rp], W, dt=float(DT), seed=3)
    rng = np.random.default_rng(5)
    poses = np.stack([rng.uniform(-100, 100, W), rng.uniform(-100, 100, W), rng.uniform(-3.1, 3.1, W)], 1).astype(np.float32)
    vel = np.zeros((W, 3), np.float32)
    if kind != "ut_jackal":             # RobotModel::SetVel does not reach the diff-drive model's own doubles
        vel[:, 0] = rng.uniform(-1, 1, W)
    if kind == "cobot":
        vel[:, 1] = rng.uniform(-1, 1, W)
        vel[:, 2] = rng.uniform(-1, 1, W)
    world.SetPose(poses.reshape(W, 1, 3), vel.reshape(W, 1, 3))
    cmds = np.zeros((W, 1, 3), np.float32)
    cmds[..., 0] = rng.uniform(-2, 2, (W, 1))
    cmds[..., 1] = rng.uniform(-1.2, 1.2, (W, 1))
    cmds[..., 2] = rng.uniform(-1, 1, (W, 1))
    world.Step(cmds)
    gp, gv = world.GetPose()
    exact = 0
    for w in range(W):
        m = ref.robot_model(rp)
        m.SetPose(poses[w])
        m.SetVel(vel[w])
        m.DriveCallback(cmds[w, 0])
        m.Step(float(DT))
        rpose, rvel = m.GetPose(), m.GetVel()
        assert np.all(np.abs(gp[w, 0].astype(np.float64) - rpose) <= 2e-6 * (1 + np.abs(rpose))), (w, gp[w, 0], rpose)
        assert np.array_equal(gv[w, 0].view(np.uint32), rvel.view(np.uint32)), (w, gv[w, 0], rvel)
        exact += int(np.array_equal(gp[w, 0].view(np.uint32), rpose.view(np.uint32)))
    assert exact >= 0.9 * W, "only %d/%d poses bit-identical to the compiled model" % (exact, W)


@pytest.mark.parametrize("kind", ["ut_automata", "ut_jackal", "cobot"])
def test_single_step_from_identical_state(oracle, kind):
    """One step from bit-identical state: 2e-6 * (1 + |pose|)."""
    rp = umrs.default_robot(kind)
    rp.laser.angle_increment = np.float32(0.1)
    W = 256
    vmap = umrs.VectorMap(lines=np.array([[-50, -50, 50, -50]], np.float32))
    world = umrs.World(vmap, [rp], W, dt=float(DT), seed=3)
    rng = np.random.default_rng(5)
    poses = np.stack([rng.uniform(-100, 100, W), rng.uniform(-100, 100, W), rng.uniform(-3.1, 3.1, W)], 1).astype(np.float32)
    vel = np.zeros((W, 3), np.float32)
    vel[:, 0] = rng.uniform(-1, 1, W)
    if kind == "cobot":
        vel[:, 1] = rng.uniform(-1, 1, W)
    if kind != "ut_automata":
        vel[:, 2] = rng.uniform(-1, 1, W)
    world.SetPose(poses.reshape(W, 1, 3), vel.reshape(W, 1, 3))
    cmds = np.zeros((W, 1, 3), np.float32)
    cmds[..., 0] = rng.uniform(-2, 2, (W, 1))
    cmds[..., 1] = rng.uniform(-1.2, 1.2, (W, 1))
    cmds[..., 2] = rng.uniform(-1, 1, (W, 1))
    world.Step(cmds)
    gp, _ = world.GetPose()
    exact = 0
    for w in range(W):
        s = oracle_lib.RobotState()
        oracle.lib.orc_robot_init(C.byref(s), oracle_lib._fp(np.ascontiguousarray(poses[w])))
        for k in range(3):
            s.vel[k] = vel[w, k]
        s.linear_vel, s.angular_vel = float(vel[w, 0]), float(vel[w, 2])
        _oracle_step(oracle, s, rp, cmds[w, 0])
        op = np.array(list(s.pose), np.float32)
        assert np.all(np.abs(gp[w, 0].astype(np.float64) - op) <= 2e-6 * (1 + np.abs(op))), (w, gp[w, 0], op)
        exact += int(np.array_equal(gp[w, 0].view(np.uint32), op.view(np.uint32)))
    assert exact >= 0.9 * W, "only %d/%d poses bit-identical" % (exact, W)


def _world_lines_oracle(oracle, statics, human_poses, hps, robot_poses, rps):
    """[statics][humans][robot bodies] in the C-ABI's documented order."""
    lines = [oracle.transform_lines(oracle.square_template(), p) for p in statics]
    lines += [oracle.transform_lines(oracle.human_template(hp), p) for hp, p in zip(hps, human_poses)]
    offs = []
    n = sum(len(l) for l in lines)
    for rp, p in zip(rps, robot_poses):
        if rp.footprint_enabled:
            lines.append(oracle.transform_lines(oracle.robot_template(rp), p))
            offs.append((n, n + 4))
            n += 4
        else:
            offs.append((n, n))
    return (np.concatenate(lines) if lines else np.zeros((0, 4), np.float32)), offs


def _config2_roster():
    """BASELINE config 2: 3 ut_jackal + 3 bwibot + 2 cobot, 16 humans."""
    rps = [umrs.default_robot("ut_jackal") for _ in range(3)]
    rps += [umrs.default_robot("bwibot") for _ in range(3)]
    rps += [umrs.default_robot("cobot") for _ in range(2)]
    hps = [umrs.default_human() for _ in range(16)]
    return rps, hps


def test_config2_crowded_world(oracle):
    """One world, 8 heterogeneous robots + 16 humans + 1 static object on the
    campus floor plan; every scan sees the other robots' bodies and the humans
    (and not its own body).  20 steps; object lines, human poses and every scan
    checked against the oracle each step."""
    extent = 60.0
    segs = umrs.map_cleanup(synth.floorplan_segments(300, extent=extent, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    rps, hps = _config2_roster()
    R, H = len(rps), len(hps)
    world = umrs.World(vmap, rps, 1, humans=hps, static_objects=1, dt=float(DT), seed=9)
    assert world.rays_per_world == 3 * 3601 + 5 * 1081
    assert world.object_lines_per_world == 4 + 20 * H + 4 * R
    free = synth.sample_free_poses(segs, R + 2 * H + 1, extent, seed=21, clearance=0.6)
    rposes = free[:R].copy()
    hstart, hgoal = free[R:R + H].copy(), free[R + H:R + 2 * H].copy()
    # keep everybody close so that occlusion happens in every scan
    centre = rposes[0, :2]
    for arr in (rposes, hstart, hgoal):
        arr[:, :2] = centre + (arr[:, :2] - centre) * 0.15
    spose = free[-1:].copy()
    spose[:, :2] = centre + (spose[:, :2] - centre) * 0.15
    world.SetPose(rposes.reshape(1, R, 3))
    world.SetHumans(hstart.reshape(1, H, 3), hgoal.reshape(1, H, 3))
    world.SetStaticObjects(spose.reshape(1, 1, 3))
    rstates = _robot_states(oracle, rposes)
    hstates = []
    for h in range(H):
        hp = umrs.default_human()
        hp.start_x, hp.start_y, hp.start_theta = hstart[h]
        hp.goal_x, hp.goal_y, hp.goal_theta = hgoal[h]
        s = oracle_lib.HumanState()
        oracle.lib.orc_human_init(C.byref(s), C.byref(hp))
        hstates.append(s)
    rng = np.random.default_rng(1)
    occluded_by_objects = 0
    for step in range(20):
        cmds = rng.uniform(-1, 1, (1, R, 3)).astype(np.float32)
        ranges, idx = world.Step(cmds, want_idx=True)
        for r in range(R):
            _oracle_step(oracle, rstates[r], rps[r], cmds[0, r])
        for h in range(H):
            oracle.lib.orc_human_step(C.byref(hstates[h]), C.byref(hps[h]), DT)
        gp, _ = world.GetPose()
        ghp = world.GetHumanPoses()
        for h in range(H):
            assert np.allclose(ghp[0, h], list(hstates[h].pose), atol=1e-5)
        # object lines: built by the oracle from the GPU's own poses (so pose
        # rounding does not leak), compared bit for bit
        lines_o, offs = _world_lines_oracle(oracle, spose, ghp[0], hps, gp[0], rps)
        lines_g = world.GetLines()[0]
        assert lines_g.shape == lines_o.shape
        assert np.abs(lines_g - lines_o).max() <= 2e-6 * (1 + np.abs(lines_o).max())
        fans = world.fan_params()[0]
        for r in range(R):
            hdr = world.scan_header(r)
            sl = slice(hdr.ranges_offset, hdr.ranges_offset + hdr.num_rays)
            # laser origin, simulator.cpp:398-400
            o = oracle.laser_origin(gp[0, r], rps[r].laser.loc_x, rps[r].laser.loc_y)
            assert np.allclose(fans[r, :2], o, atol=2e-6 * (1 + np.abs(o).max()))
            assert fans[r, 2] == np.float32(rps[r].laser.angle_min) + gp[0, r, 2]
            check_against_tierb(oracle, segs, lines_g, fans[r], rps[r].laser.range_max, hdr.num_rays,
                                ranges[0, sl], idx[0, sl], skip=offs[r])
            occluded_by_objects += int((idx[0, sl] >= len(segs)).sum())
            own = (idx[0, sl] >= len(segs) + offs[r][0]) & (idx[0, sl] < len(segs) + offs[r][1])
            assert not own.any(), "robot %d sees its own body" % r
    assert occluded_by_objects > 1000


def test_scan_header_is_laserscan_shaped():
    """simulator.cpp:150-159."""
    vmap = umrs.VectorMap(lines=np.array([[0, 0, 1, 0]], np.float32))
    rps = [umrs.default_robot("ut_automata"), umrs.default_robot("ut_jackal")]
    world = umrs.World(vmap, rps, 2)
    h0, h1 = world.scan_header(0), world.scan_header(1)
    assert (h0.num_rays, h0.ranges_offset) == (1081, 0)
    assert (h1.num_rays, h1.ranges_offset) == (3601, 1081)
    assert h0.time_increment == 0.0 and h0.scan_time == np.float32(0.05)
    assert h0.angle_min == rps[0].laser.angle_min and h0.range_max == 100.0
    assert h1.range_max == 150.0 and h1.angle_increment == rps[1].laser.angle_increment


def _make_car_world(W, M=256, seed=42, base=0, noise=False, extent=200.0, pose_seed=11, total=None):
    segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=extent, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    world = umrs.World(vmap, [car], W, dt=float(DT), seed=seed, laser_noise=noise, motion_noise=noise,
                       world_index_base=base)
    total = total or W
    poses = synth.sample_free_poses_fast(segs, total, extent, seed=pose_seed)[base:base + W]
    world.SetPose(poses.reshape(W, 1, 3))
    return world, segs, car, vmap


def test_noise_off_fused_step_equals_oracle_scan(oracle):
    """BASELINE config 3 shape at a checker-friendly size: Ackermann step fused
    with the scan, 5 steps, every world checked."""
    W = 48
    world, segs, car, _ = _make_car_world(W, M=256)
    for step in range(5):
        cmds = synth.ackermann_commands(W, 1, step)
        ranges, idx = world.Step(cmds, want_idx=True)
        fans = world.fan_params()
        for w in range(W):
            check_against_tierb(oracle, segs, None, fans[w, 0], 100.0, 1081, ranges[w], idx[w])


def test_laser_noise_dropout_rule_and_distribution(oracle):
    """simulator.cpp:412-418 with Philox noise: r > range_max - 0.1 -> 0 exactly;
    otherwise (out - clean) / sigma ~ N(0, 1); independent across rays / worlds /
    steps; reproducible from (seed, world, ray, step)."""
    from scipy import stats
    W = 256
    clean_w, segs, car, _ = _make_car_world(W, M=64, noise=False)
    noisy_w, _, _, _ = _make_car_world(W, M=64, noise=True)
    # motion noise off for the noisy world would be needed for identical poses:
    # use zero curvature commands (linear branch draws nothing)
    cmds = np.zeros((W, 1, 3), np.float32)
    cmds[..., 0] = 0.5
    clean = clean_w.Step(cmds)
    noisy = noisy_w.Step(cmds)
    assert np.array_equal(clean_w.fan_params(), noisy_w.fan_params())
    drop = clean.astype(np.float64) > np.float64(np.float32(100.0)) - 0.1
    assert drop.any() and (~drop).sum() > 100000
    assert np.all(noisy[drop] == 0.0)
    z = (noisy[~drop].astype(np.float64) - clean[~drop]) / float(car.laser.noise_stddev)
    # float rounding of out = r + sigma*g at r ~ 50 m is ~4e-6 m = 4e-4 sigma
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1.0) < 0.01
    assert stats.kstest(z[::7], "norm").pvalue > 1e-3
    assert abs(stats.skew(z)) < 0.02 and abs(stats.kurtosis(z)) < 0.05
    # independence: neighbouring rays and neighbouring worlds uncorrelated
    zz = np.where(drop, 0.0, (noisy.astype(np.float64) - clean) / float(car.laser.noise_stddev))
    m = ~drop[:, :-1] & ~drop[:, 1:]
    assert abs(np.corrcoef(zz[:, :-1][m], zz[:, 1:][m])[0, 1]) < 0.01
    m = ~drop[:-1] & ~drop[1:]
    assert abs(np.corrcoef(zz[:-1][m], zz[1:][m])[0, 1]) < 0.01
    # next step draws fresh noise
    noisy2 = noisy_w.Step(cmds)
    clean2 = clean_w.Step(cmds)
    keep = ~drop & ~(clean2.astype(np.float64) > np.float64(np.float32(100.0)) - 0.1)
    z2 = (noisy2.astype(np.float64) - clean2) / float(car.laser.noise_stddev)
    assert abs(np.corrcoef(zz[keep], z2[keep])[0, 1]) < 0.01
    # reproducible: same seed -> identical bits
    again, _, _, _ = _make_car_world(W, M=64, noise=True)
    assert np.array_equal(again.Step(cmds).view(np.uint32), noisy.view(np.uint32))
    # a different seed -> different noise
    other, _, _, _ = _make_car_world(W, M=64, noise=True, seed=43)
    assert not np.array_equal(other.Step(cmds).view(np.uint32), noisy.view(np.uint32))


def test_noise_is_the_documented_philox_stream(oracle):
    """out = max(0, r + sigma * g): ray i = 128*s + 32*q + l takes normal q of
    Philox4x32-10(counter = {32*s + l, global fan, step_lo, step_hi ^ 'LASR'},
    key = seed) + Box-Muller, restated on the host by the oracle.  The kernel uses
    the hardware log/sin/cos approximations (abs. error ~4e-7 in g, 4e-9 m in the
    range): tolerance 2e-6 * (1 + |range|)."""
    W = 4
    clean_w, segs, car, _ = _make_car_world(W, M=64, noise=False)
    noisy_w, _, _, _ = _make_car_world(W, M=64, noise=True, seed=0x1234567800000042, base=7, total=7 + W)
    clean_w.SetPose(noisy_w.GetPose()[0])
    cmds = np.zeros((W, 1, 3), np.float32)
    clean = clean_w.Step(cmds)
    noisy = noisy_w.Step(cmds)
    sigma = np.float32(car.laser.noise_stddev)
    for w in range(W):
        for i in (0, 1, 2, 3, 4, 517, 1080):
            if float(clean[w, i]) > float(np.float32(100.0)) - 0.1:
                continue
            blk, q = 32 * (i >> 7) + (i & 31), (i >> 5) & 3
            words = oracle.philox([blk, 7 + w, 0, 0 ^ 0x4C415352], [0x00000042, 0x12345678])
            pair = oracle.box_muller(words[2], words[3]) if (q & 2) else oracle.box_muller(words[0], words[1])
            g = pair[1] if (q & 1) else pair[0]
            want = np.float32(max(np.float32(0), clean[w, i] + sigma * g))
            assert abs(float(noisy[w, i]) - float(want)) <= 2e-6 * (1 + abs(float(want))), (w, i)


def test_sharded_run_equals_unsharded(oracle):
    """world_index_base keeps the Philox counters global: two shards of 96 worlds
    produce exactly the ranges of one 192-world job (multi-GPU weak scaling has
    no effect on results)."""
    W = 192
    full, _, _, _ = _make_car_world(W, noise=True)
    a, _, _, _ = _make_car_world(96, noise=True, base=0, total=W)
    b, _, _, _ = _make_car_world(96, noise=True, base=96, total=W)
    for step in range(3):
        cmds = synth.ackermann_commands(W, 1, step)
        rf = full.Step(cmds)
        ra = a.Step(cmds[:96])
        rb = b.Step(cmds[96:])
        assert np.array_equal(rf[:96].view(np.uint32), ra.view(np.uint32))
        assert np.array_equal(rf[96:].view(np.uint32), rb.view(np.uint32))
    sf, sa, sb = full.stats(), a.stats(), b.stats()
    assert sf["hits"] == sa["hits"] + sb["hits"]
    assert sf["min"] == min(sa["min"], sb["min"]) and sf["max"] == max(sa["max"], sb["max"])
    assert abs(sf["sum"] - (sa["sum"] + sb["sum"])) < 1e-6 * sf["sum"]


def test_stats_kernel_matches_numpy():
    W = 64
    world, _, _, _ = _make_car_world(W, noise=True)
    r = world.Step(synth.ackermann_commands(W, 1, 0))
    s = world.stats()
    pos = r[r > 0].astype(np.float64)
    assert s["hits"] == len(pos)
    assert s["min"] == pos.min() and s["max"] == pos.max()
    assert abs(s["sum"] - pos.sum()) < 1e-9 * pos.sum()


def test_full_size_config3_properties(oracle):
    """BASELINE config 3 at full size (4096 worlds x 1081 rays, 1024-segment map):
    size-independent properties + a sampled oracle check."""
    W = 4096
    world, segs, car, vmap = _make_car_world(W, M=1024)
    cmds = synth.ackermann_commands(W, 1, 0)
    ranges, idx = world.Step(cmds, want_idx=True)
    assert ranges.shape == (W, 1081)
    assert np.isfinite(ranges).all() and (ranges > 0).all()
    assert ranges.max() <= 100.0 * np.sqrt(2.0) * (1 + 1e-5)
    assert np.all(ranges[idx < 0] == np.float32(100.0))
    assert idx.max() < len(segs)
    fans = world.fan_params()
    for w in range(0, W, 128):
        check_against_tierb(oracle, segs, None, fans[w, 0], 100.0, 1081, ranges[w], idx[w])
    # the fused step's scan == the standalone scan entry point on the same fans
    rb = vmap.ScanBatch(fans.reshape(-1, 4), 100.0, 1081)
    assert np.array_equal(rb.view(np.uint32), ranges.view(np.uint32))
    # idempotence of the scan w.r.t. state: stepping with zero velocity command from
    # rest keeps the pose (v stays 0 after the accel clamp) -> identical scan
    world2, _, _, _ = _make_car_world(W, M=1024)
    z = np.zeros((W, 1, 3), np.float32)
    r1 = world2.Step(z)
    r2 = world2.Step(z)
    assert np.array_equal(r1.view(np.uint32), r2.view(np.uint32))
    # linearity-like property of the geometry: the hit point lies on the hit segment
    w = 17
    a = np.array([oracle.lib.orc_ray_angle(float(fans[w, 0, 2]), float(fans[w, 0, 3]), 1081, i) for i in range(1081)])
    hit = idx[w] >= 0
    px = fans[w, 0, 0] + ranges[w] * np.cos(a)
    py = fans[w, 0, 1] + ranges[w] * np.sin(a)
    s = segs[idx[w][hit]].astype(np.float64)
    d = s[:, 2:4] - s[:, 0:2]
    q = np.stack([px[hit], py[hit]], 1) - s[:, 0:2]
    cross = np.abs(d[:, 0] * q[:, 1] - d[:, 1] * q[:, 0]) / np.hypot(d[:, 0], d[:, 1])
    assert cross.max() < 2e-3  # metres off the wall line (float32 angles at up to 141 m)


def test_world_argument_validation():
    from ut_multirobot_sim_b200 import _lib
    vmap = umrs.VectorMap(lines=np.array([[0, 0, 1, 0]], np.float32))
    with pytest.raises(_lib.UmrsError):
        umrs.World(vmap, [umrs.default_robot("ut_automata")], 0)
    bad = umrs.default_robot("ut_automata")
    bad.laser.angle_max = bad.laser.angle_min
    with pytest.raises(_lib.UmrsError):
        umrs.World(vmap, [bad], 1)
    w = umrs.World(vmap, [umrs.default_robot("ut_automata")], 2, write_hit_idx=False)
    with pytest.raises(_lib.UmrsError):
        w.Step(np.zeros((2, 1, 3), np.float32), want_idx=True)


def test_repeated_runs_are_bit_identical():
    """Stand-in for racecheck (compute-sanitizer is closed on this pool): the fan kernel hands out chunks by
    ticket, compacts with warp ballots + shared atomics and lowers per-chunk bounds with atomicMin -- all
    order-dependent INSIDE a phase but, by design, not in their results.  Six identical worlds on the
    busiest plan (1024-segment map, pre-cull on, noise on), five ticks each: every output bit equal."""
    outs = []
    for rep in range(6):
        world, segs, car, vmap = _make_car_world(1500, M=1024, noise=True)
        o = []
        for step in range(5):
            r, i = world.Step(synth.ackermann_commands(1500, 1, step), want_idx=True)
            o += [r.view(np.uint32).copy(), i.copy()]
        outs.append(o)
    for rep in range(1, 6):
        for a, b in zip(outs[0], outs[rep]):
            assert np.array_equal(a, b), rep


def test_map_outlives_its_worlds():
    """a world keeps using the map's device buffers, so umrs_map_destroy refuses
    (UMRS_ERR_INVALID, nothing freed) while a world created from the map exists."""
    lib = _lib.load()
    segs = umrs.map_cleanup(synth.floorplan_segments(64, extent=50.0, seed=5))
    h = C.c_void_p()
    _lib.check(lib.umrs_map_create(segs.ctypes.data_as(C.POINTER(C.c_float)), len(segs), C.byref(h)), "map_create")
    car = umrs.default_robot("ut_automata")
    desc = _lib.WorldDesc(num_worlds=4, robots_per_world=1, humans_per_world=0, static_objects_per_world=0,
                          robots=C.pointer(car), humans=None, dt=0.025, seed=1, laser_noise_enabled=0,
                          motion_noise_enabled=0, world_index_base=0, write_hit_idx=0, reference_render=0)
    wh = C.c_void_p()
    _lib.check(lib.umrs_world_create(h, C.byref(desc), C.byref(wh)), "world_create")
    assert lib.umrs_map_destroy(h) == _lib.ERR_INVALID
    out = np.empty((4, 1081), np.float32)
    cmds = np.zeros((4, 1, 3), np.float32)
    _lib.check(lib.umrs_world_step(wh, cmds.ctypes.data, None, out.ctypes.data, None), "step")   # the map is intact
    assert np.isfinite(out).all() and lib.umrs_world_check(wh) == _lib.UMRS_OK
    _lib.check(lib.umrs_world_destroy(wh), "world_destroy")
    assert lib.umrs_map_destroy(h) == _lib.UMRS_OK


def test_pipelined_async_ticks_equal_synchronous_ticks():
    """umrs_world_step_async: 12 pipelined ticks into three rotating pinned buffers give, tick by
    tick, the bits of 12 synchronous ticks (noise on: the step counter advances the same way)."""
    import torch
    W = 2048
    segs = umrs.map_cleanup(synth.floorplan_segments(256, extent=100.0, seed=5))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    poses = synth.sample_free_poses_fast(segs, W, 100.0, seed=11).reshape(W, 1, 3)
    worlds = [umrs.World(vmap, [car], W, seed=3, laser_noise=True, motion_noise=True, write_hit_idx=False)
              for _ in range(2)]
    for wd in worlds:
        wd.SetPose(poses)
    n = worlds[0].rays_per_world
    cmds = [torch.from_numpy(synth.ackermann_commands(W, 1, s)).pin_memory() for s in range(12)]
    bufs = [torch.empty((W, n), dtype=torch.float32).pin_memory() for _ in range(3)]
    want = [worlds[0].Step(cmds[s].numpy()).copy() for s in range(12)]
    got = []
    for s in range(12):
        worlds[1].step_async_raw(cmds[s].data_ptr(), None, bufs[s % 3].data_ptr(), None)
        if s >= 1:
            worlds[1].wait_outputs(1)                      # tick s-1 is in host memory
            got.append(bufs[(s - 1) % 3].numpy().copy())
    worlds[1].wait()
    got.append(bufs[11 % 3].numpy().copy())
    for s in range(12):
        assert np.array_equal(got[s].view(np.uint32), want[s].view(np.uint32)), s
    # a synchronous tick after pipelined ones is still correct
    nxt = synth.ackermann_commands(W, 1, 12)
    assert np.array_equal(worlds[1].Step(nxt).view(np.uint32), worlds[0].Step(nxt).view(np.uint32))


@pytest.mark.parametrize("segments", [64, 1024])
def test_cta_shapes_give_the_same_bits(segments):
    """The fan kernels exist in seven CTA shapes (256x4, 192x5, 192x6, 128x8, 96x10, 512x2, 320x3 threads x
    CTAs per SM) and a launch picks one from its shared-memory plan and
    grid size.  The "cta_shape" tuning hook forces one: ranges, hit indices and poses of three
    noisy ticks, and a scan batch over the same map, must be bit-identical in all seven and
    in the picked one (a 1024-world grid on the 64-segment map picks 96x10, the
    1024-segment map 256x4)."""
    W = 1200
    runs = {}
    for shape in (-1, 0, 1, 2, 3, 4, 5, 6):
        with _lib.tuning(cta_shape=shape):
            world, segs, car, vmap = _make_car_world(W, M=segments, noise=True)
            out = []
            for step in range(3):
                r, i = world.Step(synth.ackermann_commands(W, 1, step), want_idx=True)
                out += [r.view(np.uint32).copy(), i.copy()]
            pose, vel = world.GetPose()
            fans = world.fan_params().reshape(-1, 4)[:160]
            rb, ib = vmap.ScanBatch(fans, 100.0, 1081, return_index=True)
        runs[shape] = out + [pose.view(np.uint32), vel.view(np.uint32), rb.view(np.uint32), ib]
    for shape in (0, 1, 2, 3, 4, 5, 6):
        for a, b in zip(runs[-1], runs[shape]):
            assert np.array_equal(a, b), "shape %s differs from the picked shape" % shape


def test_two_launch_step_gives_the_fused_step_bits():
    """A tick can run fused (one kernel: entity steps + scans) or split (k_world_entities, then the
    fan kernel on the stepped poses; "split_step" tuning, auto for large batches).  Same bits either
    way: ranges, hit indices, poses, velocities, human poses, object lines and fan parameters, on a
    crowded roster with every model, humans, a static object, laser + motion noise, over 4 ticks --
    and on a car-only roster, blocking and pipelined ticks."""
    segs = umrs.map_cleanup(synth.floorplan_segments(300, extent=60.0, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    rps, hps = _config2_roster()
    W, R, H = 40, len(rps), len(hps)
    free = synth.sample_free_poses_fast(segs, W * (R + 2 * H + 1), 60.0, seed=5, clearance=0.5).reshape(W, -1, 3)
    cmds = np.random.default_rng(3).uniform(-1, 1, (4, W, R, 3)).astype(np.float32)
    runs = {}
    for split in (0, 1):
        with _lib.tuning(split_step=split):
            wd = umrs.World(vmap, rps, W, humans=hps, static_objects=1, dt=float(DT), seed=77,
                            laser_noise=True, motion_noise=True)
            wd.SetPose(free[:, :R])
            wd.SetHumans(free[:, R:R + H], free[:, R + H:R + 2 * H])
            wd.SetStaticObjects(free[:, R + 2 * H:])
            out = []
            for k in range(4):
                r, i = wd.Step(cmds[k], want_idx=True)
                out += [r.view(np.uint32).copy(), i.copy(), wd.GetLines().view(np.uint32).copy(),
                        wd.fan_params().view(np.uint32).copy()]
            pose, vel = wd.GetPose()
            out += [pose.view(np.uint32), vel.view(np.uint32), np.frombuffer(wd.get_state(), np.uint8).copy()]
        runs[split] = out
    for a, b in zip(runs[0], runs[1]):
        assert np.array_equal(a, b)
    # car-only roster (the instance without the other models), device-resident and pipelined ticks
    Wc = 700
    outs = {}
    for split in (0, 1):
        with _lib.tuning(split_step=split):
            world, _, _, _ = _make_car_world(Wc, M=256, noise=True)
            o = []
            for step in range(3):
                r, i = world.Step(synth.ackermann_commands(Wc, 1, step), want_idx=True)
                o += [r.view(np.uint32).copy(), i.copy()]
            pose, vel = world.GetPose()
            o += [pose.view(np.uint32), vel.view(np.uint32)]
        outs[split] = o
    for a, b in zip(outs[0], outs[1]):
        assert np.array_equal(a, b)


def test_checkpoint_resume_is_bit_reproducible():
    """umrs_world_get_state / set_state (the reference only has SetPose, simulator.cpp:191-196):
    50 ticks, snapshot, 50 more; restore the snapshot into the same world AND into a fresh one, 50 more
    => identical ranges, noise included, identical poses / velocities / human poses.  The crowded
    roster covers all three robot models (the diff-drive double velocities, command ages) and
    humans that swap start and goal."""
    segs = umrs.map_cleanup(synth.floorplan_segments(300, extent=60.0, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    rps, hps = _config2_roster()
    W, R, H = 6, len(rps), len(hps)

    def make():
        wd = umrs.World(vmap, rps, W, humans=hps, static_objects=1, dt=float(DT), seed=77,
                        laser_noise=True, motion_noise=True)
        free = synth.sample_free_poses_fast(segs, W * (R + 2 * H + 1), 60.0, seed=5, clearance=0.5).reshape(W, -1, 3)
        wd.SetPose(free[:, :R])
        # short walks so that humans reach their goals and swap them inside the run
        hs, hg = free[:, R:R + H].copy(), free[:, R + H:R + 2 * H].copy()
        hg[..., :2] = hs[..., :2] + 0.02 * (hg[..., :2] - hs[..., :2])
        wd.SetHumans(hs, hg)
        wd.SetStaticObjects(free[:, -1:])
        return wd

    rng = np.random.default_rng(3)
    cmds = rng.uniform(-1, 1, (150, W, R, 3)).astype(np.float32)
    masks = (rng.uniform(0, 1, (150, W, R)) < 0.7).astype(np.uint8)   # some robots go un-commanded: ages matter
    a = make()
    for s in range(50):
        a.Step(cmds[s], cmd_mask=masks[s])
    snap = a.get_state()
    want = [a.Step(cmds[s], cmd_mask=masks[s]).copy() for s in range(50, 100)]
    want_pose, want_vel = a.GetPose()
    want_h = a.GetHumanPoses()
    b = make()                       # a fresh world of the same shape and seed
    for wd in (a, b):
        wd.set_state(snap)
        for s in range(50, 100):
            got = wd.Step(cmds[s], cmd_mask=masks[s])
            assert np.array_equal(got.view(np.uint32), want[s - 50].view(np.uint32)), s
        pose, vel = wd.GetPose()
        assert np.array_equal(pose.view(np.uint32), want_pose.view(np.uint32))
        assert np.array_equal(vel.view(np.uint32), want_vel.view(np.uint32))
        assert np.array_equal(wd.GetHumanPoses().view(np.uint32), want_h.view(np.uint32))
    # a snapshot of another shape or seed is refused
    other = umrs.World(vmap, rps[:2], W, dt=float(DT), seed=77)
    with pytest.raises(_lib.UmrsError):
        other.set_state(snap)
    other_seed = umrs.World(vmap, rps, W, humans=hps, static_objects=1, dt=float(DT), seed=78)
    with pytest.raises(_lib.UmrsError):
        other_seed.set_state(snap)


def test_device_step_after_pipelined_tick_does_not_clobber_its_outputs():
    """a 4-slice pipelined host-buffer tick followed at once by (a) a device-resident step,
    (b) a tick without outputs (one launch), (c) a tick sliced differently -- each overwrites d_ranges
    and must first wait for EVERY copy of the earlier tick.  The earlier tick's host buffer must hold
    that tick's bits, not the next one's."""
    import torch
    W = 4096
    segs = umrs.map_cleanup(synth.floorplan_segments(256, extent=100.0, seed=5))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    poses = synth.sample_free_poses_fast(segs, W, 100.0, seed=11).reshape(W, 1, 3)
    ref_w = umrs.World(vmap, [car], W, seed=3, laser_noise=True, motion_noise=True, write_hit_idx=False)
    ref_w.SetPose(poses)
    cmds = [torch.from_numpy(synth.ackermann_commands(W, 1, s)).pin_memory() for s in range(8)]
    d_cmds = [c.cuda() for c in cmds]
    want = [ref_w.Step(cmds[s].numpy()).copy() for s in range(8)]
    n = ref_w.rays_per_world
    for variant in ("device", "no_outputs", "other_slicing"):
        wd = umrs.World(vmap, [car], W, seed=3, laser_noise=True, motion_noise=True, write_hit_idx=False)
        wd.SetPose(poses)
        bufs = [torch.empty((W, n), dtype=torch.float32).pin_memory() for _ in range(2)]
        for rep in range(3):                      # ticks 2*rep (async, outputs) and 2*rep + 1 (the intruder)
            s = 2 * rep
            bufs[0].fill_(-1.0)
            wd.step_async_raw(cmds[s].data_ptr(), None, bufs[0].data_ptr(), None)
            if variant == "device":
                wd.step_device(d_cmds[s + 1].data_ptr(), None, 0)
            elif variant == "no_outputs":
                wd.step_async_raw(cmds[s + 1].data_ptr(), None, None, None)
            else:
                with _lib.tuning(step_slices=7):
                    wd.step_async_raw(cmds[s + 1].data_ptr(), None, bufs[1].data_ptr(), None)
            wd.wait()
            torch.cuda.synchronize()
            assert np.array_equal(bufs[0].numpy().view(np.uint32), want[s].view(np.uint32)), (variant, s)
            if variant == "other_slicing":
                assert np.array_equal(bufs[1].numpy().view(np.uint32), want[s + 1].view(np.uint32)), (variant, s)
