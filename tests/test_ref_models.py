"""Pins the oracle's restated kinematic Step()s on the UNMODIFIED
reference models: ackermann_model.cpp, diff_drive_model.cpp, omnidirectional_model.cpp and
human_object.cpp are compiled in place into oracle/_ref/libumrs_ref.so behind data-only
stand-ins for roscpp / config_reader (oracle/shim, oracle/ref_models_bridge.cpp), driven with
the same command stream as the oracle (oracle/umrs_oracle.c), and compared BIT FOR BIT on
pose and velocity after every step.

The chain is then: reference model == oracle (here, CPU) == fused step kernel
(tests/test_gpu_world.py, GPU).  dt is the float the reference reads from CONFIG_FLOAT(DT)
(simulator.cpp:74,479), widened to the double Step() takes."""
import ctypes as C
import time

import numpy as np
import pytest

import ut_multirobot_sim_b200 as umrs
from tests import oracle_lib
from tests.oracle_lib import HumanState, RobotState

_F = C.POINTER(C.c_float)
DT = float(np.float32(0.025))


@pytest.fixture(scope="module")
def ref():
    r = oracle_lib.load_ref()
    if r is None or not r.has_models:
        pytest.skip("oracle/_ref/libumrs_ref.so without the model bridge (no /root/reference here)")
    return r


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


class Pair:
    """A reference model and the oracle state stepped side by side."""

    def __init__(self, oracle, ref, rp, pose):
        self.o, self.rp = oracle.lib, rp
        self.m = ref.robot_model(rp)
        pose = np.ascontiguousarray(pose, np.float32)
        self.m.SetPose(pose)
        self.s = RobotState()
        self.o.orc_robot_init(C.byref(self.s), pose.ctypes.data_as(_F))

    def command(self, cmd):
        c = np.zeros(3, np.float32)
        c[:len(cmd)] = cmd
        self.m.DriveCallback(c)
        self.o.orc_robot_command(C.byref(self.s), C.byref(self.rp), c.ctypes.data_as(_F))

    def step(self, n1=0.0, n2=0.0, dt=DT):
        self.m.Step(dt)
        self.o.orc_robot_step(C.byref(self.s), C.byref(self.rp), C.c_float(dt), C.c_float(n1), C.c_float(n2))

    def same(self):
        return (np.array_equal(_bits(self.m.GetPose()), _bits(self.s.pose[:])) and
                np.array_equal(_bits(self.m.GetVel()), _bits(self.s.vel[:])))

    def check(self, what=""):
        assert self.same(), "%s: reference pose %r vel %r, oracle pose %r vel %r" % (
            what, self.m.GetPose(), self.m.GetVel(), list(self.s.pose), list(self.s.vel))


def _random_commands(rng, n, scale=(3.0, 1.5, 2.0)):
    c = np.stack([rng.uniform(-s, s, n) for s in scale], 1).astype(np.float32)
    c[3::7, 1] = 0.0          # straight-line branch of the Ackermann step / pure translation
    c[5::11] = 0.0
    return c


@pytest.mark.parametrize("name", ["ut_automata", "ut_jackal", "bwibot", "cobot"])
@pytest.mark.parametrize("seed", [0, 1])
def test_step_matches_reference_model(oracle, ref, name, seed):
    """400 random commands, one Step each, from a random pose: every pose and velocity equal."""
    rng = np.random.default_rng(100 + seed)
    rp = umrs.default_robot(name)
    rp.ak_angular_error_rate = 0.0            # the draws are covered by the noise test below
    pair = Pair(oracle, ref, rp, [rng.uniform(-20, 20), rng.uniform(-20, 20), rng.uniform(-3.1, 3.1)])
    for t, cmd in enumerate(_random_commands(rng, 400)):
        pair.command(cmd)
        pair.step()
        pair.check("%s step %d" % (name, t))


@pytest.mark.parametrize("name", ["ut_automata", "ut_jackal", "cobot"])
def test_saturating_commands(oracle, ref, name):
    """Commands far beyond the speed / curvature / acceleration limits, held for many ticks
    (the clamps in DriveCallback and Step), then a hard reversal."""
    rp = umrs.default_robot(name)
    rp.ak_angular_error_rate = 0.0
    pair = Pair(oracle, ref, rp, [0.0, 0.0, 3.0])
    for cmd in ([50.0, 40.0, 30.0], [-50.0, -40.0, -30.0], [1e-4, 1e-4, 1e-4], [0.0, 1e6, 0.0]):
        for t in range(120):
            pair.command(cmd)
            pair.step()
            pair.check("%s cmd %r tick %d" % (name, cmd, t))


def test_ackermann_noise_draws(oracle, ref):
    """ackermann_model.cpp:87-89 draws angular_error_(rng_) twice per arc step: with the model's
    own stream (default-seeded default_random_engine, normal_distribution<float>) fed to the
    oracle's explicit noise arguments, first draw -> bias term, second -> rate term, the poses
    stay equal bit for bit -- with the shipped parameters (bias 0) and with a bias."""
    for bias in (0.0, 0.05):
        rng = np.random.default_rng(7)
        rp = umrs.default_robot("ut_automata")
        rp.ak_angular_error_bias = bias
        assert rp.ak_angular_error_rate > 0
        pair = Pair(oracle, ref, rp, [1.5, -2.25, 0.7])
        draws, k = ref.normal_stream(1000), 0
        max_curv = np.float32(1.0 / rp.ak_min_turn_radius)
        for t, cmd in enumerate(_random_commands(rng, 400)):
            pair.command(cmd)
            curv = min(max(cmd[1], -max_curv), max_curv)
            n1 = n2 = 0.0
            if abs(curv) >= np.float32(1.0 / 1e3):      # the arc branch is the one that draws
                n1, n2 = float(draws[k]), float(draws[k + 1])
                k += 2
            pair.step(n1, n2)
            pair.check("bias %g step %d" % (bias, t))
        assert k > 500


@pytest.mark.parametrize("name", ["ut_automata", "ut_jackal", "cobot"])
def test_command_watchdog(oracle, ref, name):
    """A command older than 0.1 s is dropped (ackermann_model.cpp:55-58, diff_drive_model.cpp:85-88,
    omnidirectional_model.cpp:85-90).  The reference ages it on the wall clock, the oracle in
    ticks of dt: 5 ticks of 0.025 s here == a 0.13 s pause there."""
    rp = umrs.default_robot(name)
    rp.ak_angular_error_rate = 0.0
    pair = Pair(oracle, ref, rp, [2.0, 1.0, -1.0])
    for _ in range(20):
        pair.command([1.5, 0.4, 0.3])
        pair.step()
    pair.check("moving")
    assert abs(pair.m.GetVel()[0]) > 0.1
    # no new command: the oracle keeps it while age * dt <= 0.1, i.e. for ticks of age 1..3 (4 ticks
    # of float(0.025) already exceed 0.1), and so does the reference model stepped without waiting ...
    for t in range(3):
        pair.step()
        pair.check("within the watchdog window, tick %d" % t)
    # ... and once the wall clock passes 0.1 s both have dropped it
    time.sleep(0.13)
    for t in range(60):
        pair.step()
        pair.check("after the watchdog, tick %d" % t)
    assert abs(pair.m.GetVel()[0]) < abs(1.5)


def test_never_commanded_stays_put(oracle, ref):
    for name in ("ut_automata", "ut_jackal", "cobot"):
        rp = umrs.default_robot(name)
        pair = Pair(oracle, ref, rp, [3.0, 4.0, 0.5])
        for _ in range(10):
            pair.step()
        pair.check(name)
        assert np.array_equal(pair.m.GetPose(), np.array([3.0, 4.0, 0.5], np.float32))


def test_non_finite_commands(oracle, ref, capfd):
    """The Ackermann callback ignores a non-finite command (ackermann_model.cpp:41-47); the
    omnidirectional one only prints and stores it anyway (omnidirectional_model.cpp:50-57)."""
    rp = umrs.default_robot("ut_automata")
    rp.ak_angular_error_rate = 0.0
    pair = Pair(oracle, ref, rp, [0.0, 0.0, 0.0])
    pair.command([1.0, 0.2])
    pair.step()
    for bad in ([np.nan, 0.1], [1.0, np.inf], [-np.inf, np.nan]):
        pair.command(bad)
        pair.step()
        pair.check("ackermann %r" % (bad,))
    assert np.isfinite(pair.m.GetPose()).all()
    capfd.readouterr()


def test_set_vel_and_pose_midway(oracle, ref):
    """SetPose / SetVel between steps (entity_base.h:49, robot_model.h:36) carry into the next Step."""
    rng = np.random.default_rng(5)
    for name in ("ut_automata", "cobot"):
        rp = umrs.default_robot(name)
        rp.ak_angular_error_rate = 0.0
        pair = Pair(oracle, ref, rp, [0.0, 0.0, 0.0])
        for t, cmd in enumerate(_random_commands(rng, 120)):
            if t % 30 == 10:
                p = np.array([rng.uniform(-9, 9), rng.uniform(-9, 9), rng.uniform(-3, 3)], np.float32)
                v = np.array([rng.uniform(-1, 1), rng.uniform(-1, 1) if name == "cobot" else 0.0,
                              rng.uniform(-1, 1) if name == "cobot" else 0.0], np.float32)
                pair.m.SetPose(p)
                pair.m.SetVel(v)
                pair.s.pose[:] = p.tolist()
                pair.s.vel[:] = v.tolist()
            pair.command(cmd)
            pair.step()
            pair.check("%s step %d" % (name, t))


def test_diff_drive_parameter_corners(oracle, ref):
    """Inverted commands, zero (= unlimited) acceleration limits, zero (= unlimited) speed caps:
    the branches of diff_drive_model.cpp:91-130,176-198."""
    rng = np.random.default_rng(11)
    for inv_x, inv_z, lim, cap in ((1, 0, 0.0, 0.0), (0, 1, 0.02, 0.0), (1, 1, 0.0, 1.0), (0, 0, 0.5, 0.3)):
        rp = umrs.default_robot("ut_jackal")
        rp.invert_linear_vel_cmds, rp.invert_angular_vel_cmds = inv_x, inv_z
        rp.linear_pos_accel_limit = rp.angular_neg_accel_limit = lim
        rp.max_linear_vel = rp.max_angular = cap
        pair = Pair(oracle, ref, rp, [1.0, 1.0, 1.0])
        for t, cmd in enumerate(_random_commands(rng, 150)):
            pair.command(cmd)
            pair.step()
            pair.check("inv %d%d lim %g cap %g step %d" % (inv_x, inv_z, lim, cap, t))


def test_other_tick_lengths(oracle, ref):
    for dt in (0.01, 0.05, 1.0 / 30.0):
        dtf = float(np.float32(dt))
        rng = np.random.default_rng(13)
        for name in ("ut_automata", "ut_jackal", "cobot"):
            rp = umrs.default_robot(name)
            rp.ak_angular_error_rate = 0.0
            pair = Pair(oracle, ref, rp, [0.5, -0.5, 2.0])
            for t, cmd in enumerate(_random_commands(rng, 100)):
                pair.command(cmd)
                pair.step(dt=dtf)
                pair.check("%s dt %g step %d" % (name, dt, t))


def test_human_matches_reference_object(oracle, ref):
    """HumanObject: template, Step (goal seeking, speed clip, goal swap / single shot) and the
    transformed lines, 20 random humans x 600 ticks."""
    rng = np.random.default_rng(3)
    for trial in range(20):
        hp = umrs.default_human()
        hp.start_x, hp.start_y, hp.start_theta = [float(x) for x in rng.uniform(-5, 5, 3)]
        hp.goal_x, hp.goal_y, hp.goal_theta = [float(x) for x in rng.uniform(-5, 5, 3)]
        hp.avg_speed = float(rng.uniform(0.5, 3))
        hp.max_speed = float(rng.uniform(0.5, 3))
        hp.mode = trial % 2
        hp.reach_goal_threshold = float(rng.uniform(0.05, 0.5))
        if trial % 5 == 4:
            hp.num_segments = int(rng.integers(3, 30))
        h = ref.human(hp)
        s = HumanState()
        oracle.lib.orc_human_init(C.byref(s), C.byref(hp))
        tmpl = h.GetTemplateLines()
        assert len(tmpl) == hp.num_segments
        assert np.array_equal(_bits(tmpl), _bits(oracle.human_template(hp)).reshape(-1, 4))
        rot = float(rng.uniform(-1, 1)) if trial % 3 == 0 else 0.0
        h.SetVel(0.0, 0.0, rot)
        s.rot_vel = rot
        for t in range(600):
            h.Step(DT)
            oracle.lib.orc_human_step(C.byref(s), C.byref(hp), C.c_float(DT))
            pose = np.array(s.pose[:], np.float32)
            assert np.array_equal(_bits(h.GetPose()), _bits(pose)), (trial, t)
            if t % 50 == 0:
                assert np.array_equal(_bits(h.GetLines()), _bits(oracle.transform_lines(tmpl, pose)).reshape(-1, 4))
