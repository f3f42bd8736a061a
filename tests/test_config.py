"""Config ingestion without Lua: the C++ evaluator behind
umrs_config_* against (i) fixtures carrying the reference's keys and values, (ii) the
reference's own config/*.lua when the tree is present (this container only), and
(iii) the umrs_default_* structs the rest of the tests use.  Host-only: no GPU."""
import math
import os

import numpy as np
import pytest

import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import _lib

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "golden", "config")
REF = "/root/reference"


def _same_struct(a, b, skip=()):
    for name, _ in a._fields_:
        if name in skip:
            continue
        va, vb = getattr(a, name), getattr(b, name)
        if hasattr(va, "_fields_"):
            _same_struct(va, vb, skip)
        else:
            assert va == vb, "%s: %r != %r" % (name, va, vb)


def test_expression_subset():
    c = umrs.Config("config/exprs.lua", ROOT)
    # Lua precedence: ^ right-assoc above unary minus above * / %, then + -
    assert c.get_double("a") == 1 + 2 * 3 ** (2 ** 0.5) - (-4 / 8) % 3
    assert c.get_double("b") == 9
    assert c.get_string("c") == "n=2.5!"
    assert c.get_bool("d") is True
    assert c.list_length("e") == 4 and c.get_double("e.k.z") == -1.5e-3
    assert c.get_double("f") == 4 + 20 - 1.5e-3
    assert c.get_double("g") == 16 + 0.5 + 3.0
    assert c.get_double("h") == 7 - 1 + 4
    assert c.get_string("i") == "dflt"
    assert not c.has("j")
    assert c.get_bool("k") is True
    assert c.get_int("t.v.w") == 10
    # typed getters fail on a missing key or a wrong type, never default silently
    with pytest.raises(_lib.UmrsError):
        c.get_float("nope")
    with pytest.raises(_lib.UmrsError):
        c.get_float("c")
    with pytest.raises(_lib.UmrsError):
        c.get_string("a")


def test_unsupported_constructs_and_io_errors_are_statuses():
    with pytest.raises(_lib.UmrsError) as e:
        umrs.Config("config/bad_control_flow.lua", ROOT)
    assert e.value.status == _lib.ERR_UNSUPPORTED and "bad_control_flow.lua:2" in str(e.value)
    with pytest.raises(_lib.UmrsError) as e:
        umrs.Config("config/does_not_exist.lua", ROOT)
    assert e.value.status == _lib.ERR_IO


def test_sim_config_keys_and_lists():
    c = umrs.Config(["config/sim.lua"], ROOT)
    assert c.get_float("delta_t") == np.float32(0.025)
    assert c.get_bool("publish_tfs") and c.get_float("laser_loc.x") == np.float32(0.2)
    assert c.get_string_list("robot_types") == ["ACKERMANN_DRIVE", "DIFF_DRIVE"]
    assert c.get_string("map_name") == "maps/test.vectormap.txt"
    sp = c.get_vector3_list("start_poses")
    assert sp.shape == (2, 3)
    assert np.array_equal(sp[0], np.array([33.016267, 18.534157, 0], np.float32))
    assert np.array_equal(sp[1], np.array([1.5, -2, math.pi * 90 / 180], np.float32))
    assert c.get_string_list("human_config_list") == ["config/humans/h0.lua", "config/humans/h1.lua"]
    lp, miss = c.laser_params()
    assert miss == 0
    _same_struct(lp, umrs.default_laser())
    assert umrs.laser_num_rays(lp) == 1081


def test_robot_structs_equal_the_defaults():
    """ConfigReader({env_config, robot_config}) order: sim first, robot file overrides."""
    car, miss = umrs.Config(["config/car.lua"], ROOT).robot_params("ACKERMANN_DRIVE")   # require() pulls sim in
    assert miss == 0
    _same_struct(car, umrs.default_robot("ut_automata"), skip=(
        "footprint_enabled", "linear_pos_accel_limit", "linear_neg_accel_limit", "angular_pos_accel_limit",
        "angular_neg_accel_limit", "max_angular", "max_linear_vel", "co_max_accel", "co_max_angle_accel",
        "co_max_speed", "co_max_angle_vel"))
    jk, miss = umrs.Config(["config/sim.lua", "config/jackal.lua"], ROOT).robot_params("DIFF_DRIVE")
    assert miss == 0
    _same_struct(jk, umrs.default_robot("ut_jackal"), skip=(
        "footprint_enabled", "ak_min_turn_radius", "ak_max_speed", "ak_max_accel", "ak_angular_error_bias",
        "ak_angular_error_rate", "co_max_accel", "co_max_angle_accel", "co_max_speed", "co_max_angle_vel"))
    assert umrs.laser_num_rays(jk.laser) == 3601
    om, miss = umrs.Config(["config/sim.lua", "config/omni.lua"], ROOT).robot_params("OMNIDIRECTIONAL_DRIVE")
    assert miss == 0
    d = umrs.default_robot("cobot")
    for k in ("co_max_accel", "co_max_angle_accel", "co_max_speed", "co_max_angle_vel"):
        assert getattr(om, k) == getattr(d, k)
    # the jackal file alone leaves the noise key undefined: counted, value 0 like CONFIG_FLOAT
    jk2, miss = umrs.Config(["config/jackal.lua"], ROOT).robot_params("DIFF_DRIVE")
    assert miss == 1 and jk2.laser.noise_stddev == 0.0
    with pytest.raises(_lib.UmrsError):
        umrs.Config(["config/sim.lua"], ROOT).robot_params("HOVERCRAFT")


def test_human_struct_equals_the_default():
    hp, miss = umrs.Config("config/humans/h0.lua", ROOT).human_params()
    assert miss == 0
    _same_struct(hp, umrs.default_human())


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "config")), reason="reference tree not present")
def test_reference_config_files_parse_to_the_defaults():
    """The reference's own files (read in place, this container only)."""
    sim = "config/sim_config.lua"
    for f in sorted(os.listdir(os.path.join(REF, "config"))):
        if f.endswith(".lua"):
            # every shipped file is inside the subset (after the simulator config, as in
            # the reference: bwibot_config.lua uses Vector3 without defining it)
            umrs.Config([sim, "config/" + f], REF)
    pairs = (("config/ut_automata_config.lua", "ACKERMANN_DRIVE", "ut_automata"),
             ("config/ut_jackal_config.lua", "DIFF_DRIVE", "ut_jackal"),
             ("config/bwibot_config.lua", "DIFF_DRIVE", "bwibot"))
    groups = {"ACKERMANN_DRIVE": ("ak_",), "DIFF_DRIVE": ("invert", "linear_", "angular_", "max_angular", "max_linear")}
    for path, rtype, name in pairs:
        rp, miss = umrs.Config([sim, path], REF).robot_params(rtype)
        assert miss == 0
        d = umrs.default_robot(name)
        for field, _ in rp._fields_:
            if field in ("footprint_enabled",):
                continue
            if field in ("laser", "model_type", "car_length", "car_width", "rear_axle_offset") or \
                    field.startswith(groups[rtype]):
                va, vb = getattr(rp, field), getattr(d, field)
                if field == "laser":
                    _same_struct(va, vb)
                else:
                    assert va == vb, (path, field, va, vb)
    om, miss = umrs.Config([sim, "config/cobot_config.lua"], REF).robot_params("OMNIDIRECTIONAL_DRIVE")
    d = umrs.default_robot("cobot")
    assert miss == 0 and all(getattr(om, k) == getattr(d, k) for k in
                             ("co_max_accel", "co_max_angle_accel", "co_max_speed", "co_max_angle_vel"))
    hp, miss = umrs.Config("config/human_config.lua", REF).human_params()
    assert miss == 0
    _same_struct(hp, umrs.default_human())
    init = umrs.Config("config/default_init_config.lua", REF)
    assert init.get_string("map_name") == "UT_Campus/UT_Campus.vectormap.txt"
    assert np.array_equal(init.get_vector3_list("start_poses"),
                          np.array([[33.016267, 18.534157, 0]], np.float32))
    assert umrs.Config(sim, REF).get_string_list("robot_types") == ["DIFF_DRIVE"]
