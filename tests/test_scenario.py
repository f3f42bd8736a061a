"""Scenario generator parity: ut_multirobot_sim_b200.scenario against the
files the UNMODIFIED reference scripts/generate_config.py wrote for tests/golden/scenario/crowd.yml
(tests/golden/scenario, made by tests/golden/make_scenario_golden.py), the round trip through
the config evaluator, and the batched world descriptors (seeding, sharding); one GPU test
steps a scenario batch and checks humans and scans against the oracle."""
import ctypes as C
import os

import numpy as np
import pytest

from ut_multirobot_sim_b200 import scenario

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scenario")
EXAMPLE = os.path.join(GOLD, "crowd.yml")


def _tree(root):
    out = {}
    for d, _, files in os.walk(root):
        for f in files:
            p = os.path.join(d, f)
            out[os.path.relpath(p, root)] = open(p, "rb").read()
    return out


def test_generated_files_equal_the_reference_scripts_output(tmp_path):
    init = scenario.generate_config(EXAMPLE, str(tmp_path))
    want = _tree(os.path.join(GOLD, "config"))
    got = _tree(str(tmp_path / "config"))
    assert sorted(got) == sorted(want) and len(want) == 7
    for name in want:
        assert got[name] == want[name], name          # byte for byte, "\n\r" separators included
    assert init.endswith("config/human_crowd_scenario_configs/corridor_crowd/init_config.lua")


def test_yaml_scalars_are_formatted_like_the_reference(tmp_path):
    """The reference str()-formats the YAML scalars: an int stays '12', a float '-4.5'."""
    scn = scenario.load_yaml(EXAMPLE)
    assert scn.human_start_goal_positions[1][1][0] == 12 and isinstance(scn.human_start_goal_positions[1][1][0], int)
    assert scn.start_y == -11 and isinstance(scn.start_y, int)
    scenario.generate_config(scn, str(tmp_path))
    txt = open(tmp_path / "config/human_crowd_scenario_configs/corridor_crowd/human/human_config_1.lua").read()
    assert "hu_goal_x = 12\n" in txt and "hu_start_y = -4.5\n" in txt
    assert "start_y = -11\n" in open(tmp_path / "config/human_crowd_scenario_configs/corridor_crowd/init_config.lua").read()


def test_missing_keys_are_reported(tmp_path):
    p = tmp_path / "bad.yml"
    p.write_text('prefix: "x"\nmap_name: "m.txt"\n')
    with pytest.raises(KeyError):
        scenario.load_yaml(str(p))
    p.write_text('prefix: "x"\nmap_name: "m"\nstart_x: 0\nstart_y: 0\nstart_angle: 0\n'
                 'human_start_goal_positions: [[[1, 2]]]\n')
    with pytest.raises(ValueError):
        scenario.load_yaml(str(p))


def test_generated_files_load_through_the_config_evaluator(tmp_path):
    """What Simulator::init/loadObject read back (simulator.cpp:93-97,230-233): the golden files
    written by the reference and ours give the same parameters through umrs_config_*."""
    scn = scenario.load_yaml(EXAMPLE)
    ours = scenario.load_scenario_configs(scenario.generate_config(scn, str(tmp_path)), str(tmp_path))
    gold = scenario.load_scenario_configs(
        os.path.join(GOLD, "config/human_crowd_scenario_configs/corridor_crowd/init_config.lua"), GOLD)
    for ls in (ours, gold):
        assert ls.map_name == "maps/AHG2.txt"
        assert np.array_equal(ls.start_pose, np.array([4.25, -11, 1.5707963], np.float32))
        assert len(ls.humans) == 6 and ls.human_files[3].endswith("human/human_config_3.lua")
        for hp, (s, g) in zip(ls.humans, scn.human_start_goal_positions):
            assert (hp.start_x, hp.start_y) == (np.float32(s[0]), np.float32(s[1]))
            assert (hp.goal_x, hp.goal_y) == (np.float32(g[0]), np.float32(g[1]))
            assert hp.mode == 1 and hp.num_segments == 20 and hp.radius == np.float32(0.1)
            assert hp.max_speed == np.float32(1.5) and hp.avg_speed == np.float32(1.0)
            assert hp.reach_goal_threshold == np.float32(0.3)
    a = scenario.world_batch(ours, 3, seed=5)
    b = scenario.world_batch(scn, 3, seed=5)
    assert np.array_equal(a.human_start, b.human_start) and np.array_equal(a.robot_pose, b.robot_pose)


def test_world_batch_seeding_and_sharding(tmp_path):
    scn = scenario.load_yaml(EXAMPLE)
    full = scenario.world_batch(scn, 64, seed=7)
    assert full.robot_pose.shape == (64, 1, 3) and full.human_start.shape == (64, 6, 3)
    # world 0 is the scenario itself
    assert np.array_equal(full.robot_pose[0, 0], np.array([4.25, -11, 1.5707963], np.float32))
    assert np.array_equal(full.human_goal[0, 2, :2], np.array([2.5, -14.0], np.float32))
    # the others are perturbed within the stated bounds, and differ from each other
    d = full.robot_pose[1:, 0, :2] - full.robot_pose[0, 0, :2]
    assert (np.abs(d) <= 0.5 + 1e-5).all() and np.abs(d).min() > 0
    assert len({tuple(p) for p in full.robot_pose[:, 0]}) == 64
    assert (np.abs(full.human_start[1:, :, :2] - full.human_start[0, :, :2]) <= 0.5 + 1e-5).all()
    # deterministic, and a shard of the job holds exactly the worlds of the unsharded job
    again = scenario.world_batch(scn, 64, seed=7)
    assert np.array_equal(full.human_goal, again.human_goal)
    lo = scenario.world_batch(scn, 32, seed=7, world_index_base=0)
    hi = scenario.world_batch(scn, 32, seed=7, world_index_base=32)
    assert np.array_equal(np.concatenate([lo.robot_pose, hi.robot_pose]), full.robot_pose)
    assert np.array_equal(np.concatenate([lo.human_start, hi.human_start]), full.human_start)
    assert not np.array_equal(scenario.world_batch(scn, 64, seed=8).robot_pose[1:], full.robot_pose[1:])
    # file round trip + the command line
    full.save(str(tmp_path / "b.npz"))
    back = scenario.WorldBatch.load(str(tmp_path / "b.npz"))
    assert back.map_name == "maps/AHG2.txt" and np.array_equal(back.human_goal, full.human_goal)
    scenario.main(["--config-file", EXAMPLE, "--out-root", str(tmp_path), "--worlds", "8", "--seed", "7"])
    cli = scenario.WorldBatch.load(
        str(tmp_path / "config/human_crowd_scenario_configs/corridor_crowd/world_batch.npz"))
    assert np.array_equal(cli.robot_pose, full.robot_pose[:8])


@pytest.mark.gpu
def test_scenario_batch_steps_on_the_gpu_like_the_oracle(oracle, tmp_path):
    """16 variants of the golden scenario (its map file is not in the repo: a 30 x 30 m hall with
    pillars stands in), one ut_automata car + the 6 walking humans per world, 60 steps:
    human poses follow the restated HumanObject::Step, every scan equals the tier-B oracle over
    map + that world's human footprints."""
    import ut_multirobot_sim_b200 as umrs
    from tests import oracle_lib
    from tests.test_gpu_scan import check_against_tierb
    ls = scenario.load_scenario_configs(scenario.generate_config(EXAMPLE, str(tmp_path)), str(tmp_path))
    walls = [[-13, -18, 17, -18], [17, -18, 17, 12], [17, 12, -13, 12], [-13, 12, -13, -18]]
    for cx, cy in [(-4, -9), (8, -9), (-4, 5), (8, 5), (14, -1)]:
        walls += [[cx - 1, cy - 1, cx + 1, cy - 1], [cx + 1, cy - 1, cx + 1, cy + 1],
                  [cx + 1, cy + 1, cx - 1, cy + 1], [cx - 1, cy + 1, cx - 1, cy - 1]]
    segs = umrs.map_cleanup(np.asarray(walls, np.float32))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    W, H = 16, len(ls.humans)
    batch = scenario.world_batch(ls, W, seed=3)
    world = scenario.make_world(batch, vmap, car, human=ls.humans[0], dt=0.025, seed=1)
    hstates = []
    for w in range(W):
        row = []
        for h in range(H):
            hp = umrs.default_human()
            C.memmove(C.byref(hp), C.byref(ls.humans[h]), C.sizeof(hp))
            hp.start_x, hp.start_y, hp.start_theta = batch.human_start[w, h]
            hp.goal_x, hp.goal_y, hp.goal_theta = batch.human_goal[w, h]
            s = oracle_lib.HumanState()
            oracle.lib.orc_human_init(C.byref(s), C.byref(hp))
            row.append((s, hp))
        hstates.append(row)
    n = umrs.laser_num_rays(car.laser)
    # object lines of a world: [humans][robot bodies]; a scan does not see its own body
    assert world.object_lines_per_world == 20 * H + (4 if car.footprint_enabled else 0)
    own = (20 * H, world.object_lines_per_world)
    seen_humans = 0
    rng = np.random.default_rng(0)
    for step in range(60):
        cmds = np.zeros((W, 1, 3), np.float32)
        cmds[..., 0] = rng.uniform(0, 1.2, (W, 1))
        cmds[..., 1] = rng.uniform(-0.5, 0.5, (W, 1))
        ranges, idx = world.Step(cmds, want_idx=True)
        ghp = world.GetHumanPoses()
        for w in range(W):
            for h in range(H):
                s, hp = hstates[w][h]
                oracle.lib.orc_human_step(C.byref(s), C.byref(hp), np.float32(0.025))
                assert np.allclose(ghp[w, h], list(s.pose), atol=2e-5), (step, w, h)
        if step % 10 == 9:
            lines, fans = world.GetLines(), world.fan_params()
            for w in range(W):
                check_against_tierb(oracle, segs, lines[w], fans[w, 0], car.laser.range_max, n,
                                    ranges[w], idx[w], skip=own)
                seen_humans += int(((idx[w] >= len(segs)) & (idx[w] < len(segs) + 20 * H)).sum())
    assert seen_humans > 100       # the humans do occlude the scans
    # 60 steps at 1 m/s = 1.5 m: everybody moved, nobody reached a 20 m goal yet
    assert np.all(np.linalg.norm(ghp[..., :2] - batch.human_start[..., :2], axis=-1) > 1.0)
