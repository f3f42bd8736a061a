"""include/umrs_vector_map.hpp: the C++ host adapter with the reference's VectorMap
signatures.  CPU: it compiles as C++11 and fails loudly without a
device.  GPU: its GetPredictedScan equals the C-ABI call made from Python."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    exe = str(tmp_path / "adapter_demo")
    subprocess.run(["g++", "-std=c++11", "-Wall", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "cpp", "adapter_demo.cpp"), "-o", exe,
                    "-L", os.path.join(ROOT, "ut_multirobot_sim_b200"), "-lumrs",
                    "-Wl,-rpath," + os.path.join(ROOT, "ut_multirobot_sim_b200")], check=True)
    return exe


def _write_map(tmp_path):
    from ut_multirobot_sim_b200 import synth
    raw = synth.floorplan_segments(96, extent=40.0, seed=2)
    path = str(tmp_path / "m.vectormap.txt")
    synth.write_map(path, raw)
    return path, raw


def test_adapter_compiles_and_reports_missing_device(tmp_path):
    import torch
    exe = _build(tmp_path)
    path, _ = _write_map(tmp_path)
    r = subprocess.run([exe, path, "20", "20", "30", "-1", "1", "16"], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0
    else:
        assert r.returncode == 3 and "device" in r.stderr.lower()
    # a missing map is an error status, not the reference's exit(1)
    r = subprocess.run([exe, "/nonexistent.txt", "0", "0", "30", "-1", "1", "16"], capture_output=True, text=True)
    assert r.returncode == 3 and "I/O" in r.stderr


@pytest.mark.gpu
def test_adapter_scan_equals_capi_scan(tmp_path):
    import ut_multirobot_sim_b200 as umrs
    exe = _build(tmp_path)
    path, raw = _write_map(tmp_path)
    obj = [[18.0, 19.0, 18.0, 21.0], [22.5, 18.0, 22.5, 22.0]]
    args = [exe, path, "20.25", "20.5", "30", "-2.356", "2.356", "1081"] + [str(v) for o in obj for v in o]
    r = subprocess.run(args, capture_output=True, text=True, check=True)
    out = r.stdout.split("\n")
    vmap = umrs.VectorMap(file=path)
    assert out[0] == "lines %d" % len(vmap.lines)
    got = np.array([float(x) for x in out[1:] if x], np.float32)
    vmap.object_lines = np.array(obj, np.float32)
    want = vmap.GetPredictedScan((np.float32(20.25), np.float32(20.5)), 0.4, 30.0, np.float32(-2.356),
                                 np.float32(2.356), 1081)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert (got < 30.0).any()


def test_robot_model_adapter_reports_missing_device(tmp_path):
    import torch
    exe = _build(tmp_path)
    path, _ = _write_map(tmp_path)
    r = subprocess.run([exe, "model", path, "ut_automata", "20", "20", "0.3", "3", "1.0", "0.2", "0"],
                       capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0
    else:
        assert r.returncode == 3 and "device" in r.stderr.lower()


@pytest.mark.gpu
@pytest.mark.parametrize("kind,cmd", [("ut_automata", (1.0, 0.4, 0.0)), ("ut_jackal", (0.8, -0.5, 0.0)),
                                      ("cobot", (0.5, 0.3, 0.7))])
def test_robot_model_adapter_equals_the_batched_world(tmp_path, kind, cmd):
    """include/umrs_robot_model.hpp (EntityBase/RobotModel-shaped, one robot) stepped from C++ gives
    the poses, velocities and the LaserScan of a 1-world World driven from Python with the same
    commands and command gaps."""
    import ut_multirobot_sim_b200 as umrs
    exe = _build(tmp_path)
    path, raw = _write_map(tmp_path)
    steps = 9
    r = subprocess.run([exe, "model", path, kind, "20.25", "20.5", "0.3", str(steps)] + [repr(c) for c in cmd],
                       capture_output=True, text=True, check=True)
    out = r.stdout.strip().split("\n")
    poses = np.array([[float(x) for x in l.split()[1:]] for l in out[:steps]], np.float32)
    body, rays = int(out[steps].split()[1]), int(out[steps].split()[3])
    scan = np.array([float(x) for x in out[steps + 1:]], np.float32)
    vmap = umrs.VectorMap(file=path)
    rp = umrs.default_robot(kind)
    world = umrs.World(vmap, [rp], 1, dt=0.025, seed=0)
    world.SetPose(np.array([[[20.25, 20.5, 0.3]]], np.float32))
    assert rays == world.rays_per_world and body == world.object_lines_per_world
    for s in range(steps):
        mask = np.array([[0 if s % 3 == 2 else 1]], np.uint8)
        want_scan = world.Step(np.array([[cmd]], np.float32), cmd_mask=mask)
        p, v = world.GetPose()
        assert np.array_equal(poses[s, :3].view(np.uint32), p[0, 0].view(np.uint32)), (s, poses[s], p)
        assert np.array_equal(poses[s, 3:].view(np.uint32), v[0, 0].view(np.uint32)), (s, poses[s], v)
    assert np.array_equal(scan.view(np.uint32), want_scan[0].view(np.uint32))
