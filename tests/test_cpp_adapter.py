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
