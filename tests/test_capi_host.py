"""C-ABI library: loads, exports every symbol include/umrs.h declares, host-only
entry points work without a GPU, compute entry points fail loudly without one
(no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "umrs.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(umrs_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 45
    for n in names:
        assert hasattr(lib, n), "libumrs.so does not export %s" % n
    # the ctypes table binds all of them too (signature drift shows up here)
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_header_compiles_as_plain_c(tmp_path):
    import subprocess
    src = tmp_path / "t.c"
    src.write_text('#include "umrs.h"\nint main(void){umrs_laser_params l; umrs_default_laser(&l); '
                   'return umrs_laser_num_rays(&l) == 1081 ? 0 : 1;}\n')
    exe = tmp_path / "t"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src),
                    "-o", str(exe), "-L", os.path.join(ROOT, "ut_multirobot_sim_b200"), "-lumrs",
                    "-Wl,-rpath," + os.path.join(ROOT, "ut_multirobot_sim_b200")], check=True)
    assert subprocess.run([str(exe)]).returncode == 0


def test_version_and_status_strings():
    lib = _lib.load()
    assert lib.umrs_version() == 100
    assert lib.umrs_status_string(0) == b"ok"
    for s in range(1, 7):
        assert lib.umrs_status_string(s) not in (b"ok", b"unknown status")


def test_num_rays_matches_reference_formula():
    """simulator.cpp:402-404 in float32: 1081 for the default laser, 3601 for the jackal."""
    assert umrs.laser_num_rays(umrs.default_laser()) == 1081
    assert umrs.laser_num_rays(umrs.default_laser(jackal=True)) == 3601
    assert umrs.laser_num_rays(umrs.default_robot("bwibot").laser) == 1081


def test_model_type_strings():
    lib = _lib.load()
    assert lib.umrs_model_type_from_string(b"ACKERMANN_DRIVE") == umrs.ACKERMANN_DRIVE
    assert lib.umrs_model_type_from_string(b"OMNIDIRECTIONAL_DRIVE") == umrs.OMNIDIRECTIONAL_DRIVE
    assert lib.umrs_model_type_from_string(b"DIFF_DRIVE") == umrs.DIFF_DRIVE
    assert lib.umrs_model_type_from_string(b"HOVERCRAFT") == -1  # the reference exit(1)s (simulator.cpp:139-141)
    assert lib.umrs_model_type_from_string(None) == -1


def test_defaults_restate_lua_configs():
    car = umrs.default_robot("ut_automata")
    assert car.model_type == umrs.ACKERMANN_DRIVE
    assert (car.ak_min_turn_radius, car.ak_max_speed, car.ak_max_accel) == (
        np.float32(0.98), np.float32(1.2), np.float32(3.0))
    assert car.laser.loc_x == np.float32(0.2) and car.laser.range_max == 100.0
    jk = umrs.default_robot("ut_jackal")
    assert jk.model_type == umrs.DIFF_DRIVE and jk.laser.range_max == 150.0
    assert umrs.default_robot("cobot").model_type == umrs.OMNIDIRECTIONAL_DRIVE
    hp = umrs.default_human()
    assert hp.num_segments == 20 and hp.radius == np.float32(0.1) and hp.mode == 1


def test_philox_host_restatement():
    from tests.test_oracle import PHILOX_KAT
    lib = _lib.load()
    for ctr, key, want in PHILOX_KAT:
        out = (C.c_uint32 * 4)()
        lib.umrs_philox4x32((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
        assert list(out) == want


def test_map_cleanup_errors():
    lib = _lib.load()
    n = C.c_int32(0)
    assert lib.umrs_map_cleanup(None, -1, None, 0, C.byref(n)) == _lib.ERR_INVALID
    raw = np.array([[0, 0, 2, 0], [1, -1, 1, 1]], np.float32)
    out = np.zeros((1, 4), np.float32)
    # too small an output: INVALID, and the needed count is reported
    st = lib.umrs_map_cleanup(raw.ctypes.data_as(C.POINTER(C.c_float)), 2,
                              out.ctypes.data_as(C.POINTER(C.c_float)), 1, C.byref(n))
    assert st == _lib.ERR_INVALID and n.value == 3
    assert len(umrs.map_cleanup(raw)) == 3
    assert len(umrs.map_cleanup(np.zeros((0, 4), np.float32))) == 0


def test_map_file_io_error_is_a_status_not_exit():
    """The reference exit(1)s on a missing map (vector_map.cpp:399-402)."""
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.umrs_map_load_file(b"/nonexistent/map.txt", C.byref(h)) == _lib.ERR_IO
    assert not h


def _no_gpu():
    import torch
    return not torch.cuda.is_available()


@pytest.mark.skipif(not _no_gpu(), reason="checks the no-device error path")
def test_compute_entry_points_fail_loudly_without_a_device():
    """No CPU fallback: map creation (the first device touch) reports NO_DEVICE."""
    lib = _lib.load()
    segs = np.array([[0, 0, 1, 0]], np.float32)
    h = C.c_void_p()
    st = lib.umrs_map_create(segs.ctypes.data_as(C.POINTER(C.c_float)), 1, C.byref(h))
    assert st in (_lib.ERR_NO_DEVICE, _lib.ERR_CUDA)
    assert not h
    with pytest.raises(_lib.UmrsError):
        umrs.VectorMap(lines=segs)


def test_product_does_not_reference_the_oracle():
    """The product path must never route through oracle/ (or any CPU fallback)."""
    pkg = os.path.join(ROOT, "ut_multirobot_sim_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                text = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle_lib" not in text and "liboracle" not in text and "umrs_ref" not in text, f
                assert "oracle/" not in text.replace("tests/oracle", ""), f
