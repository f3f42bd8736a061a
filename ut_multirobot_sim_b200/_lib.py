"""ctypes binding of libumrs.so (the C-ABI in include/umrs.h).

The library is the product: if it is missing this module raises -- there is no
Python or CPU fallback for any compute entry point.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# UMRS_LIB: alternative build of the same library (kernel-variant A/B runs)
LIB_PATH = os.environ.get("UMRS_LIB") or os.path.join(_HERE, "libumrs.so")

UMRS_OK = 0
ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_UNSUPPORTED, ERR_NO_DEVICE, ERR_IO = 1, 2, 3, 4, 5, 6
ACKERMANN_DRIVE, OMNIDIRECTIONAL_DRIVE, DIFF_DRIVE = 0, 1, 2


class LaserParams(C.Structure):
    _fields_ = [("loc_x", C.c_float), ("loc_y", C.c_float),
                ("angle_min", C.c_float), ("angle_max", C.c_float),
                ("angle_increment", C.c_float),
                ("range_min", C.c_float), ("range_max", C.c_float),
                ("noise_stddev", C.c_float)]


class RobotParams(C.Structure):
    _fields_ = [("model_type", C.c_int32),
                ("ak_min_turn_radius", C.c_float), ("ak_max_speed", C.c_float),
                ("ak_max_accel", C.c_float), ("ak_angular_error_bias", C.c_float),
                ("ak_angular_error_rate", C.c_float),
                ("invert_linear_vel_cmds", C.c_int32), ("invert_angular_vel_cmds", C.c_int32),
                ("linear_pos_accel_limit", C.c_float), ("linear_neg_accel_limit", C.c_float),
                ("angular_pos_accel_limit", C.c_float), ("angular_neg_accel_limit", C.c_float),
                ("max_angular", C.c_float), ("max_linear_vel", C.c_float),
                ("co_max_accel", C.c_float), ("co_max_angle_accel", C.c_float),
                ("co_max_speed", C.c_float), ("co_max_angle_vel", C.c_float),
                ("car_length", C.c_float), ("car_width", C.c_float),
                ("rear_axle_offset", C.c_float), ("footprint_enabled", C.c_int32),
                ("laser", LaserParams)]


class HumanParams(C.Structure):
    _fields_ = [("radius", C.c_float), ("num_segments", C.c_int32),
                ("start_x", C.c_float), ("start_y", C.c_float), ("start_theta", C.c_float),
                ("goal_x", C.c_float), ("goal_y", C.c_float), ("goal_theta", C.c_float),
                ("max_speed", C.c_float), ("avg_speed", C.c_float),
                ("max_omega", C.c_float), ("avg_omega", C.c_float),
                ("reach_goal_threshold", C.c_float), ("mode", C.c_int32)]


class LaserScanHeader(C.Structure):
    _fields_ = [("angle_min", C.c_float), ("angle_max", C.c_float),
                ("angle_increment", C.c_float), ("time_increment", C.c_float),
                ("scan_time", C.c_float), ("range_min", C.c_float),
                ("range_max", C.c_float), ("num_rays", C.c_int32),
                ("ranges_offset", C.c_int32)]


class WorldDesc(C.Structure):
    _fields_ = [("num_worlds", C.c_int32), ("robots_per_world", C.c_int32),
                ("humans_per_world", C.c_int32), ("static_objects_per_world", C.c_int32),
                ("robots", C.POINTER(RobotParams)), ("humans", C.POINTER(HumanParams)),
                ("dt", C.c_float), ("seed", C.c_uint64),
                ("laser_noise_enabled", C.c_int32), ("motion_noise_enabled", C.c_int32),
                ("world_index_base", C.c_int32), ("write_hit_idx", C.c_int32),
                ("reference_render", C.c_int32)]


_P = C.c_void_p
_F = C.POINTER(C.c_float)
_I = C.POINTER(C.c_int32)

# name -> (restype, argtypes); every symbol include/umrs.h declares
SIGNATURES = {
    "umrs_version": (C.c_int, []),
    "umrs_status_string": (C.c_char_p, [C.c_int]),
    "umrs_last_cuda_error": (C.c_char_p, []),
    "umrs_kernel_launch_count": (C.c_uint64, []),
    "umrs_set_device": (C.c_int, [C.c_int32]),
    "umrs_get_device": (C.c_int, [_I]),
    "umrs_device_count": (C.c_int, [_I]),
    "umrs_model_type_from_string": (C.c_int, [C.c_char_p]),
    "umrs_laser_num_rays": (C.c_int, [C.POINTER(LaserParams)]),
    "umrs_default_laser": (None, [C.POINTER(LaserParams)]),
    "umrs_default_jackal_laser": (None, [C.POINTER(LaserParams)]),
    "umrs_default_ut_automata": (None, [C.POINTER(RobotParams)]),
    "umrs_default_ut_jackal": (None, [C.POINTER(RobotParams)]),
    "umrs_default_bwibot": (None, [C.POINTER(RobotParams)]),
    "umrs_default_cobot": (None, [C.POINTER(RobotParams)]),
    "umrs_default_human": (None, [C.POINTER(HumanParams)]),
    "umrs_map_create": (C.c_int, [_F, C.c_int32, C.POINTER(_P)]),
    "umrs_map_load_file": (C.c_int, [C.c_char_p, C.POINTER(_P)]),
    "umrs_map_destroy": (C.c_int, [_P]),
    "umrs_map_num_segments": (C.c_int, [_P, _I]),
    "umrs_map_get_segments": (C.c_int, [_P, _F, C.c_int32]),
    "umrs_map_cleanup": (C.c_int, [_F, C.c_int32, _F, C.c_int32, _I]),
    "umrs_map_enable_counters": (C.c_int, [_P, C.c_int32]),
    "umrs_map_read_counters": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "umrs_map_cleanup_ex": (C.c_int, [_F, C.c_int32, _F, C.c_int32, _I, C.c_int32]),
    "umrs_predicted_scan": (C.c_int, [_P, _F, C.c_int32, C.c_float, C.c_float, C.c_float,
                                      C.c_float, C.c_float, C.c_float, C.c_int32, _F, _I]),
    "umrs_scan_batch_reference_render": (C.c_int, [_P, _F, C.c_int32, _F, C.c_int32, C.c_float, C.c_int32, _F]),
    "umrs_scan_batch_reference_render_device": (C.c_int, [_P, _P, C.c_int32, _P, C.c_int32, C.c_float,
                                                          C.c_int32, _P, _P]),
    "umrs_scan_batch_reference_render_noalloc": (C.c_int, [_P, _P, C.c_int32, _P, C.c_int32, C.c_float,
                                                           C.c_int32, _P, _P]),
    "umrs_scan_batch": (C.c_int, [_P, _F, C.c_int32, _F, C.c_int32, C.c_float, C.c_int32, _F, _I]),
    "umrs_scan_batch_device": (C.c_int, [_P, _P, C.c_int32, _P, C.c_int32, C.c_float,
                                         C.c_int32, _P, _P, _P]),
    "umrs_map_set_scan_kernel": (C.c_int, [_P, C.c_int32]),
    "umrs_world_create": (C.c_int, [_P, C.POINTER(WorldDesc), C.POINTER(_P)]),
    "umrs_world_destroy": (C.c_int, [_P]),
    "umrs_world_rays_per_world": (C.c_int, [_P, _I]),
    "umrs_world_scan_header": (C.c_int, [_P, C.c_int32, C.POINTER(LaserScanHeader)]),
    "umrs_world_object_lines_per_world": (C.c_int, [_P, _I]),
    "umrs_world_set_robot_state": (C.c_int, [_P, _F, _F]),
    "umrs_world_get_robot_state": (C.c_int, [_P, _F, _F]),
    "umrs_world_set_humans": (C.c_int, [_P, _F, _F]),
    "umrs_world_get_human_poses": (C.c_int, [_P, _F]),
    "umrs_world_set_static_objects": (C.c_int, [_P, _F]),
    "umrs_world_get_object_lines": (C.c_int, [_P, _F]),
    "umrs_world_step": (C.c_int, [_P, _P, _P, _P, _P]),
    "umrs_world_step_async": (C.c_int, [_P, _P, _P, _P, _P]),
    "umrs_world_wait_outputs": (C.c_int, [_P, C.c_int32]),
    "umrs_world_wait": (C.c_int, [_P]),
    "umrs_world_step_device": (C.c_int, [_P, _P, _P, _P]),
    "umrs_world_device_buffers": (C.c_int, [_P, C.POINTER(_P), C.POINTER(_P), C.POINTER(_P)]),
    "umrs_world_get_fan_params": (C.c_int, [_P, _F]),
    "umrs_world_sync": (C.c_int, [_P]),
    "umrs_world_step_count": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "umrs_world_enable_counters": (C.c_int, [_P, C.c_int32]),
    "umrs_world_read_counters": (C.c_int, [_P, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                           C.POINTER(C.c_uint64)]),
    "umrs_world_read_counters_ex": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "umrs_world_read_phase_cycles": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "umrs_world_stats": (C.c_int, [_P, C.POINTER(C.c_double)]),
    "umrs_config_load": (C.c_int, [C.POINTER(C.c_char_p), C.c_int32, C.c_char_p, C.POINTER(_P)]),
    "umrs_config_destroy": (C.c_int, [_P]),
    "umrs_config_last_error": (C.c_char_p, []),
    "umrs_config_has": (C.c_int, [_P, C.c_char_p]),
    "umrs_config_get_double": (C.c_int, [_P, C.c_char_p, C.POINTER(C.c_double)]),
    "umrs_config_get_float": (C.c_int, [_P, C.c_char_p, _F]),
    "umrs_config_get_int": (C.c_int, [_P, C.c_char_p, _I]),
    "umrs_config_get_bool": (C.c_int, [_P, C.c_char_p, _I]),
    "umrs_config_get_string": (C.c_int, [_P, C.c_char_p, C.c_char_p, C.c_int32]),
    "umrs_config_list_length": (C.c_int, [_P, C.c_char_p, _I]),
    "umrs_config_get_list_string": (C.c_int, [_P, C.c_char_p, C.c_int32, C.c_char_p, C.c_int32]),
    "umrs_config_get_list_vector3": (C.c_int, [_P, C.c_char_p, C.c_int32, _F]),
    "umrs_config_laser_params": (C.c_int, [_P, C.POINTER(LaserParams), _I]),
    "umrs_config_robot_params": (C.c_int, [_P, C.c_char_p, C.POINTER(RobotParams), _I]),
    "umrs_config_human_params": (C.c_int, [_P, C.POINTER(HumanParams), _I]),
    "umrs_philox4x32": (None, [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                               C.POINTER(C.c_uint32)]),
    "umrs_world_state_size": (C.c_int, [_P, C.POINTER(C.c_uint64)]),
    "umrs_world_get_state": (C.c_int, [_P, _P, C.c_uint64]),
    "umrs_world_set_state": (C.c_int, [_P, _P, C.c_uint64]),
    "umrs_map_check": (C.c_int, [_P]),
    "umrs_world_check": (C.c_int, [_P]),
    "umrs_tuning_set": (C.c_int, [C.c_char_p, C.c_int32]),
    "umrs_tuning_get": (C.c_int, [C.c_char_p, _I]),
}

# Dev convenience for A/B scripts: UMRS_TUNE_<KEY>=<int> in the environment is handed to
# umrs_tuning_set once, when the library is loaded (the library itself never reads the environment).
TUNING_KEYS = ("cta_shape", "cap", "grid_from", "step_slices", "tma_stage", "render_cta_fans", "split_step")

_lib = None


def load():
    """Load libumrs.so and attach signatures; raises if it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libumrs.so is missing (%s): build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` or `make -C ut_multirobot_sim_b200/csrc`. There is no fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    for key in TUNING_KEYS:
        v = os.environ.get("UMRS_TUNE_" + key.upper())
        if v not in (None, ""):
            check(lib.umrs_tuning_set(key.encode(), int(v)), "umrs_tuning_set(%s)" % key)
    return lib


class tuning:
    """with tuning(cta_shape=3, cap=96): ...   -- sets the hooks, restores the previous values."""

    def __init__(self, **kv):
        self.kv, self.old = kv, {}

    def __enter__(self):
        lib = load()
        for k, v in self.kv.items():
            o = C.c_int32(0)
            check(lib.umrs_tuning_get(k.encode(), C.byref(o)), "umrs_tuning_get")
            self.old[k] = o.value
            check(lib.umrs_tuning_set(k.encode(), int(v)), "umrs_tuning_set(%s=%s)" % (k, v))
        return self

    def __exit__(self, *exc):
        lib = load()
        for k, v in self.old.items():
            lib.umrs_tuning_set(k.encode(), int(v))
        return False


class UmrsError(RuntimeError):
    def __init__(self, status, where):
        lib = load()
        msg = lib.umrs_status_string(status).decode()
        cuda = lib.umrs_last_cuda_error().decode()
        super().__init__("%s: %s%s" % (where, msg, (" [" + cuda + "]") if cuda else ""))
        self.status = status


def check(status, where):
    if status != UMRS_OK:
        raise UmrsError(status, where)
