"""ctypes access to the checkers: oracle/liboracle.so (plain-C restatement, "port")
and oracle/_ref/libumrs_ref.so (the unmodified reference scan, "reference").
TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's CPU-baseline legs use it."""
import ctypes as C
import os
import subprocess

import numpy as np

from ut_multirobot_sim_b200._lib import HumanParams, RobotParams

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_F = C.POINTER(C.c_float)
_I = C.POINTER(C.c_int32)


def _fp(a):
    return a.ctypes.data_as(_F) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(_I) if a is not None else None


def build(quiet=True):
    """(Re)build the checkers; the reference part only where /root/reference exists."""
    subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


class RobotState(C.Structure):
    _fields_ = [("pose", C.c_float * 3), ("vel", C.c_float * 3),
                ("linear_vel", C.c_double), ("angular_vel", C.c_double),
                ("cmd", C.c_float * 3), ("cmd_age", C.c_int32)]


class HumanState(C.Structure):
    _fields_ = [("pose", C.c_float * 3), ("start", C.c_float * 3), ("goal", C.c_float * 3),
                ("trans_vel", C.c_float * 2), ("rot_vel", C.c_double)]


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        lib.orc_scan_tierb.restype = None
        lib.orc_scan_tierb.argtypes = [_F, C.c_int32, _F, C.c_int32, C.c_int32, C.c_int32,
                                       C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                       C.c_int32, _F, _I]
        lib.orc_ray_angle.restype = C.c_float
        lib.orc_ray_angle.argtypes = [C.c_float, C.c_float, C.c_int32, C.c_int32]
        lib.orc_candidate_range.restype = C.c_float
        lib.orc_candidate_range.argtypes = [_F, C.c_float, C.c_float, C.c_float, C.c_float,
                                            C.c_int32, C.c_int32]
        lib.orc_laser_postprocess.restype = None
        lib.orc_laser_postprocess.argtypes = [_F, C.c_int32, C.c_float, C.c_float, _F]
        lib.orc_laser_origin.restype = None
        lib.orc_laser_origin.argtypes = [_F, C.c_float, C.c_float, _F]
        lib.orc_map_cleanup.restype = C.c_int32
        lib.orc_map_cleanup.argtypes = [_F, C.c_int32, _F, C.c_int32]
        lib.orc_robot_init.argtypes = [C.POINTER(RobotState), _F]
        lib.orc_robot_command.argtypes = [C.POINTER(RobotState), C.POINTER(RobotParams), _F]
        lib.orc_robot_step.argtypes = [C.POINTER(RobotState), C.POINTER(RobotParams), C.c_float,
                                       C.c_float, C.c_float]
        lib.orc_robots_tick.restype = None
        lib.orc_robots_tick.argtypes = [C.POINTER(RobotState), C.c_int32, C.POINTER(RobotParams), _F,
                                        C.c_float, _F]
        lib.orc_human_init.argtypes = [C.POINTER(HumanState), C.POINTER(HumanParams)]
        lib.orc_human_step.argtypes = [C.POINTER(HumanState), C.POINTER(HumanParams), C.c_float]
        lib.orc_human_template.argtypes = [C.POINTER(HumanParams), _F]
        lib.orc_transform_lines.argtypes = [_F, C.c_int32, _F, _F]
        lib.orc_square_template.argtypes = [_F]
        lib.orc_robot_template.argtypes = [C.POINTER(RobotParams), _F]
        lib.orc_philox4x32.argtypes = [C.POINTER(C.c_uint32)] * 3
        lib.orc_box_muller.argtypes = [C.c_uint32, C.c_uint32, _F]
        for n in ("orc_robot_init", "orc_robot_command", "orc_robot_step", "orc_human_init",
                  "orc_human_step", "orc_human_template", "orc_transform_lines",
                  "orc_square_template", "orc_robot_template", "orc_philox4x32", "orc_box_muller"):
            getattr(lib, n).restype = None

    def scan_tierb(self, map_segs, obj_segs, loc, range_max, angle_min, angle_max, num_rays,
                   skip=(0, 0)):
        m = np.ascontiguousarray(map_segs, np.float32).reshape(-1, 4)
        o = np.ascontiguousarray(obj_segs if obj_segs is not None else np.zeros((0, 4)), np.float32).reshape(-1, 4)
        ranges = np.empty(num_rays, np.float32)
        idx = np.empty(num_rays, np.int32)
        self.lib.orc_scan_tierb(_fp(m), len(m), _fp(o), len(o), skip[0], skip[1],
                                loc[0], loc[1], range_max, angle_min, angle_max, num_rays,
                                _fp(ranges), _ip(idx))
        return ranges, idx

    def candidate_range(self, seg, loc, angle_min, angle_max, num_rays, i):
        s = np.ascontiguousarray(seg, np.float32)
        return float(self.lib.orc_candidate_range(_fp(s), loc[0], loc[1], angle_min, angle_max, num_rays, i))

    def map_cleanup(self, segs):
        s = np.ascontiguousarray(segs, np.float32).reshape(-1, 4)
        cap = 4 * len(s) + 16
        out = np.empty((cap, 4), np.float32)
        n = self.lib.orc_map_cleanup(_fp(s), len(s), _fp(out), cap)
        assert n <= cap
        return out[:n].copy()

    def laser_postprocess(self, ranges, range_max, stddev, normals=None):
        r = np.ascontiguousarray(ranges, np.float32).copy()
        nrm = None if normals is None else np.ascontiguousarray(normals, np.float32)
        self.lib.orc_laser_postprocess(_fp(r), r.size, range_max, stddev, _fp(nrm))
        return r

    def laser_origin(self, pose, lx, ly):
        p = np.ascontiguousarray(pose, np.float32)
        out = np.empty(2, np.float32)
        self.lib.orc_laser_origin(_fp(p), lx, ly, _fp(out))
        return out

    def human_template(self, hp):
        out = np.empty((hp.num_segments, 4), np.float32)
        self.lib.orc_human_template(C.byref(hp), _fp(out))
        return out

    def square_template(self):
        out = np.empty((4, 4), np.float32)
        self.lib.orc_square_template(_fp(out))
        return out

    def robot_template(self, rp):
        out = np.empty((4, 4), np.float32)
        self.lib.orc_robot_template(C.byref(rp), _fp(out))
        return out

    def transform_lines(self, tmpl, pose):
        t = np.ascontiguousarray(tmpl, np.float32).reshape(-1, 4)
        p = np.ascontiguousarray(pose, np.float32)
        out = np.empty_like(t)
        self.lib.orc_transform_lines(_fp(t), len(t), _fp(p), _fp(out))
        return out

    def philox(self, ctr, key):
        c = (C.c_uint32 * 4)(*ctr)
        k = (C.c_uint32 * 2)(*key)
        o = (C.c_uint32 * 4)()
        self.lib.orc_philox4x32(c, k, o)
        return list(o)

    def box_muller(self, u0, u1):
        out = np.empty(2, np.float32)
        self.lib.orc_box_muller(u0, u1, _fp(out))
        return out


class Ref:
    """The reference's own VectorMap behind oracle/ref_bridge.cpp."""

    def __init__(self, lib):
        self.lib = lib
        lib.umrs_ref_map_load.restype = C.c_void_p
        lib.umrs_ref_map_load.argtypes = [C.c_char_p]
        lib.umrs_ref_map_from_segments.restype = C.c_void_p
        lib.umrs_ref_map_from_segments.argtypes = [_F, C.c_int, C.c_int]
        lib.umrs_ref_map_free.argtypes = [C.c_void_p]
        lib.umrs_ref_map_num_lines.argtypes = [C.c_void_p]
        lib.umrs_ref_map_get_lines.argtypes = [C.c_void_p, _F, C.c_int]
        lib.umrs_ref_map_set_object_lines.argtypes = [C.c_void_p, _F, C.c_int]
        lib.umrs_ref_scene_lines.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, _F, C.c_int]
        lib.umrs_ref_scene_render.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float,
                                              C.c_float, C.c_float, _F, C.c_int]
        lib.umrs_ref_predicted_scan.argtypes = [C.c_void_p] + [C.c_float] * 6 + [C.c_int, _F]
        lib.umrs_ref_predicted_scan.restype = None
        lib.umrs_ref_scan_batch.argtypes = [C.c_void_p, _F, C.c_int, _F, _I, C.c_float, C.c_float,
                                            C.c_int, C.c_int, _F]
        lib.umrs_ref_scan_batch.restype = None
        lib.umrs_ref_line_intersects.argtypes = [_F, _F]
        lib.umrs_ref_line_intersection.argtypes = [_F, _F, _F]
        lib.umrs_ref_ray_intersects.argtypes = [_F] + [C.c_float] * 4
        lib.umrs_ref_cross.restype = C.c_float
        lib.umrs_ref_cross.argtypes = [C.c_float] * 4
        lib.umrs_ref_angle_mod_f.restype = C.c_float
        lib.umrs_ref_angle_mod_f.argtypes = [C.c_float]
        lib.umrs_ref_angle_mod_d.restype = C.c_double
        lib.umrs_ref_angle_mod_d.argtypes = [C.c_double]
        lib.umrs_ref_angle_diff_f.restype = C.c_float
        lib.umrs_ref_angle_diff_f.argtypes = [C.c_float, C.c_float]
        lib.umrs_ref_heading_f.argtypes = [C.c_float, _F]
        lib.umrs_ref_unit_normal.argtypes = [_F, _F]
        lib.umrs_ref_apply_pose.argtypes = [_F, _F]
        lib.umrs_ref_short_term_object.argtypes = [_F, _F, _F]
        # the reference's kinematic models (oracle/ref_models_bridge.cpp)
        self.has_models = hasattr(lib, "umrs_ref_model_create")
        if self.has_models:
            lib.umrs_ref_config_set_float.argtypes = [C.c_char_p, C.c_float]
            lib.umrs_ref_config_set_int.argtypes = [C.c_char_p, C.c_int]
            lib.umrs_ref_config_set_bool.argtypes = [C.c_char_p, C.c_int]
            lib.umrs_ref_config_set_string.argtypes = [C.c_char_p, C.c_char_p]
            lib.umrs_ref_model_create.restype = C.c_void_p
            lib.umrs_ref_model_create.argtypes = [C.c_int]
            lib.umrs_ref_model_free.argtypes = [C.c_void_p]
            for n in ("set_pose", "get_pose", "set_vel", "get_vel", "command"):
                getattr(lib, "umrs_ref_model_" + n).argtypes = [C.c_void_p, _F]
                getattr(lib, "umrs_ref_model_" + n).restype = None
            lib.umrs_ref_model_step.argtypes = [C.c_void_p, C.c_double]
            lib.umrs_ref_model_step.restype = None
            lib.umrs_ref_normal_stream.argtypes = [C.c_int, _F]
            lib.umrs_ref_human_create.restype = C.c_void_p
            lib.umrs_ref_human_free.argtypes = [C.c_void_p]
            lib.umrs_ref_human_step.argtypes = [C.c_void_p, C.c_double]
            lib.umrs_ref_human_set_vel.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_double]
            lib.umrs_ref_human_set_pose.argtypes = [C.c_void_p, _F]
            lib.umrs_ref_human_get_pose.argtypes = [C.c_void_p, _F]
            lib.umrs_ref_human_lines.argtypes = [C.c_void_p, _F, C.c_int]
            lib.umrs_ref_human_template.argtypes = [C.c_void_p, _F, C.c_int]

    def configure_robot(self, rp):
        """Fill the ConfigReader slots the three models read with a umrs_robot_params."""
        L = self.lib
        for key, field in (("ak_min_turn_radius", "ak_min_turn_radius"), ("ak_max_accel", "ak_max_accel"),
                           ("ak_max_speed", "ak_max_speed"), ("ak_angular_error_bias", "ak_angular_error_bias"),
                           ("ak_angular_error_rate", "ak_angular_error_rate"),
                           ("linear_pos_accel_limit", "linear_pos_accel_limit"),
                           ("linear_neg_accel_limit", "linear_neg_accel_limit"),
                           ("angular_pos_accel_limit", "angular_pos_accel_limit"),
                           ("angular_neg_accel_limit", "angular_neg_accel_limit"),
                           ("max_angular", "max_angular"), ("max_linear_vel", "max_linear_vel"),
                           ("co_max_accel", "co_max_accel"), ("co_max_angle_accel", "co_max_angle_accel"),
                           ("co_max_speed", "co_max_speed"), ("co_max_angle_vel", "co_max_angle_vel")):
            L.umrs_ref_config_set_float(key.encode(), float(getattr(rp, field)))
        for key in ("linear_odom_scale", "angular_odom_scale", "co_base_radius"):
            L.umrs_ref_config_set_float(key.encode(), 1.0)
        for i in range(4):
            L.umrs_ref_config_set_float(b"co_w%d" % i, 0.7853981633974483 * (2 * i + 1))
        L.umrs_ref_config_set_bool(b"invert_linear_vel_cmds", int(rp.invert_linear_vel_cmds))
        L.umrs_ref_config_set_bool(b"invert_angular_vel_cmds", int(rp.invert_angular_vel_cmds))
        for key in ("ak_drive_callback_topic", "drive_callback_topic", "diff_drive_odom_topic",
                    "co_drive_callback_topic", "co_cobot_odom_topic"):
            L.umrs_ref_config_set_string(key.encode(), b"/" + key.encode())

    def robot_model(self, rp):
        self.configure_robot(rp)
        return RefModel(self, self.lib.umrs_ref_model_create(int(rp.model_type)))

    def human(self, hp):
        L = self.lib
        for f in ("radius", "num_segments", "start_x", "start_y", "start_theta", "goal_x", "goal_y",
                  "goal_theta", "max_speed", "avg_speed", "max_omega", "avg_omega", "reach_goal_threshold"):
            L.umrs_ref_config_set_float(("hu_" + f).encode(), float(getattr(hp, f)))
        L.umrs_ref_config_set_int(b"hu_mode", int(hp.mode))
        return RefHuman(self, L.umrs_ref_human_create())

    def normal_stream(self, n):
        out = np.empty(n, np.float32)
        self.lib.umrs_ref_normal_stream(n, _fp(out))
        return out

    def map_from_segments(self, segs, cleanup):
        s = np.ascontiguousarray(segs, np.float32).reshape(-1, 4)
        return RefMap(self, self.lib.umrs_ref_map_from_segments(_fp(s), len(s), int(cleanup)))

    def map_load(self, path):
        h = self.lib.umrs_ref_map_load(str(path).encode())
        assert h, "reference could not open %s" % path
        return RefMap(self, h)


class RefModel:
    """One of the reference's own robot models (MakeMotionModel, simulator.cpp:128-142)."""

    def __init__(self, ref, handle):
        self.ref, self.h = ref, C.c_void_p(handle)

    def _get(self, fn):
        out = np.empty(3, np.float32)
        fn(self.h, _fp(out))
        return out

    def SetPose(self, xyt):
        self.ref.lib.umrs_ref_model_set_pose(self.h, _fp(np.ascontiguousarray(xyt, np.float32)))

    def SetVel(self, xyt):
        self.ref.lib.umrs_ref_model_set_vel(self.h, _fp(np.ascontiguousarray(xyt, np.float32)))

    def GetPose(self):
        return self._get(self.ref.lib.umrs_ref_model_get_pose)

    def GetVel(self):
        return self._get(self.ref.lib.umrs_ref_model_get_vel)

    def DriveCallback(self, cmd):
        c = np.zeros(3, np.float32)
        c[:len(cmd)] = cmd
        self.ref.lib.umrs_ref_model_command(self.h, _fp(c))

    def Step(self, dt):
        self.ref.lib.umrs_ref_model_step(self.h, float(dt))

    def __del__(self):
        try:
            self.ref.lib.umrs_ref_model_free(self.h)
        except Exception:
            pass


class RefHuman:
    """The reference's HumanObject (human_object.cpp)."""

    def __init__(self, ref, handle):
        self.ref, self.h = ref, C.c_void_p(handle)

    def Step(self, dt):
        self.ref.lib.umrs_ref_human_step(self.h, float(dt))

    def SetVel(self, vx, vy, rot):
        self.ref.lib.umrs_ref_human_set_vel(self.h, float(vx), float(vy), float(rot))

    def SetPose(self, xyt):
        self.ref.lib.umrs_ref_human_set_pose(self.h, _fp(np.ascontiguousarray(xyt, np.float32)))

    def GetPose(self):
        out = np.empty(3, np.float32)
        self.ref.lib.umrs_ref_human_get_pose(self.h, _fp(out))
        return out

    def _lines(self, fn):
        out = np.empty((64, 4), np.float32)
        n = fn(self.h, _fp(out), 64)
        return out[:n].copy()

    def GetLines(self):
        return self._lines(self.ref.lib.umrs_ref_human_lines)

    def GetTemplateLines(self):
        return self._lines(self.ref.lib.umrs_ref_human_template)

    def __del__(self):
        try:
            self.ref.lib.umrs_ref_human_free(self.h)
        except Exception:
            pass


class RefMap:
    def __init__(self, ref, handle):
        self.ref, self.h = ref, C.c_void_p(handle)

    @property
    def lines(self):
        n = self.ref.lib.umrs_ref_map_num_lines(self.h)
        out = np.empty((n, 4), np.float32)
        self.ref.lib.umrs_ref_map_get_lines(self.h, _fp(out), n)
        return out

    def set_object_lines(self, segs):
        s = np.ascontiguousarray(segs, np.float32).reshape(-1, 4)
        self.ref.lib.umrs_ref_map_set_object_lines(self.h, _fp(s), len(s))

    def scene_lines(self, loc, max_range, cap=200000):
        out = np.empty((cap, 4), np.float32)
        n = self.ref.lib.umrs_ref_scene_lines(self.h, loc[0], loc[1], max_range, _fp(out), cap)
        return out[:n].copy()

    def predicted_scan(self, loc, range_min, range_max, angle_min, angle_max, num_rays):
        out = np.empty(num_rays, np.float32)
        self.ref.lib.umrs_ref_predicted_scan(self.h, loc[0], loc[1], range_min, range_max,
                                             angle_min, angle_max, num_rays, _fp(out))
        return out

    def scan_batch(self, fans, range_min, range_max, num_rays, threads=1, obj=None, obj_offset=None):
        f = np.ascontiguousarray(fans, np.float32).reshape(-1, 4)
        out = np.empty((len(f), num_rays), np.float32)
        o = None if obj is None else np.ascontiguousarray(obj, np.float32)
        oo = None if obj_offset is None else np.ascontiguousarray(obj_offset, np.int32)
        self.ref.lib.umrs_ref_scan_batch(self.h, _fp(f), len(f), _fp(o), _ip(oo), range_min,
                                         range_max, num_rays, threads, _fp(out))
        return out

    def __del__(self):
        try:
            self.ref.lib.umrs_ref_map_free(self.h)
        except Exception:
            pass


_oracle = None
_ref = None


def load_oracle():
    global _oracle
    if _oracle is None:
        p = os.path.join(ORACLE_DIR, "liboracle.so")
        srcs = [os.path.join(ORACLE_DIR, f) for f in ("umrs_oracle.c", "umrs_oracle.h")]
        if not os.path.exists(p) or any(os.path.getmtime(f) > os.path.getmtime(p) for f in srcs):
            build()
        _oracle = Oracle(C.CDLL(p))
    return _oracle


def load_ref():
    global _ref
    if _ref is None:
        p = os.path.join(ORACLE_DIR, "_ref", "libumrs_ref.so")
        have_ref = os.path.isdir("/root/reference/src/simulator")
        srcs = [os.path.join(ORACLE_DIR, f) for f in ("ref_bridge.cpp", "ref_models_bridge.cpp")]
        if have_ref and (not os.path.exists(p) or any(os.path.getmtime(f) > os.path.getmtime(p) for f in srcs)):
            build()
        if not os.path.exists(p):
            return None
        _ref = Ref(C.CDLL(p))
    return _ref
