"""Host-side mirror of the reference interface for the scan path, over the C-ABI.

Names follow the reference (ut-amrl/ut_multirobot_sim):
  VectorMap.Load / .lines / .object_lines / .GetPredictedScan
      src/simulator/vector_map.h:42-83
  World.Step  = Simulator::update + publishLaser (simulator.cpp:474-509,393-426)
      batched over independent worlds.
Everything computes on the GPU through libumrs.so; numpy only carries buffers.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (ACKERMANN_DRIVE, DIFF_DRIVE, OMNIDIRECTIONAL_DRIVE, HumanParams,
                   LaserParams, LaserScanHeader, RobotParams, WorldDesc, check)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32)) if a is not None else None


def _vp(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _segs(a):
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 4)
    return a


def default_robot(kind):
    """'ut_automata' | 'ut_jackal' | 'bwibot' | 'cobot' -> RobotParams restating config/*.lua."""
    lib = _lib.load()
    p = RobotParams()
    getattr(lib, "umrs_default_" + kind)(C.byref(p))
    return p


def default_human():
    lib = _lib.load()
    p = HumanParams()
    lib.umrs_default_human(C.byref(p))
    return p


def default_laser(jackal=False):
    lib = _lib.load()
    p = LaserParams()
    (lib.umrs_default_jackal_laser if jackal else lib.umrs_default_laser)(C.byref(p))
    return p


def laser_num_rays(laser):
    return int(_lib.load().umrs_laser_num_rays(C.byref(laser)))


def map_cleanup(segs, algorithm=0):
    """VectorMap::Cleanup (vector_map.cpp:367-395) on raw segments, host-side.
    algorithm: 0 auto, 1 the reference's quadratic search, 2 uniform grid (same output)."""
    lib = _lib.load()
    segs = _segs(segs)
    cap = 4 * len(segs) + 16
    while True:
        out = np.empty((cap, 4), np.float32)
        n = C.c_int32(0)
        st = lib.umrs_map_cleanup_ex(_fp(segs), len(segs), _fp(out), cap, C.byref(n), algorithm)
        if st == _lib.UMRS_OK:
            return out[: n.value].copy()
        if n.value > cap:
            cap = n.value
            continue
        check(st, "umrs_map_cleanup")


class VectorMap:
    """vector_map::VectorMap (vector_map.h:42-83) backed by a device-resident map."""

    def __init__(self, lines=None, file=None, cleaned=True):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self.object_lines = np.zeros((0, 4), np.float32)  # vector_map.h:81
        self.file_name = ""
        if file is not None:
            self.Load(file)
        elif lines is not None:
            segs = _segs(lines)
            if not cleaned:
                segs = map_cleanup(segs)
            check(self._lib.umrs_map_create(_fp(segs), len(segs), C.byref(self._h)),
                  "umrs_map_create")

    def Load(self, file):
        """VectorMap::Load (vector_map.cpp:397-411): parse + Cleanup."""
        self._free()
        check(self._lib.umrs_map_load_file(str(file).encode(), C.byref(self._h)),
              "umrs_map_load_file")
        self.file_name = str(file)

    @property
    def lines(self):
        n = C.c_int32(0)
        check(self._lib.umrs_map_num_segments(self._h, C.byref(n)), "umrs_map_num_segments")
        out = np.empty((n.value, 4), np.float32)
        check(self._lib.umrs_map_get_segments(self._h, _fp(out), n.value), "umrs_map_get_segments")
        return out

    def GetPredictedScan(self, loc, range_min, range_max, angle_min, angle_max, num_rays,
                         return_index=False):
        """VectorMap::GetPredictedScan (vector_map.h:66-72); reads self.object_lines."""
        obj = _segs(self.object_lines)
        ranges = np.empty(num_rays, np.float32)
        idx = np.empty(num_rays, np.int32) if return_index else None
        check(self._lib.umrs_predicted_scan(
            self._h, _fp(obj) if len(obj) else None, len(obj), float(loc[0]), float(loc[1]),
            float(range_min), float(range_max), float(angle_min), float(angle_max),
            int(num_rays), _fp(ranges), _ip(idx)), "umrs_predicted_scan")
        return (ranges, idx) if return_index else ranges

    def ScanBatch(self, fans, range_max, num_rays, return_index=False):
        """Many fans {loc_x, loc_y, angle_min, angle_max} against map + object_lines."""
        fans = np.ascontiguousarray(fans, np.float32).reshape(-1, 4)
        obj = _segs(self.object_lines)
        ranges = np.empty((len(fans), num_rays), np.float32)
        idx = np.empty((len(fans), num_rays), np.int32) if return_index else None
        check(self._lib.umrs_scan_batch(
            self._h, _fp(obj) if len(obj) else None, len(obj), _fp(fans), len(fans),
            float(range_max), int(num_rays), _fp(ranges), _ip(idx)), "umrs_scan_batch")
        return (ranges, idx) if return_index else ranges

    def ScanBatchReferenceRender(self, fans, range_max, num_rays):
        """The same fans through the reference's own render (SceneRender/TrimOcclusion +
        first covering piece): bit-identical to the reference's GetPredictedScan output."""
        fans = np.ascontiguousarray(fans, np.float32).reshape(-1, 4)
        obj = _segs(self.object_lines)
        ranges = np.empty((len(fans), num_rays), np.float32)
        check(self._lib.umrs_scan_batch_reference_render(
            self._h, _fp(obj) if len(obj) else None, len(obj), _fp(fans), len(fans),
            float(range_max), int(num_rays), _fp(ranges)), "umrs_scan_batch_reference_render")
        return ranges

    def enable_counters(self, on=True):
        check(self._lib.umrs_map_enable_counters(self._h, int(on)), "umrs_map_enable_counters")

    def read_counters(self):
        """Work accounting of the last scan call (see World.read_counters)."""
        ex = (C.c_uint64 * 8)()
        check(self._lib.umrs_map_read_counters(self._h, ex), "umrs_map_read_counters")
        return {"ray_seg_tests": ex[0], "segs_prepared": ex[1], "chunk_seg_pairs": ex[2], "passes": ex[3],
                "tie_sweeps": ex[4], "spilled_entries": ex[5], "sweep_all_fans": ex[6],
                "source_segments": ex[7] & ((1 << 40) - 1), "source_rows": ex[7] >> 40}

    def _free(self):
        if self._h:
            self._lib.umrs_map_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self._free()
        except Exception:
            pass


class World:
    """W independent worlds in lockstep; one Step = Simulator::Run's update() +
    publishLaser() (simulator.cpp:531-541) for every world."""

    def __init__(self, vmap, robots, num_worlds, humans=(), static_objects=0, dt=0.025,
                 seed=0, laser_noise=False, motion_noise=False, world_index_base=0,
                 write_hit_idx=True, reference_render=False):
        self._lib = _lib.load()
        self._map = vmap
        self.W, self.R, self.H, self.O = int(num_worlds), len(robots), len(humans), int(static_objects)
        self._robots = (RobotParams * self.R)(*robots)
        self._humans = (HumanParams * max(self.H, 1))(*humans)
        d = WorldDesc()
        d.num_worlds, d.robots_per_world = self.W, self.R
        d.humans_per_world, d.static_objects_per_world = self.H, self.O
        d.robots = C.cast(self._robots, C.POINTER(RobotParams))
        d.humans = C.cast(self._humans, C.POINTER(HumanParams))
        d.dt, d.seed = dt, seed
        d.laser_noise_enabled, d.motion_noise_enabled = int(laser_noise), int(motion_noise)
        d.world_index_base, d.write_hit_idx = int(world_index_base), int(write_hit_idx)
        d.reference_render = int(reference_render)
        self._h = C.c_void_p()
        check(self._lib.umrs_world_create(vmap._h, C.byref(d), C.byref(self._h)), "umrs_world_create")
        n = C.c_int32(0)
        check(self._lib.umrs_world_rays_per_world(self._h, C.byref(n)), "rays_per_world")
        self.rays_per_world = n.value
        check(self._lib.umrs_world_object_lines_per_world(self._h, C.byref(n)), "object_lines")
        self.object_lines_per_world = n.value
        self.has_idx = bool(write_hit_idx)

    def scan_header(self, robot):
        h = LaserScanHeader()
        check(self._lib.umrs_world_scan_header(self._h, robot, C.byref(h)), "scan_header")
        return h

    def SetPose(self, pose_xyt, vel_xyt=None):
        pose = np.ascontiguousarray(pose_xyt, np.float32).reshape(self.W, self.R, 3)
        vel = None if vel_xyt is None else np.ascontiguousarray(vel_xyt, np.float32).reshape(self.W, self.R, 3)
        check(self._lib.umrs_world_set_robot_state(self._h, _fp(pose), _fp(vel)), "set_robot_state")

    def GetPose(self):
        pose = np.empty((self.W, self.R, 3), np.float32)
        vel = np.empty((self.W, self.R, 3), np.float32)
        check(self._lib.umrs_world_get_robot_state(self._h, _fp(pose), _fp(vel)), "get_robot_state")
        return pose, vel

    def SetHumans(self, start_xyt, goal_xyt):
        s = np.ascontiguousarray(start_xyt, np.float32).reshape(self.W, self.H, 3)
        g = np.ascontiguousarray(goal_xyt, np.float32).reshape(self.W, self.H, 3)
        check(self._lib.umrs_world_set_humans(self._h, _fp(s), _fp(g)), "set_humans")

    def GetHumanPoses(self):
        p = np.empty((self.W, self.H, 3), np.float32)
        check(self._lib.umrs_world_get_human_poses(self._h, _fp(p) if self.H else None), "get_human_poses")
        return p

    def SetStaticObjects(self, pose_xyt):
        p = np.ascontiguousarray(pose_xyt, np.float32).reshape(self.W, self.O, 3)
        check(self._lib.umrs_world_set_static_objects(self._h, _fp(p)), "set_static_objects")

    def GetLines(self):
        """Current object lines [W][L][4] (EntityBase::GetLines of every entity)."""
        o = np.empty((self.W, self.object_lines_per_world, 4), np.float32)
        check(self._lib.umrs_world_get_object_lines(self._h, _fp(o) if o.size else None), "get_object_lines")
        return o

    def fan_params(self):
        f = np.empty((self.W, self.R, 4), np.float32)
        check(self._lib.umrs_world_get_fan_params(self._h, _fp(f)), "get_fan_params")
        return f

    def Step(self, cmds, cmd_mask=None, ranges_out=None, idx_out=None, want_idx=False):
        """Host-buffer step: cmds [W][R][3] in, ranges [W][rays_per_world] out."""
        cmds = None if cmds is None else np.ascontiguousarray(cmds, np.float32).reshape(self.W, self.R, 3)
        mask = None if cmd_mask is None else np.ascontiguousarray(cmd_mask, np.uint8).reshape(self.W, self.R)
        if ranges_out is None:
            ranges_out = np.empty((self.W, self.rays_per_world), np.float32)
        if idx_out is None and want_idx:
            idx_out = np.empty((self.W, self.rays_per_world), np.int32)
        check(self._lib.umrs_world_step(self._h, _vp(cmds), _vp(mask), _vp(ranges_out), _vp(idx_out)),
              "umrs_world_step")
        return (ranges_out, idx_out) if want_idx else ranges_out

    def step_raw(self, cmds_ptr, mask_ptr, ranges_ptr, idx_ptr):
        """Host-buffer step on raw addresses (e.g. pinned torch tensors' data_ptr())."""
        check(self._lib.umrs_world_step(self._h, cmds_ptr, mask_ptr, ranges_ptr, idx_ptr), "umrs_world_step")

    def step_async_raw(self, cmds_ptr, mask_ptr, ranges_ptr, idx_ptr):
        """Pipelined host-buffer tick on raw addresses (pinned buffers); see wait_outputs / wait."""
        check(self._lib.umrs_world_step_async(self._h, cmds_ptr, mask_ptr, ranges_ptr, idx_ptr),
              "umrs_world_step_async")

    def wait_outputs(self, ticks_back=0):
        check(self._lib.umrs_world_wait_outputs(self._h, int(ticks_back)), "umrs_world_wait_outputs")

    def wait(self):
        check(self._lib.umrs_world_wait(self._h), "umrs_world_wait")

    def step_device(self, cmds_dev_ptr, mask_dev_ptr=None, stream=None):
        check(self._lib.umrs_world_step_device(self._h, cmds_dev_ptr, mask_dev_ptr, stream),
              "umrs_world_step_device")

    def device_buffers(self):
        r, i, p = C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self._lib.umrs_world_device_buffers(self._h, C.byref(r), C.byref(i), C.byref(p)), "device_buffers")
        return r.value, i.value, p.value

    def sync(self):
        check(self._lib.umrs_world_sync(self._h), "umrs_world_sync")

    def enable_counters(self, on=True):
        check(self._lib.umrs_world_enable_counters(self._h, int(on)), "enable_counters")

    def read_counters(self):
        a, b, c = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        check(self._lib.umrs_world_read_counters(self._h, C.byref(a), C.byref(b), C.byref(c)), "read_counters")
        ex = (C.c_uint64 * 8)()
        check(self._lib.umrs_world_read_counters_ex(self._h, ex), "read_counters_ex")
        return {"ray_seg_tests": a.value, "segs_prepared": b.value, "chunk_seg_pairs": c.value,
                "passes": ex[3], "tie_sweeps": ex[4], "spilled_entries": ex[5], "sweep_all_fans": ex[6],
                "source_segments": ex[7] & ((1 << 40) - 1), "source_rows": ex[7] >> 40}

    def get_state(self):
        """Snapshot for a bit-reproducible resume (umrs_world_get_state): bytes."""
        n = C.c_uint64(0)
        check(self._lib.umrs_world_state_size(self._h, C.byref(n)), "umrs_world_state_size")
        buf = (C.c_ubyte * n.value)()
        check(self._lib.umrs_world_get_state(self._h, buf, n.value), "umrs_world_get_state")
        return bytes(buf)

    def set_state(self, blob):
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        check(self._lib.umrs_world_set_state(self._h, buf, len(blob)), "umrs_world_set_state")

    def check(self):
        """Raises if a device-side condition was flagged since the last check (umrs_world_check)."""
        check(self._lib.umrs_world_check(self._h), "umrs_world_check")

    def phase_cycles(self):
        """Dev builds (-DUMRS_PHASE_TIMING): summed critical-path cycles per phase; zeros otherwise."""
        ex = (C.c_uint64 * 8)()
        check(self._lib.umrs_world_read_phase_cycles(self._h, ex), "read_phase_cycles")
        names = ("phase_a", "setup", "gather", "records", "buckets", "ray_loop", "finalise")
        return {k: int(ex[i]) for i, k in enumerate(names)}

    def stats(self):
        out = (C.c_double * 4)()
        check(self._lib.umrs_world_stats(self._h, out), "umrs_world_stats")
        return {"hits": out[0], "min": out[1], "max": out[2], "sum": out[3]}

    def __del__(self):
        try:
            if self._h:
                self._lib.umrs_world_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass


class Config:
    """Mirror of config_reader::ConfigReader (a list of Lua config files loaded into
    one state, simulator.cpp:106) over umrs_config_*: typed dotted lookups and the
    parameter structs the reference fills from its CONFIG_* variables."""

    def __init__(self, files, root_dir=None):
        lib = _lib.load()
        if isinstance(files, (str, bytes)):
            files = [files]
        arr = (C.c_char_p * len(files))(*[f.encode() if isinstance(f, str) else f for f in files])
        h = C.c_void_p()
        st = lib.umrs_config_load(arr, len(files), root_dir.encode() if root_dir else None, C.byref(h))
        if st != _lib.UMRS_OK:
            raise _lib.UmrsError(st, "umrs_config_load: " + lib.umrs_config_last_error().decode())
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            _lib.load().umrs_config_destroy(self._h)
            self._h = None

    def has(self, key):
        return bool(_lib.load().umrs_config_has(self._h, key.encode()))

    def _get(self, fn, ctype, key):
        out = ctype()
        _lib.check(getattr(_lib.load(), fn)(self._h, key.encode(), C.byref(out)), "%s(%r)" % (fn, key))
        return out.value

    def get_double(self, key):
        return self._get("umrs_config_get_double", C.c_double, key)

    def get_float(self, key):
        return self._get("umrs_config_get_float", C.c_float, key)

    def get_int(self, key):
        return self._get("umrs_config_get_int", C.c_int32, key)

    def get_bool(self, key):
        return bool(self._get("umrs_config_get_bool", C.c_int32, key))

    def get_string(self, key):
        buf = C.create_string_buffer(4096)
        _lib.check(_lib.load().umrs_config_get_string(self._h, key.encode(), buf, 4096),
                   "umrs_config_get_string(%r)" % key)
        return buf.value.decode()

    def list_length(self, key):
        return self._get("umrs_config_list_length", C.c_int32, key)

    def get_string_list(self, key):
        out = []
        for i in range(self.list_length(key)):
            buf = C.create_string_buffer(4096)
            _lib.check(_lib.load().umrs_config_get_list_string(self._h, key.encode(), i, buf, 4096),
                       "umrs_config_get_list_string(%r)" % key)
            out.append(buf.value.decode())
        return out

    def get_vector3_list(self, key):
        n = self.list_length(key)
        out = np.zeros((n, 3), np.float32)
        for i in range(n):
            _lib.check(_lib.load().umrs_config_get_list_vector3(
                self._h, key.encode(), i, out[i].ctypes.data_as(C.POINTER(C.c_float))),
                "umrs_config_get_list_vector3(%r)" % key)
        return out

    def robot_params(self, robot_type):
        """(RobotParams, number of keys the files left undefined)."""
        rp, miss = _lib.RobotParams(), C.c_int32()
        _lib.check(_lib.load().umrs_config_robot_params(self._h, robot_type.encode(), C.byref(rp),
                                                        C.byref(miss)), "umrs_config_robot_params")
        return rp, miss.value

    def human_params(self):
        hp, miss = _lib.HumanParams(), C.c_int32()
        _lib.check(_lib.load().umrs_config_human_params(self._h, C.byref(hp), C.byref(miss)),
                   "umrs_config_human_params")
        return hp, miss.value

    def laser_params(self):
        lp, miss = _lib.LaserParams(), C.c_int32()
        _lib.check(_lib.load().umrs_config_laser_params(self._h, C.byref(lp), C.byref(miss)),
                   "umrs_config_laser_params")
        return lp, miss.value
