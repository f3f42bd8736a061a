#!/usr/bin/env python
"""bench.py -- lidar rays/s (+ env-steps/s) of the fused world step on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  torchrun ... bench.py --gpus N ...          (one rank per GPU, N > 1)

Workload (BASELINE.json configs):
  every N : BASELINE config 5's per-GPU share -- 8192 worlds x 1 Ackermann car x 1081
            rays per GPU, contiguous world blocks (65,536 worlds at N = 8; weak scaling);
            --worlds 4096 gives config 3
against one shared seeded floor plan of --map-segments (default 1024) cleaned
segments.  A "step" is one umrs_world_step over all worlds of the rank.

Prints ONE JSON line (rank 0).  `value` = rays/s with inputs resident in HBM;
`e2e` = the same through the host-buffer C-ABI call (pinned host buffers,
H2D commands + D2H ranges inside the timed region).  `--impl reference` times the
reference's own CPU scan (oracle/_ref, the unmodified vector_map.cpp) on the
host cores instead.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_RAYS = 1081
WORLDS_PER_GPU = 8192  # BASELINE config 5: 65,536 worlds over 8 GPUs (weak scaling at every N)
EXTENT = 200.0
FLOPS_PER_TEST = 12.0  # nominal cost of one ray-segment candidate test
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # 74.4: SMs x lanes x FMA x max SM clock
FACTS_FILE = "r2_kernel_facts.json"   # per-launch ncu facts of k_world_step on the bench workload (scripts/ncu_facts.py)


_JSON_FD = None


def emit(line):
    """The result line, on the real stdout."""
    data = (line + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line + "\n")
    else:
        os.write(_JSON_FD, data)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--map-segments", type=int, default=1024)
    ap.add_argument("--worlds", type=int, default=0, help="override total worlds")
    ap.add_argument("--cpu-sample-fans", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--no-legs", action="store_true", help="skip the config 1-4 / reference-render legs")
    ap.add_argument("--sustained-seconds", type=float, default=2.0)
    return ap.parse_args()


def total_worlds(args):
    if args.worlds:
        return args.worlds
    return WORLDS_PER_GPU * max(1, args.gpus)


def workload_name(args):
    w = total_worlds(args)
    return ("config5 shard (65,536 worlds over 8 GPUs = %d per GPU; config3 is --worlds 4096): %d worlds x 1 "
            "Ackermann x %d rays on %d GPU(s), entity step + scan per tick, floor plan %d segs" % (
                WORLDS_PER_GPU, w, N_RAYS, max(1, args.gpus), args.map_segments))


def config_dict(args, n_segs):
    """The workload, stated identically by both arms."""
    return {"workload": workload_name(args), "worlds_per_gpu": total_worlds(args) // max(1, args.gpus),
            "map_segments": int(n_segs), "laser_noise": True, "motion_noise": True,
            "l2": "b200 arm: %s; reference arm: CPU, not applicable" % (
                "not flushed" if args.no_flush else "flushed between timed steps (256 MiB memset outside the event pairs)"),
            "timing": "b200 arm: sum of per-step CUDA-event pairs on the launch stream, max over ranks; "
                      "reference arm: wall clock around the steps"}


def make_inputs(args, rank, world_size):
    """The whole job's poses and command sets (every rank builds the same arrays and takes its block)."""
    import ut_multirobot_sim_b200 as umrs
    from ut_multirobot_sim_b200 import synth
    from ut_multirobot_sim_b200.shard import shard_worlds
    W_total = total_worlds(args)
    base, per = shard_worlds(W_total, rank, world_size)
    segs = umrs.map_cleanup(synth.floorplan_segments(args.map_segments, extent=EXTENT, seed=1234))
    poses = synth.sample_free_poses_fast(segs, W_total, EXTENT, seed=11)
    cmds = [synth.ackermann_commands(W_total, 1, s, seed=42) for s in range(8)]
    return segs, poses, cmds, per, base


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag = index, False
        self.sm, self.reasons, self.sm_max = [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.sm_max = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            }
            while not self.stop_flag:
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for bit, name in names.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    pass
                time.sleep(0.01)
        except Exception as e:  # NVML unavailable: report that, never fake a clock
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons),
                "samples": len(self.sm)}


def cpu_baseline(segs, fans, range_max, threads, budget_s):
    """oracle/_ref (the unmodified reference scan, vector_map.cpp compiled in place) on
    `threads` host cores: whole passes over `fans` until about budget_s of wall time."""
    from tests import oracle_lib
    ref = oracle_lib.load_ref()
    if ref is None:
        return None
    rm = ref.map_from_segments(segs, False)
    f = np.ascontiguousarray(fans, np.float32)
    rm.scan_batch(f[:max(1, threads)], 0.4, range_max, N_RAYS, threads=threads)  # warm
    t0 = time.perf_counter()
    reps = 0
    while True:
        rm.scan_batch(f, 0.4, range_max, N_RAYS, threads=threads)
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or reps >= 1000:
            break
    return {"value": reps * len(f) * N_RAYS / dt, "unit": "rays/s", "cores": threads,
            "kind": "reference", "seconds": dt, "fans": len(f), "passes": reps}


def ut_automata_params():
    """The UT-Automata car for the reference arm, from the reference's own config files
    (config/ut_automata_config.lua:4-10, config/sim_config.lua:25-31,39-44) -- restated here so
    that the reference arm never maps the product library."""
    from ut_multirobot_sim_b200._lib import RobotParams      # the POD struct only; nothing is loaded
    p = RobotParams()
    p.model_type = 0                                         # ACKERMANN_DRIVE
    p.ak_min_turn_radius, p.ak_max_speed, p.ak_max_accel = 0.98, 1.2, 3.0
    p.ak_angular_error_bias, p.ak_angular_error_rate = 0.0, 0.1
    p.car_width, p.car_length, p.rear_axle_offset, p.footprint_enabled = 0.281, 0.535, -0.162, 1
    p.laser.loc_x, p.laser.loc_y = 0.2, 0.0
    p.laser.angle_min, p.laser.angle_max = np.float32(np.pi * -135.0 / 180.0), np.float32(np.pi * 135.0 / 180.0)
    p.laser.angle_increment = np.float32(np.pi * 0.25 / 180.0)
    p.laser.range_min, p.laser.range_max, p.laser.noise_stddev = 0.4, 100.0, 0.01
    return p


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on all host cores --
    the map cleaned by the reference's own Load/Cleanup (vector_map.cpp:397-411), the restated
    Ackermann tick (oracle port) + the UNMODIFIED GetPredictedScan (oracle/_ref) -- on the same
    workload: every world of the workload, every step.  Nothing of the product is loaded."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import ctypes as C
    from ut_multirobot_sim_b200 import synth                 # pure numpy
    from tests import oracle_lib
    cores = os.cpu_count() or 1
    ref = oracle_lib.load_ref()
    if ref is None:
        emit(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libumrs_ref.so not built"}))
        return
    orc = oracle_lib.load_oracle()
    rm = ref.map_from_segments(synth.floorplan_segments(args.map_segments, extent=EXTENT, seed=1234), True)
    segs = rm.lines
    car = ut_automata_params()
    W_total = total_worlds(args)
    S = min(W_total, args.cpu_sample_fans or W_total)
    poses = synth.sample_free_poses_fast(segs, W_total, EXTENT, seed=11)[:S]
    states = (oracle_lib.RobotState * S)()
    for k in range(S):
        orc.lib.orc_robot_init(C.byref(states[k]), oracle_lib._fp(np.ascontiguousarray(poses[k], np.float32)))
    cmd_sets = [np.ascontiguousarray(synth.ackermann_commands(W_total, 1, s, seed=42)[:S, 0]) for s in range(8)]
    fans = np.empty((S, 4), np.float32)

    # publishLaser's dropout + Gaussian noise (simulator.cpp:412-418): the normals are drawn once,
    # outside the timed region (the reference draws them from std::default_random_engine per ray)
    normals = np.random.default_rng(7).standard_normal(S * N_RAYS, dtype=np.float32)

    def one_step(step):
        orc.lib.orc_robots_tick(states, S, C.byref(car), oracle_lib._fp(cmd_sets[step % 8]),
                                0.025, oracle_lib._fp(fans))
        out = rm.scan_batch(fans, car.laser.range_min, car.laser.range_max, N_RAYS, threads=cores)
        orc.lib.orc_laser_postprocess(oracle_lib._fp(out), out.size, car.laser.range_max,
                                      car.laser.noise_stddev, oracle_lib._fp(normals))
        return out

    steps, warm = args.steps, args.warmup
    # keep the whole run within a few minutes whatever --steps says (the warm-up is kept as asked
    # unless it alone would take more than a minute)
    t0 = time.perf_counter()
    one_step(0)
    per = time.perf_counter() - t0
    warm_run = max(0, min(warm, int(60.0 / max(per, 1e-6))))
    steps = max(1, min(steps, int(150.0 / max(per, 1e-6))))
    for s in range(max(0, warm_run - 1)):     # the probe step above was warm-up step 1
        one_step(1 + s)
    t0 = time.perf_counter()
    for s in range(steps):
        one_step(warm_run + s)
    dt = time.perf_counter() - t0
    value = S * N_RAYS * steps / dt
    sample = "%d of %d worlds per step, %d steps (requested %d), warm-up %d (requested %d), all %d host threads" % (
        S, W_total, steps, args.steps, warm_run, warm, cores)
    emit(json.dumps({
        "impl": "reference", "metric": "lidar_rays_per_sec", "value": value, "unit": "rays/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm_run, "ms_per_step": dt / steps * 1e3,
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, len(segs)),
        "reference_run": {"sample": sample, "map_cleanup": "the reference's own VectorMap::Load/Cleanup (oracle/_ref)",
                          "tick": "restated Ackermann step (oracle port), motion-noise draws zero",
                          "scan": "unmodified GetPredictedScan, one private VectorMap per thread",
                          "post": "dropout + noise (simulator.cpp:412-418) with pre-drawn normals"},
        "env_steps_per_sec": S * steps / dt,
        "cpu_baseline": {"value": value, "unit": "rays/s", "cores": cores, "kind": "reference",
                         "sample": sample},
        "e2e": {"value": value, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def source_hash():
    """sha256 over the kernel sources the committed ncu facts describe (scripts/ncu_facts.py writes the same)."""
    import hashlib
    h = hashlib.sha256()
    for f in ("umrs_kernels.cu", "umrs_device.cuh"):
        with open(os.path.join(ROOT, "ut_multirobot_sim_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


def run_b200(args):
    import torch
    import torch.distributed as dist
    import ut_multirobot_sim_b200 as umrs
    from ut_multirobot_sim_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world_size != args.gpus and world_size > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world_size))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    from ut_multirobot_sim_b200 import shard
    cpus = None if os.environ.get("UMRS_NO_BIND") else shard.bind_host_to_gpu(local_rank)
    lib = _lib.load()
    _lib.check(lib.umrs_set_device(local_rank), "umrs_set_device")
    if world_size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world_size == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    segs, poses_all, cmds_all, W, base = make_inputs(args, rank, world_size)
    W_total = total_worlds(args)
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")

    def make_world(b, n):
        wd = umrs.World(vmap, [car], n, dt=0.025, seed=42, laser_noise=True, motion_noise=True,
                        world_index_base=b, write_hit_idx=False)
        wd.SetPose(np.ascontiguousarray(poses_all[b:b + n]).reshape(n, 1, 3))
        return wd

    world = make_world(base, W)
    cmd_sets = [np.ascontiguousarray(c[base:base + W]) for c in cmds_all]
    rays_per_step = W * N_RAYS

    stream = torch.cuda.Stream()
    sp = stream.cuda_stream
    d_cmds = [torch.from_numpy(c).cuda() for c in cmd_sets]
    flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    K, Wm = args.steps, args.warmup

    def device_leg():
        ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
        with torch.cuda.stream(stream):
            for s in range(Wm):
                world.step_device(d_cmds[s % 8].data_ptr(), None, sp)
            barrier()
            l0 = lib.umrs_kernel_launch_count()
            for s in range(K):
                if flush is not None:
                    flush.zero_()          # L2 flush, outside the event pair
                ev0[s].record(stream)
                world.step_device(d_cmds[(Wm + s) % 8].data_ptr(), None, sp)
                ev1[s].record(stream)
            barrier()
            l1 = lib.umrs_kernel_launch_count()
        ms = sum(a.elapsed_time(b) for a, b in zip(ev0, ev1))
        return ms, l1 - l0

    def e2e_leg(world, b, W, pipelined):
        # host buffers (pinned), through the public host-buffer C-ABI calls.  pipelined: successive
        # ticks through umrs_world_step_async into two output buffers in turn, the previous tick's
        # output waited for (and read) while the next one runs; else one synchronous call per tick.
        h_cmds = [torch.from_numpy(np.ascontiguousarray(c[b:b + W])).pin_memory() for c in cmds_all]
        h_out = [torch.empty((W, N_RAYS), dtype=torch.float32).pin_memory() for _ in range(2)]
        # warm-up through the very call that is timed (the pipelined path has its own slicing and events)
        for s in range(Wm):
            if pipelined:
                world.step_async_raw(h_cmds[s % 8].data_ptr(), None, h_out[s & 1].data_ptr(), None)
                if s >= 1:
                    world.wait_outputs(1)
            else:
                world.step_raw(h_cmds[s % 8].data_ptr(), None, h_out[0].data_ptr(), None)
        if pipelined:
            world.wait()
        barrier()
        chk = 0.0
        t0 = time.perf_counter()
        if pipelined:
            for s in range(K):
                world.step_async_raw(h_cmds[s % 8].data_ptr(), None, h_out[s & 1].data_ptr(), None)
                if s >= 1:
                    world.wait_outputs(1)
                    chk += float(h_out[(s - 1) & 1][0, 0])      # the step's result is read on the host
            world.wait()
            chk += float(h_out[(K - 1) & 1][0, 0])
        else:
            for s in range(K):
                world.step_raw(h_cmds[s % 8].data_ptr(), None, h_out[0].data_ptr(), None)
                chk += float(h_out[0][0, 0])
        barrier()
        dt = time.perf_counter() - t0
        return dt * 1e3, h_cmds[0].numel() * 4, h_out[0].numel() * 4, chk

    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_dev, launches = device_leg()
    clocks = sampler.result()
    ms_dev = max_over_ranks(ms_dev)
    # plain pinned D2H of one step's output, timed in this run on every rank at once: the PCIe
    # roofline of `e2e`, and (N > 1) the weights of the host-buffer sharding
    def pcie_peak():
        n_bytes = W * N_RAYS * 4
        dev = torch.empty(n_bytes, dtype=torch.uint8, device="cuda")
        host = torch.empty(n_bytes, dtype=torch.uint8).pin_memory()
        for _ in range(3):
            host.copy_(dev, non_blocking=True)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            host.copy_(dev, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        return reps * n_bytes / (e0.elapsed_time(e1) * 1e-3) / 1e9

    pcie_all = shard.exchange_rates(pcie_peak())      # one all-gather, before the timed e2e region
    # Host-buffer stepping is bound by each GPU's link to host memory, and on this pool's 8-GPU box the
    # links are not equal (12.6 vs 22 GB/s with all ranks copying): the e2e legs shard the SAME job in
    # proportion to the measured rates (contiguous blocks, global world indices: same bits), so that
    # every rank's copy takes the same time.  The device-resident `value` keeps the uniform blocks.
    if world_size > 1:
        base_e, W_e = shard.shard_worlds_weighted(W_total, rank, pcie_all)
        world_e = make_world(base_e, W_e)
    else:
        base_e, W_e, world_e = base, W, world
    ms_sync, _, _, _ = e2e_leg(world_e, base_e, W_e, False)
    ms_sync = max_over_ranks(ms_sync)
    ms_e2e, _, _, _ = e2e_leg(world_e, base_e, W_e, True)
    ms_e2e = max_over_ranks(ms_e2e)
    ms_e2e_uniform = None
    if world_size > 1:
        ms_e2e_uniform, _, _, _ = e2e_leg(world, base, W, True)
        ms_e2e_uniform = max_over_ranks(ms_e2e_uniform)
    h2d, d2h = W_total * 3 * 4, W_total * N_RAYS * 4      # whole job, per step

    # sustained leg: back-to-back steps for >= --sustained-seconds, no flush, clocks sampled
    def sustained_leg():
        secs = max(0.0, args.sustained_seconds)
        if secs <= 0:
            return None
        chunk = 200
        smp = ClockSampler(local_rank)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 0
        with torch.cuda.stream(stream):
            barrier()
            smp.start()
            t0 = time.perf_counter()
            e0.record(stream)
            while True:
                for s in range(chunk):
                    world.step_device(d_cmds[(n + s) % 8].data_ptr(), None, sp)
                n += chunk
                stream.synchronize()
                if time.perf_counter() - t0 >= secs:
                    break
            e1.record(stream)
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        return {"steps": n, "seconds": ms * 1e-3, "ms_per_step": ms / n,
                "value": rays_per_step * n / (ms * 1e-3), "unit": "rays/s (this rank)",
                "l2": "not flushed (back-to-back steps)", "clocks": smp.result()}

    sustained = sustained_leg()

    # work accounting for the roofline: one untimed step with device counters on
    world.enable_counters(True)
    world.step_device(d_cmds[0].data_ptr(), None, sp)
    cnt = world.read_counters()
    world.enable_counters(False)
    stats = shard.reduce_stats(world.stats())   # the only cross-GPU exchange (NCCL), after the timed region

    total_rays = rays_per_step * world_size
    value = total_rays * K / (ms_dev * 1e-3)
    e2e_value = total_rays * K / (ms_e2e * 1e-3)
    tests = cnt["ray_seg_tests"]
    kernel_s = ms_dev * 1e-3 / K
    achieved_tflops = FLOPS_PER_TEST * tests / kernel_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    alg_bytes = rays_per_step * 4 + W * (64 * 2 + 12 + 16 + 12)  # ranges out + state r/w + cmd + fan + pose
    # per-launch facts of the committed `ncu --set full` capture of this workload (profiles/):
    # DRAM traffic and warp instructions executed (both properties of the launch, not of a clock)
    traffic, issue, pipes = None, None, None
    facts_note = {"stale": False, "source": "profiles/%s" % FACTS_FILE}
    try:
        facts = json.load(open(os.path.join(ROOT, "profiles", FACTS_FILE)))
        stale = facts.get("source_sha256") != source_hash()
        if stale:
            facts_note = {"stale": True, "why": "csrc/umrs_kernels.cu or umrs_device.cuh changed since the capture in profiles/%s" % FACTS_FILE}
        elif facts.get("worlds") == W and args.map_segments == 1024:
            traffic = facts["dram_bytes_read"] + facts["dram_bytes_write"]
            ginst = facts["inst_executed"] / kernel_s / 1e9
            peak_ginst = 148 * 4 * 1.965            # 4 schedulers per SM, one warp instruction per clock each
            pipes = {k: facts[k] for k in ("fma_pipe_pct", "alu_pipe_pct", "xu_pipe_pct", "lsu_pipe_pct",
                                           "fp64_pipe_pct", "smem_wavefronts_pct_of_peak", "issue_active_pct")
                     if k in facts}
            issue = {"achieved_ginst_per_s": ginst, "peak_ginst_per_s": peak_ginst, "frac": ginst / peak_ginst,
                     "source": "smsp__inst_executed.sum of the committed capture / live launch time; peak = 148 SM x 4 x 1.965 GHz"}
    except Exception as e:
        facts_note = {"stale": True, "why": "profiles/%s missing or unreadable (%s)" % (FACTS_FILE, type(e).__name__)}
    out = {
        "metric": "lidar_rays_per_sec", "value": value, "unit": "rays/s",
        "n_gpus": world_size, "steps": K, "warmup": Wm, "ms_per_step": ms_dev / K,
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, len(segs)),
        "path": "exact nearest hit per ray (tier B, SURVEY 8c): the reference's per-line formulae, minimum over all "
                "candidates instead of its occlusion-trimmed render; the reference's own bits are the opt-in "
                "reference render (legs.reference_render_*)",
        "env_steps_per_sec": W * world_size * K / (ms_dev * 1e-3),
        "e2e": {"value": e2e_value, "unit": "rays/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / K,
                "env_steps_per_sec": W * world_size * K / (ms_e2e * 1e-3),
                "call": "umrs_world_step_async + umrs_world_wait_outputs (ticks pipelined, two pinned output buffers)",
                "pcie": {"achieved_gbs": (d2h + h2d) / (ms_e2e / K * 1e-3) / 1e9,
                         "peak_gbs": sum(pcie_all), "peak_gbs_per_rank": pcie_all,
                         "frac": (d2h + h2d) / (ms_e2e / K * 1e-3) / 1e9 / sum(pcie_all),
                         "peak_source": "sum over ranks of a plain pinned device->host copy of a rank's ranges "
                                        "(%d bytes, 20 back to back), all ranks copying at once, timed in this run" % (W * N_RAYS * 4)},
                "sharding": {"policy": "uniform contiguous blocks" if world_size == 1 else
                             "contiguous blocks in proportion to each rank's measured device->host rate "
                             "(shard_worlds_weighted; global world indices, same bits as the uniform blocks)",
                             "worlds_this_rank": W_e,
                             "uniform_blocks": None if ms_e2e_uniform is None else {
                                 "value": total_rays * K / (ms_e2e_uniform * 1e-3), "ms_per_step": ms_e2e_uniform / K}},
                "synchronous": {"value": total_rays * K / (ms_sync * 1e-3), "ms_per_step": ms_sync / K,
                                "call": "umrs_world_step"}},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "fp32", "achieved": achieved_tflops, "peak": FP32_PEAK_TFLOPS,
                     "unit": "TFLOP/s", "frac": achieved_tflops / FP32_PEAK_TFLOPS, "traffic": traffic,
                     "issue": issue, "ncu_pipe_utilisation": pipes, "ncu_facts": facts_note,
                     "brute_force_equivalent_tflops": FLOPS_PER_TEST * cnt["segs_prepared"] * N_RAYS / kernel_s / 1e12,
                     "kernel": "k_world_step", "tests_per_launch": int(tests),
                     "tests_per_ray": tests / rays_per_step,
                     "segs_prepared_per_fan": cnt["segs_prepared"] / W,
                     "flops_per_test": FLOPS_PER_TEST,
                     "peak_source": "nominal FP32 pipe: 148 SM x 128 lanes x 2 x 1.965 GHz (MEASURED_PEAKS.json has no FP32 figure)",
                     "hbm": {"achieved_gbs": alg_bytes / kernel_s / 1e9, "peak_gbs": hbm_peak,
                             "frac": alg_bytes / kernel_s / 1e9 / hbm_peak,
                             "peak_source": "measured" if "hbm_gbs" in peaks else "fallback"}},
        "sustained": sustained,
        "stats": stats,
        "host": {"cpus_bound": None if cpus is None else len(cpus), "cpu_count": os.cpu_count()},
    }
    if rank == 0 and world_size == 1 and not args.no_legs:
        # BASELINE configs 1-4 and the reference render, short forms (scripts/bench_configs.py)
        from scripts import bench_configs
        t0 = time.perf_counter()
        out["legs"] = bench_configs.bench_legs()
        out["legs"]["seconds"] = time.perf_counter() - t0
    if rank == 0 and world_size == 1 and not args.no_cpu_baseline:
        fans = world.fan_params().reshape(-1, 4)
        cores = os.cpu_count() or 1
        n_fans = min(W, args.cpu_sample_fans or W)
        cb = cpu_baseline(segs, fans[:n_fans], car.laser.range_max, cores, 12.0)
        if cb is not None:
            one = cpu_baseline(segs, fans[:min(n_fans, 512)], car.laser.range_max, 1, 4.0)
            cb["sample"] = ("%d passes over %d of the %d fans of the last step (%.1f s), unmodified reference "
                            "GetPredictedScan, %d threads" % (cb["passes"], n_fans, W, cb["seconds"], cores))
            cb["single_thread_value"] = one["value"]
            out["cpu_baseline"] = cb
    if rank == 0:
        emit(json.dumps(out))
    if world_size > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    # stdout carries exactly ONE JSON line: anything a library prints there while we run
    # (NCCL's version banner, ...) is sent to stderr instead
    global _JSON_FD
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
