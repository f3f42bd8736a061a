"""BASELINE.json configs 1-4 and the opt-in reference render, timed on one GPU (the bench.py
headline is config 5's shard).  Each function returns one dict; `python scripts/bench_configs.py
[1 2 3 4 R]` prints them as JSON lines, and bench.py runs the short forms as its `legs`.
Device-resident timings use CUDA events on the launch stream; host-buffer timings use the
synchronous C-ABI calls.  Where the reference's own routine can run on the same inputs
(tests/oracle_lib: oracle/_ref, the unmodified vector_map.cpp) a `cpu_baseline` is timed beside
it on a bounded sample, as the checker -- never as the thing measured."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import synth, _lib

lib = _lib.load()
d135 = np.float32(np.pi * 135.0 / 180.0)


def ev_time(fn, stream, iters, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(iters):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def _ref():
    from tests import oracle_lib
    return oracle_lib.load_ref()


def _cpu_scan_baseline(segs, fans, range_max, n, threads=None, budget_s=1.0):
    """The unmodified reference GetPredictedScan over `fans` on the host cores: whole passes until
    about budget_s.  None when oracle/_ref was not built."""
    ref = _ref()
    if ref is None:
        return None
    threads = threads or (os.cpu_count() or 1)
    rm = ref.map_from_segments(segs, False)
    f = np.ascontiguousarray(fans, np.float32)
    rm.scan_batch(f[:max(1, min(len(f), threads))], 0.4, range_max, n, threads=threads)   # warm
    t0 = time.perf_counter()
    reps = 0
    while True:
        rm.scan_batch(f, 0.4, range_max, n, threads=threads)
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or reps >= 1000:
            break
    return {"value": reps * len(f) * n / dt, "unit": "rays/s", "cores": threads, "kind": "reference",
            "sample": "%d passes over %d fans x %d rays, unmodified GetPredictedScan" % (reps, len(f), n)}


def config1(cpu=True):
    segs = umrs.map_cleanup(synth.floorplan_segments(1500, 200.0, 1234))
    vmap = umrs.VectorMap(lines=segs)
    poses = synth.sample_free_poses(segs, 64, 200.0, 1234)
    # the first calls pay one-time costs (kernel image load, shared-memory opt-in, pinned staging):
    # reported apart, not averaged into the per-scan figure
    t0 = time.perf_counter()
    for p in poses[:3]:
        vmap.GetPredictedScan(p[:2], 0.4, 100.0, p[2] - d135, p[2] + d135, 1081)
    first3 = time.perf_counter() - t0
    per = []
    for p in poses:
        t0 = time.perf_counter()
        vmap.GetPredictedScan(p[:2], 0.4, 100.0, p[2] - d135, p[2] + d135, 1081)
        per.append(time.perf_counter() - t0)
    dt = float(np.median(per))
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - d135, poses[:, 2] + d135], 1).astype(np.float32)
    vmap.ScanBatch(fans, 100.0, 1081)        # first call grows the handle's device buffers
    t0 = time.perf_counter()
    vmap.ScanBatch(fans, 100.0, 1081)
    db = time.perf_counter() - t0
    out = {"config": 1, "what": "single 1081-ray scan through umrs_predicted_scan, host buffers, 1500-seg campus map, noise off",
           "us_per_scan": dt * 1e6, "rays_per_s": 1081 / dt, "us_per_scan_mean": float(np.mean(per)) * 1e6,
           "first_3_calls_ms": first3 * 1e3, "batch64_ms": db * 1e3, "batch64_rays_per_s": 64 * 1081 / db}
    if cpu:
        # the reference's execution model: one scan at a time on one thread
        cb = _cpu_scan_baseline(segs, fans, 100.0, 1081, threads=1, budget_s=0.5)
        if cb is not None:
            cb["us_per_scan"] = 1081 / cb["value"] * 1e6
        out["cpu_baseline"] = cb
    return out


def config2(worlds=(1, 1024), iters=50, cpu=True):
    """SURVEY 8(d) config 2 on the 1500-segment campus map ("map as (1)")."""
    extent = 200.0
    segs = umrs.map_cleanup(synth.floorplan_segments(1500, extent=extent, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    rps = [umrs.default_robot("ut_jackal") for _ in range(3)] + [umrs.default_robot("bwibot") for _ in range(3)] + \
          [umrs.default_robot("cobot") for _ in range(2)]
    hps = [umrs.default_human() for _ in range(16)]
    outs = []
    for W in worlds:
        world = umrs.World(vmap, rps, W, humans=hps, static_objects=1, seed=9, laser_noise=True, write_hit_idx=False)
        free = synth.sample_free_poses_fast(segs, W * 8, extent, seed=21, clearance=0.6).reshape(W, 8, 3)
        hs = synth.sample_free_poses_fast(segs, W * 32, extent, seed=22).reshape(W, 32, 3)
        so = synth.sample_free_poses_fast(segs, W, extent, seed=23).reshape(W, 1, 3)
        for arr in (hs, so, free):      # everybody within a few metres of robot 0: bodies occlude every scan
            c = free[:, :1, :2].copy()
            arr[..., :2] = c + (arr[..., :2] - c) * 0.04
        world.SetPose(free)
        world.SetHumans(hs[:, :16], hs[:, 16:])
        world.SetStaticObjects(so)
        cmds = torch.from_numpy(np.random.default_rng(1).uniform(-1, 1, (W, 8, 3)).astype(np.float32)).cuda()
        st = torch.cuda.Stream()
        ms = ev_time(lambda: world.step_device(cmds.data_ptr(), None, st.cuda_stream), st, iters)
        rays = W * world.rays_per_world
        world.enable_counters(True)
        world.step_device(cmds.data_ptr(), None, st.cuda_stream)
        cnt = world.read_counters()
        world.enable_counters(False)
        o = {"config": 2, "what": "8 robots (3 jackal 3601 rays, 3 bwibot, 2 cobot) + 16 humans + 1 object per world, "
                                  "1500-seg campus map, step + scan (fused for 1 world, two launches for 1024), noise on",
             "worlds": W, "rays_per_world": world.rays_per_world, "ms_per_step": ms, "rays_per_s": rays / ms * 1e3,
             "env_steps_per_s": W / ms * 1e3, "tests_per_ray": cnt["ray_seg_tests"] / rays,
             "segs_prepared_per_fan": cnt["segs_prepared"] / (W * 8)}
        if cpu and W == worlds[-1]:
            # the reference routine on the 8 scans of world 0 (its own object lines minus each robot's body);
            # jackal scans (150 m) run into the reference's 2000-candidate cap on this map
            lines, fp = world.GetLines()[0], world.fan_params()[0]
            ref = _ref()
            if ref is not None:
                rm = ref.map_from_segments(segs, False)
                t0 = time.perf_counter()
                nr = 0
                bodies = [k for k, q in enumerate(rps) if q.footprint_enabled]
                for r_i, rp in enumerate(rps):
                    ol = lines
                    if rp.footprint_enabled:
                        b0 = 4 + 20 * 16 + 4 * bodies.index(r_i)
                        ol = np.concatenate([lines[:b0], lines[b0 + 4:]])
                    rm.set_object_lines(ol)
                    n = umrs.laser_num_rays(rp.laser)
                    rm.predicted_scan(fp[r_i, :2], 0.4, rp.laser.range_max, fp[r_i, 2], fp[r_i, 3], n)
                    nr += n
                dt = time.perf_counter() - t0
                o["cpu_baseline"] = {"value": nr / dt, "unit": "rays/s", "cores": 1, "kind": "reference",
                                     "sample": "the 8 scans of world 0, one thread (the reference's execution model)"}
        outs.append(o)
    return outs


def config3(maps=(64, 256, 1024), noises=(False, True), iters=200, cpu=True, W=4096):
    outs = []
    for M in maps:
        for noise in noises:
            segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
            vmap = umrs.VectorMap(lines=segs)
            car = umrs.default_robot("ut_automata")
            world = umrs.World(vmap, [car], W, seed=42, laser_noise=noise, motion_noise=noise, write_hit_idx=False)
            world.SetPose(synth.sample_free_poses_fast(segs, W, 200.0, seed=11).reshape(W, 1, 3))
            cmds = torch.from_numpy(synth.ackermann_commands(W, 1, 0)).cuda()
            st = torch.cuda.Stream()
            ms = ev_time(lambda: world.step_device(cmds.data_ptr(), None, st.cuda_stream), st, iters)
            world.enable_counters(True)
            world.step_device(cmds.data_ptr(), None, st.cuda_stream)
            cnt = world.read_counters()
            world.enable_counters(False)
            o = {"config": 3, "map_segments": M, "noise": noise, "worlds": W, "ms_per_step": ms,
                 "rays_per_s": W * 1081 / ms * 1e3, "env_steps_per_s": W / ms * 1e3,
                 "tests_per_ray": cnt["ray_seg_tests"] / (W * 1081),
                 "segs_prepared_per_fan": cnt["segs_prepared"] / W}
            if cpu and noise == noises[-1]:
                fans = world.fan_params().reshape(-1, 4)
                o["cpu_baseline"] = _cpu_scan_baseline(segs, fans[:512], 100.0, 1081, budget_s=0.5)
            outs.append(o)
    return outs


def config4(iters=100):
    segs = synth.cityblock_segments(100000, extent=2000.0, seed=99)
    vmap = umrs.VectorMap(lines=segs)
    rng = np.random.default_rng(4)
    n_fans = 925
    poses = np.stack([rng.uniform(100, 1900, n_fans), rng.uniform(100, 1900, n_fans),
                      rng.uniform(-np.pi, np.pi, n_fans)], 1).astype(np.float32)
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - d135, poses[:, 2] + d135], 1).astype(np.float32)
    d_fans = torch.from_numpy(fans).cuda()
    d_out = torch.empty((n_fans, 1081), dtype=torch.float32, device="cuda")
    st = torch.cuda.Stream()

    def run():
        _lib.check(lib.umrs_scan_batch_device(vmap._h, None, 0, d_fans.data_ptr(), n_fans, 100.0, 1081,
                                              d_out.data_ptr(), None, st.cuda_stream), "scan_batch_device")
    ms = ev_time(run, st, iters)
    vmap.ScanBatch(fans, 100.0, 1081)        # first call grows the handle's device buffers
    dhs = []
    for _ in range(5):
        t0 = time.perf_counter()
        vmap.ScanBatch(fans, 100.0, 1081)
        dhs.append(time.perf_counter() - t0)
    dh = sorted(dhs)[2]                      # median of 5 (pageable host buffers, wall clock)
    vmap.enable_counters(True)       # rows visited and segment bytes fetched per ray (grid walk)
    run()
    cnt = vmap.read_counters()
    vmap.enable_counters(False)
    rays = n_fans * 1081
    return {"config": 4, "what": "100k-segment city-block map (2 km), 925 fans x 1081 rays, uniform-grid rows (K2)",
            "ms_per_step": ms, "rays_per_s": rays / ms * 1e3,
            "host_buffer_ms": dh * 1e3, "host_buffer_rays_per_s": rays / dh,
            "grid_rows_per_fan": cnt["source_rows"] / n_fans,
            "source_segments_per_fan": cnt["source_segments"] / n_fans,
            "segment_bytes_fetched_per_ray": cnt["source_segments"] * 20.0 / rays,
            "segs_prepared_per_fan": cnt["segs_prepared"] / n_fans,
            "tests_per_ray": cnt["ray_seg_tests"] / rays,
            "passes_per_fan": cnt["passes"] / n_fans,
            "cpu_baseline": None,
            "cpu_baseline_note": "the reference cannot run this map: SceneRender stops at 2000 candidates "
                                 "(vector_map.cpp:223,229,248); the map alone is 100k segments"}


def reference_render(maps=(256, 1024), batches=(1, 64, 4096, 16384), iters=5, cpu=True):
    """The opt-in reference render (bit-identical to the reference's GetPredictedScan), device buffers."""
    outs = []
    for M in maps:
        segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
        vmap = umrs.VectorMap(lines=segs)
        for W in batches:
            poses = synth.sample_free_poses_fast(segs, W, 200.0, seed=11)
            fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - d135, poses[:, 2] + d135], 1).astype(np.float32)
            d_fans = torch.from_numpy(fans).cuda()
            d_out = torch.empty((W, 1081), dtype=torch.float32, device="cuda")
            st = torch.cuda.Stream()

            def run():
                _lib.check(lib.umrs_scan_batch_reference_render_device(
                    vmap._h, None, 0, d_fans.data_ptr(), W, 100.0, 1081, d_out.data_ptr(), st.cuda_stream), "render")
            ms = ev_time(run, st, iters, warm=2)
            o = {"config": "reference_render", "map_segments": int(len(segs)), "fans": W,
                 "ms_per_call": ms, "rays_per_s": W * 1081 / ms * 1e3}
            if cpu and W == batches[-1]:
                o["cpu_baseline"] = _cpu_scan_baseline(segs, fans[:512], 100.0, 1081, budget_s=0.5)
            outs.append(o)
    return outs


def bench_legs():
    """The short forms bench.py puts into its JSON line (about 10 s in all)."""
    legs = {}
    c1 = config1()
    legs["config1_single_scan"] = c1
    c2 = config2(worlds=(1024,), iters=20)[0]
    legs["config2_crowded_1024_worlds"] = c2
    for o in config3(noises=(True,), iters=50):
        legs["config3_4096_worlds_%d_segs" % o["map_segments"]] = o
    legs["config4_100k_segments"] = config4(iters=30)
    legs["reference_render_4096_fans_1024_segs"] = reference_render(maps=(1024,), batches=(4096,), iters=3)[0]
    return legs


if __name__ == "__main__":
    which = sys.argv[1:] or ["1", "2", "3", "4", "R"]
    for c in which:
        res = {"1": config1, "2": config2, "3": config3, "4": config4, "R": reference_render}[c]()
        for o in (res if isinstance(res, list) else [res]):
            print(json.dumps(o))
