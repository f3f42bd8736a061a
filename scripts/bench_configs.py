"""Times BASELINE.json configs 1-4 (the bench.py headline is config 5's shard) on one GPU and
prints one JSON line per config.  Device-resident timings use CUDA events on the launch
stream; host-buffer timings use the synchronous C-ABI calls."""
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


def config1():
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
    t0 = time.perf_counter()
    vmap.ScanBatch(fans, 100.0, 1081)
    db = time.perf_counter() - t0
    print(json.dumps({"config": 1, "what": "single 1081-ray scan through umrs_predicted_scan, host buffers, 1500-seg map",
                      "ms_per_scan": dt * 1e3, "rays_per_s": 1081 / dt, "ms_per_scan_mean": float(np.mean(per)) * 1e3,
                      "first_3_calls_ms": first3 * 1e3,
                      "batch64_ms": db * 1e3, "batch64_rays_per_s": 64 * 1081 / db}))


def config2():
    extent = 60.0
    segs = umrs.map_cleanup(synth.floorplan_segments(300, extent=extent, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    rps = [umrs.default_robot("ut_jackal") for _ in range(3)] + [umrs.default_robot("bwibot") for _ in range(3)] + \
          [umrs.default_robot("cobot") for _ in range(2)]
    hps = [umrs.default_human() for _ in range(16)]
    for W in (1, 1024):
        world = umrs.World(vmap, rps, W, humans=hps, static_objects=1, seed=9, laser_noise=True, write_hit_idx=False)
        free = synth.sample_free_poses_fast(segs, W * 8, extent, seed=21, clearance=0.6)
        world.SetPose(free.reshape(W, 8, 3))
        hs = synth.sample_free_poses_fast(segs, W * 32, extent, seed=22).reshape(W, 32, 3)
        world.SetHumans(hs[:, :16], hs[:, 16:])
        world.SetStaticObjects(synth.sample_free_poses_fast(segs, W, extent, seed=23).reshape(W, 1, 3))
        cmds = torch.from_numpy(np.random.default_rng(1).uniform(-1, 1, (W, 8, 3)).astype(np.float32)).cuda()
        st = torch.cuda.Stream()
        ms = ev_time(lambda: world.step_device(cmds.data_ptr(), None, st.cuda_stream), st, 50)
        rays = W * world.rays_per_world
        world.enable_counters(True)
        world.step_device(cmds.data_ptr(), None, st.cuda_stream)
        print("config2 counters", world.read_counters(), file=sys.stderr)
        world.enable_counters(False)
        print(json.dumps({"config": 2, "what": "8 robots (3 jackal 3601 rays, 3 bwibot, 2 cobot) + 16 humans + 1 object per world, fused step",
                          "worlds": W, "ms_per_step": ms, "rays_per_s": rays / ms * 1e3, "env_steps_per_s": W / ms * 1e3}))


def config3():
    for M in (64, 256, 1024):
        for noise in (False, True):
            segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
            vmap = umrs.VectorMap(lines=segs)
            car = umrs.default_robot("ut_automata")
            W = 4096
            world = umrs.World(vmap, [car], W, seed=42, laser_noise=noise, motion_noise=noise, write_hit_idx=False)
            world.SetPose(synth.sample_free_poses_fast(segs, W, 200.0, seed=11).reshape(W, 1, 3))
            cmds = torch.from_numpy(synth.ackermann_commands(W, 1, 0)).cuda()
            st = torch.cuda.Stream()
            ms = ev_time(lambda: world.step_device(cmds.data_ptr(), None, st.cuda_stream), st, 200)
            world.enable_counters(True)
            world.step_device(cmds.data_ptr(), None, st.cuda_stream)
            cnt = world.read_counters()
            print(json.dumps({"config": 3, "map_segments": M, "noise": noise, "worlds": W, "ms_per_step": ms,
                              "rays_per_s": W * 1081 / ms * 1e3, "env_steps_per_s": W / ms * 1e3,
                              "tests_per_ray": cnt["ray_seg_tests"] / (W * 1081),
                              "segs_prepared_per_fan": cnt["segs_prepared"] / W}))


def config4():
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
    ms = ev_time(run, st, 100)
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
    print(json.dumps({"config": 4, "what": "100k-segment city-block map (2 km), 925 fans x 1081 rays, uniform-grid rows (K2)",
                      "ms_per_step": ms, "rays_per_s": rays / ms * 1e3,
                      "host_buffer_ms": dh * 1e3, "host_buffer_rays_per_s": rays / dh,
                      "grid_rows_per_fan": cnt["source_rows"] / n_fans,
                      "source_segments_per_fan": cnt["source_segments"] / n_fans,
                      "segment_bytes_fetched_per_ray": cnt["source_segments"] * 20.0 / rays,
                      "segs_prepared_per_fan": cnt["segs_prepared"] / n_fans,
                      "tests_per_ray": cnt["ray_seg_tests"] / rays,
                      "passes_per_fan": cnt["passes"] / n_fans}))


def reference_render():
    """The opt-in reference render (bit-identical to the reference's GetPredictedScan), device buffers."""
    for M in (256, 1024):
        segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
        vmap = umrs.VectorMap(lines=segs)
        for W in (1, 64, 4096):
            poses = synth.sample_free_poses_fast(segs, W, 200.0, seed=11)
            fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - d135, poses[:, 2] + d135], 1).astype(np.float32)
            d_fans = torch.from_numpy(fans).cuda()
            d_out = torch.empty((W, 1081), dtype=torch.float32, device="cuda")
            st = torch.cuda.Stream()

            def run():
                _lib.check(lib.umrs_scan_batch_reference_render_device(
                    vmap._h, None, 0, d_fans.data_ptr(), W, 100.0, 1081, d_out.data_ptr(), st.cuda_stream), "render")
            ms = ev_time(run, st, 5, warm=2)
            print(json.dumps({"config": "reference_render", "map_segments": int(len(segs)), "fans": W,
                              "ms_per_call": ms, "rays_per_s": W * 1081 / ms * 1e3}))


if __name__ == "__main__":
    which = sys.argv[1:] or ["1", "2", "3", "4", "R"]
    for c in which:
        {"1": config1, "2": config2, "3": config3, "4": config4, "R": reference_render}[c]()
