"""Dev helper: correctness + rough timing of the fused step on one GPU."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import synth
from tests import oracle_lib

orc = oracle_lib.load_oracle()
W = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
MAPS = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [64, 256, 1024]
for M in MAPS:
    segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    world = umrs.World(vmap, [car], W, dt=0.025, seed=1)
    poses = synth.sample_free_poses_fast(segs, W, 200.0, seed=11)
    world.SetPose(poses.reshape(W, 1, 3))
    cmds = synth.ackermann_commands(W, 1, 0)
    ranges, idx = world.Step(cmds, want_idx=True)
    fans = world.fan_params()
    n = umrs.laser_num_rays(car.laser)
    bad = 0; exact = 0
    for w in range(0, W, max(1, W // 64)):
        f = fans[w, 0]
        r_o, i_o = orc.scan_tierb(segs, None, f[:2], car.laser.range_max, f[2], f[3], n)
        rel = np.abs(ranges[w] - r_o) / np.maximum(np.abs(r_o), 1e-9)
        bad += int((rel > 1e-5).sum()) + int((idx[w] != i_o).sum())
        exact += int((ranges[w] == r_o).sum())
    world.enable_counters(True)
    world.Step(cmds)
    cnt = world.read_counters()
    world.enable_counters(False)
    dcmd = torch.from_numpy(cmds).cuda()
    ts = torch.cuda.Stream()
    st = ts.cuda_stream
    torch.cuda.synchronize()
    for _ in range(5):
        world.step_device(dcmd.data_ptr(), None, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 50
    e0.record(ts)
    for _ in range(K):
        world.step_device(dcmd.data_ptr(), None, st)
    e1.record(ts); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    rays = W * n
    print("M=%d W=%d: bad=%d exact=%d/%d  %.3f ms/step  %.3e rays/s  tests/ray=%.1f pairs/chunk=%.1f segs=%d" % (
        M, W, bad, exact, 64 * n, ms, rays / ms * 1e3, cnt["ray_seg_tests"] / rays,
        cnt["chunk_seg_pairs"] / (W * ((n + 31) // 32)), cnt["segs_prepared"] // W))
