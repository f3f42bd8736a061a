"""Dev helper: CUDA-event timing of the reference render for several batch sizes (1024-seg map)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import synth, _lib
lib = _lib.load()
d135 = np.float32(np.pi * 135.0 / 180.0)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
vmap = umrs.VectorMap(lines=segs)
poses_all = synth.sample_free_poses_fast(segs, 16384, 200.0, seed=11)
for W in (1, 512, 4096, 8192, 16384):
    poses = poses_all[:W]
    fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - d135, poses[:, 2] + d135], 1).astype(np.float32)
    d_fans = torch.from_numpy(fans).cuda()
    d_out = torch.empty((W, 1081), dtype=torch.float32, device="cuda")
    st = torch.cuda.Stream()
    def run():
        _lib.check(lib.umrs_scan_batch_reference_render_device(vmap._h, None, 0, d_fans.data_ptr(), W, 100.0, 1081, d_out.data_ptr(), st.cuda_stream), "render")
    for _ in range(2): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(4): run()
    e1.record(st); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 4
    print("M=%d fans=%d  %.3f ms  %.3e rays/s" % (M, W, ms, W * 1081 / ms * 1e3))
