"""Dev helper: a few launches of the reference render (4096 fans x 1081 rays, 1024-seg map) for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import synth, _lib
lib = _lib.load()
M = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
W = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
d135 = np.float32(np.pi * 135.0 / 180.0)
segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
vmap = umrs.VectorMap(lines=segs)
poses = synth.sample_free_poses_fast(segs, W, 200.0, seed=11)
fans = np.stack([poses[:, 0], poses[:, 1], poses[:, 2] - d135, poses[:, 2] + d135], 1).astype(np.float32)
d_fans = torch.from_numpy(fans).cuda()
d_out = torch.empty((W, 1081), dtype=torch.float32, device="cuda")
for _ in range(3):
    _lib.check(lib.umrs_scan_batch_reference_render_device(vmap._h, None, 0, d_fans.data_ptr(), W, 100.0, 1081, d_out.data_ptr(), None), "render")
torch.cuda.synchronize()
