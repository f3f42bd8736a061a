"""Dev helper: per-phase critical-path cycles of the fan CTAs (needs a library built with
-DUMRS_PHASE_TIMING: UMRS_LIB=build/libumrs_timing.so python scripts/phase_cycles.py [W] [M])."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import synth

W = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
for M in ([int(sys.argv[2])] if len(sys.argv) > 2 else [256, 1024]):
    segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    world = umrs.World(vmap, [car], W, dt=0.025, seed=1, laser_noise=True, motion_noise=True)
    world.SetPose(synth.sample_free_poses_fast(segs, W, 200.0, seed=11).reshape(W, 1, 3))
    cmds = synth.ackermann_commands(W, 1, 0)
    for _ in range(3):
        world.Step(cmds)
    world.enable_counters(True)
    world.Step(cmds)
    pc = world.phase_cycles()
    world.enable_counters(False)
    tot = sum(pc.values()) or 1
    print("M=%d W=%d: %d cycles per fan CTA" % (M, W, tot // W))
    for k, v in pc.items():
        print("  %-9s %8d cycles/fan  %5.1f %%" % (k, v // W, 100.0 * v / tot))
