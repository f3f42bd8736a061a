"""Dev helper: histogram of the candidates S2 prepares per fan (library built with -DUMRS_HIST:
UMRS_LIB=build/libumrs_hist.so python scripts/surv_hist.py [W] [M,M,...])."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ut_multirobot_sim_b200 as umrs
from ut_multirobot_sim_b200 import synth

W = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
MAPS = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [512, 768, 1024, 1500]
edges = ["<=192", "<=256", "<=320", "<=384", "<=448", "<=512", "<=608", ">608"]
for M in MAPS:
    segs = umrs.map_cleanup(synth.floorplan_segments(M, extent=200.0, seed=1234))
    vmap = umrs.VectorMap(lines=segs)
    car = umrs.default_robot("ut_automata")
    world = umrs.World(vmap, [car], W, dt=0.025, seed=1, laser_noise=True, motion_noise=True)
    world.SetPose(synth.sample_free_poses_fast(segs, W, 200.0, seed=11).reshape(W, 1, 3))
    cmds = synth.ackermann_commands(W, 1, 0)
    for _ in range(2):
        world.Step(cmds)
    world.enable_counters(True)
    world.Step(cmds)
    pc = list(world.phase_cycles().values())
    import ctypes as C
    ex = (C.c_uint64 * 8)()
    umrs._lib.load().umrs_world_read_phase_cycles(world._h, ex)
    cnt = world.read_counters()
    world.enable_counters(False)
    print("M=%d (%d cleaned) W=%d prepared/fan=%.1f passes=%d" % (M, len(segs), W, cnt["segs_prepared"] / W, cnt["passes"]))
    print("   " + "  ".join("%s:%d" % (e, ex[i]) for i, e in enumerate(edges)))
