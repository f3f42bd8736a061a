"""Scenario generator: the reference's scripts/generate_config.py
(YAML -> init_config.lua + one human_config_<i>.lua per human, :40-62) and its extension to
batched world descriptors -- thousands of seeded variants of one scenario, ready for
`World.SetPose / SetHumans`.

`generate_config` writes byte-for-byte what the reference script writes (tests/golden/scenario
holds the reference script's own output for crowd.yml); `load_scenario_configs` reads those
files back through the C-ABI config evaluator (umrs_config_*), i.e. the way
Simulator::init/loadObject consume them (simulator.cpp:93-97,230-233).

Host-side only; no device work here.
"""
import os
from dataclasses import dataclass, field

import numpy as np
import yaml

# The text the reference's templates expand to (scripts/init_config_template.txt and
# scripts/human_config_template.txt after str.format); a data format, restated so that the
# output files are identical and the reference's Lua loader accepts them unchanged.
_INIT_FMT = (
    'map_name =  "{0}"\n'
    "-- Simulator starting location.\n"
    "start_x = {1}\n"
    "start_y = {2}\n"
    "start_angle = {3}\n"
    "\n"
    "human_config_list = {{ {4} }}\n"
    "\n"
)
_HUMAN_FMT = (
    "-- Human shape information\n"
    "hu_radius = 0.1\n"
    "hu_num_segments = 20\n"
    "\n"
    "-- Human start position\n"
    "hu_start_x = {0}\n"
    "hu_start_y = {1}\n"
    "hu_start_theta = 0.\n"
    "\n"
    "-- Human goal position\n"
    "hu_goal_x = {2}\n"
    "hu_goal_y = {3}\n"
    "hu_goal_theta = 0.\n"
    "\n"
    "-- Human speed information\n"
    "hu_max_speed = 1.5\n"
    "hu_avg_speed = 1.\n"
    "hu_max_omega = 0.2\n"
    "hu_avg_omega = 0.\n"
    "hu_reach_goal_threshold = 0.3\n"
    "\n"
    "-- Human walking mode\n"
    "local HumanMode = {{\n"
    "    Singleshot=0,\n"
    "    Repeat=1\n"
    "}}\n"
    "\n"
    "hu_mode = HumanMode.Repeat\n"
)


@dataclass
class Scenario:
    """The keys of the reference's scenario YAML (scripts/example.yml)."""
    prefix: str
    map_name: str
    start_x: float
    start_y: float
    start_angle: float
    # [[ [sx, sy], [gx, gy] ], ...] exactly as parsed (ints stay ints: the reference formats
    # the YAML scalars with str(), so "29" and "29." give different files)
    human_start_goal_positions: list = field(default_factory=list)

    @property
    def num_humans(self):
        return len(self.human_start_goal_positions)


def load_yaml(path):
    """generate_config.py:17-18 (yaml.FullLoader)."""
    with open(path, "rb") as f:
        data = yaml.load(f, yaml.FullLoader)
    missing = [k for k in ("prefix", "map_name", "start_x", "start_y", "start_angle",
                           "human_start_goal_positions") if k not in data]
    if missing:
        # the reference dies with KeyError at generate_config.py:29,40-44
        raise KeyError("scenario file %s lacks %s" % (path, ", ".join(missing)))
    pos = data["human_start_goal_positions"]
    for i, sg in enumerate(pos):
        if len(sg) != 2 or len(sg[0]) < 2 or len(sg[1]) < 2:
            raise ValueError("human %d: expected [[start_x, start_y], [goal_x, goal_y]]" % i)
    return Scenario(data["prefix"], data["map_name"], data["start_x"], data["start_y"],
                    data["start_angle"], pos)


def scenario_dir(prefix):
    """generate_config.py:30, relative to the simulator's working directory."""
    return "config/human_crowd_scenario_configs/%s" % prefix


def generate_config(scn, out_root="."):
    """generate_config.py:29-62: writes <out_root>/config/human_crowd_scenario_configs/<prefix>/
    init_config.lua and human/human_config_<i>.lua; the paths inside init_config.lua stay
    relative to the working directory, as in the reference.  Returns the init file's path."""
    if isinstance(scn, (str, bytes, os.PathLike)):
        scn = load_yaml(scn)
    rel = scenario_dir(scn.prefix)
    os.makedirs(os.path.join(out_root, rel, "human"), exist_ok=True)
    listing = ""
    for i, (start, goal) in enumerate(scn.human_start_goal_positions):
        name = "%s/human/human_config_%d.lua" % (rel, i)
        with open(os.path.join(out_root, name), "w") as f:
            f.write(_HUMAN_FMT.format(start[0], start[1], goal[0], goal[1]))
        listing += '"%s" ' % name
        if i != scn.num_humans - 1:
            listing += ",\n\r"          # sic, generate_config.py:56
    init_path = os.path.join(out_root, rel, "init_config.lua")
    with open(init_path, "w") as f:
        f.write(_INIT_FMT.format(scn.map_name, scn.start_x, scn.start_y, scn.start_angle, listing))
    return init_path


@dataclass
class LoadedScenario:
    """What Simulator::init / loadObject read from the generated files."""
    map_name: str
    start_pose: np.ndarray          # [3] x, y, angle
    humans: list                    # HumanParams per human_config_list entry
    human_files: list


def load_scenario_configs(init_config, root_dir="."):
    """Reads init_config.lua (map_name, start_x/y/angle, human_config_list: simulator.cpp:93-97)
    and every listed human file (one fresh config state per human, as HumanObject's ctor does,
    human_object.cpp:70-84) through umrs_config_*."""
    from . import api
    init = api.Config([init_config])
    files = init.get_string_list("human_config_list") if init.has("human_config_list") else []
    humans = []
    for f in files:
        hp, missing = api.Config([os.path.join(root_dir, f)]).human_params()
        if missing:
            raise ValueError("%s: %d human keys undefined" % (f, missing))
        humans.append(hp)
    pose = np.array([init.get_float("start_x"), init.get_float("start_y"),
                     init.get_float("start_angle")], np.float32)
    return LoadedScenario(init.get_string("map_name"), pose, humans, files)


# ---- batched world descriptors -------------------------------------------------------------
@dataclass
class WorldBatch:
    """W seeded variants of one scenario: what World.SetPose / World.SetHumans take.
    Variant 0 of the whole job (world_index_base + w == 0) is the scenario itself."""
    map_name: str
    robot_pose: np.ndarray      # [W][1][3] x, y, angle
    human_start: np.ndarray     # [W][H][3]
    human_goal: np.ndarray      # [W][H][3]
    world_index_base: int = 0

    @property
    def num_worlds(self):
        return self.robot_pose.shape[0]

    @property
    def num_humans(self):
        return self.human_start.shape[1]

    def save(self, path):
        np.savez_compressed(path, map_name=np.array(self.map_name), robot_pose=self.robot_pose,
                            human_start=self.human_start, human_goal=self.human_goal,
                            world_index_base=np.array(self.world_index_base))

    @staticmethod
    def load(path):
        z = np.load(path)
        return WorldBatch(str(z["map_name"]), z["robot_pose"], z["human_start"], z["human_goal"],
                          int(z["world_index_base"]))


def _as_scenario_arrays(scn):
    if isinstance(scn, LoadedScenario):
        pose = np.asarray(scn.start_pose, np.float32)
        s = np.array([[h.start_x, h.start_y, h.start_theta] for h in scn.humans], np.float32).reshape(-1, 3)
        g = np.array([[h.goal_x, h.goal_y, h.goal_theta] for h in scn.humans], np.float32).reshape(-1, 3)
        return scn.map_name, pose, s, g
    pose = np.array([scn.start_x, scn.start_y, scn.start_angle], np.float32)
    sg = scn.human_start_goal_positions
    s = np.array([[p[0][0], p[0][1], 0.0] for p in sg], np.float32).reshape(-1, 3)   # hu_start_theta = 0.
    g = np.array([[p[1][0], p[1][1], 0.0] for p in sg], np.float32).reshape(-1, 3)   # hu_goal_theta = 0.
    return scn.map_name, pose, s, g


def world_batch(scn, num_worlds, seed=0, pos_jitter=0.5, angle_jitter=np.pi, human_jitter=0.5,
                world_index_base=0):
    """`num_worlds` variants of a Scenario / LoadedScenario.  World g = world_index_base + w
    draws from its own generator keyed by (seed, g), so a shard of a job gets exactly the
    worlds the unsharded job would (the same rule as the Philox counters of the step kernel,
    DESIGN.md 6).  g == 0 is the unperturbed scenario; the others move the robot start by
    U[-pos_jitter, pos_jitter]^2 and U[-angle_jitter, angle_jitter], every human start and
    goal by U[-human_jitter, human_jitter]^2."""
    name, pose, hs, hg = _as_scenario_arrays(scn)
    W, H = int(num_worlds), hs.shape[0]
    rp = np.empty((W, 1, 3), np.float32)
    s = np.empty((W, H, 3), np.float32)
    g = np.empty((W, H, 3), np.float32)
    for w in range(W):
        gi = world_index_base + w
        rp[w, 0], s[w], g[w] = pose, hs, hg
        if gi == 0:
            continue
        rng = np.random.default_rng([int(seed), gi])
        rp[w, 0, :2] += rng.uniform(-pos_jitter, pos_jitter, 2).astype(np.float32)
        rp[w, 0, 2] += np.float32(rng.uniform(-angle_jitter, angle_jitter))
        s[w, :, :2] += rng.uniform(-human_jitter, human_jitter, (H, 2)).astype(np.float32)
        g[w, :, :2] += rng.uniform(-human_jitter, human_jitter, (H, 2)).astype(np.float32)
    return WorldBatch(name, rp, s, g, int(world_index_base))


def make_world(batch, vmap, robot, human=None, **world_kwargs):
    """A batched World (device) holding `batch`: one robot of type `robot` per world at the
    scenario's start pose (simulator.cpp:177-181) and H humans walking start <-> goal
    (human_object.cpp:131-161).  `human`: HumanParams template (default config/human_config.lua)."""
    from . import api
    hp = human if human is not None else api.default_human()
    world = api.World(vmap, [robot], batch.num_worlds, humans=[hp] * batch.num_humans,
                      world_index_base=batch.world_index_base, **world_kwargs)
    world.SetPose(batch.robot_pose)
    if batch.num_humans:
        world.SetHumans(batch.human_start, batch.human_goal)
    return world


def main(argv=None):
    """python -m ut_multirobot_sim_b200.scenario --config-file example.yml [--out-root DIR]
    [--worlds W --seed S --batch-out FILE.npz]   (generate_config.py's command line + batching)"""
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--config-file", default="example.yml", type=str)
    ap.add_argument("--out-root", default=".")
    ap.add_argument("--worlds", type=int, default=0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--batch-out", default=None)
    a = ap.parse_args(argv)
    scn = load_yaml(a.config_file)
    print(generate_config(scn, a.out_root))
    if a.worlds > 0:
        b = world_batch(scn, a.worlds, a.seed)
        out = a.batch_out or os.path.join(a.out_root, scenario_dir(scn.prefix), "world_batch.npz")
        b.save(out)
        print(out)


if __name__ == "__main__":
    main()
