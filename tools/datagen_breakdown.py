"""Where a recycled data-generation round spends its time: wall seconds inside env.reset / sort_colours / env.step
(torch.cuda.synchronize around each), for tuning the generator.  python tools/datagen_breakdown.py [envs] [episodes]"""
import sys, time, collections
sys.path.insert(0, ".")
import torch
from mujoco_robot_environments_b200 import config as C
from mujoco_robot_environments_b200.env import BatchedRearrangementEnv
from mujoco_robot_environments_b200 import data_generation as dg

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
episodes = int(sys.argv[2]) if len(sys.argv) > 2 else n
env = BatchedRearrangementEnv(cfg=C.compose(dg.COLOUR_SPLITTER_OVERRIDES), num_envs=n, strict=False)
acc, cnt, launches = collections.Counter(), collections.Counter(), collections.Counter()


def timed(name, fn):
    def wrap(*a, **k):
        torch.cuda.synchronize(); t0 = time.perf_counter(); l0 = env.physics.launches
        r = fn(*a, **k)
        torch.cuda.synchronize(); acc[name] += time.perf_counter() - t0; cnt[name] += 1; launches[name] += env.physics.launches - l0
        return r
    return wrap


env.reset = timed("reset", env.reset)
env.sort_colours = timed("sort_colours", env.sort_colours)
env.step = timed("step", env.step)
env._place_props = timed("reset/place_props", env._place_props) if hasattr(env, "_place_props") else None
t0 = time.perf_counter()
out = dg.rollout_recycled(env, episodes, 10)
torch.cuda.synchronize()
print("episodes", out["success"].shape[0], "wall", round(time.perf_counter() - t0, 1))
for k in acc:
    print("%-20s calls %4d  launches %6d  seconds %7.1f" % (k, cnt[k], launches[k], acc[k]))
