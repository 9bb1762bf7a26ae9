"""Per-stage cycle breakdown of the step kernel (developer aid).  Needs a library built with -DB2P_PROF:

  python -c "from mujoco_robot_environments_b200 import ffi; ffi.build(force=True, extra=['-DB2P_PROF'])"
  python tools/stage_profile.py [--workload c3|c2] [--envs N] [--ticks T]      # on the GPU box
  (rebuild without the flag afterwards: the counters cost a few percent)
"""
import argparse, ctypes as C, os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import scenarios as sc
from mujoco_robot_environments_b200 import ffi
from mujoco_robot_environments_b200.physics import BatchedPhysics

NAMES = ["fk", "dyn(crb,rne,M chol)", "collide", "make_rows", "osc", "smooth", "solve_init", "assemble_H", "chol_H", "solve_H", "line_search", "update", "integrate", "barrier", "misc", "niter", "nstep", "  collide: broad phase", "  collide: narrow phase"]
ap = argparse.ArgumentParser(); ap.add_argument("--workload", default="c3"); ap.add_argument("--envs", type=int, default=16384); ap.add_argument("--ticks", type=int, default=20)
a = ap.parse_args()
dt = torch.float32 if a.workload == "c3" else torch.float64
n = a.envs if a.workload == "c3" else min(a.envs, 4096)
st, size, active = sc.make_states(n, seed=1234, settled=True)
p = BatchedPhysics(n, dtype=dt); p.state.copy_(torch.tensor(st, dtype=dt)); p.set_env_params(torch.tensor(size), torch.tensor(active))
tg, grip = sc.random_targets(n, seed=0)
def go():
    if a.workload == "c3": p.run_controller(torch.tensor(tg), torch.tensor(grip), a.ticks)
    else: p.step(5 * a.ticks)
go(); go()
buf = (C.c_ulonglong * 24)()
ffi.check(p._L.b2p_prof_read(p._h, buf, 1))
go()
ffi.check(p._L.b2p_prof_read(p._h, buf, 1))
v = list(buf); nstep = max(1, v[16]); tot = sum(v[:15])
print("run_env total cycles/env-step: %.0f" % (v[19] / nstep))
print("env-steps profiled: %d   mean Newton iterations/step: %.2f   cycles/env-step (warp timeline): %.0f" % (nstep, v[15] / nstep, tot / nstep))
NAMES += ["total", "  sub0", "  sub1", "  sub2", "  sub3"]
for i in list(range(15)) + [17, 18, 20, 21, 22, 23]:
    print("  %-22s %9.0f cycles/step  %5.1f %%" % (NAMES[i], v[i] / nstep, 100.0 * v[i] / max(1, tot)))
