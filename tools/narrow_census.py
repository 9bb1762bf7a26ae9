"""Narrow-phase census of c3-like scenes on the host build of the kernel templates (-DB2P_HOST_COUNTERS): candidates, box-box calls,
MPR queries, support calls and hill-climb steps per physics step.  Developer aid (python tools/narrow_census.py [n_scenes] [ticks])."""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenarios as sc
import hostemu

so = os.path.join(ROOT, "tests", "_hostemu", "libb2p_hostemu_cnt.so")
subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-DB2P_HOST_COUNTERS", "-o", so, os.path.join(ROOT, "tests", "hostemu", "b2p_hostemu.cpp")])
L = C.CDLL(so)
L.emu_create.restype = C.c_void_p; L.emu_create.argtypes = [C.c_char_p, C.c_size_t, C.c_int]
L.emu_run.argtypes = [C.c_void_p, C.c_int, C.POINTER(hostemu.EmuArgs)]
L.emu_destroy.argtypes = [C.c_void_p, C.c_int]
hostemu._lib = L
n, ticks = int(sys.argv[1]) if len(sys.argv) > 1 else 8, int(sys.argv[2]) if len(sys.argv) > 2 else 40
st, size, active = sc.make_states(n, seed=1234)
tg, grip = sc.random_targets(n, seed=0)
e = hostemu.HostEmu(sc.default_blob(), np.float64)
cnt = (C.c_long * 6)()
L.emu_counters(cnt, 1)
state = st.copy()
e.run(hostemu.MODE_OSC, state, n_steps=ticks, control_steps=5, target=tg, grip=grip, prop_size=size, prop_active=active, want_out=False)
L.emu_counters(cnt, 0)
names = ["support calls", "climb steps", "mpr queries", "candidates", "box-box calls", "steps"]
steps = max(1, cnt[5])
for k, v in zip(names, cnt):
    print("%-14s %10d  per step %.2f" % (k, v, v / steps))
pm = (C.c_long * 1024)(); L.emu_pair_mpr(pm)
from mujoco_robot_environments_b200.physics import load_default_blob
from mujoco_robot_environments_b200.compile import blob as B
blob, names = load_default_blob(); _, _, arr = B.unpack(blob)
top = sorted(range(1024), key=lambda i: -pm[i])[:16]
print("MPR queries per step by pair:")
for i in top:
    if pm[i]:
        print("  %.3f  %s  |  %s" % (pm[i] / steps, names["geom"][arr["pair_geom1"][i]], names["geom"][arr["pair_geom2"][i]]))
