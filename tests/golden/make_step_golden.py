"""Generate tests/golden/step_golden.npz: regression vectors of the CPU oracle for the physics half of the path.

These are NOT MuJoCo outputs (MuJoCo is not installable here; parity stays "unpinned"): they freeze
the oracle's current behaviour on seeded scenes so that any later change to oracle/b2p_oracle.c shows up as a diff
of committed numbers, and they give the CUDA tests a fixture that does not need the oracle at run time.
tools/dump_mujoco_golden.py writes the same schema from real MuJoCo where the wheel exists.

Schema (S scenes, K recorded steps):  state0 [S,128], prop_size [S,4,3], prop_active [S,4], ctrl [S,8],
  qpos [S,K,43], qvel [S,K,39], qacc [S,K,39], ncon [S,K], nefc [S,K], contact_geom [S,K,48,2] (-1 padded),
  contact_dist [S,K,48];  OSC rollout: target [S,13], grip [S], osc_qpos [S,43] after 20 ticks, osc_converged [S].

  python tests/golden/make_step_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import scenarios as sc  # noqa: E402
from orc import Oracle  # noqa: E402

S, K, MAXCON = 6, 25, 48


def main():
    o = Oracle(sc.default_blob())
    st, size, active = sc.make_states(S, seed=77, settled=False)
    # scene 4/5: coupled solver blocks (gripper squeezing a prop; stacked props)
    for e, kind in ((4, "grasp"), (5, "stack")):
        a, b, c = sc.coupled_states(kind)
        st[e], size[e], active[e] = a[0], b[0], c[0]
    sc.oracle_load(o, st[0], size[0], active[0]); o.step1(); bias = o.qfrc_bias[:7].copy()
    ctrl = sc.halton_ctrl(S, 0, bias)
    for e in (4, 5):
        sc.oracle_load(o, st[e], size[e], active[e]); o.step1(); ctrl[e, :7] = o.qfrc_bias[:7]; ctrl[e, 7] = 255.0
    out = dict(state0=st, prop_size=size, prop_active=active, ctrl=ctrl, qpos=np.zeros((S, K, sc.NQ)), qvel=np.zeros((S, K, sc.NV)),
               qacc=np.zeros((S, K, sc.NV)), ncon=np.zeros((S, K), np.int32), nefc=np.zeros((S, K), np.int32),
               contact_geom=-np.ones((S, K, MAXCON, 2), np.int32), contact_dist=np.zeros((S, K, MAXCON)))
    for e in range(S):
        sc.oracle_load(o, st[e], size[e], active[e]); o.set_option("full_contacts", 0); o.ctrl[:] = ctrl[e]
        for k in range(K):
            o.step()
            out["qpos"][e, k] = o.qpos[:sc.NQ]; out["qvel"][e, k] = o.qvel[:sc.NV]; out["qacc"][e, k] = o.qacc[:sc.NV]
            c = o.contacts(); n = len(c["dist"])
            out["ncon"][e, k] = n; out["nefc"][e, k] = o.dim("nefc")
            out["contact_geom"][e, k, :n] = c["geom"]; out["contact_dist"][e, k, :n] = c["dist"]
    tg, grip = sc.random_targets(S, seed=78)
    out["target"], out["grip"] = tg, grip
    out["osc_qpos"] = np.zeros((S, sc.NQ)); out["osc_converged"] = np.zeros(S, np.uint8)
    for e in range(S):
        sc.oracle_load(o, st[e], size[e], active[e]); o.set_option("full_contacts", 0)
        out["osc_converged"][e] = o.run_controller(tg[e], int(grip[e]), 20)
        out["osc_qpos"][e] = o.qpos[:sc.NQ]
    np.savez_compressed(os.path.join(HERE, "step_golden.npz"), **out)
    print("wrote step_golden.npz: ncon range %d..%d, nefc range %d..%d" % (out["ncon"].min(), out["ncon"].max(), out["nefc"].min(), out["nefc"].max()))


if __name__ == "__main__":
    main()
