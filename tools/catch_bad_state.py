"""Catch the rare fp32 blow-ups of the recycled data generator (profiles/r01_datagen_recycle_8192.json: badqpos/badqvel/badqacc = 2
of 16 978 episodes) and bisect them to the physics step.

Every run_controller phase of the scripted pick / place is preceded by a copy of the state batch; when a scene raises a bad-state
flag during a phase, the phase is replayed from the copy for that scene alone (the kernels are bitwise reproducible and
independent of batch position) tick by tick and then step by step, and the last good state + the command are written to
gpurun_out/bad_state_<k>.json together with the contact / solver figures around the event.

  python tools/catch_bad_state.py [--envs 8192] [--rounds 12] [--max-catch 6]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from mujoco_robot_environments_b200 import config as cfglib  # noqa: E402
from mujoco_robot_environments_b200.data_generation import COLOUR_SPLITTER_OVERRIDES  # noqa: E402
from mujoco_robot_environments_b200.env import BatchedRearrangementEnv  # noqa: E402
from mujoco_robot_environments_b200.physics import BatchedPhysics  # noqa: E402

BAD = 7


def replay(env, state, size, active, target, grip, ticks, k):
    """one scene, fp32, from `state`: find the first tick, then the first step, that raises a bad-state flag"""
    p = BatchedPhysics(1, dtype=torch.float32, device=env.device)
    p.set_env_params(size[None], active[None])
    p.state.copy_(state[None]); p.status(clear=True)
    tg, gr = target[None].contiguous(), grip[None].contiguous()
    hist = []
    bad_tick = None
    for t in range(ticks):
        before = p.state.clone()
        p.run_controller(tg, gr, 1)
        f = int(p.status().item())
        hist.append(dict(tick=t, nefc=int(p.data.nefc.item()), ncon=int(p.data.ncon.item()), niter=int(p.derived("NITER")[0, 0].item()), flags=f,
                         max_qvel=float(p.qvel.abs().max().item())))
        if f & BAD:
            bad_tick = t
            break
    rec = dict(k=k, ticks=ticks, bad_tick=bad_tick, history=hist[-12:])
    if bad_tick is not None:
        # step by step inside the bad tick: the OSC command of that tick, then 5 x step(1)
        p.state.copy_(before); p.status(clear=True)
        p.run_controller(tg, gr, 0)
        ctrl = p.derived("CTRL")[:, :8].clone()
        p.set_control(ctrl)
        steps = []
        for s in range(5):
            good = p.state.clone()
            p.forward(full_contacts=False)
            con = {kk: v[0, :int(p.data.ncon.item())].cpu().numpy().tolist() for kk, v in p.data.contact.items() if kk in ("dist", "geom1", "geom2")}
            p.state.copy_(good)
            p.step(1)
            f = int(p.status().item())
            steps.append(dict(step=s, flags=f, nefc=int(p.data.nefc.item()), niter=int(p.derived("NITER")[0, 0].item()), max_qacc=float(p.data.qacc.abs().max().item()),
                              max_qvel=float(p.qvel.abs().max().item()), contacts_before=con))
            if f & BAD:
                rec["last_good_state"] = good[0].cpu().numpy().astype(np.float64).tolist()
                rec["ctrl"] = ctrl[0].cpu().numpy().astype(np.float64).tolist()
                break
        rec["steps"] = steps
        rec["tick_start_state"] = before[0].cpu().numpy().astype(np.float64).tolist()
    rec.update(size=size.cpu().numpy().astype(np.float64).tolist(), active=active.cpu().numpy().tolist(), target=target.cpu().numpy().astype(np.float64).tolist(), grip=int(grip.item()),
               phase_start_state=state.cpu().numpy().astype(np.float64).tolist())
    p.close()
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=8192); ap.add_argument("--rounds", type=int, default=12); ap.add_argument("--max-catch", type=int, default=6); ap.add_argument("--seconds", type=float, default=1e9)
    a = ap.parse_args()
    cfg = cfglib.compose(COLOUR_SPLITTER_OVERRIDES)
    env = BatchedRearrangementEnv(cfg=cfg, num_envs=a.envs, dtype=torch.float32, strict=False, seed=1)
    env.reset()
    p, arm, osc, mm = env.physics, env._robot, env._osc, env._minmax
    caught = []
    seen = torch.zeros(a.envs, dtype=torch.bool, device=env.device)
    orig = arm.run_controller

    def hooked(duration):
        before = p.state.clone(); tg = osc.target13.clone(); gr = mm.bits.clone()
        r = orig(duration)
        new = ((p.status() & BAD) != 0) & ~seen
        if bool(new.any()):
            for e in torch.nonzero(new).flatten().tolist():
                seen[e] = True
                if len(caught) < a.max_catch:
                    rec = replay(env, before[e], p.prop_size[e].clone(), p.prop_active[e].clone(), tg[e], gr[e], arm.ticks_for(duration), len(caught))
                    rec["scene"] = e
                    caught.append(rec)
                    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
                    json.dump(rec, open(os.path.join(ROOT, "gpurun_out", "bad_state_%d.json" % rec["k"]), "w"))
                    print("caught scene %d: bad tick %s, steps %s" % (e, rec["bad_tick"], [(s["step"], s["flags"], s["nefc"], s["niter"], "%.3g" % s["max_qacc"]) for s in rec.get("steps", [])]), flush=True)
        return r
    arm.run_controller = hooked
    episodes = 0
    import time
    t_end = time.time() + a.seconds
    for rnd in range(a.rounds):
        if time.time() > t_end or len(caught) >= a.max_catch:
            break
        in_progress, pick, place = env.sort_colours()
        act = in_progress & ~env.failed
        ended = ~act
        if bool(ended.any()):
            episodes += int(ended.sum()); seen[ended] = False
            env.reset(mask=ended)
            continue
        env.step({"pose": pick.clone()}); env.step({"pose": place.clone()})
        print("round %d: flagged so far %d, caught %d" % (rnd, int(((p.status() & BAD) != 0).sum()), len(caught)), flush=True)
    print(json.dumps(dict(episodes_finished=episodes, caught=len(caught))))


if __name__ == "__main__":
    main()
