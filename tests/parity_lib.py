"""Shared machinery of the configuration-level parity tests (tests/test_gpu_config_parity.py) and of
tools/parity_report.py: take scene states out of a large GPU batch (BASELINE configs[1] / configs[2], or an env in the
middle of a pick / place), replay a seeded sample of them on the GPU and on the CPU oracle, and MEASURE the differences.

Test infrastructure: the oracle is the checker.  Error definitions (what DESIGN.md section 2 quotes):
  abs   = max_i |gpu_i - oracle_i|                       over the dofs of one scene
  rel   = abs / max(floor, max_i |oracle_i|)              floor = 1 (qpos rad/m, qvel, qacc in SI units)
and the figure reported per field is the maximum (and the median) over the sampled scenes.
"""
import json
import os

import numpy as np

import scenarios as sc
from orc import Oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def torch_():
    import torch
    return torch


def pick_sample(nefc, niter, k, seed=0):
    """k scene indices: the k/4 with the most constraint rows, the k/4 with the most Newton iterations in the last step, the
    rest drawn uniformly (seeded) from the others."""
    n = len(nefc)
    k = min(k, n)
    top = list(np.argsort(-nefc, kind="stable")[:k // 4]) + list(np.argsort(-niter, kind="stable")[:k // 4])
    chosen = list(dict.fromkeys(int(i) for i in top))
    rest = np.setdiff1d(np.arange(n), np.array(chosen, dtype=np.int64))
    rng = np.random.default_rng(seed)
    chosen += [int(i) for i in rng.choice(rest, size=k - len(chosen), replace=False)]
    return np.array(chosen, dtype=np.int64)


def _err(got, want, floor=1.0):
    a = float(np.abs(got - want).max(initial=0.0))
    return a, a / max(floor, float(np.abs(want).max(initial=0.0)))


def contact_key(geom, excl=None):
    """multiset of (geom1, geom2) pairs as a sorted list"""
    return sorted((int(a), int(b)) for a, b in geom)


def compare_sample(dtype, state, size, active, ctrl=None, target=None, grip=None, horizon=100, gains=None):
    """state [k,128] (values of `dtype`), per-scene prop params, and either ctrl [k,8] (physics only) or target [k,13] + grip [k]
    (OSC in the loop).  Returns a dict of measured differences GPU vs oracle:
      single step: qacc / qvel / qpos (abs, rel per scene), contact sets at the snapshot state;
      `horizon` steps: max |dqpos| at steps 1, 10, 25, 50, 100 (physics) or after horizon/5 control ticks (OSC)."""
    torch = torch_()
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    k = state.shape[0]
    tdt = torch.float64 if dtype == "f64" else torch.float32
    st64 = np.asarray(state, np.float64)
    p = BatchedPhysics(k, dtype=tdt)
    p.set_env_params(torch.tensor(size), torch.tensor(active))

    def load():
        p.state.copy_(torch.tensor(st64, dtype=tdt)); p.status(clear=True)

    o = Oracle(sc.default_blob())
    osc = target is not None
    # ---- the command held over the single step: given ctrl, or the OSC torque evaluated at the snapshot state
    load()
    if osc:
        tg = torch.tensor(np.asarray(target, np.float64), dtype=tdt, device=p.device); gr = torch.tensor(np.asarray(grip, np.uint8), device=p.device)
        p.run_controller(tg, gr, 0)                      # zero-tick launch: OSC.compute_control_output() for the current state
        ctrl_gpu = p.derived("CTRL").cpu().numpy().astype(np.float64)[:, :sc.NU]
    else:
        ctrl_gpu = np.asarray(ctrl, np.float64)
    # ---- contact list at the snapshot state + one step
    load()
    p.forward(full_contacts=False); torch.cuda.synchronize()
    ncon = p.data.ncon.cpu().numpy(); con = {kk: v.cpu().numpy() for kk, v in p.data.contact.items()}
    nefc_gpu = p.data.nefc.cpu().numpy()
    load()
    p.set_control(torch.tensor(ctrl_gpu)); p.step(1); torch.cuda.synchronize()
    after = p.state.cpu().numpy().astype(np.float64); qacc = p.data.qacc.cpu().numpy().astype(np.float64)
    niter_gpu = p.derived("NITER")[:, 0].cpu().numpy()
    flags1 = p.status().cpu().numpy().copy()
    rows = []
    tau_err = []
    for e in range(k):
        sc.oracle_load(o, st64[e], size[e], active[e]); o.set_option("full_contacts", 0)
        o.step1()
        oc = o.contacts()
        if osc:
            tau = o.osc_torque(np.asarray(target[e], np.float64))
            tau_err.append(_err(ctrl_gpu[e, :7], tau))
            # the oracle steps with ITS OWN command below (end-to-end tick parity); the single step isolates mj_step: same command
        o.ctrl[:sc.NU] = ctrl_gpu[e]
        o.step2()
        want_qacc = o.qacc[:sc.NV].copy()
        act = np.concatenate([np.ones(15, bool), np.repeat(active[e].astype(bool), 6)])
        a_qacc = _err(qacc[e, :sc.NV][act], want_qacc[act])
        dq = np.abs(qacc[e, :sc.NV] - want_qacc) * act
        grp = dict(arm=float(dq[:7].max()), gripper=float(dq[7:15].max()), props=float(dq[15:].max()))
        a_qvel = _err(after[e, sc.REC_QVEL:sc.REC_QVEL + sc.NV][act], o.qvel[:sc.NV][act])
        a_qpos = _err(after[e, :sc.NQ], o.qpos[:sc.NQ])
        g_keys = contact_key(np.stack([con["geom1"][e, :ncon[e]], con["geom2"][e, :ncon[e]]], 1)) if ncon[e] else []
        o_keys = contact_key(oc["geom"])
        same = g_keys == o_keys
        ddist = 0.0
        if same and ncon[e]:
            ddist = float(np.abs(np.sort(con["dist"][e, :ncon[e]]) - np.sort(oc["dist"])).max())
        rows.append(dict(qacc_abs=a_qacc[0], qacc_rel=a_qacc[1], qvel_abs=a_qvel[0], qvel_rel=a_qvel[1], qpos_abs=a_qpos[0], qpos_rel=a_qpos[1],
                         ncon=int(ncon[e]), ncon_oracle=len(oc["dist"]), nefc=int(nefc_gpu[e]), nefc_oracle=int(o.dim("nefc")), niter=int(niter_gpu[e]),
                         niter_oracle=int(o.dim("solver_iter")), contacts_identical=bool(same), contact_dist_abs=ddist, qacc_scale=float(np.abs(want_qacc).max()),
                         qacc_abs_arm=grp["arm"], qacc_abs_gripper=grp["gripper"], qacc_abs_props=grp["props"], worst_dof=int(np.argmax(dq))))
        if not same:      # keep what is needed to replay the scene on the CPU (tests/hostemu vs oracle)
            rows[-1]["replay"] = dict(state=st64[e].tolist(), size=np.asarray(size[e]).tolist(), active=np.asarray(active[e]).tolist(), ctrl=ctrl_gpu[e].tolist(),
                                      gpu_pairs=g_keys, oracle_pairs=o_keys, gpu_dist=con["dist"][e, :ncon[e]].tolist(), oracle_dist=oc["dist"].tolist())
    # ---- horizon
    marks = [m for m in (1, 10, 25, 50, 100, horizon) if m <= horizon]
    marks = sorted(set(marks))
    growth = {}
    if osc:
        load()
        ticks = max(1, horizon // 5)
        conv = p.run_controller(tg, gr, ticks).cpu().numpy(); torch.cuda.synchronize()
        fin = p.state.cpu().numpy().astype(np.float64)
        d, conv_same = [], 0
        for e in range(k):
            sc.oracle_load(o, st64[e], size[e], active[e]); o.set_option("full_contacts", 0)
            c = o.run_controller(np.asarray(target[e], np.float64), int(grip[e]), ticks)
            d.append(float(np.abs(fin[e, :sc.NQ] - o.qpos[:sc.NQ]).max())); conv_same += int(bool(conv[e]) == c)
        growth[str(5 * ticks)] = dict(max=float(np.max(d)), median=float(np.median(d)))
        last_d = d
        extra = dict(converged_bits_equal=conv_same, osc_tau_abs_max=float(max(t[0] for t in tau_err)), osc_tau_rel_max=float(max(t[1] for t in tau_err)))
    else:
        load(); p.set_control(torch.tensor(ctrl_gpu))
        snaps, prev = {}, 0
        for m in marks:
            p.step(m - prev); prev = m
            snaps[m] = p.state.cpu().numpy().astype(np.float64)[:, :sc.NQ]
        d = {m: [] for m in marks}
        for e in range(k):
            sc.oracle_load(o, st64[e], size[e], active[e]); o.set_option("full_contacts", 0); o.ctrl[:sc.NU] = ctrl_gpu[e]
            done = 0
            for m in marks:
                for _ in range(m - done):
                    o.step()
                done = m
                d[m].append(float(np.abs(snaps[m][e] - o.qpos[:sc.NQ]).max()))
        growth = {str(m): dict(max=float(np.max(v)), median=float(np.median(v))) for m, v in d.items()}
        extra = {}
        last_d = d[marks[-1]]
    flagsH = p.status().cpu().numpy()
    for e in range(k):      # scenes that drift apart keep what is needed to replay them on the CPU
        rows[e]["drift"] = last_d[e]
        if last_d[e] > (1e-9 if dtype == "f64" else 1e-3) and "replay" not in rows[e]:
            rows[e]["replay"] = dict(state=st64[e].tolist(), size=np.asarray(size[e]).tolist(), active=np.asarray(active[e]).tolist(), ctrl=ctrl_gpu[e].tolist(),
                                     target=None if target is None else np.asarray(target[e]).tolist(), grip=None if grip is None else int(grip[e]))
    p.close()
    agg = lambda key: dict(max=float(max(r[key] for r in rows)), median=float(np.median([r[key] for r in rows])))
    res = dict(dtype=dtype, scenes=k, single_step={f: agg(f) for f in ("qacc_abs", "qacc_rel", "qvel_abs", "qvel_rel", "qpos_abs", "qpos_rel", "contact_dist_abs", "qacc_abs_arm", "qacc_abs_gripper", "qacc_abs_props")},
               contact_sets_identical=int(sum(r["contacts_identical"] for r in rows)), nefc_equal=int(sum(r["nefc"] == r["nefc_oracle"] for r in rows)),
               nefc_max=int(max(r["nefc"] for r in rows)), niter_max=int(max(r["niter"] for r in rows)),
               qpos_drift=growth, drift_scenes_above=dict(thresh=(1e-9 if dtype == "f64" else 1e-3), count=int(sum(x > (1e-9 if dtype == "f64" else 1e-3) for x in last_d))), flags_single_step=int(np.bitwise_or.reduce(flags1)), flags_horizon=int(np.bitwise_or.reduce(flagsH)), rows=rows)
    res.update(extra)
    return res


def save_report(name, res):
    """best effort: gpurun_out/ travels back from the GPU box"""
    try:
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_%s.json" % name), "w") as f:
            json.dump(res, f, indent=1)
            f.write("\n")
    except OSError:
        pass


def summary_line(name, r):
    s = r["single_step"]
    return ("%s [%s, %d scenes, nefc<=%d, niter<=%d]: single step qacc abs %.3e rel %.3e | qvel rel %.3e | qpos rel %.3e | contact sets identical %d/%d | "
            "qpos drift %s" % (name, r["dtype"], r["scenes"], r["nefc_max"], r["niter_max"], s["qacc_abs"]["max"], s["qacc_rel"]["max"], s["qvel_rel"]["max"],
                               s["qpos_rel"]["max"], r["contact_sets_identical"], r["scenes"], {k: "%.2e" % v["max"] for k, v in r["qpos_drift"].items()}))


# ------------------------------------------------------------------------------------------ where the states come from
def c2_batch(n=4096, launches=5, steps=100, seed=1234):
    """BASELINE configs[1] exactly as bench.py --workload c2 drives it: fp64, physics only, Halton torque noise around the home
    gravity load, `launches` launches of `steps` steps -> (physics handle, last ctrl [n,8] numpy, size, active)"""
    torch = torch_()
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    st, size, active = sc.make_states(n, seed=seed, settled=True)
    p = BatchedPhysics(n, dtype=torch.float64)
    p.state.copy_(torch.tensor(st)); p.set_env_params(torch.tensor(size), torch.tensor(active))
    p.forward(full_contacts=False); torch.cuda.synchronize()
    bias = p.data.qfrc_bias[0, :7].cpu().numpy().astype(np.float64)
    ctrl = None
    for i in range(launches):
        ctrl = np.tile(sc.halton_ctrl(min(n, 256), i, bias), ((n + 255) // 256, 1))[:n]
        p.set_control(torch.tensor(ctrl)); p.step(steps)
    torch.cuda.synchronize()
    return p, ctrl, size, active


def c3_batch(n=16384, launches=20, ticks=20, seed=1234):
    """BASELINE configs[2] as bench.py drives it: fp32, OSC in the loop, a new target per launch"""
    torch = torch_()
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    st, size, active = sc.make_states(n, seed=seed, settled=True)
    p = BatchedPhysics(n, dtype=torch.float32)
    p.state.copy_(torch.tensor(st, dtype=torch.float32)); p.set_env_params(torch.tensor(size), torch.tensor(active))
    tg = grip = None
    for i in range(launches):
        tg, grip = sc.random_targets(n, seed=i)
        p.run_controller(torch.tensor(tg), torch.tensor(grip), ticks)
    torch.cuda.synchronize()
    return p, tg, grip, size, active


def snapshot(p, idx):
    return p.state.cpu().numpy().astype(np.float64)[idx]


def batch_sample(p, k, seed=0):
    nefc = p.data.nefc.cpu().numpy(); niter = p.derived("NITER")[:, 0].cpu().numpy()
    return pick_sample(nefc, niter, k, seed)


def pick_place_snapshots(dtype, n=32, seed=3):
    """An env stepped through the reference's scripted pick and place (rearrangement.py:358-440) and stopped three times:
    'grasp' 0.5 s into closing the gripper on the cube, 'lift' 1.0 s into the pre-grasp ascent with the cube in hand,
    'release' 0.15 s into opening the gripper above the place pose.  -> {name: (state, target13, grip, size, active)}"""
    torch = torch_()
    from mujoco_robot_environments_b200.controllers import _mulquat
    from mujoco_robot_environments_b200.env import BatchedRearrangementEnv
    tdt = torch.float64 if dtype == "f64" else torch.float32
    env = BatchedRearrangementEnv(num_envs=n, dtype=tdt, strict=False, seed=seed)
    ts = env.reset()
    p, osc, arm, mm = env.physics, env._osc, env._robot, env._minmax
    pos0 = ts.observation["prop_pos"][:, 0].clone(); quat0 = ts.observation["prop_quat"][:, 0]
    half = torch.atan2(quat0[:, 3], quat0[:, 0])
    half = torch.where(half > math_pi() / 4, half - math_pi() / 2, half)          # prop_pick's min(theta, theta - 90 deg) rule
    down = torch.tensor([0.0, 0.0, 1.0, 0.0], device=p.device, dtype=tdt)
    qz = torch.stack([torch.cos(half), torch.zeros_like(half), torch.zeros_like(half), torch.sin(half)], 1)
    grasp = _mulquat(qz, down[None].expand(n, 4))
    zeros = torch.zeros(3, device=p.device, dtype=tdt)
    pose = torch.cat([pos0, grasp], 1); pose[:, 2] = 0.575
    pre = pose.clone(); pre[:, 2] = 0.9
    out = {}

    def snap(name):
        torch.cuda.synchronize()
        out[name] = (p.state.cpu().numpy().astype(np.float64), osc.target13.cpu().numpy().astype(np.float64), mm.bits.cpu().numpy().copy(),
                     env.prop_half_size.cpu().numpy().astype(np.float64), env.has_prop.cpu().numpy().astype(np.uint8))
    osc.set_target(position=pre[:, :3], velocity=zeros, quat=pre[:, 3:], angular_velocity=zeros); arm.run_controller(2.0)
    osc.set_target(position=pose[:, :3]); arm.run_controller(2.0)
    mm.status = "max"; arm.run_controller(0.5); snap("grasp"); arm.run_controller(0.5)
    osc.set_target(position=pre[:, :3]); arm.run_controller(1.0); snap("lift"); arm.run_controller(1.0)
    osc.set_target(position=env.eef_home_pose, quat=down); arm.run_controller(2.0)
    place = pose.clone(); place[:, 0] = 0.45; place[:, 1] = torch.linspace(-0.3, 0.3, n, device=p.device, dtype=tdt)
    ppre = place.clone(); ppre[:, 2] = 0.9
    osc.set_target(position=ppre[:, :3], quat=ppre[:, 3:]); arm.run_controller(2.0)
    osc.set_target(position=place[:, :3]); arm.run_controller(2.0)
    mm.status = "min"; arm.run_controller(0.15); snap("release")
    lifted = int((out["lift"][0][:, 15 + 2] > 0.45).sum())
    env.close()
    return out, lifted


def math_pi():
    return float(np.pi)
