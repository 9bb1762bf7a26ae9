"""Seeded synthetic scenes shared by the CPU and GPU parity tests (C2 / C3 inputs)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ARM_HOME = np.array([0, -0.785, 0, -2.356, 0, 1.571, 0.785])
NQ, NV, NU, NP = 43, 39, 8, 4
REC_QPOS, REC_QVEL, REC_QACC, REC_TIME, REC_SIZE = 0, 44, 84, 124, 128
DOWN_QUAT = np.array([0.0, 0.0, 1.0, 0.0])   # mju_mat2Quat(R_xyz(0, 180deg, 0))  (rearrangement.py:391-393)


def default_blob(tag="rearrangement_n4"):
    with open(os.path.join(ROOT, "mujoco_robot_environments_b200", "assets", tag + ".b2pm"), "rb") as f:
        return f.read()


def halton(i, base):
    f, r = 1.0, 0.0
    while i > 0:
        f /= base
        r += f * (i % base)
        i //= base
    return r


def make_states(n_env, seed=1234, settled=True, jitter=0.15):
    """-> (state [n,128], prop_size [n,4,3], prop_active [n,4]).  Arm near home, props on (or above) the table,
    >= 0.19 m apart (well outside the 0.15 margin), per-env N ~ U{2,3,4}, cube half-size ~ U(.015,.016)."""
    rng = np.random.default_rng(seed)
    st = np.zeros((n_env, REC_SIZE))
    size = np.zeros((n_env, NP, 3))
    active = np.zeros((n_env, NP), np.uint8)
    slots = [(x, y) for x in (0.36, 0.55) for y in (-0.36, -0.12, 0.12, 0.36)]
    for e in range(n_env):
        st[e, :7] = ARM_HOME + jitter * (rng.random(7) - 0.5)
        n = rng.integers(2, 5)
        pick = rng.permutation(len(slots))[:NP]
        for p in range(NP):
            hs = rng.uniform(0.015, 0.016)
            size[e, p] = hs
            active[e, p] = p < n
            x, y = slots[pick[p]]
            x += 0.02 * (rng.random() - 0.5); y += 0.02 * (rng.random() - 0.5)
            z = 0.4 + hs - (1.05e-4 if settled else -0.02)
            yaw = np.pi * rng.random()
            if not active[e, p]:
                x, y, z = 5.0 + p, 5.0, 0.2
            q = 15 + 7 * p
            st[e, q:q + 3] = (x, y, z)
            st[e, q + 3:q + 7] = (np.cos(yaw / 2), 0, 0, np.sin(yaw / 2))
    return st, size, active


def halton_ctrl(n_env, step, bias):
    """ctrl = qfrc_bias_home + 0.1*ctrlrange*(2*Halton(t, j+2)-1); gripper bit from Halton(t, 11) (model_test.py:39-54 pattern)."""
    rngs = np.array([87, 87, 87, 87, 12, 12, 12.0])
    primes = [2, 3, 5, 7, 11, 13, 17, 19]
    c = np.zeros((n_env, NU))
    for e in range(n_env):
        t = step + 1 + 977 * e
        for j in range(7):
            c[e, j] = bias[j] + 0.1 * rngs[j] * (2 * halton(t, primes[j]) - 1)
        c[e, 7] = 255.0 * (halton(step // 1000 + 1 + e, primes[7]) > 0.5)
    return c


def random_targets(n_env, seed=0):
    """OSC targets of the reference's own demo distribution (mujoco_controllers/osc.py:272-286)."""
    rng = np.random.default_rng(seed)
    t = np.zeros((n_env, 13))
    t[:, 0] = rng.uniform(0.25, 0.5, n_env); t[:, 1] = rng.uniform(-0.3, 0.3, n_env); t[:, 2] = rng.uniform(0.6, 0.7, n_env)
    t[:, 6:10] = DOWN_QUAT
    grip = (rng.random(n_env) > 0.5).astype(np.uint8)
    return t, grip


def oracle_load(o, st_row, size_row, active_row):
    o.reset()
    for p in range(NP):
        o.set_prop(p, size_row[p], bool(active_row[p]))
    o.qpos[:NQ] = st_row[REC_QPOS:REC_QPOS + NQ]
    o.qvel[:NV] = st_row[REC_QVEL:REC_QVEL + NV]
    o.qacc_warmstart[:NV] = st_row[REC_QACC:REC_QACC + NV]
    o.time[0] = st_row[REC_TIME]


def oracle_state(o):
    r = np.zeros(REC_SIZE)
    r[REC_QPOS:REC_QPOS + NQ] = o.qpos[:NQ]; r[REC_QVEL:REC_QVEL + NV] = o.qvel[:NV]
    r[REC_QACC:REC_QACC + NV] = o.qacc_warmstart[:NV]; r[REC_TIME] = o.time[0]
    return r


def coupled_states(kind="grasp", seed=0):
    """Scenes that exercise the coupled solver blocks (test infrastructure; uses the oracle as scene generator).
    grasp: gripper pre-closed around prop 0 hovering between the pads (tree <-> prop rows, box-box pad contacts);
    stack: prop 1 resting on prop 0 (prop <-> prop rows); grasp_stack: both (fill-in between the tree and prop 1)."""
    from orc import Oracle
    st, size, active = make_states(1, seed=seed, jitter=0.0)
    active[0, :] = (1, 1, 1, 0)
    o = Oracle(default_blob())
    park = st.copy()
    for p in range(NP):
        park[0, 15 + 7 * p:15 + 7 * p + 3] = (5.0 + p, 5.0, 0.2)
    oracle_load(o, park[0], size[0], np.zeros(NP, np.uint8)); o.set_option("full_contacts", 0)
    o.step1(); bias = o.qfrc_bias[:7].copy()
    pads = [18, 24]     # right_pad1, left_pad1 geoms (names.json)
    for _ in range(1500):
        o.ctrl[:7] = o.qfrc_bias[:7]; o.ctrl[7] = 255.0
        o.step()
        if abs(o.geom_xpos[pads[1]][0] - o.geom_xpos[pads[0]][0]) < 2 * 0.004 + 2 * size[0, 0, 0] + 0.004:
            break
    st[0, :15] = o.qpos[:15]; st[0, REC_QVEL:REC_QVEL + 15] = o.qvel[:15]
    mid = 0.5 * (o.geom_xpos[pads[0]] + o.geom_xpos[pads[1]])
    if kind in ("grasp", "grasp_stack"):
        st[0, 15:18] = mid; st[0, 18:22] = (1, 0, 0, 0)
    if kind == "stack":
        st[0, 22:25] = st[0, 15:18] + (0, 0, 2 * size[0, 0, 2] - 1e-4); st[0, 25:29] = st[0, 18:22]
    if kind == "grasp_stack":
        st[0, 22:25] = mid + (0, 0, size[0, 0, 2] + size[0, 1, 2] - 1e-4); st[0, 25:29] = (1, 0, 0, 0)
    return st, size, active
