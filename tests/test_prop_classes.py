"""The sphere (GalaApple) and cylinder (condim 6) prop classes of environment/props.py:307-489 (include/b2p_prop_types.h): the kernel
templates (host build here, CUDA in the gpu-marked test) against the oracle on a scene that holds one prop of every class -- a
resting apple, a spinning upright cylinder (torsional friction), a rolling cylinder on its side (rolling friction) and a box -- plus
analytic facts of the classes themselves."""
import numpy as np
import pytest

import scenarios as sc
from orc import Oracle
from mujoco_robot_environments_b200.compile import blob as B

P = lambda k: 15 + 7 * k          # qpos address of prop k
V = lambda k: 15 + 6 * k          # dof address of prop k
TYPES = np.array([B.GEOM_SPHERE, B.GEOM_CYLINDER, B.GEOM_CYLINDER, B.GEOM_BOX], np.uint8)
SIZES = np.array([[0.02, 0, 0], [0.02, 0.03, 0], [0.02, 0.03, 0], [0.02, 0.02, 0.02]])


def mixed_scene():
    st, size, active = sc.make_states(1, seed=3)
    q = st[0].copy()
    q[P(0):P(0) + 7] = [0.55, 0.36, 0.4 + 0.02 + 0.002, 1, 0, 0, 0]
    q[P(1):P(1) + 7] = [0.55, -0.30, 0.4 + 0.03 + 0.002, 1, 0, 0, 0]
    q[P(2):P(2) + 7] = [0.45, 0.10, 0.4 + 0.02 + 0.002, np.cos(np.pi / 4), np.sin(np.pi / 4), 0, 0]      # axis horizontal
    q[P(3):P(3) + 7] = [0.45, -0.10, 0.4 + 0.02 + 0.002, 1, 0, 0, 0]
    v = np.zeros(sc.NV); v[V(0)] = 0.3; v[V(1) + 5] = 5.0; v[V(2)] = 0.2; v[V(2) + 4] = 3.0
    state = np.zeros(128); state[:sc.NQ] = q[:sc.NQ]; state[sc.REC_QVEL:sc.REC_QVEL + sc.NV] = v
    return state, SIZES.copy(), np.ones(4, np.uint8), TYPES.copy()


def load_oracle(o, state, size, active, types):
    sc.oracle_load(o, state, size, active)
    for p in range(4):
        o.set_prop_type(p, int(types[p]))
    for p in range(4):
        o.set_prop(p, size[p], bool(active[p]))
    o.qvel[:sc.NV] = state[sc.REC_QVEL:sc.REC_QVEL + sc.NV]
    o.set_option("full_contacts", 0)


def test_classes_in_the_oracle():
    """mass / inertia of the classes, condim-6 rows, and the steady state: the apple's normal force is its weight (density 1000 sphere),
    the upright cylinder rests on one condim-6 contact (6 rows) carrying m g, torsional friction has stopped its spin, rolling friction
    the rolling cylinder."""
    o = Oracle(sc.default_blob())
    state, size, active, types = mixed_scene()
    load_oracle(o, state, size, active, types)
    for _ in range(400):
        o.step()
    con = o.contacts(); g = con["geom"]
    prop_g = [int(x) for x in B.unpack(sc.default_blob())[2]["prop_geomid"]]
    n_per = [int((g[:, 1] == pg).sum()) for pg in prop_g]
    assert n_per == [1, 1, 1, 4]                  # sphere: 1 point; cylinders against the table box: one MPR contact each (mjc_Convex); box: 4 corners
    nefc = o.dim("nefc")
    ncr = 3 * 1 + 6 * 1 + 6 * 1 + 3 * 4           # contact rows: condim 3 for the apple and the box, 6 for the cylinders
    assert nefc - ncr in (7, 8)                   # 7 equality rows (+ a joint-limit row of the gripper)
    f = o.efc_force[nefc - ncr:nefc]
    m_apple = 1000 * 4 / 3 * np.pi * 0.02 ** 3
    assert abs(f[0] - m_apple * 9.8) < 1e-4 and abs(f[3] - 0.1 * 9.8) < 1e-4
    assert abs(o.qvel[V(1) + 5]) < 1e-6 and np.abs(o.qvel[V(2):V(2) + 6]).max() < 1e-3


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-10), (np.float32, 2e-3)])
def test_host_build_matches_oracle(dtype, tol):
    from hostemu import MODE_PHYSICS, HostEmu, out_slices
    S = out_slices()
    blob = sc.default_blob()
    o = Oracle(blob); e = HostEmu(blob, dtype)
    state, size, active, types = mixed_scene()
    load_oracle(o, state, size, active, types)
    got = state[None].astype(dtype)
    r = e.run(MODE_PHYSICS, got, n_steps=200, ctrl=np.zeros((1, 8)), prop_size=size[None], prop_active=active[None], prop_type=types[None])
    for _ in range(200):
        o.step()
    assert r["status"][0] == 0
    assert int(r["out"][0][S["nefc"]][0]) == o.dim("nefc") and int(r["out"][0][S["ncon"]][0]) == o.dim("ncon")
    assert np.abs(got[0, :sc.NQ] - o.qpos[:sc.NQ]).max() < tol
    assert np.abs(got[0, sc.REC_QVEL:sc.REC_QVEL + sc.NV] - o.qvel[:sc.NV]).max() < 100 * tol


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [("float64", 1e-9), ("float32", 2e-3)])
def test_gpu_matches_oracle(dtype, tol):
    torch = pytest.importorskip("torch")
    from mujoco_robot_environments_b200.physics import BatchedPhysics
    dt = getattr(torch, dtype)
    n = 8
    state, size, active, types = mixed_scene()
    states = np.repeat(state[None], n, 0)
    rng = np.random.default_rng(0)
    states[1:, sc.REC_QVEL + V(0)] += rng.uniform(-0.2, 0.2, n - 1)           # a different push per scene
    p = BatchedPhysics(n, dtype=dt)
    p.state.copy_(torch.tensor(states, dtype=dt))
    p.set_env_params(torch.tensor(np.repeat(size[None], n, 0)), torch.tensor(np.repeat(active[None], n, 0)), torch.tensor(np.repeat(types[None], n, 0)))
    p.step(200); torch.cuda.synchronize()
    got = p.state.cpu().numpy().astype(np.float64)
    assert int(p.status().cpu().max()) == 0
    o = Oracle(sc.default_blob())
    for i in range(n):
        load_oracle(o, states[i], size, active, types)
        for _ in range(200):
            o.step()
        assert np.abs(got[i, :sc.NQ] - o.qpos[:sc.NQ]).max() < tol, i
    # mixed batch: scene 0 keeps its classes, scene 1 is all boxes -- per-scene classes do not leak between scenes
    types2 = np.repeat(types[None], n, 0); types2[1] = B.GEOM_BOX
    size2 = np.repeat(size[None], n, 0); size2[1] = 0.02
    q = BatchedPhysics(n, dtype=dt); q.state.copy_(torch.tensor(states, dtype=dt))
    q.set_env_params(torch.tensor(size2), torch.tensor(np.repeat(active[None], n, 0)), torch.tensor(types2))
    q.step(200); torch.cuda.synchronize()
    got2 = q.state.cpu().numpy().astype(np.float64)
    assert np.abs(got2[0] - got[0]).max() == 0
    sc.oracle_load(o, states[1], size2[1], active)
    for k in range(4):
        o.set_prop_type(k, B.GEOM_BOX); o.set_prop(k, size2[1][k], True)
    o.qvel[:sc.NV] = states[1, sc.REC_QVEL:sc.REC_QVEL + sc.NV]; o.set_option("full_contacts", 0)
    for _ in range(200):
        o.step()
    assert np.abs(got2[1, :sc.NQ] - o.qpos[:sc.NQ]).max() < tol
