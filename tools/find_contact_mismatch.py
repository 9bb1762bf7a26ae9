"""CPU hunt for scenes whose contact list differs between the kernel logic (host build of the templates, tests/hostemu) and the
oracle: drives BASELINE configs[1]'s workload (Halton torque noise) on a slice of scenes and compares the lists every `every`
steps.  Test infrastructure; prints and saves offending states to gpurun_out/contact_mismatch.json.

  python tools/find_contact_mismatch.py [--scenes 256] [--steps 600] [--every 20] [--procs 8] [--dtype f64]
"""
import argparse
import json
import multiprocessing as mp
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def work(a):
    lo, hi, steps, every, dtype = a
    import numpy as np
    import scenarios as sc
    from hostemu import MODE_FORWARD, MODE_PHYSICS, HostEmu, out_slices
    from orc import Oracle
    S = out_slices()
    npdt = np.float64 if dtype == "f64" else np.float32
    blob = sc.default_blob()
    e = HostEmu(blob, npdt); o = Oracle(blob)
    n_all = 4096
    st, size, active = sc.make_states(n_all, seed=1234, settled=True)
    st, size, active = st[lo:hi].astype(npdt), size[lo:hi], active[lo:hi]
    sc.oracle_load(o, st[0].astype(np.float64), size[0], active[0]); o.step1(); bias = o.qfrc_bias[:7].copy()
    found = []
    for k0 in range(0, steps, every):
        launch = k0 // 100
        ctrl = np.tile(sc.halton_ctrl(256, launch, bias), (16, 1))[lo:hi]
        e.run(MODE_PHYSICS, st, n_steps=every, ctrl=ctrl, prop_size=size, prop_active=active, want_out=False)
        r = e.run(MODE_FORWARD, st.copy(), prop_size=size, prop_active=active, full_contacts=False)
        for i in range(hi - lo):
            out = r["out"][i].astype(np.float64)
            nc = int(out[S["ncon"]][0]); c = out[S["contact"]].reshape(-1, 10)[:nc]
            sc.oracle_load(o, st[i].astype(np.float64), size[i], active[i]); o.set_option("full_contacts", 0); o.step1()
            oc = o.contacts()
            gk = sorted((int(x[7]), int(x[8])) for x in c); ok = sorted((int(x), int(y)) for x, y in oc["geom"])
            if gk != ok:
                found.append(dict(scene=lo + i, step=k0 + every, state=st[i].astype(np.float64).tolist(), size=size[i].tolist(), active=active[i].tolist(),
                                  emu_pairs=gk, oracle_pairs=ok, emu_dist=c[:, 0].tolist(), oracle_dist=oc["dist"].tolist()))
    return found


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", type=int, default=256); ap.add_argument("--steps", type=int, default=600); ap.add_argument("--every", type=int, default=20)
    ap.add_argument("--procs", type=int, default=os.cpu_count()); ap.add_argument("--dtype", default="f64")
    a = ap.parse_args()
    per = (a.scenes + a.procs - 1) // a.procs
    jobs = [(lo, min(a.scenes, lo + per), a.steps, a.every, a.dtype) for lo in range(0, a.scenes, per)]
    with mp.get_context("spawn").Pool(a.procs) as pool:
        res = sum(pool.map(work, jobs), [])
    print("%d mismatching (scene, step) samples out of %d x %d" % (len(res), a.scenes, a.steps // a.every))
    for r in res[:20]:
        d = [p for p in r["emu_pairs"] if p not in r["oracle_pairs"]], [p for p in r["oracle_pairs"] if p not in r["emu_pairs"]]
        print(" scene %d step %d: only-kernel %s only-oracle %s" % (r["scene"], r["step"], d[0], d[1]))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "contact_mismatch.json"), "w"))
