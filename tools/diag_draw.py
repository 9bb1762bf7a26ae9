"""Diagnostic: identical drawing scenes (tool on the table, position servos) must give identical results on the GPU."""
import sys, os
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import blobtools as bt
from mujoco_robot_environments_b200.physics import BatchedPhysics, load_default_blob
TOUCH = [-0.59998, 0.56896, 0.44407, -2.42968, -0.90239, 2.84152, -1.69566]
blob0 = load_default_blob("lasa_draw_position")[0]
for label, blob in (("pyramidal", blob0), ("elliptic", bt.surgery(blob0, opts={"cone": 1}))):
    for dtype in (torch.float32, torch.float64):
        for rep in range(2):
            n = 6
            p = BatchedPhysics(n, dtype=dtype, blob=blob)
            p.reset(); p.qpos[:, :7] = torch.tensor(TOUCH, dtype=dtype, device=p.device)
            first = None
            for t in range(30):
                target = np.array(TOUCH) + np.array([0.004 * t, 0.002, 0, 0, 0, 0.001, 0])
                p.set_control(torch.tensor(target).expand(n, 7)); p.step(5)
                torch.cuda.synchronize()
                q = p.qpos[:, :7].cpu().numpy()
                if first is None and np.abs(q - q[0]).max() > 0:
                    first = (t, float(np.abs(q - q[0]).max()), p.data.nefc.cpu().numpy().tolist(), p.data.niter.cpu().numpy().tolist() if hasattr(p.data, "niter") else None)
            print(label, dtype, "rep", rep, "spread", float(np.abs(q - q[0]).max()), "q0", q[:, 0].tolist(), "first divergence", first, flush=True)
            p.close()
