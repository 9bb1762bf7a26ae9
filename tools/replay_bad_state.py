"""Replay tests/golden/bad_state_fp32_*.json on the GPU: one physics step from the last good state (fp32), then the OSC tick it came
from; prints the status bits (bits 8.. = first NaN probe of a -DB2P_NANTRACE build selected with B2P_LIB=...)."""
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from mujoco_robot_environments_b200.physics import BatchedPhysics  # noqa: E402

for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "bad_state_fp32_*.json"))):
    r = json.load(open(f))
    for dtype in (torch.float32, torch.float64):
        p = BatchedPhysics(1, dtype=dtype)
        p.set_env_params(torch.tensor(r["size"])[None], torch.tensor(r["active"], dtype=torch.uint8)[None])
        p.state.copy_(torch.tensor(r["last_good_state"], dtype=dtype)[None]); p.status(clear=True)
        p.set_control(torch.tensor(r["ctrl"])[None]); p.step(1); torch.cuda.synchronize()
        st = int(p.status().item())
        print(os.path.basename(f), str(dtype), "one step: flags", st & 255, "probe", st >> 8, "(iteration %d, site %d)" % ((st >> 8) // 16, (st >> 8) % 16), "nefc", int(p.data.nefc.item()), "niter", int(p.derived("NITER")[0, 0].item()),
              "max|qacc| %.4g" % float(p.data.qacc.abs().max().item()), "finite", bool(torch.isfinite(p.state).all()))
        p.state.copy_(torch.tensor(r["tick_start_state"], dtype=dtype)[None]); p.status(clear=True)
        p.run_controller(torch.tensor(r["target"], dtype=dtype)[None], torch.tensor([r["grip"]], dtype=torch.uint8), 3); torch.cuda.synchronize()
        st = int(p.status().item())
        print("   3 ticks from the tick start: flags", st & 255, "probe", st >> 8, "finite", bool(torch.isfinite(p.state).all()))
        p.close()
