"""Regenerate include/b2p_model_layout.h from compile/blob.py."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from mujoco_robot_environments_b200.compile import blob
out = os.path.join(os.path.dirname(__file__), "..", "include", "b2p_model_layout.h")
open(out, "w").write(blob.c_header())
print("wrote", os.path.normpath(out))
