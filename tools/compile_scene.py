"""Compile the canonical rearrangement scene(s) from the MJCF sources into B2PM blobs.

The MJCF / mesh inputs are the menagerie files the reference loads (SURVEY.md 0.4:
MEN-B franka_emika_panda/panda_nohand.xml + robotiq_2f85/2f85.xml); they are third-party
data that cannot travel to the GPU box, so the compiled blobs (hull vertices, inertias,
pair table, mj_setConst quantities) are committed under
mujoco_robot_environments_b200/assets/ together with this generating script.

  python tools/compile_scene.py [--menagerie PATH]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from mujoco_robot_environments_b200.compile import scene  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--menagerie", default=scene.DEFAULT_MENAGERIE)
args = ap.parse_args()
out = os.path.join(os.path.dirname(__file__), "..", "mujoco_robot_environments_b200", "assets")
os.makedirs(out, exist_ok=True)
for tag, fuse in (("rearrangement_n4", True), ("rearrangement_n4_unfused", False)):
    blob, names = scene.compile_rearrangement(4, fuse=fuse, menagerie=args.menagerie)
    open(os.path.join(out, tag + ".b2pm"), "wb").write(blob)
    json.dump(names, open(os.path.join(out, tag + ".names.json"), "w"), indent=0)
    print(tag, len(blob), "bytes")
    if fuse:    # render model of the same scene (overhead camera + target zones): assets/<tag>.b2pr
        from mujoco_robot_environments_b200 import config as cfglib
        from mujoco_robot_environments_b200.compile import render
        rb = render.build_render_blob(blob, names["geom"], cfglib.CAMERAS[0], cfglib.TASK_GROUPS["rearrangement_w_targets"]["target_locations"])
        open(os.path.join(out, tag + ".b2pr"), "wb").write(rb)
        print(tag + ".b2pr", len(rb), "bytes")
# the drawing scenes of tasks/lasa_draw.py (arm without the gripper, cylinder tool piece, pyramidal cone): torque motors / position servos
for tag, kw in (("lasa_draw", {}), ("lasa_draw_position", {"actuators": "position"})):
    blob, names = scene.compile_lasa_draw(menagerie=args.menagerie, **kw)
    open(os.path.join(out, tag + ".b2pm"), "wb").write(blob)
    json.dump(names, open(os.path.join(out, tag + ".names.json"), "w"), indent=0)
    print(tag, len(blob), "bytes")
    if tag == "lasa_draw":      # render model: main_camera, the draw plane as a translucent zone, the tool piece as a 64-gon prism
        from mujoco_robot_environments_b200 import config as cfglib
        from mujoco_robot_environments_b200.compile import render
        rb = render.build_render_blob(blob, names["geom"], cfglib.LASA_CAMERAS[0], cfglib.LASA_ZONES, colours=cfglib.LASA_COLOURS)
        open(os.path.join(out, tag + ".b2pr"), "wb").write(rb)
        print(tag + ".b2pr", len(rb), "bytes")
