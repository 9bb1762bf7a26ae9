"""The committed scene blobs are what the scene compiler makes of the MJCF sources (mujoco_robot_environments_b200/compile/scene.py).
The sources (the menagerie snapshot the reference vendors) cannot travel, so this check runs only where they exist; everywhere else
the blobs themselves are checked for the model class they claim."""
import os

import numpy as np
import pytest

from mujoco_robot_environments_b200.compile import blob as B
from mujoco_robot_environments_b200.compile import scene

ASSETS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mujoco_robot_environments_b200", "assets")


def _read(tag):
    with open(os.path.join(ASSETS, tag + ".b2pm"), "rb") as f:
        return f.read()


def test_drawing_scene_blobs_describe_the_reference_scene():
    """tasks/lasa_draw.py:60-176: arm without gripper (7 hinge dofs, 7 actuators), no props, no equalities, table + ground + arm hulls
    + ONE cylinder (the tool piece, radius 0.015, half height 0.05), MuJoCo's default pyramidal cone, config/lasa.yaml's 10 ms step."""
    for tag, servo in (("lasa_draw", False), ("lasa_draw_position", True)):
        c, o, a = B.unpack(_read(tag))
        assert (c["nq"], c["nv"], c["nu"], c["nprop"], c["neq"]) == (7, 7, 7, 0, 0)
        assert c["narm"] == 7 and c["ngrip"] == 0 and c["site_pinch"] == c["site_attachment"]
        assert o["cone"] == B.CONE_PYRAMIDAL and o["integrator"] == B.INT_IMPLICITFAST and o["impratio"] == 10 and o["timestep"] == 0.01
        types = list(a["geom_type"])
        assert types.count(B.GEOM_CYLINDER) == 1 and types.count(B.GEOM_PLANE) == 1 and types.count(B.GEOM_BOX) == 1
        g = types.index(B.GEOM_CYLINDER)
        assert np.allclose(a["geom_size"][g][:2], (0.015, 0.05)) and np.allclose(a["geom_aabb"][g][3:], (0.015, 0.015, 0.05))
        gain, bias = np.asarray(a["act_gain"]), np.asarray(a["act_bias"])
        if servo:      # config/robots/arm/actuator_config/position.yaml: kp, (0, -kp, -kv)
            assert np.allclose(gain, (4500, 4500, 3500, 3500, 2000, 2000, 2000)) and np.allclose(bias[:, 1], -gain) and np.allclose(bias[:, 2], -gain / 10)
        else:
            assert np.allclose(gain, 1) and not bias.any()
        # every pair of the table pairs a static geom with a geom of the moving tree; the tool piece meets the table
        table, pairs = c["geom_table"], set(zip(a["pair_geom1"], a["pair_geom2"]))
        assert (g, table) in pairs or (table, g) in pairs


@pytest.mark.skipif(not os.path.isdir(scene.DEFAULT_MENAGERIE), reason="MJCF sources (menagerie snapshot) not on this machine")
@pytest.mark.parametrize("tag,kw", [("lasa_draw", {}), ("lasa_draw_position", {"actuators": "position"})])
def test_drawing_scene_blobs_are_reproducible_from_the_mjcf_sources(tag, kw):
    blob, names = scene.compile_lasa_draw(**kw)
    c0, o0, a0 = B.unpack(_read(tag)); c1, o1, a1 = B.unpack(blob)
    assert c0 == c1 and o0 == o1
    for k in a0:
        assert np.array_equal(np.asarray(a0[k]), np.asarray(a1[k])), k
