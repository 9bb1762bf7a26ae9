"""Hydra-style overrides of the config slice the hot path reads (reference: config/*.yaml, tasks/rearrangement.py:34-49)."""
import pytest

from mujoco_robot_environments_b200 import config as C


def test_defaults_match_reference_yaml():
    cfg = C.compose()
    assert cfg["physics_dt"] == 0.001 and cfg["control_dt"] == 0.005 and cfg["gravity"] == [0.0, 0.0, -9.8]
    assert cfg["arena"]["props"]["min_objects"] == 2 and cfg["arena"]["props"]["max_objects"] == 5
    g = cfg["robots"]["arm"]["controller_config"]["controller_params"]["gains"]
    assert (g["position"]["kp"], g["position"]["kd"], g["orientation"]["kp"], g["orientation"]["kd"], g["nullspace"]["kp"], g["nullspace"]["kd"]) == (350.0, 20.0, 500.0, 100.0, 200.0, 30.0)
    assert cfg["task"]["initializers"]["workspace"]["min_pose"] == [0.35, -0.4, 0.43]


def test_group_and_dotted_overrides():
    cfg = C.compose(["arena/props=single_block", "robots.arm.controller_config.controller_params.gains.position.kp=300", "control_dt=0.01"])
    assert cfg["arena"]["props"]["max_objects"] == 1
    cp = cfg["robots"]["arm"]["controller_config"]["controller_params"]
    assert cp["gains"]["position"]["kp"] == 300 and cp["control_dt"] == 0.01
    assert C.compose()["arena"]["props"]["max_objects"] == 5          # defaults untouched


def test_bad_overrides_fail_loudly():
    with pytest.raises(ValueError):
        C.compose(["arena/props=apple"])
    with pytest.raises(KeyError):
        C.compose(["arena/props=nope"])
    with pytest.raises(KeyError):
        C.compose(["no.such.key=1"])
