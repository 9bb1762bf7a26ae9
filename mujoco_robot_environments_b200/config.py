"""Configuration of the batched rearrangement env: the slice of the reference's hydra tree that the hot path reads,
with hydra-style overrides (``"arena/props=colour_splitter"``, ``"control_dt=0.005"``,
``"robots.arm.controller_config.controller_params.gains.position.kp=300"``).

hydra / omegaconf are not installed in this image, so this module re-creates the two features the
reference's entry points rely on (``initialize`` + ``compose(config_name="rearrangement", overrides=[...])``,
tasks/rearrangement.py:34-49): config groups selected with ``group/sub=name`` and dotted scalar overrides.  Values
are transcribed from the reference YAML files cited next to each block.
"""
from __future__ import annotations

import copy
from typing import Any, Dict, Iterable, Optional

import yaml

# config/arena/props/*.yaml  (only box-prop sets are compiled into the kernels' model class; others raise)
PROP_GROUPS: Dict[str, Dict[str, Any]] = {
    "colour_splitter": dict(min_objects=2, max_objects=5, min_object_size=0.015, max_object_size=0.016, sample_size=True, shapes=["cube"],
                            colours=["green", "blue", "red", "yellow", "cyan", "magenta"]),                    # colour_splitter.yaml:1-15
    "single_block": dict(min_objects=1, max_objects=1, min_object_size=0.02, max_object_size=0.025, sample_size=True, shapes=["cube"],
                         colours=["red", "green", "blue", "yellow", "magenta", "cyan", "grey"]),                # single_block.yaml:1-16
}
UNSUPPORTED_PROP_GROUPS = ("apple", "itl_agent_demo", "itl_rearrangement")   # sphere / cylinder / condim-6 colliders: out of the model class

# config/task/*.yaml
TASK_GROUPS: Dict[str, Dict[str, Any]] = {
    "rearrangement": dict(initializers=dict(seed=1, workspace=dict(min_pose=[0.35, -0.4, 0.43], max_pose=[0.55, 0.4, 0.435]))),   # rearrangement.yaml:2-7
    # rearrangement_w_targets.yaml:2-56 (zones are visual-only boxes on the table, rearrangement.py:104-116)
    "rearrangement_w_targets": dict(
        initializers=dict(seed=1, workspace=dict(min_pose=[0.35, -0.4, 0.43], max_pose=[0.55, 0.4, 0.435])),
        target_locations={
            "top_left": dict(name="top_left", location=[0.55, 0.4, 0.4], size=[0.075, 0.15, 0.01], rgba=[1.0, 0.0, 0.0, 0.1]),
            "bottom_left": dict(name="bottom_left", location=[0.35, 0.4, 0.4], size=[0.075, 0.15, 0.01], rgba=[0.0, 1.0, 0.0, 0.1]),
            "top_right": dict(name="top_right", location=[0.55, -0.4, 0.4], size=[0.075, 0.15, 0.01], rgba=[0.0, 0.0, 1.0, 0.1]),
            "bottom_right": dict(name="bottom_right", location=[0.35, -0.4, 0.4], size=[0.075, 0.15, 0.01], rgba=[1.0, 1.0, 0.0, 0.1]),
            "top_middle": dict(name="top_middle", location=[0.55, 0.0, 0.4], size=[0.075, 0.15, 0.01], rgba=[1.0, 0.0, 1.0, 0.1]),
            "bottom_middle": dict(name="bottom_middle", location=[0.35, 0.0, 0.4], size=[0.075, 0.15, 0.01], rgba=[0.0, 1.0, 1.0, 0.1]),
        },
        colour_target_map=dict(red="top_left", green="bottom_left", blue="top_right", yellow="bottom_right", magenta="top_middle", cyan="bottom_middle")),
}

# config/arena/cameras/transporter_data_collection.yaml:1-6 (the `rearrangement_table` arena's camera list)
CAMERAS = [dict(name="overhead_camera", pos=[0.7, 0.0, 1.3], quat=[0.707, 0.0, 0.0, -0.707], height=480, width=640, fovy=61)]
# config/dataset/default.yaml
DATASET = dict(num_episodes=1000, max_steps=10, max_episodes_per_file=10)

DEFAULT_CONFIG: Dict[str, Any] = {
    # config/rearrangement.yaml:2-6
    "physics_dt": 0.001, "control_dt": 0.005, "gravity": [0.0, 0.0, -9.8], "nconmax": 1000, "njmax": 2000,
    "arena": {"props": copy.deepcopy(PROP_GROUPS["colour_splitter"]), "cameras": copy.deepcopy(CAMERAS)},
    "dataset": copy.deepcopy(DATASET),
    "task": copy.deepcopy(TASK_GROUPS["rearrangement"]),
    "robots": {
        "arm": {
            "configuration": {"home": [0, -0.785, 0, -2.356, 0, 1.571, 0.785]},            # robots/arm/franka_emika_panda.yaml:8
            "controller_config": {"controller_params": {                                    # controller_config/osc.yaml:1-22
                "name": "osc", "physics_dt": 0.001, "control_dt": 0.005,
                "gains": {"position": {"kp": 350.0, "kd": 20.0}, "orientation": {"kp": 500.0, "kd": 100.0}, "nullspace": {"kp": 200.0, "kd": 30.0}},
                "nullspace": {"joint_config": [0, -0.785, 0, -2.356, 0, 1.571, 0.785]},
                "convergence": {"position_threshold": 5e-3, "orientation_threshold": 68e-3}}},
        },
        "end_effector": {"controller_config": {"max_val": 255.0, "min_val": 0.0}},           # controller_config/min_max.yaml:1-4
    },
    # PropPlacer defaults (environment/prop_initializer.py:60-75)
    "prop_placer": {"max_attempts_per_prop": 1000, "max_qvel_tol": 1e-3, "max_qacc_tol": 1e-2, "min_settle_physics_time": 0.3,
                    "max_settle_physics_time": 2.0, "max_settle_physics_attempts": 10},
}

_GROUPS = {"arena/props": ("arena", "props", PROP_GROUPS), "task": ("task", None, TASK_GROUPS)}


def compose(overrides: Optional[Iterable[str]] = None) -> Dict[str, Any]:
    """hydra ``compose(config_name="rearrangement", overrides=[...])`` for the keys above."""
    cfg = copy.deepcopy(DEFAULT_CONFIG)
    for ov in overrides or ():
        key, _, val = ov.partition("=")
        add = key.strip().startswith("+")              # hydra "+key=value": append a new key (e.g. "+name=colour_splitter")
        key = key.strip().lstrip("+")
        if key in _GROUPS:
            top, sub, table = _GROUPS[key]
            if val in UNSUPPORTED_PROP_GROUPS:
                raise ValueError("prop set %r needs sphere / cylinder / condim-6 colliders, which the B200 kernels' model class does not cover" % val)
            if val not in table:
                raise KeyError("unknown option %r for config group %r (have: %s)" % (val, key, ", ".join(sorted(table))))
            if sub is None:
                cfg[top] = copy.deepcopy(table[val])
            else:
                cfg[top][sub] = copy.deepcopy(table[val])
            continue
        node = cfg
        parts = key.split(".")
        for p in parts[:-1]:
            if p not in node or not isinstance(node[p], dict):
                raise KeyError("override %r: no config node %r" % (ov, p))
            node = node[p]
        if parts[-1] not in node and not add:
            raise KeyError("override %r: unknown key (prefix with a config group or use an existing dotted path)" % ov)
        node[parts[-1]] = yaml.safe_load(val)
    # ${physics_dt} / ${control_dt} interpolation of osc.yaml:3-4
    cp = cfg["robots"]["arm"]["controller_config"]["controller_params"]
    cp["physics_dt"], cp["control_dt"] = cfg["physics_dt"], cfg["control_dt"]
    return cfg
