"""Configuration of the batched rearrangement env: the slice of the reference's hydra tree that the hot path reads,
with hydra-style overrides (``"arena/props=colour_splitter"``, ``"control_dt=0.005"``,
``"robots.arm.controller_config.controller_params.gains.position.kp=300"``).

hydra / omegaconf are not installed in this image, so this module re-creates the features the reference's entry
points rely on (``initialize`` + ``compose(config_name="rearrangement", overrides=[...])``, tasks/rearrangement.py:34-49): a
``defaults:`` list per file, config groups selected with ``group/sub=name``, dotted scalar overrides (``+key=value`` appends) and
``${a.b.c}`` interpolation.

Where the reference's YAML tree is on disk (``$B2P_REFERENCE_CONFIG``, else ``/root/reference/mujoco_robot_environments/config``)
``compose`` READS it (``compose_from_tree``), plus this package's own ``config/`` directory (``osc_b200.yaml``).  Where it is
not (the GPU box), the table below -- transcribed from the same files, cited per block, and checked against the tree by
tests/test_config.py -- is the fallback.  Both give the same dictionary for the keys the hot path reads.
"""
from __future__ import annotations

import copy
import os
import re
from typing import Any, Dict, Iterable, Optional

import yaml

# config/arena/props/*.yaml.  shapes -> collider classes: cube / rectangle = the compiled box, cylinder = condim-6 cylinder,
# apple = GalaApple's sphere collider (include/b2p_prop_types.h; environment/props.py:520-581)
PROP_GROUPS: Dict[str, Dict[str, Any]] = {
    "colour_splitter": dict(min_objects=2, max_objects=5, min_object_size=0.015, max_object_size=0.016, sample_size=True, sample_colour=True, color_noise=0.1,
                            shapes=["cube"], colours=["green", "blue", "red", "yellow", "cyan", "magenta"], textures=["plain", "stary", "dotted"]),   # colour_splitter.yaml:1-19
    "single_block": dict(min_objects=1, max_objects=1, min_object_size=0.02, max_object_size=0.025, sample_size=True, sample_colour=True, color_noise=0.1,
                         shapes=["cube"], colours=["red", "green", "blue", "yellow", "magenta", "cyan", "grey"], textures=["plain", "stary", "dotted"]),   # single_block.yaml:1-20
    "itl_agent_demo": dict(min_objects=2, max_objects=2, min_object_size=0.02, max_object_size=0.025, sample_size=True, sample_colour=True, color_noise=0.1,
                           shapes=["cube"], colours=["green", "blue"], textures=["plain", "stary", "dotted"]),                                          # itl_agent_demo.yaml:1-15
    "itl_rearrangement": dict(min_objects=4, max_objects=6, min_object_size=0.02, max_object_size=0.025, sample_size=True, sample_colour=True, color_noise=0.1,
                              shapes=["cube", "cylinder", "rectangle"], colours=["red", "green", "blue", "yellow", "magenta", "cyan", "grey"],
                              textures=["plain", "stary", "dotted"]),                                                                           # itl_rearrangement.yaml:1-22
    "apple": dict(min_objects=2, max_objects=3, min_object_size=0.015, max_object_size=0.016, sample_size=True, sample_colour=True, color_noise=0.1,
                  shapes=["apple"], colours=["green", "blue", "red", "yellow", "cyan", "magenta"], textures=["plain", "stary", "dotted"]),              # apple.yaml:1-20
}
SHAPES = ("cube", "rectangle", "cylinder", "apple")          # add_object (props.py:520-581); "sphere" is defined there but no prop group lists it
APPLE_RADIUS = 0.02                                           # assets/mjcf/gala_apple/Apple.xml:7 (class "collision": sphere size 0.02)

# config/task/*.yaml
TASK_GROUPS: Dict[str, Dict[str, Any]] = {
    "rearrangement": dict(initializers=dict(seed=1, workspace=dict(min_pose=[0.35, -0.4, 0.43], max_pose=[0.55, 0.4, 0.435]),
                                            gripper=dict(min_pose=[0.4, -0.0, 0.7, 3.14, 0.0, 0.0], max_pose=[0.5, 0.0, 0.7, 3.14, 0.0, 0.0]))),   # rearrangement.yaml:2-10
    "apple": dict(initializers=dict(seed=1, workspace=dict(min_pose=[0.35, -0.4, 0.43], max_pose=[0.55, 0.4, 0.435]),
                                    gripper=dict(min_pose=[0.4, -0.0, 0.7, 3.14, 0.0, 0.0], max_pose=[0.5, 0.0, 0.7, 3.14, 0.0, 0.0]))),           # apple.yaml:2-10
    # rearrangement_w_targets.yaml:2-56 (zones are visual-only boxes on the table, rearrangement.py:104-116)
    "rearrangement_w_targets": dict(
        initializers=dict(seed=1, workspace=dict(min_pose=[0.35, -0.4, 0.43], max_pose=[0.55, 0.4, 0.435]),
                          gripper=dict(min_pose=[0.4, -0.0, 0.7, 3.14, 0.0, 0.0], max_pose=[0.5, 0.0, 0.7, 3.14, 0.0, 0.0])),
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
# config/arena/cameras/lasa.yaml:1-6 (the drawing task's observation camera)
LASA_CAMERAS = [dict(name="main_camera", pos=[0.363, -0.919, 1.663], quat=[0.94215352, 0.3349722, -0.00401909, -0.01114498], height=640, width=640, fovy=61)]
# the drawing task's visual-only geoms: the draw plane (tasks/lasa_draw.py:147-163: box half (0.3, 0.4, 0.001) at (0.4, 0, 0.55), rgba
# (0, 0, 1, 0.05), contype = conaffinity = 0) is a translucent zone for the ray-caster; the tool piece and the table carry their rgba
LASA_ZONES = {"draw_plane": dict(location=[0.4, 0.0, 0.55], size=[0.3, 0.4, 0.001], rgba=[0.0, 0.0, 1.0, 0.05])}
LASA_COLOURS = [("panda nohand/tool", (0.02, 0.302, 0.4)), ("table/table", (1.0, 1.0, 1.0))]      # lasa_draw.py:118, :97
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
            "default_configurations": {"home": [0, -0.785, 0, -2.356, 0, 1.571, 0.785]},     # robots/arm/franka_emika_panda.yaml:7-8
            "configuration": {"home": [0, -0.785, 0, -2.356, 0, 1.571, 0.785]},             # (this package's older name for the same block)
            "controller_config": {"controller_params": {                                    # controller_config/osc.yaml:1-22
                "name": "osc", "physics_dt": 0.001, "control_dt": 0.005,
                "gains": {"position": {"kp": 350.0, "kd": 20.0}, "orientation": {"kp": 500.0, "kd": 100.0}, "nullspace": {"kp": 200.0, "kd": 30.0}},
                "nullspace": {"joint_config": [0, -0.785, 0, -2.356, 0, 1.571, 0.785]},
                "convergence": {"position_threshold": 5e-3, "orientation_threshold": 68e-3}}},
        },
        "end_effector": {"controller_config": {"controller": {"min_val": 0.0, "max_val": 255.0}, "max_val": 255.0, "min_val": 0.0}},   # controller_config/min_max.yaml:1-4
    },
    # PropPlacer defaults (environment/prop_initializer.py:60-75)
    "prop_placer": {"max_attempts_per_prop": 1000, "max_qvel_tol": 1e-3, "max_qacc_tol": 1e-2, "min_settle_physics_time": 0.3,
                    "max_settle_physics_time": 2.0, "max_settle_physics_attempts": 10},
}

_GROUPS = {"arena/props": ("arena", "props", PROP_GROUPS), "task": ("task", None, TASK_GROUPS)}


REFERENCE_CONFIG_DEFAULT = "/root/reference/mujoco_robot_environments/config"
PACKAGE_CONFIG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "config")
# the reference's own DEFAULT_CONFIG passes this override (tasks/rearrangement.py:44-49); arena/rearrangement_table.yaml:3 names
# `props: default`, a file the tree does not contain
MISSING_GROUP_DEFAULTS = {"arena/props": "colour_splitter"}
_INTERP = re.compile(r"^\$\{([\w\.]+)\}$")


def reference_config_root() -> Optional[str]:
    root = os.environ.get("B2P_REFERENCE_CONFIG") or REFERENCE_CONFIG_DEFAULT
    return root if os.path.isfile(os.path.join(root, "rearrangement.yaml")) else None


_FLOAT = re.compile(r"^[+-]?(\d+\.?\d*|\.\d+)([eE][+-]?\d+)?$")


def _numbers(node):
    """PyYAML follows YAML 1.1 and reads `5e-3` as a string; omegaconf (what hydra uses) reads it as a float"""
    if isinstance(node, dict):
        return {k: _numbers(v) for k, v in node.items()}
    if isinstance(node, list):
        return [_numbers(v) for v in node]
    if isinstance(node, str) and _FLOAT.match(node):
        return float(node)
    return node


def _group_file(roots, group: str, name: str) -> Optional[str]:
    for r in roots:
        f = os.path.join(r, group, name + ".yaml")
        if os.path.isfile(f):
            return f
    return None


def _load_node(roots, group: str, name: str, choices: Dict[str, str]):
    """One config file + everything its ``defaults:`` list pulls in (group paths are relative to the file's own group)."""
    f = _group_file(roots, group, name)
    if f is None:
        raise KeyError("config group %r has no option %r (searched %s)" % (group, name, ", ".join(roots)))
    with open(f) as fh:
        node = _numbers(yaml.safe_load(fh))
    if not isinstance(node, dict):
        return node if node is not None else {}
    out: Dict[str, Any] = {}
    for d in node.pop("defaults", None) or []:
        if isinstance(d, str):
            continue                                   # `_self_`
        (g, n), = d.items()
        key = (group + "/" + g) if group else g
        n = choices.get(key, n)
        if _group_file(roots, key, n) is None and key in MISSING_GROUP_DEFAULTS and key not in choices:
            n = MISSING_GROUP_DEFAULTS[key]
        out[g] = _load_node(roots, key, n, choices)
    for k, v in node.items():
        out[k] = v
    return out


def _resolve(cfg):
    """``${a.b.c}`` -> the node at that dotted path (only whole-value interpolations occur in the tree)."""
    def lookup(path, depth=0):
        node = cfg
        for p in path.split("."):
            node = node[p]
        if isinstance(node, str) and _INTERP.match(node) and depth < 16:
            return lookup(_INTERP.match(node).group(1), depth + 1)
        return node

    def walk(node):
        if isinstance(node, dict):
            for k in list(node):
                v = node[k]
                if isinstance(v, str) and _INTERP.match(v):
                    try:
                        node[k] = copy.deepcopy(lookup(_INTERP.match(v).group(1)))
                    except KeyError:
                        pass                            # points outside what was composed (wandb etc.): left as written
                else:
                    walk(v)
        elif isinstance(node, list):
            for v in node:
                walk(v)
    walk(cfg); walk(cfg)
    return cfg


def _apply_scalar_overrides(cfg, scalar):
    for ov, key, val, add in scalar:
        node = cfg
        parts = key.split(".")
        for p in parts[:-1]:
            if p not in node or not isinstance(node[p], dict):
                raise KeyError("override %r: no config node %r" % (ov, p))
            node = node[p]
        if parts[-1] not in node and not add:
            raise KeyError("override %r: unknown key (prefix with a config group or use an existing dotted path)" % ov)
        node[parts[-1]] = yaml.safe_load(val)


def _split_overrides(overrides, groups):
    choices, scalar = {}, []
    for ov in overrides or ():
        key, _, val = ov.partition("=")
        add = key.strip().startswith("+")
        key = key.strip().lstrip("+")
        if "/" in key or key in groups:
            choices[key] = val.strip()
        else:
            scalar.append((ov, key, val, add))
    return choices, scalar


def _finish(cfg):
    """keys this package adds on top of the reference's: PropPlacer's constructor defaults and two aliases of older key names"""
    cfg.setdefault("prop_placer", copy.deepcopy(DEFAULT_CONFIG["prop_placer"]))
    arm, ee = cfg["robots"]["arm"], cfg["robots"]["end_effector"]
    if "default_configurations" in arm:
        arm["configuration"] = copy.deepcopy(arm["default_configurations"])
    cc = ee["controller_config"]
    if isinstance(cc.get("controller"), dict):
        cc["max_val"], cc["min_val"] = cc["controller"]["max_val"], cc["controller"]["min_val"]
    cp = arm["controller_config"]["controller_params"]
    cp["physics_dt"], cp["control_dt"] = cfg["physics_dt"], cfg["control_dt"]           # ${physics_dt} / ${control_dt} of osc.yaml:3-4, after overrides
    return cfg


def compose_from_tree(root: str, overrides: Optional[Iterable[str]] = None) -> Dict[str, Any]:
    """``hydra.compose(config_name="rearrangement", overrides=[...])`` over the reference's YAML tree at ``root`` (+ this package's
    ``config/`` for ``osc_b200``)."""
    roots = [root, PACKAGE_CONFIG]
    choices, scalar = _split_overrides(overrides, ("arena", "robots", "task", "dataset"))
    cfg = _load_node(roots, "", "rearrangement", choices)
    _apply_scalar_overrides(cfg, scalar)
    _resolve(cfg)
    return _finish(cfg)


def compose_from_table(overrides: Optional[Iterable[str]] = None) -> Dict[str, Any]:
    """The same composition from the transcribed table (no YAML tree on this machine)."""
    cfg = copy.deepcopy(DEFAULT_CONFIG)
    choices, scalar = _split_overrides(overrides, ("arena", "robots", "task", "dataset"))
    for key, val in choices.items():
        if key not in _GROUPS:
            raise KeyError("config group %r is not in the transcribed table (have: %s); point B2P_REFERENCE_CONFIG at the reference's config/ tree" % (key, ", ".join(sorted(_GROUPS))))
        top, sub, table = _GROUPS[key]
        if val not in table:
            raise KeyError("unknown option %r for config group %r (have: %s)" % (val, key, ", ".join(sorted(table))))
        if sub is None:
            cfg[top] = copy.deepcopy(table[val])
        else:
            cfg[top][sub] = copy.deepcopy(table[val])
    _apply_scalar_overrides(cfg, scalar)
    return _finish(cfg)


def compose(overrides: Optional[Iterable[str]] = None, root: Optional[str] = None) -> Dict[str, Any]:
    """hydra ``compose(config_name="rearrangement", overrides=[...])``: from the reference's YAML tree when it is on disk, else from
    the transcribed table."""
    root = root or reference_config_root()
    return compose_from_tree(root, overrides) if root else compose_from_table(overrides)
