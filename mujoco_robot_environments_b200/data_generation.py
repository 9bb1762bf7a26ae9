"""Batched transporter-network data generation (BASELINE configs[3]).

Mirrors the episode loop of the reference's ``transporter_network_data_generation.py:97-143`` -- ``env.reset()``, then up to
``dataset.max_steps`` rounds of ``sort_colours() -> step(pick_action) -> step(place_action)`` with the action dict
``{"pose", "pixel_coords" = world_2_pixel(pose[:3]), "gripper_rot" = 0.0}`` -- for ``num_envs`` scenes at once on every GPU,
then gathers the demonstrations to rank 0 over NCCL (``dist.gather_demos``) and sums the episode statistics
(``dist.reduce_stats``).  envlogger / TFDS (the reference's on-disk format) are not installed in this image: rank 0 writes the
same fields as one ``.npz`` per call.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 -m mujoco_robot_environments_b200.data_generation --envs 1024
"""
from __future__ import annotations

import argparse
import json
import os
import time
from typing import Dict, Optional

import numpy as np
import torch

from . import config as cfglib
from . import dist as distlib

# the reference's COLOR_SEPERATOR_TASK_CONFIG (transporter_network_data_generation.py:26-33)
COLOUR_SPLITTER_OVERRIDES = ["+name=colour_splitter", "task=rearrangement_w_targets", "arena/props=colour_splitter"]
ACTION_WIDTH = 7 + 2 + 1          # pose f64[7] | pixel_coords i32[2] | gripper_rot f64


def pack_action(pose: torch.Tensor, pixel: torch.Tensor) -> torch.Tensor:
    """action dict -> one row [n, 10] (pose, u, v, gripper_rot = 0.0) in float64 (the reference's action dtypes fit exactly)."""
    return torch.cat([pose.to(torch.float64), pixel.to(torch.float64), torch.zeros(pose.shape[0], 1, dtype=torch.float64, device=pose.device)], 1)


def rollout_episode(env, max_steps: int, render: bool = False, resample_props: bool = False) -> Dict[str, torch.Tensor]:
    """One episode per scene.  -> dict of device tensors with leading dim n_env:
    ``actions`` [n, T, 2, 10] (pick, place), ``valid`` [n, T] (the scene acted in round t), ``prop_pos0`` / ``prop_quat0`` (state
    observation at reset), ``success`` [n] (every prop inside its colour zone at the end), ``failed`` [n], ``rounds`` [n];
    with ``render`` also ``rgb`` u8 [n, T, 2, H, W, 3] / ``depth`` f32 [n, T, 2, H, W] (the observation ``step`` returns: the
    frame BEFORE acting, rearrangement.py:343)."""
    n, dev = env.num_envs, env.device
    ts = env.reset(resample_props=resample_props)
    obs0 = ts.observation
    actions = torch.zeros(n, max_steps, 2, ACTION_WIDTH, dtype=torch.float64, device=dev)
    valid = torch.zeros(n, max_steps, dtype=torch.bool, device=dev)
    frames = {"rgb": [], "depth": []}
    alive = torch.ones(n, dtype=torch.bool, device=dev)
    for t in range(max_steps):
        in_progress, pick, place = env.sort_colours()
        act = in_progress & alive
        if not bool(act.any()):
            break                                                    # "Task demonstration is complete"
        pick_action = {"pose": pick, "pixel_coords": env.world_2_pixel("overhead_camera/overhead_camera", pick[:, :3]), "gripper_rot": 0.0}
        place_action = {"pose": place, "pixel_coords": env.world_2_pixel("overhead_camera/overhead_camera", place[:, :3]), "gripper_rot": 0.0}
        actions[:, t, 0] = pack_action(pick, pick_action["pixel_coords"]) * act[:, None]
        actions[:, t, 1] = pack_action(place, place_action["pixel_coords"]) * act[:, None]
        valid[:, t] = act
        for a in (pick_action, place_action):
            if render:
                o = env.render()
                frames["rgb"].append(o["overhead_camera/rgb"]); frames["depth"].append(o["overhead_camera/depth"])
            env.step({"pose": a["pose"].clone(), "pixel_coords": a["pixel_coords"], "gripper_rot": a["gripper_rot"]}, active=act)
        alive = alive & ~env.failed                                  # "Task demonstration failed with exception": episode over
    in_progress = env.has_prop & ~env.expert.within_target()
    out = {"actions": actions, "valid": valid, "prop_pos0": obs0["prop_pos"].clone(), "prop_quat0": obs0["prop_quat"].clone(),
           "prop_colour": env.prop_colour, "has_prop": env.has_prop, "success": ~in_progress.any(1) & alive, "failed": env.failed.clone(),
           "rounds": valid.sum(1), "status": env.physics.status().clone()}     # status: OR of the kernel's per-scene flags over the episode
    if render and frames["rgb"]:
        T = len(frames["rgb"]) // 2
        out["rgb"] = torch.stack(frames["rgb"], 1).view(n, T, 2, *frames["rgb"][0].shape[1:])
        out["depth"] = torch.stack(frames["depth"], 1).view(n, T, 2, *frames["depth"][0].shape[1:])
    return out


def rollout_recycled(env, episodes: int, max_steps: int) -> Dict[str, torch.Tensor]:
    """At least ``episodes`` episodes on the scene batch with finished scenes recycled: after every pick/place round the scenes
    whose demonstration is complete, has failed or has used ``max_steps`` rounds are recorded, reset (new props every 10th
    episode of a scene, transporter_network_data_generation.py:99-101,118) and start their next episode while the others carry
    on, so every round steps a full batch (a late round of ``rollout_episode`` steps only the stragglers, and one scene-step
    costs the same ~0.2 ms whatever the batch).  Same record layout as ``rollout_episode`` (leading dim = episodes, in
    completion order), state observations only."""
    n, dev = env.num_envs, env.device
    obs0 = env.reset().observation
    actions = torch.zeros(n, max_steps, 2, ACTION_WIDTH, dtype=torch.float64, device=dev)
    valid = torch.zeros(n, max_steps, dtype=torch.bool, device=dev)
    pos0, quat0 = obs0["prop_pos"].clone(), obs0["prop_quat"].clone()
    t = torch.zeros(n, dtype=torch.int64, device=dev)
    played = torch.zeros(n, dtype=torch.int64, device=dev)           # episodes finished per scene
    status = torch.zeros(n, dtype=torch.int32, device=dev)           # reset() clears the kernel's flags for the whole batch
    e = torch.arange(n, device=dev)
    chunks, done = [], 0
    while done < episodes:
        in_progress, pick, place = env.sort_colours()                # flags scenes without a collision-free place pose as failed
        act = in_progress & ~env.failed & (t < max_steps)
        ended = ~act
        if bool(ended.any()):
            status |= env.physics.status()
            unsorted = (env.has_prop & ~env.expert.within_target()).any(1)
            chunks.append({"actions": actions[ended].clone(), "valid": valid[ended].clone(), "prop_pos0": pos0[ended].clone(),
                           "prop_quat0": quat0[ended].clone(), "prop_colour": env.prop_colour[ended].clone(), "has_prop": env.has_prop[ended].clone(),
                           "success": (~unsorted & ~env.failed)[ended], "failed": env.failed[ended].clone(), "rounds": valid[ended].sum(1),
                           "status": status[ended].clone()})
            done += int(ended.sum())
            played[ended] += 1
            fresh = ended & (played % 10 == 0)                       # a "new env" in the reference: props re-drawn
            if bool((ended & ~fresh).any()):
                env.reset(mask=ended & ~fresh)
            if bool(fresh.any()):
                env.reset(mask=fresh, resample_props=True)
            obs = env._compute_observation()
            pos0[ended], quat0[ended] = obs["prop_pos"][ended], obs["prop_quat"][ended]
            actions[ended] = 0; valid[ended] = False; t[ended] = 0; status[ended] = 0
            continue
        pick_px = env.world_2_pixel("overhead_camera/overhead_camera", pick[:, :3])
        place_px = env.world_2_pixel("overhead_camera/overhead_camera", place[:, :3])
        actions[e, t, 0] = pack_action(pick, pick_px); actions[e, t, 1] = pack_action(place, place_px)
        valid[e, t] = True
        env.step({"pose": pick.clone(), "pixel_coords": pick_px, "gripper_rot": 0.0})
        env.step({"pose": place.clone(), "pixel_coords": place_px, "gripper_rot": 0.0})
        t += 1
    return {k: torch.cat([c[k] for c in chunks], 0) for k in chunks[0]}


def generate(num_envs: int = 1024, episodes_per_env: int = 1, cfg: Optional[dict] = None, device=None, dtype=torch.float32,
             seed: Optional[int] = None, render: bool = False, max_steps: Optional[int] = None, recycle: bool = False):
    """Run ``episodes_per_env`` episodes on ``num_envs`` scenes of THIS rank, gather to rank 0.
    -> (demos dict on rank 0 / None elsewhere, stats dict summed over ranks)."""
    from .env import BatchedRearrangementEnv
    rank, world, local = distlib.init()
    cfg = cfg if cfg is not None else cfglib.compose(COLOUR_SPLITTER_OVERRIDES)
    T = int(cfg["dataset"]["max_steps"] if max_steps is None else max_steps)
    base_seed = int(cfg["task"]["initializers"]["seed"]) if seed is None else int(seed)
    env = BatchedRearrangementEnv(cfg=cfg, num_envs=num_envs, device=device if device is not None else "cuda:%d" % local, dtype=dtype,
                                  seed=base_seed + 100003 * rank, strict=False)
    demos, t0 = [], time.perf_counter()
    launches0 = env.physics.launches
    if recycle:
        if render:
            raise ValueError("recycle=True records state observations only")
        demos.append(rollout_recycled(env, num_envs * episodes_per_env, T))
    for ep in range(0 if recycle else episodes_per_env):
        # the reference builds a fresh env (new props) every 10 episodes (transporter_network_data_generation.py:99-101,118)
        demos.append(rollout_episode(env, T, render, resample_props=ep > 0 and ep % 10 == 0))
    torch.cuda.synchronize(env.device)
    wall = time.perf_counter() - t0
    local = {k: torch.cat([d[k] for d in demos], 0) for k in demos[0] if all(k in d for d in demos)}
    stats = torch.tensor([float(local["success"].numel()), float(local["success"].sum()), float(local["rounds"].sum()), float(local["failed"].sum()),
                          float(env.physics.time.sum().item() / env.physics.model.opt.timestep), float(env.physics.launches - launches0)]
                         + [float(((local["status"] & (1 << b)) != 0).sum()) for b in range(len(env.physics.STATUS_FLAGS))],
                         dtype=torch.float64, device=env.device)
    wall_t = distlib.max_over_ranks(torch.tensor([wall], dtype=torch.float64, device=env.device))
    distlib.reduce_stats(stats)
    torch.cuda.synchronize(env.device)
    tg0 = time.perf_counter()
    gathered = {k: distlib.gather_demos(v.to(torch.uint8) if v.dtype == torch.bool else v) for k, v in local.items()}
    torch.cuda.synchronize(env.device)
    gather_s = distlib.max_over_ranks(torch.tensor([time.perf_counter() - tg0], dtype=torch.float64, device=env.device)).item()
    nbytes = torch.tensor([float(sum(v.numel() * v.element_size() for v in local.values()))], dtype=torch.float64, device=env.device)
    distlib.reduce_stats(nbytes)                           # bytes of demonstration records over all ranks (what rank 0 ends up holding)
    s = stats.tolist()
    stat = {"episodes": s[0], "successes": s[1], "pick_place_rounds": s[2], "failed": s[3], "physics_env_steps_last_episode": s[4], "kernel_launches": s[5],
            "wall_s": wall_t.item(), "world_size": world, "envs_per_gpu": num_envs, "episodes_per_s": s[0] / max(wall_t.item(), 1e-9),
            "gather_s": gather_s, "gathered_bytes": nbytes.item(), "frames_rendered": bool(render),
            "flagged_episodes": {k: s[6 + i] for i, k in enumerate(env.physics.STATUS_FLAGS)}}
    env.close()
    return (gathered if rank == 0 else None), stat


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=1024, help="scenes per GPU")
    ap.add_argument("--episodes", type=int, default=1, help="episodes per scene")
    ap.add_argument("--max-steps", type=int, default=None)
    ap.add_argument("--render", action="store_true")
    ap.add_argument("--recycle", action="store_true", help="finished scenes start their next episode at once (full batch every round)")
    ap.add_argument("--out", default=None, help="rank 0 writes <out>.npz")
    a = ap.parse_args()
    demos, stat = generate(a.envs, a.episodes, render=a.render, max_steps=a.max_steps, recycle=a.recycle)
    if demos is not None:
        if a.out:
            os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
            np.savez_compressed(a.out, **{k: v.cpu().numpy() for k, v in demos.items()})
        stat["demo_rows_on_rank0"] = int(demos["actions"].shape[0])
        print(json.dumps(stat), flush=True)
    if distlib._world() > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
