"""Batched controller tuning: CMA-ES over the six OSC gains, one (candidate, target) pair per scene.

Mirrors ``automated_controller_tuning/rearrangement_controller_tuning.py``:
  * ``p2p_fitness`` (:43-70): target position ~ U(workspace) + (0, 0, U(0.15, 0.55)), orientation = mju_mat2Quat(R_xyz(0, 180,
    U(-2 pi, 2 pi)) in degrees), zero target velocities, ``run_controller(4.0)``; fitness = 1e3 * position error + orientation
    error + 1e6 if the controller did not converge;
  * the evolution loop (:145-198): CMA-ES, popsize 20, 6 dims, sigma_init 500, mean initialised at 500, candidates mapped
    through ``abs`` ("only consider positive params"), gains written to ``controller_gains`` before every roll-out.
The reference evaluates the 20 candidates one after the other on one scene; here every candidate is evaluated on
``targets_per_candidate`` targets at once -- popsize x targets scenes in ONE ``b2p_step_osc`` launch per generation, the
candidate's gains entering the kernel through ``b2p_set_env_gains``.  evosax / jax are not installed: the strategy below is the
standard (mu/mu_w, lambda)-CMA-ES (Hansen's tutorial defaults, which are also evosax's).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import numpy as np
import torch

from . import config as cfglib
from .controllers import BatchedOSC
from .expert import euler_xyz_matrix, mat2quat
from .physics import ARM_HOME, BatchedPhysics
from .robot_arm import BatchedRobotArm


class CMAES:
    """Minimal CMA-ES (ask / tell), minimisation."""

    def __init__(self, num_dims: int, popsize: int, sigma_init: float, mean_init: float, seed: int = 0):
        self.n, self.lam = num_dims, popsize
        self.mu = popsize // 2
        w = math.log(self.mu + 0.5) - np.log(np.arange(1, self.mu + 1))
        self.w = w / w.sum()
        self.mueff = 1.0 / (self.w ** 2).sum()
        n = num_dims
        self.cc = (4 + self.mueff / n) / (n + 4 + 2 * self.mueff / n)
        self.cs = (self.mueff + 2) / (n + self.mueff + 5)
        self.c1 = 2 / ((n + 1.3) ** 2 + self.mueff)
        self.cmu = min(1 - self.c1, 2 * (self.mueff - 2 + 1 / self.mueff) / ((n + 2) ** 2 + self.mueff))
        self.damps = 1 + 2 * max(0.0, math.sqrt((self.mueff - 1) / (n + 1)) - 1) + self.cs
        self.chi_n = math.sqrt(n) * (1 - 1 / (4 * n) + 1 / (21 * n * n))
        self.mean = np.full(n, float(mean_init))
        self.sigma = float(sigma_init)
        self.C, self.pc, self.ps = np.eye(n), np.zeros(n), np.zeros(n)
        self.rng = np.random.default_rng(seed)
        self.gen = 0
        self.best_fitness, self.best_member = np.inf, self.mean.copy()

    def ask(self) -> np.ndarray:
        d2, self.B = np.linalg.eigh(self.C)
        self.D = np.sqrt(np.maximum(d2, 1e-20))
        self.z = self.rng.standard_normal((self.lam, self.n))
        self.y = (self.z * self.D) @ self.B.T
        return self.mean + self.sigma * self.y

    def tell(self, x: np.ndarray, fitness: np.ndarray):
        order = np.argsort(fitness)
        if fitness[order[0]] < self.best_fitness:
            self.best_fitness, self.best_member = float(fitness[order[0]]), x[order[0]].copy()
        ysel = self.y[order[:self.mu]]
        yw = self.w @ ysel
        self.mean = self.mean + self.sigma * yw
        inv_sqrt = self.B @ np.diag(1 / self.D) @ self.B.T
        self.ps = (1 - self.cs) * self.ps + math.sqrt(self.cs * (2 - self.cs) * self.mueff) * (inv_sqrt @ yw)
        self.gen += 1
        hsig = np.linalg.norm(self.ps) / math.sqrt(1 - (1 - self.cs) ** (2 * self.gen)) / self.chi_n < 1.4 + 2 / (self.n + 1)
        self.pc = (1 - self.cc) * self.pc + hsig * math.sqrt(self.cc * (2 - self.cc) * self.mueff) * yw
        rank_mu = (ysel * self.w[:, None]).T @ ysel
        self.C = (1 - self.c1 - self.cmu) * self.C + self.c1 * (np.outer(self.pc, self.pc) + (1 - hsig) * self.cc * (2 - self.cc) * self.C) + self.cmu * rank_mu
        self.C = (self.C + self.C.T) / 2
        self.sigma *= math.exp((self.cs / self.damps) * (np.linalg.norm(self.ps) / self.chi_n - 1))


class BatchedTuner:
    def __init__(self, popsize: int = 20, targets_per_candidate: int = 32, cfg: Optional[dict] = None, device=None,
                 dtype: torch.dtype = torch.float32, seed: int = 0, sigma_init: float = 500.0, mean_init: float = 500.0):
        self.cfg = cfg if cfg is not None else cfglib.compose(["arena/props=single_block"])     # the reference's TUNING_CONFIG (:133-137)
        self.popsize, self.tpc = int(popsize), int(targets_per_candidate)
        n = self.popsize * self.tpc
        self.physics = BatchedPhysics(n, dtype=dtype, device=device)
        p = self.physics
        p.set_out_fields("SITE")        # the fitness functions read the end-effector site only
        self.osc = BatchedOSC(p, None, self.cfg["robots"]["arm"]["controller_config"]["controller_params"])
        self.robot = BatchedRobotArm(p, self.osc, None, self.cfg["control_dt"])
        self.strategy = CMAES(6, self.popsize, sigma_init, mean_init, seed)
        self.gen = torch.Generator(device=p.device); self.gen.manual_seed(seed + 1)
        self.best_params = np.abs(self.strategy.best_member)
        self._env, self._seed = None, seed
        # point-to-point motion needs no props: all parked with collisions off (the reference's single block plays no part in p2p_fitness)
        self._home = torch.zeros_like(p.state[0])
        self._home[:p.model.nq] = p.model.qpos0.to(p.device)
        self._home[0:7] = torch.tensor(ARM_HOME, device=p.device, dtype=p.dtype)
        for k, a in enumerate(p.model.prop_qposadr):
            self._home[a:a + 7] = torch.tensor([5.0 + k, 5.0, 0.2, 1, 0, 0, 0], device=p.device, dtype=p.dtype)
        p.set_env_params(None, torch.zeros(n, p.maxprop, dtype=torch.uint8))

    def sample_targets(self):
        """p2p_fitness target distribution (:46-53), one per scene."""
        p = self.physics
        w = self.cfg["task"]["initializers"]["workspace"]
        lo, hi = (torch.tensor(w[k], device=p.device, dtype=torch.float64) for k in ("min_pose", "max_pose"))
        u = lambda *shape: torch.rand(*shape, device=p.device, dtype=torch.float64, generator=self.gen)
        pos = lo + (hi - lo) * u(p.n_env, 3)
        pos[:, 2] += 0.15 + 0.4 * u(p.n_env)
        yaw_deg = -2 * math.pi + 4 * math.pi * u(p.n_env)
        z = torch.zeros_like(yaw_deg)
        quat = mat2quat(euler_xyz_matrix(z, z + math.pi, torch.deg2rad(yaw_deg)))
        return pos.to(p.dtype), quat.to(p.dtype)

    def p2p_fitness(self, params: np.ndarray, duration: float = 4.0) -> np.ndarray:
        """``params`` [popsize, 6] (already positive) -> fitness [popsize] = mean over the candidate's targets."""
        p = self.physics
        p.state[:] = self._home
        gains = torch.tensor(params, device=p.device, dtype=p.dtype).repeat_interleave(self.tpc, dim=0)
        self.osc.set_gains_per_scene(gains)
        pos, quat = self.sample_targets()
        zeros = torch.zeros(3, device=p.device, dtype=p.dtype)
        self.osc.set_target(position=pos, velocity=zeros, quat=quat, angular_velocity=zeros)
        ok = self.robot.run_controller(duration)
        f = 1e3 * self.osc.current_position_error() + self.osc.current_orientation_error().norm(dim=1) + (~ok).to(p.dtype) * 1e6
        bad = (p.status(clear=True) & 7) != 0                      # a diverged roll-out (BADQPOS / BADQVEL / BADQACC) is a failure too
        f = torch.where(bad | ~torch.isfinite(f), torch.full_like(f, 2e6), f)
        return f.view(self.popsize, self.tpc).mean(1).double().cpu().numpy()

    def circle_fitness(self, params: np.ndarray, radius: float = 0.1, speed: float = 0.1) -> np.ndarray:
        """:73-111: fixed orientation R_xyz(0, 180, 0), the first 20 of 100 way-points of a circle of ``radius`` around
        (0.25, 0, 0.9); ``run_controller(5.0)`` to the first, ``run_controller(1.0)`` to each; fitness = 0, or 1e6 for a candidate
        that missed a way-point (the reference raises there).  ``speed`` is unused, as in the reference."""
        p = self.physics
        p.state[:] = self._home
        self.osc.set_gains_per_scene(torch.tensor(params, device=p.device, dtype=p.dtype).repeat_interleave(self.tpc, dim=0))
        z = torch.zeros(1, device=p.device, dtype=torch.float64)
        quat = mat2quat(euler_xyz_matrix(z, z + math.pi, z)).to(p.dtype).expand(p.n_env, 4)
        zeros = torch.zeros(3, device=p.device, dtype=p.dtype)
        ang = torch.arange(20, device=p.device, dtype=torch.float64) * (2 * math.pi / 100)
        way = torch.stack([radius * torch.cos(ang) + 0.25, radius * torch.sin(ang), torch.full_like(ang, 0.9)], 1).to(p.dtype)
        goto = lambda w, dur: (self.osc.set_target(position=w.expand(p.n_env, 3), velocity=zeros, quat=quat, angular_velocity=zeros),
                               self.robot.run_controller(dur))[1]
        goto(way[0], 5.0)
        ok = torch.ones(p.n_env, dtype=torch.bool, device=p.device)
        for w in way:
            ok &= goto(w, 1.0)
        ok &= (p.status(clear=True) & 7) == 0
        return ((~ok).double() * 1e6).view(self.popsize, self.tpc).mean(1).cpu().numpy()

    def pick_place_fitness(self, params: np.ndarray) -> np.ndarray:
        """Second tuning stage (:113-129,161-171): ``random_pick_and_place`` on a single-block scene, pick then place through the
        env's phase tables, fitness = 1000 x distance of the block to the place target (whose z the place phase table has set to
        0.575, as in the reference), 1e6 for a scene in which a phase did not converge.  One (candidate, place target) per scene."""
        env = self._pick_place_env()
        env.reset()
        env._osc.set_gains_per_scene(torch.tensor(params, device=env.device, dtype=env.dtype).repeat_interleave(self.tpc, dim=0))
        pick, place = env.random_pick_and_place()
        env.pick(pick); env.place(place)
        pos = env.props_info["position"][torch.arange(env.num_envs, device=env.device), env.has_prop.to(torch.int64).argmax(1)]
        f = 1000.0 * (place[:, :3] - pos).norm(dim=1)
        bad = env.failed | ((env.physics.status(clear=True) & 7) != 0) | ~torch.isfinite(f)
        f = torch.where(bad, torch.full_like(f, 1e6), f)
        return f.view(self.popsize, self.tpc).mean(1).double().cpu().numpy()

    def _pick_place_env(self):
        if self._env is None:
            from .env import BatchedRearrangementEnv
            self._env = BatchedRearrangementEnv(cfg=self.cfg, num_envs=self.popsize * self.tpc, device=self.physics.device, dtype=self.physics.dtype,
                                                seed=self._seed, strict=False, camera_obs=False)
        return self._env

    def run(self, generations: int = 50, duration: float = 4.0, method: str = "p2p") -> List[Dict[str, float]]:
        """One CMA-ES stage (:183-192).  The reference runs ``p2p`` and then ``pick_place`` with the same strategy state."""
        evaluate = {"p2p": lambda x: self.p2p_fitness(x, duration), "circle": self.circle_fitness, "pick_place": self.pick_place_fitness}[method]
        hist = []
        for g in range(generations):
            x = self.strategy.ask()
            fit = evaluate(np.abs(x))                                                         # "only consider positive params" (:187)
            self.strategy.tell(x, fit)
            self.best_params = np.abs(self.strategy.best_member)
            hist.append({"generation": g, "best_fitness": self.strategy.best_fitness, "mean_fitness": float(fit.mean()), "sigma": self.strategy.sigma})
        return hist
