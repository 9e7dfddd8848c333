"""Host staging between a vectorised env and the batched shield.

In the reference every env owns its controller and calls the shield itself, three times per control tick
(HRG/environments/manipulation/human_env.py:503-524 -> failsafe_controller.py:300 `newLongTermTrajectory`, :310
`humanMeasurement`, :329 `step`), inside the worker processes of SB3's `SubprocVecEnv`
(HRG/utils/training_utils_SB3.py:45-75). With one shield for all envs the order of a tick becomes

    workers:  adapter.set_human_measurement(i, site_xpos, t)      # what _set_human_measurement passed on
              adapter.set_goal(i, goal_qpos)                       # policy steps only (every 25th tick)
    once:     adapter.step()                                       # ONE hrg_shield_tick_host for all envs
    workers:  q, dq, ddq = adapter.desired_motion(i); adapter.get_safety(i)

The adapter owns the page-locked buffers the C ABI copies from / to, so gathering is a row write and scattering a row
read; nothing is allocated per tick. Worker processes that cannot share the address space send their rows through the
VecEnv's pipes and the main process files them with `put_rows`.
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch


class VecEnvShieldAdapter:
    """`shield`: a BatchedSafetyShield or PipelinedHostShield (anything with n_envs, J, K and
    tick_host(kp, t, goal_mask, goal_q, q, dq, ddq, flags) on page-locked tensors, plus synchronize() or a device)."""

    def __init__(self, shield, pin_memory: Optional[bool] = None):
        self.shield = shield
        self.n_envs, self.J, self.K = int(shield.n_envs), int(shield.J), int(shield.K)
        pin = torch.cuda.is_available() if pin_memory is None else bool(pin_memory)
        f64 = dict(dtype=torch.float64, pin_memory=pin)
        self._kp = torch.zeros(self.n_envs, self.K, 3, **f64)
        self._t = torch.zeros(self.n_envs, **f64)
        self._goal = torch.zeros(self.n_envs, self.J, **f64)
        self._goal_mask = torch.zeros(self.n_envs, dtype=torch.uint8, pin_memory=pin)
        self._q, self._dq, self._ddq = (torch.zeros(self.n_envs, self.J, **f64) for _ in range(3))
        self._flags = torch.zeros(self.n_envs, dtype=torch.uint8, pin_memory=pin)
        # numpy views: workers in this process write rows without going through torch
        self.kp, self.t, self.goal, self.goal_mask = (x.numpy() for x in (self._kp, self._t, self._goal, self._goal_mask))
        self.q, self.dq, self.ddq, self.flags = (x.numpy() for x in (self._q, self._dq, self._ddq, self._flags))
        self._fresh = np.zeros(self.n_envs, dtype=bool)  # measurement received since the last step
        self.ticks = 0

    # ---- worker-facing: the FailsafeController methods that touched the shield ---------------------------------------
    def set_human_measurement(self, env: int, human_measurement, time: float):
        """failsafe_controller.py:302-310 (list of K [x, y, z] positions in HumanObject.joint_elements order)."""
        m = np.asarray(human_measurement, dtype=np.float64)
        if m.shape != (self.K, 3):
            raise ValueError(f"env {env}: expected {self.K} keypoints x 3, got shape {m.shape}")
        self.kp[env] = m
        self.t[env] = time
        self._fresh[env] = True

    def set_goal(self, env: int, goal_qpos):
        """failsafe_controller.py:300: safety_shield.newLongTermTrajectory(goal_qpos, zeros)."""
        g = np.asarray(goal_qpos, dtype=np.float64)
        if g.shape != (self.J,):
            raise ValueError(f"env {env}: expected {self.J} joint positions, got shape {g.shape}")
        self.goal[env] = g
        self.goal_mask[env] = 1

    def put_rows(self, envs: Sequence[int], human_measurements, times, goal_qpos=None, goal_envs=None):
        """Rows that arrived over the VecEnv's pipes: measurements [n,K,3], times [n]; goals [m,J] for goal_envs [m]."""
        envs = np.asarray(envs, dtype=np.int64)
        self.kp[envs] = np.asarray(human_measurements, dtype=np.float64).reshape(len(envs), self.K, 3)
        self.t[envs] = np.asarray(times, dtype=np.float64)
        self._fresh[envs] = True
        if goal_qpos is not None:
            ge = envs if goal_envs is None else np.asarray(goal_envs, dtype=np.int64)
            self.goal[ge] = np.asarray(goal_qpos, dtype=np.float64).reshape(len(ge), self.J)
            self.goal_mask[ge] = 1

    # ---- once per control tick ------------------------------------------------------------------------------------------
    def step(self, require_all: bool = True):
        """humanMeasurement + [newLongTermTrajectory] + step for every env: one batched call, then the setpoints are
        readable. require_all: every env must have delivered a measurement since the last step (the reference calls
        humanMeasurement before every step, human_env.py:505)."""
        if require_all and not self._fresh.all():
            missing = np.flatnonzero(~self._fresh)
            raise RuntimeError(f"{missing.size} env(s) delivered no human measurement this tick (first: {missing[:8].tolist()})")
        any_goal = bool(self.goal_mask.any())
        self.shield.tick_host(self._kp, self._t, self._goal_mask if any_goal else None, self._goal if any_goal else None,
                              self._q, self._dq, self._ddq, self._flags)
        if hasattr(self.shield, "synchronize"):
            self.shield.synchronize()
        else:
            torch.cuda.current_stream(self.shield.device).synchronize()
        self.goal_mask[:] = 0
        self._fresh[:] = False
        self.ticks += 1

    # ---- worker-facing: results ------------------------------------------------------------------------------------------
    def desired_motion(self, env: int):
        """(q, dq, ddq) of `safety_shield.step(time)` for this env (failsafe_controller.py:329-335)."""
        return self.q[env].copy(), self.dq[env].copy(), self.ddq[env].copy()

    def get_safety(self, env: int) -> bool:
        """failsafe_controller.py:375-381."""
        return bool(self.flags[env] & 1)

    def branch(self, env: int) -> int:
        return int(self.flags[env] >> 1)
