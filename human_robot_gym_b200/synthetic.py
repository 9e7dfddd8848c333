"""Synthetic shield inputs: a moving human next to each robot and a stream of goal poses.

Used by tests (identical inputs to the oracle and to the CUDA path) and by bench.py. Pure numpy on the host; the
caller uploads what it needs. Nothing here is on the hot path.

Human: a standing 23-keypoint skeleton (the keypoint set of SS/config/mujoco_mocap.yaml:7-13, z-up, metres) placed at
polar offset r ~ U[r_min, r_max], theta ~ U[0, 2pi) from the robot base with yaw ~ U[0, 2pi); every keypoint
coordinate oscillates, p_k(t) = p_k0 + A_k sin(2 pi f_k t + phi_k), A ~ U[0, 0.15] m, f ~ U[0.2, 1.5] Hz (speeds stay
below the model's v_max = 2 m/s). Goals: a clamped random walk, goal_k = clamp(goal_{k-1} + U[-0.2, 0.2]^J), mirroring
`FailsafeController.set_goal` with output_max 0.2 (failsafe_controller.py:285-300, config/failsafe.json).
"""
from __future__ import annotations

import numpy as np

# name -> rest position (x forward, y left, z up) of a 1.75 m person standing at the origin, arms hanging
_SKELETON = {
    "L_Hip": (0.0, 0.09, 0.93), "R_Hip": (0.0, -0.09, 0.93), "Torso": (0.0, 0.0, 1.05),
    "L_Knee": (0.0, 0.10, 0.50), "R_Knee": (0.0, -0.10, 0.50), "Spine": (0.0, 0.0, 1.18),
    "L_Ankle": (0.0, 0.10, 0.08), "R_Ankle": (0.0, -0.10, 0.08), "Chest": (0.0, 0.0, 1.30),
    "L_Toe": (0.13, 0.10, 0.02), "R_Toe": (0.13, -0.10, 0.02), "Neck": (0.0, 0.0, 1.50),
    "L_Thorax": (0.0, 0.08, 1.42), "R_Thorax": (0.0, -0.08, 1.42), "Head": (0.0, 0.0, 1.65),
    "L_ShouLder": (0.0, 0.19, 1.43), "R_ShouLder": (0.0, -0.19, 1.43),
    "L_Elbow": (0.0, 0.22, 1.15), "R_Elbow": (0.0, -0.22, 1.15),
    "L_Wrist": (0.05, 0.23, 0.90), "R_Wrist": (0.05, -0.23, 0.90),
    "L_Hand": (0.08, 0.23, 0.82), "R_Hand": (0.08, -0.23, 0.82),
}


def rest_pose(joint_names) -> np.ndarray:
    """[K,3] rest positions for the given keypoint names; unknown names get a torso-height default spread."""
    out = np.zeros((len(joint_names), 3))
    for i, n in enumerate(joint_names):
        if n in _SKELETON:
            out[i] = _SKELETON[n]
        else:  # other mocap sets (cmu_mocap*.yaml): spread them over a 1.75 m column deterministically
            out[i] = (0.03 * ((i * 7) % 5 - 2), 0.05 * ((i * 3) % 7 - 3), 0.1 + 1.6 * ((i * 11) % 17) / 16.0)
    return out


class SyntheticScenario:
    def __init__(self, n_envs: int, joint_names, q_min, q_max, seed: int, init_q=None, r_range=(0.3, 2.0),
                 amp_max=0.15, f_range=(0.2, 1.5), goal_step=0.2):
        rng = np.random.default_rng(seed)
        self.n, self.K = int(n_envs), len(joint_names)
        self.q_min, self.q_max = np.asarray(q_min, float), np.asarray(q_max, float)
        self.J = self.q_min.shape[0]
        base = rest_pose(joint_names)  # [K,3]
        r = rng.uniform(r_range[0], r_range[1], self.n)
        th = rng.uniform(0, 2 * np.pi, self.n)
        yaw = rng.uniform(0, 2 * np.pi, self.n)
        c, s = np.cos(yaw), np.sin(yaw)
        x = base[None, :, 0] * c[:, None] - base[None, :, 1] * s[:, None] + (r * np.cos(th))[:, None]
        y = base[None, :, 0] * s[:, None] + base[None, :, 1] * c[:, None] + (r * np.sin(th))[:, None]
        z = np.broadcast_to(base[None, :, 2], x.shape)
        self.p0 = np.stack([x, y, z], axis=-1)  # [n,K,3]
        self.amp = rng.uniform(0, amp_max, (self.n, self.K, 3))
        self.freq = rng.uniform(f_range[0], f_range[1], (self.n, self.K, 3))
        self.phase = rng.uniform(0, 2 * np.pi, (self.n, self.K, 3))
        self.goal_step = goal_step
        self._rng = rng
        if init_q is None:
            init_q = np.zeros(self.J)
        self.init_q = np.clip(np.broadcast_to(np.asarray(init_q, float), (self.n, self.J)), self.q_min, self.q_max).copy()
        self._goal = self.init_q.copy()

    def keypoints(self, t: float) -> np.ndarray:
        """[n,K,3] FP64 keypoint positions at time t."""
        return self.p0 + self.amp * np.sin(2 * np.pi * self.freq * t + self.phase)

    def next_goal(self, mask=None) -> np.ndarray:
        """Advance the goal random walk (for the envs in `mask`, default all) and return the goals [n,J]."""
        step = self._rng.uniform(-self.goal_step, self.goal_step, (self.n, self.J))
        new = np.clip(self._goal + step, self.q_min, self.q_max)
        if mask is None:
            self._goal = new
        else:
            m = np.asarray(mask, bool)
            self._goal[m] = new[m]
        return self._goal.copy()
