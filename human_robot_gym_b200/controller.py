"""Batched FailsafeController: the shield's immediate caller, over N envs on the device.

Mirrors HRG/controllers/failsafe_controller/failsafe_controller/failsafe_controller.py: `set_goal` (:252-300),
`set_human_measurement` (:302-310), `run_controller` (:312-373), `get_safety` (:375-381). Mass matrix, gravity /
Coriolis compensation and the joint state still come from the simulator (MuJoCo, out of scope): the caller passes them as
[N, ...] CUDA tensors. Everything else of the tick - goal from the policy action, the shield step, the PD+ torque law -
runs in three kernel launches for all envs, instead of five pybind crossings and three vector copies per env.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import cabi
from .batched import BatchedSafetyShield, _ptr


def _per_joint(x, J, name):
    a = np.asarray(x, dtype=np.float64)
    if a.ndim == 0:
        a = np.full(J, float(a))  # robosuite's nums2array: a scalar applies to every joint
    if a.shape != (J,):
        raise ValueError(f"{name}: expected a scalar or {J} values, got shape {a.shape}")
    return a


class BatchedFailsafeController:
    """Arguments follow JointPositionController / FailsafeController: `input_max/min`, `output_max/min` (action scaling),
    `kp`, `damping_ratio` (kd = 2 sqrt(kp) damping_ratio), `qpos_limits` ([2, J] or None), `actuator_range` ((low, high))."""

    def __init__(self, shield: BatchedSafetyShield, actuator_range, input_max=1, input_min=-1, output_max=0.05,
                 output_min=-0.05, kp=50, damping_ratio=1, qpos_limits=None):
        self.shield = shield
        J = shield.J
        k = cabi.HrgControllerConfig()
        k.nb_joints = J
        tab = {"input_min": input_min, "input_max": input_max, "output_min": output_min, "output_max": output_max,
               "torque_min": actuator_range[0], "torque_max": actuator_range[1]}
        self.kp = _per_joint(kp, J, "kp")
        self.kd = 2 * np.sqrt(self.kp) * _per_joint(damping_ratio, J, "damping_ratio")  # joint_pos.py: kd from damping
        tab["kp"], tab["kd"] = self.kp, self.kd
        if qpos_limits is not None:
            lim = np.asarray(qpos_limits, dtype=np.float64)
            tab["position_min"], tab["position_max"] = lim[0], lim[1]
            k.has_position_limits = 1
        for name, v in tab.items():
            arr = getattr(k, name)
            for j, x in enumerate(_per_joint(v, J, name)):
                arr[j] = x
        self._k = k
        f64 = dict(dtype=torch.float64, device=shield.device)
        self.goal_qpos = torch.zeros(shield.n_envs, J, **f64)
        self.torques = torch.zeros(shield.n_envs, J, **f64)

    def _t(self, x, shape):
        t = torch.as_tensor(x, dtype=torch.float64, device=self.shield.device).contiguous()
        if tuple(t.shape) != tuple(shape):
            raise ValueError(f"expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return t

    def set_goal(self, action, joint_pos, set_qpos=None, env_mask=None):
        """failsafe_controller.py:252-300 for the envs of `env_mask` (those whose policy step starts on this tick)."""
        s, N, J = self.shield, self.shield.n_envs, self.shield.J
        a = None if action is None else self._t(action, (N, J))
        q = None if joint_pos is None else self._t(joint_pos, (N, J))
        sp = None if set_qpos is None else self._t(set_qpos, (N, J))
        m = s._mask(env_mask)
        with torch.cuda.device(s.device):
            cabi.check(s.lib.hrg_controller_set_goal(s._h, C.byref(self._k), _ptr(m), _ptr(a), _ptr(q), _ptr(sp),
                                                     _ptr(self.goal_qpos), s._stream()), "hrg_controller_set_goal")
        return self.goal_qpos

    def set_human_measurement(self, human_measurement, time, env_mask=None):
        self.shield.humanMeasurement(human_measurement, time, env_mask=env_mask)

    def run_controller(self, joint_pos, joint_vel, mass_matrix, torque_compensation, current_time):
        """failsafe_controller.py:312-373: shield step, then tau = M (kp e + kd de + ddq_des) + compensation, clipped."""
        s, N, J = self.shield, self.shield.n_envs, self.shield.J
        q, dq = self._t(joint_pos, (N, J)), self._t(joint_vel, (N, J))
        M, tc = self._t(mass_matrix, (N, J, J)), self._t(torque_compensation, (N, J))
        s.step(current_time)
        with torch.cuda.device(s.device):
            cabi.check(s.lib.hrg_controller_torques(s._h, C.byref(self._k), _ptr(s.q), _ptr(s.dq), _ptr(s.ddq), _ptr(q),
                                                    _ptr(dq), _ptr(M), _ptr(tc), _ptr(self.torques), s._stream()),
                       "hrg_controller_torques")
        return self.torques

    def get_safety(self):
        return self.shield.getSafety()
