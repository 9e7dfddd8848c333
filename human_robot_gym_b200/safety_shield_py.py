"""Drop-in for the reference's pybind module `safety_shield_py` (SS/bindings/safety_shield_bindings.cpp:26-116):
`from human_robot_gym_b200.safety_shield_py import Motion, LongTermTraj, SafetyShield, ShieldType`.

`SafetyShield` keeps the reference's constructor / method names, keyword arguments, list-in / list-out conventions and
error behaviour, but its arithmetic runs in the CUDA library behind include/hrg_shield.h: one `SafetyShield` object is a
batch of ONE env on a `BatchedSafetyShield` (there is no CPU path for step / humanMeasurement; constructing it without a
CUDA device raises). `FailsafeController` (failsafe_controller.py:23,173,232-250) uses it unchanged; a vectorised env uses
`BatchedSafetyShield` directly and gets the same per-env results.

`Motion` and `LongTermTraj` are the value types that cross the boundary (the reference copies them by value, §8(b)); they
are plain host-side containers. `LongTermTraj.interpolate` & co. are only the reference's *accessors* for user-built
trajectories (tests/bindings/test_ltt.py) and are not on the step path.
"""
from __future__ import annotations

import enum
import math
from typing import List, Optional, Sequence

import numpy as np


class ShieldType(enum.IntEnum):
    """SS/include/safety_shield/safety_shield.h:55 (py::arithmetic enum)."""
    OFF = 0
    SSM = 1
    PFL = 2


def _vec(x) -> List[float]:
    return [float(v) for v in np.asarray(x, dtype=np.float64).ravel()]


class Motion:
    """SS/include/safety_shield/motion.h:30-275 / SS/src/motion.cc (all six constructors). Defaults are the BINDING's
    (SS/bindings/safety_shield_bindings.cpp:41-44: every threshold 1e-4; the C++ header default of hasSamePos is 1e-3)."""

    def __init__(self, *args, **kw):
        # pybind overloads (bindings :30-40): (), (nb_modules), (time, q[, dq[, ddq[, dddq]]], s=0.0); keywords as named there
        self._time, self._s = 0.0, 0.0
        unknown = set(kw) - {"nb_modules", "time", "q", "dq", "ddq", "dddq", "s"}
        if unknown:
            raise TypeError(f"Motion(): incompatible constructor arguments {sorted(unknown)}")
        if not args and not kw:
            self._q, self._dq, self._ddq, self._dddq = [], [], [], []
            return
        if "nb_modules" in kw or (len(args) == 1 and not kw and isinstance(args[0], (int, np.integer))
                                  and not isinstance(args[0], bool)):
            if len(args) + len(kw) != 1:
                raise TypeError("Motion(nb_modules)")
            n = int(kw["nb_modules"] if kw else args[0])
            self._q, self._dq, self._ddq, self._dddq = ([0.0] * n for _ in range(4))
            return
        a = list(args)
        if a:
            self._time = float(a.pop(0))
        elif "time" in kw:
            self._time = float(kw["time"])
        else:
            raise TypeError("Motion(time, q[, dq[, ddq[, dddq]]][, s])")
        if len(a) >= 2 and np.isscalar(a[-1]):   # trailing positional s
            self._s = float(a.pop())
        for name in ("q", "dq", "ddq", "dddq")[len(a):]:
            if name in kw:
                a.append(kw[name])
            else:
                break
        if not 1 <= len(a) <= 4:
            raise TypeError("Motion(time, q[, dq[, ddq[, dddq]]][, s])")
        vs = [_vec(v) for v in a]
        n = len(vs[0])
        while len(vs) < 4:
            vs.append([0.0] * n)
        self._q, self._dq, self._ddq, self._dddq = vs
        if "s" in kw:
            self._s = float(kw["s"])

    # motion.h:142-179
    def isStopped(self, threshold: float = 1e-4) -> bool:
        return all(abs(a) <= threshold and abs(b) <= threshold and abs(c) <= threshold
                   for a, b, c in zip(self._dq, self._ddq, self._dddq))

    def hasSamePos(self, motion: "Motion", threshold: float = 1e-4) -> bool:
        return all(abs(a - b) <= threshold for a, b in zip(self._q, motion._q))

    def hasSameVel(self, motion: "Motion", threshold: float = 1e-4) -> bool:
        return all(abs(a - b) <= threshold for a, b in zip(self._dq, motion._dq))

    def hasSameAcc(self, motion: "Motion", threshold: float = 1e-4) -> bool:
        return all(abs(a - b) <= threshold for a, b in zip(self._ddq, motion._ddq))

    def getNbModules(self): return len(self._q)
    def getTime(self): return self._time
    def getS(self): return self._s
    def getAngle(self): return list(self._q)
    def getVelocity(self): return list(self._dq)
    def getAcceleration(self): return list(self._ddq)
    def getJerk(self): return list(self._dddq)
    def setTime(self, new_time): self._time = float(new_time)
    def setS(self, new_s): self._s = float(new_s)
    def setAngle(self, new_q): self._q = _vec(new_q)
    def setVelocity(self, new_dq): self._dq = _vec(new_dq)
    def setAcceleration(self, new_ddq): self._ddq = _vec(new_ddq)
    def setJerk(self, new_dddq): self._dddq = _vec(new_dddq)

    def __repr__(self):
        return f"Motion(time={self._time}, s={self._s}, q={self._q})"


class LongTermTraj:
    """SS/include/safety_shield/long_term_traj.h:38-260, SS/src/long_term_traj.cc:41-159 (SSM constructor)."""

    def __init__(self, long_term_traj: Optional[Sequence[Motion]] = None, sample_time: float = 0.01,
                 starting_index: int = 0, sliding_window_k: int = 10, alpha_i_max: float = 1.0):
        self._cur = 0
        self._start = int(starting_index)
        self._max_acc, self._max_jerk = [], []
        if long_term_traj is None:                      # default ctor: one zero Motion(1), alpha_i = {1, 1}
            self._traj = [Motion(1)]
            self._ts = 0.01
            self._alpha = [1.0, 1.0]
        else:
            self._traj = list(long_term_traj)
            self._ts = float(sample_time)
            self.calculate_max_acc_jerk_window(self._traj, sliding_window_k)
            self._alpha = [float(alpha_i_max)] * self._traj[0].getNbModules()

    def setLongTermTrajectory(self, long_term_traj, sample_time: Optional[float] = None):
        self._traj = list(long_term_traj)
        if sample_time is not None:
            self._ts = float(sample_time)

    def getLength(self): return len(self._traj)
    def getCurrentPos(self): return self._cur
    def getCurrentMotion(self): return self._traj[self._cur]
    def getNextMotion(self): return self._traj[min(self._cur + 1, len(self._traj) - 1)]
    def getNextMotionAtIndex(self, index): return self._traj[self.getTrajectoryIndex(index)]
    def getAlphaI(self): return list(self._alpha)
    def getMaxAccelerationWindow(self, index): return list(self._max_acc[self.getTrajectoryIndex(index)])
    def getMaxJerkWindow(self, index): return list(self._max_jerk[self.getTrajectoryIndex(index)])

    def getTrajectoryIndex(self, index: int) -> int:
        return max(min(int(index) - self._start, len(self._traj) - 1), self._cur)

    def increasePosition(self):
        self._cur = min(self._cur + 1, len(self._traj) - 1)

    def interpolate(self, s, ds, dds, ddds, v_max_allowed, a_max_allowed, j_max_allowed) -> Motion:
        """long_term_traj.cc:41-77 (the jerk is not clamped; j_max_allowed is unused there too)."""
        ind = s / self._ts
        ind_mod, intpart = math.modf(ind)
        m = self.getNextMotionAtIndex(int(intpart))
        dt = ind_mod * self._ts
        q, dq, ddq, dddq = [], [], [], []
        for i, (q1, dq1, ddq1, dddq1) in enumerate(zip(m._q, m._dq, m._ddq, m._dddq)):
            q.append(q1 + dt * dq1 + 1.0 / 2 * dt * dt * ddq1 + 1.0 / 6 * dt * dt * dt * dddq1)
            v_max_int = dq1 + dt * ddq1 + 1.0 / 2 * dt * dt * dddq1
            dq.append(min(max(v_max_int * ds, -v_max_allowed[i]), v_max_allowed[i]))
            a_max_int = ddq1 + dt * dddq1
            ddq.append(min(max(v_max_int * dds + ds * ds * a_max_int, -a_max_allowed[i]), a_max_allowed[i]))
            dddq.append(dddq1 * ds * ds * ds + 3.0 * a_max_int * dds * ds + v_max_int * ddds)
        return Motion(0.0, q, dq, ddq, dddq, s)

    def calculate_max_acc_jerk_window(self, long_term_traj, k: int):
        """Forward-looking sliding-window maximum of acceleration and jerk per joint (long_term_traj.cc:79-159):
        entry i = max over samples [i, i + k). The reference appends to its member vectors; so does this."""
        n = len(long_term_traj)
        k = min(n, int(k))
        acc = np.array([m._ddq for m in long_term_traj], dtype=np.float64)
        jerk = np.array([m._dddq for m in long_term_traj], dtype=np.float64)
        for i in range(n):
            hi = min(n, i + k)
            self._max_acc.append([float(v) for v in acc[i:hi].max(axis=0)])
            self._max_jerk.append([float(v) for v in jerk[i:hi].max(axis=0)])

    # helpers for the batched boundary (not part of the reference surface)
    def _as_array(self) -> np.ndarray:
        return np.array([[m._q, m._dq, m._ddq, m._dddq] for m in self._traj], dtype=np.float64)  # [L, 4, J]


class RobotMovementException(RuntimeError):
    """SS/include/safety_shield/exceptions.h:24-41 (surfaces as RuntimeError through pybind)."""


class TrajectoryException(RuntimeError):
    """SS/include/safety_shield/exceptions.h:43-62."""


class SafetyShield:
    """`safety_shield_py.SafetyShield` (bindings :82-115): one env = a batch of one on the CUDA library."""

    def __init__(self, sample_time: Optional[float] = None, trajectory_config_file: Optional[str] = None,
                 robot_config_file: Optional[str] = None, mocap_config_file: Optional[str] = None,
                 init_x: float = 0.0, init_y: float = 0.0, init_z: float = 0.0, init_roll: float = 0.0,
                 init_pitch: float = 0.0, init_yaw: float = 0.0, init_qpos=None,
                 shield_type: ShieldType = ShieldType.SSM, device=None, ltt_capacity: Optional[int] = None):
        self._b = None
        self._type = ShieldType(int(shield_type))
        self._safe = False
        if sample_time is None:          # SafetyShield() default ctor of the bindings: an empty shell
            return
        from .batched import BatchedSafetyShield  # raises without the CUDA extension / a CUDA device
        pose = np.array([[init_x, init_y, init_z, init_roll, init_pitch, init_yaw]], dtype=np.float64)
        q0 = np.asarray(init_qpos, dtype=np.float64).reshape(1, -1)
        self._b = BatchedSafetyShield(1, float(sample_time), trajectory_config_file, robot_config_file,
                                      mocap_config_file, base_pose=pose, init_qpos=q0, shield_type=int(self._type),
                                      device=device, ltt_capacity=ltt_capacity, store_capsules=True)
        self._J = self._b.J
        self._safe = self._type == ShieldType.OFF

    def _need(self):
        if self._b is None:
            raise RuntimeError("SafetyShield was default-constructed; no configuration loaded")
        return self._b

    def reset(self, init_x, init_y, init_z, init_roll, init_pitch, init_yaw, init_qpos, current_time,
              shield_type: ShieldType = ShieldType.SSM):
        b = self._need()
        self._type = ShieldType(int(shield_type))
        pose = np.array([[init_x, init_y, init_z, init_roll, init_pitch, init_yaw]], dtype=np.float64)
        b.reset(pose, np.asarray(init_qpos, dtype=np.float64).reshape(1, -1), float(current_time), int(self._type))
        self._safe = self._type == ShieldType.OFF

    def step(self, cycle_begin_time: float) -> Motion:
        b = self._need()
        q, dq, ddq, dddq = b.step(float(cycle_begin_time))
        out = [x[0].cpu().numpy() for x in (q, dq, ddq, dddq)]      # the caller consumes the setpoint now: one sync
        self._safe = bool(b.is_safe[0].item())
        return Motion(float(cycle_begin_time), out[0], out[1], out[2], out[3])

    def newLongTermTrajectory(self, goal_position, goal_velocity):
        b = self._need()
        b.newLongTermTrajectory(np.asarray(goal_position, dtype=np.float64).reshape(1, -1),
                                np.asarray(goal_velocity, dtype=np.float64).reshape(1, -1))

    def setLongTermTrajectory(self, traj: LongTermTraj):
        """safety_shield.cc:699-730: raises when the robot moves or the trajectory does not start / end as required."""
        b = self._need()
        verdict = b.set_ltt(0, traj._as_array(), traj._ts, traj._start, traj._alpha[: self._J] if len(traj._alpha) >= self._J
                            else [traj._alpha[0]] * self._J)
        if verdict == 1:
            raise RobotMovementException("Robot must be at standstill when setting a new long-term trajectory")
        if verdict in (2, 3, 4):
            raise TrajectoryException({2: "Given LTT does not start at the same position as the current robot position",
                                       3: "Given LTT does not start with the same velocity as the current robot velocity",
                                       4: "Given LTT does not end in a complete stop of the robot"}[verdict])

    def humanMeasurement(self, human_measurement, time: float):
        b = self._need()
        kp = np.asarray([np.asarray(p, dtype=np.float64).ravel()[:3] for p in human_measurement], dtype=np.float64)
        if kp.shape != (b.K, 3):
            raise ValueError(f"expected {b.K} keypoints of 3 coordinates, got {kp.shape}")
        b.humanMeasurement(kp[None], float(time))

    def getRobotReachCapsules(self) -> List[List[float]]:
        robot, _, _ = self._need().capsules()
        return robot[0].cpu().numpy().tolist()

    def getHumanReachCapsules(self, type: int = 1) -> List[List[float]]:  # noqa: A002 (the binding's keyword)
        b = self._need()
        _, human, _ = b.capsules()
        h = human[0].cpu().numpy()
        n_ext, n_bod = b.cfg.n_ext, b.cfg.n_bodies
        lo, hi = {0: (0, n_ext), 1: (n_ext, n_ext + n_bod), 2: (n_ext + n_bod, n_ext + 2 * n_bod)}[int(type)]
        return h[lo:hi].tolist()

    def getSafety(self) -> bool:
        return bool(self._safe)

    def getShieldType(self) -> ShieldType:
        return self._type
