"""ctypes view of oracle/_ref/libsara_ref.so — the UNMODIFIED reference C++ behind oracle/ref_driver.cc.

TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product (human_robot_gym_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libsara_ref.so")
_lib = None
_dp = C.POINTER(C.c_double)
_u8p = C.POINTER(C.c_uint8)


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_int, C.c_double, C.c_char_p, C.c_char_p, C.c_char_p, _dp, _dp, C.c_int, C.c_int, C.c_int]
        L.ref_create_types.restype = C.c_void_p
        L.ref_create_types.argtypes = [C.c_int, C.c_double, C.c_char_p, C.c_char_p, C.c_char_p, _dp, _dp, C.c_int, C.c_int,
                                       C.c_void_p, C.c_int]
        L.ref_reset_types.argtypes = [C.c_void_p, _u8p, _dp, _dp, _dp, C.c_int, C.c_void_p]
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_set_threads.argtypes = [C.c_void_p, C.c_int]
        L.ref_num_keypoints.argtypes = [C.c_void_p]
        L.ref_num_joints.argtypes = [C.c_void_p]
        L.ref_reset.argtypes = [C.c_void_p, _u8p, _dp, _dp, _dp, C.c_int]
        L.ref_human_measurement.argtypes = [C.c_void_p, _u8p, _dp, _dp]
        L.ref_new_goal.argtypes = [C.c_void_p, _u8p, _dp, _dp]
        L.ref_step.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, _dp, _u8p, _u8p]
        L.ref_get_capsules.argtypes = [C.c_void_p, _dp, _dp, C.c_int, _dp]
        L.ref_get_state.argtypes = [C.c_void_p, C.c_int, _dp]
        L.ref_get_paths.argtypes = [C.c_void_p, _dp]
        L.ref_stage_reach_verify.argtypes = [C.c_void_p, C.c_int] + [_dp] * 10 + [C.c_int, _dp, _u8p]
        L.ref_capsule_dist.argtypes = [C.c_int, _dp, _dp, _dp]
        L.ref_plan_safety_shield.argtypes = [C.c_void_p, C.c_int, _dp, _u8p, _dp, _dp]
        L.ref_ltp_plan.restype = C.c_int
        L.ref_ltp_plan.argtypes = [C.c_int, C.c_double] + [_dp] * 10 + [C.c_int]
        L.ref_ltp_opt_braking.argtypes = [C.c_double, _dp, C.c_double, C.c_double, _dp, _dp, _dp]
        L.ref_ltp_opt_switch_times.argtypes = [C.c_double, _dp] + [C.c_double] * 5 + [_dp, _dp, C.POINTER(C.c_int)]
        L.ref_ltp_time_scaling.argtypes = [C.c_double, _dp] + [C.c_double] * 6 + [_dp, _dp, C.POINTER(C.c_int)]
        L.ref_ltt_interpolate.argtypes = [C.c_int, C.c_int, _dp, C.c_double, C.c_int, C.c_int, _dp, _dp, _dp, _dp, _dp]
        L.ref_human_reach.restype = C.c_int
        L.ref_ltt_max_window.argtypes = [C.c_int, C.c_int, _dp, C.c_double, C.c_int, _dp, _dp]
        L.ref_live_max_window.argtypes = [C.c_void_p, C.c_void_p, _dp, _dp]
        L.ref_human_reach.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, _dp, C.c_int]
        L.ref_poly_roots.restype = C.c_double
        L.ref_poly_roots.argtypes = [C.c_int, _dp, _dp, _dp]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _u8(a):
    return None if a is None else a.ctypes.data_as(_u8p)


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


class RefBatch:
    """N independent reference `SafetyShield` objects (one per env), SSM=1 / PFL=2 / OFF=0 (a scalar or one per env)."""

    def __init__(self, n, sample_time, traj_yaml, robot_yaml, mocap_yaml, base_pose, init_q, shield_type=1, threads=1):
        init_q = _f64(init_q)
        if init_q.ndim == 1:
            init_q = np.tile(init_q, (n, 1))
        base_pose = _f64(base_pose)
        if base_pose.ndim == 1:
            base_pose = np.tile(base_pose, (n, 1))
        self.n, self.J = n, init_q.shape[1]
        self._bp, self._iq = np.ascontiguousarray(base_pose), np.ascontiguousarray(init_q)
        ty = np.ascontiguousarray(np.broadcast_to(np.asarray(shield_type, np.int32), (n,)))
        self.h = lib().ref_create_types(n, float(sample_time), traj_yaml.encode(), robot_yaml.encode(), mocap_yaml.encode(),
                                        _d(self._bp), _d(self._iq), self.J, 1, ty.ctypes.data, int(threads))
        self.K = lib().ref_num_keypoints(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_destroy(self.h)
            self.h = None

    def set_threads(self, t):
        lib().ref_set_threads(self.h, int(t))

    def _mask(self, mask):
        return None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)

    def reset(self, base_pose, init_q, t, shield_type=1, mask=None):
        bp, iq, tt, m = _f64(base_pose, (self.n, 6)), _f64(init_q, (self.n, self.J)), _f64(t, (self.n,)), self._mask(mask)
        ty = np.ascontiguousarray(np.broadcast_to(np.asarray(shield_type, np.int32), (self.n,)))
        lib().ref_reset_types(self.h, _u8(m), _d(bp), _d(iq), _d(tt), 1, ty.ctypes.data)

    def human_measurement(self, kp, t, mask=None):
        kp, tt, m = _f64(kp, (self.n, self.K, 3)), _f64(np.broadcast_to(t, (self.n,))), self._mask(mask)
        lib().ref_human_measurement(self.h, _u8(m), _d(kp), _d(tt))

    def new_goal(self, goal_q, goal_dq=None, mask=None):
        gq = _f64(goal_q, (self.n, self.J))
        gd = np.zeros_like(gq) if goal_dq is None else _f64(goal_dq, (self.n, self.J))
        m = self._mask(mask)
        lib().ref_new_goal(self.h, _u8(m), _d(gq), _d(gd))

    def step(self, t):
        tt = _f64(np.broadcast_to(t, (self.n,)))
        out = {k: np.zeros((self.n, self.J)) for k in ("q", "dq", "ddq", "dddq")}
        out["is_safe"] = np.zeros(self.n, dtype=np.uint8)
        out["branch"] = np.zeros(self.n, dtype=np.uint8)
        lib().ref_step(self.h, _d(tt), _d(out["q"]), _d(out["dq"]), _d(out["ddq"]), _d(out["dddq"]),
                       _u8(out["is_safe"]), _u8(out["branch"]))
        return out

    def capsules(self, H):
        robot = np.zeros((self.n, self.J, 7))
        human = np.zeros((self.n, H, 7))
        md = np.full((self.n, 3), np.inf)
        lib().ref_get_capsules(self.h, _d(robot), _d(human), int(H), _d(md))
        return robot, human, md

    def live_max_window(self, index=None):
        """getMaxAccelerationWindow / getMaxJerkWindow(index) of every env's live LTT -> ([n,J], [n,J]);
        index [n] int32 or None (each env's path_s_discrete_)."""
        J = lib().ref_num_joints(self.h)
        acc, jerk = np.zeros((self.n, J)), np.zeros((self.n, J))
        idx = None if index is None else np.ascontiguousarray(index, dtype=np.int32)
        lib().ref_live_max_window(self.h, None if idx is None else idx.ctypes.data, _d(acc), _d(jerk))
        return acc, jerk

    def paths(self):
        """[n, 5, 11]: recovery, failsafe, failsafe_2, safe, potential path x (phases[6], pos, vel, acc, jerk, is_current)."""
        out = np.zeros((self.n, 5, 11))
        lib().ref_get_paths(self.h, _d(out))
        return out

    def stage_reach_verify(self, base_pose, q_start, q_goal, s_diff, t_reach, alpha_i, kp_pos, kp_vel, H):
        """reach + humanReachabilityAnalysis + verify_human_reach on explicit inputs (n cases, on the objects of env 0)
        -> robot [n,J,7], human [n,H,7], min_dist [n,3], is_safe [n]."""
        n = np.asarray(s_diff).shape[0]
        a = [_f64(base_pose, (n, 6)), _f64(q_start, (n, self.J)), _f64(q_goal, (n, self.J)), _f64(s_diff, (n,)),
             _f64(t_reach, (n,)), _f64(alpha_i, (n, self.J)), _f64(kp_pos, (n, self.K, 3)), _f64(kp_vel, (n, self.K, 3))]
        robot, human, md = np.zeros((n, self.J, 7)), np.zeros((n, H, 7)), np.full((n, 3), np.inf)
        safe = np.zeros(n, dtype=np.uint8)
        lib().ref_stage_reach_verify(self.h, n, *[_d(x) for x in a], _d(robot), _d(human), int(H), _d(md), _u8(safe))
        return robot, human, md, safe

    def state(self, i):
        out = np.zeros(63)
        lib().ref_get_state(self.h, int(i), _d(out))
        return out

    def human_reach(self, i, t_command, t_brake, max_caps=128):
        """HumanReach::humanReachabilityAnalysis of env i -> capsules [H,7] ordered Pos, Vel, Accel."""
        out = np.zeros((max_caps, 7))
        n = lib().ref_human_reach(self.h, int(i), float(t_command), float(t_brake), _d(out), max_caps)
        return out[:n]

    def plan_safety_shield(self, cases):
        cases = _f64(cases).reshape(-1, 6)
        n = cases.shape[0]
        ok, ph, fin = np.zeros(n, dtype=np.uint8), np.zeros((n, 6)), np.zeros((n, 3))
        lib().ref_plan_safety_shield(self.h, n, _d(cases), _u8(ok), _d(ph), _d(fin))
        return ok, ph, fin


def capsule_dist(c1, c2):
    c1, c2 = _f64(c1).reshape(-1, 7), _f64(c2).reshape(-1, 7)
    out = np.zeros(c1.shape[0])
    lib().ref_capsule_dist(c1.shape[0], _d(c1), _d(c2), _d(out))
    return out


def ltp_plan(ts, q_min, q_max, v_max, a_max, j_max, q_goal, q0, v0, a0, max_len=4096):
    J = len(q0)
    a = [_f64(x) for x in (q_min, q_max, v_max, a_max, j_max, q_goal, q0, v0, a0)]
    traj = np.zeros((max_len, 4, J))
    L = lib().ref_ltp_plan(J, float(ts), *[_d(x) for x in a], _d(traj), max_len)
    return L, traj[: min(L, max_len)]


def poly_roots(coef):
    coef = _f64(coef)
    n = coef.shape[0]
    re, im = np.zeros(n - 1), np.zeros(n - 1)
    s = lib().ref_poly_roots(n, _d(coef), _d(re), _d(im))
    return s, re, im


def ltp_opt_braking(ts, lim, v0, a0):
    """LongTermPlanner::optBraking for one joint; lim = (q_min, q_max, v_max, a_max, j_max) -> ok, q, t_rel[7], dir."""
    lim, t = _f64(lim), np.zeros(7)
    q, d = np.zeros(1), np.zeros(1)
    ok = lib().ref_ltp_opt_braking(float(ts), _d(lim), float(v0), float(a0), _d(q), _d(t), _d(d))
    return bool(ok), q[0], t, d[0]


def ltp_opt_switch_times(ts, lim, q_goal, q0, v0, a0, v_drive):
    lim, t, d, m = _f64(lim), np.zeros(7), np.zeros(1), C.c_int(0)
    ok = lib().ref_ltp_opt_switch_times(float(ts), _d(lim), float(q_goal), float(q0), float(v0), float(a0), float(v_drive),
                                        _d(t), _d(d), C.byref(m))
    return bool(ok), t, d[0], m.value


def ltp_time_scaling(ts, lim, q_goal, q0, v0, a0, direction, t_required):
    lim, t, vd, m = _f64(lim), np.zeros(7), np.zeros(1), C.c_int(0)
    ok = lib().ref_ltp_time_scaling(float(ts), _d(lim), float(q_goal), float(q0), float(v0), float(a0), float(direction),
                                    float(t_required), _d(t), _d(vd), C.byref(m))
    return bool(ok), t, vd[0], m.value


def ltt_interpolate(traj, sample_time, sdd, v_max, a_max, j_max, starting_index=0):
    """LongTermTraj::interpolate on traj [L,4,J] at the rows of sdd [n,4] = (s, ds, dds, ddds) -> [n,4,J]."""
    traj = _f64(traj)
    L, _, J = traj.shape
    sdd = _f64(sdd).reshape(-1, 4)
    out = np.zeros((sdd.shape[0], 4, J))
    lib().ref_ltt_interpolate(J, L, _d(traj), float(sample_time), int(starting_index), sdd.shape[0], _d(sdd),
                              _d(_f64(v_max)), _d(_f64(a_max)), _d(_f64(j_max)), _d(out))
    return out


def ltt_max_window(traj, sample_time, k):
    """LongTermTraj(traj, sample_time, 0, k).max_acceleration_window_ / max_jerk_window_ -> ([L,J], [L,J])."""
    traj = _f64(traj)
    L, _, J = traj.shape
    acc, jerk = np.zeros((L, J)), np.zeros((L, J))
    lib().ref_ltt_max_window(J, L, _d(traj), float(sample_time), int(k), _d(acc), _d(jerk))
    return acc, jerk
