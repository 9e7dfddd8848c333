"""ctypes view of the TEST-ONLY host emulation (tests/host_emul/emul.cc). Never imported by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

from human_robot_gym_b200.cabi import HrgConfig

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libhrg_emul.so")
_lib = None
_p = C.c_void_p


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(_HERE, "emul.cc")
        csrc = os.path.join(_HERE, "..", "..", "human_robot_gym_b200", "csrc")
        newest = max([os.path.getmtime(src)] + [os.path.getmtime(os.path.join(csrc, f)) for f in os.listdir(csrc)])
        if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < newest:
            subprocess.check_call([os.path.join(_HERE, "build.sh")])
        L = C.CDLL(_LIB)
        L.emul_create.restype = C.c_void_p
        L.emul_create.argtypes = [C.POINTER(HrgConfig), C.c_int, _p, _p, _p]
        L.emul_destroy.argtypes = [_p]
        L.emul_reset.argtypes = [_p] * 5
        L.emul_human_measurement.argtypes = [_p] * 4
        L.emul_new_goal.argtypes = [_p] * 3
        L.emul_step.argtypes = [_p] * 9
        L.emul_get_capsules.argtypes = [_p] * 4
        L.emul_get_state.argtypes = [_p, C.c_int, _p]
        L.emul_capsule_dist.argtypes = [C.c_int, _p, _p, _p]
        L.emul_plan_safety_shield.argtypes = [C.c_double, C.c_int, _p, _p, _p, _p]
        L.emul_ltp_plan.restype = C.c_int
        L.emul_ltp_plan.argtypes = [C.POINTER(HrgConfig), _p, _p, _p, _p, _p, C.c_int]
        D = C.c_double
        L.emul_ltp_opt_braking.argtypes = [D, _p, D, D, _p, _p, _p]
        L.emul_ltp_opt_switch_times.argtypes = [D, _p, D, D, D, D, D, _p, _p, _p]
        L.emul_ltp_time_scaling.argtypes = [D, _p, D, D, D, D, D, D, _p, _p, _p]
        L.emul_ltt_interpolate.argtypes = [C.c_int, C.c_int, _p, D, C.c_int, C.c_int, _p, _p, _p, _p]
        L.emul_human_capsules.argtypes = [C.POINTER(HrgConfig), _p, _p, D, _p]
        L.emul_capsule_markers.argtypes = [C.c_int, _p, _p, _p, _p]
        L.emul_human_keypoints.restype = C.c_int
        L.emul_human_keypoints.argtypes = [C.c_int, _p, _p, C.c_int, _p, _p, _p, _p]
        L.emul_poly_root.restype = C.c_double
        L.emul_poly_root.argtypes = [_p, C.c_int]
        L.emul_poly_root_reg.restype = C.c_double
        L.emul_poly_root_reg.argtypes = [_p, C.c_int]
        L.emul_poly_root_batch.argtypes = [C.c_int, C.c_int, _p, _p, _p]
        L.emul_poly_root_fast_batch.argtypes = [C.c_longlong, C.c_int, _p, _p, _p, _p]
        L.emul_set_root_mode.argtypes = [C.c_int]
        _lib = L
    return _lib


def _a(x, dt=np.float64):
    return np.ascontiguousarray(x, dtype=dt)


def _ptr(a):
    return None if a is None else a.ctypes.data


class EmulBatch:
    def __init__(self, cfg: HrgConfig, n, base_pose, init_q, shield_type=1):
        self.cfg, self.n, self.J, self.K = cfg, n, cfg.nb_joints, cfg.n_keypoints
        self.H = cfg.n_ext + 2 * cfg.n_bodies
        bp = _a(np.broadcast_to(_a(base_pose), (n, 6)))
        q0 = _a(np.broadcast_to(_a(init_q), (n, self.J)))
        ty = _a(np.broadcast_to(np.asarray(shield_type), (n,)), np.int32)
        self.h = lib().emul_create(C.byref(cfg), n, _ptr(bp), _ptr(q0), _ptr(ty))

    def __del__(self):
        if getattr(self, "h", None):
            lib().emul_destroy(self.h)
            self.h = None

    def reset(self, base_pose, init_q, shield_type=1, mask=None):
        bp = _a(np.broadcast_to(_a(base_pose), (self.n, 6)))
        q0 = _a(np.broadcast_to(_a(init_q), (self.n, self.J)))
        ty = _a(np.broadcast_to(np.asarray(shield_type), (self.n,)), np.int32)
        m = None if mask is None else _a(mask, np.uint8)
        lib().emul_reset(self.h, _ptr(m), _ptr(bp), _ptr(q0), _ptr(ty))

    def human_measurement(self, kp, t, mask=None):
        kp = _a(kp).reshape(self.n, self.K, 3)
        tt = _a(np.broadcast_to(t, (self.n,)))
        m = None if mask is None else _a(mask, np.uint8)
        lib().emul_human_measurement(self.h, _ptr(m), _ptr(kp), _ptr(tt))

    def new_goal(self, goal_q, mask=None):
        g = _a(goal_q).reshape(self.n, self.J)
        m = None if mask is None else _a(mask, np.uint8)
        lib().emul_new_goal(self.h, _ptr(m), _ptr(g))

    def step(self, t):
        tt = _a(np.broadcast_to(t, (self.n,)))
        o = {k: np.zeros((self.n, self.J)) for k in ("q", "dq", "ddq", "dddq")}
        o["is_safe"] = np.zeros(self.n, np.uint8)
        o["branch"] = np.zeros(self.n, np.uint8)
        o["status"] = np.zeros(self.n, np.uint32)
        lib().emul_step(self.h, _ptr(tt), _ptr(o["q"]), _ptr(o["dq"]), _ptr(o["ddq"]), _ptr(o["dddq"]),
                        _ptr(o["is_safe"]), _ptr(o["branch"]), _ptr(o["status"]))
        return o

    def capsules(self):
        r, h, md = np.zeros((self.n, self.J, 7)), np.zeros((self.n, self.H, 7)), np.zeros((self.n, 3))
        lib().emul_get_capsules(self.h, _ptr(r), _ptr(h), _ptr(md))
        return r, h, md

    def state(self, e):
        o = np.zeros(52)
        lib().emul_get_state(self.h, e, _ptr(o))
        return o


def capsule_dist(c1, c2):
    c1, c2 = _a(c1).reshape(-1, 7), _a(c2).reshape(-1, 7)
    d = np.zeros(c1.shape[0])
    lib().emul_capsule_dist(c1.shape[0], _ptr(c1), _ptr(c2), _ptr(d))
    return d


def plan_safety_shield(ts, cases):
    cases = _a(cases).reshape(-1, 6)
    n = cases.shape[0]
    ok, ph, fin = np.zeros(n, np.uint8), np.zeros((n, 6)), np.zeros((n, 3))
    lib().emul_plan_safety_shield(ts, n, _ptr(cases), _ptr(ok), _ptr(ph), _ptr(fin))
    return ok, ph, fin


def ltp_plan(cfg, goal, q0, v0, a0, cap=4096):
    J = cfg.nb_joints
    traj = np.zeros((cap, 4, J))
    L = lib().emul_ltp_plan(C.byref(cfg), _ptr(_a(goal)), _ptr(_a(q0)), _ptr(_a(v0)), _ptr(_a(a0)), _ptr(traj), cap)
    return L, traj[: max(L, 0)]


def poly_root(coef):
    c = _a(coef)
    return lib().emul_poly_root(_ptr(c), c.shape[0])


def poly_root_batch(coefs):
    """(matrix-in-memory solver, register-resident solver) smallest positive real roots of coefs [n, ncoef]."""
    c = _a(coefs)
    n, ncoef = c.shape
    a, b = np.zeros(n), np.zeros(n)
    lib().emul_poly_root_batch(n, ncoef, _ptr(c), _ptr(a), _ptr(b))
    return a, b


def poly_root_fast_batch(coefs):
    """Derivative-chain root finder (csrc/ltp_roots_fast.cuh) next to the exact QR restatement for coefs [n, ncoef], ncoef
    5..7: (fast root, margin in units of the QR iteration's noise, exact root). margin == -1 marks a polynomial on which the
    mixed-degree form (k_replan_deep) and the per-degree form returned different bits."""
    c = _a(coefs)
    n, ncoef = c.shape
    f, m, r = np.zeros(n), np.zeros(n), np.zeros(n)
    lib().emul_poly_root_fast_batch(n, ncoef, _ptr(c), _ptr(f), _ptr(m), _ptr(r))
    return f, m, r


def set_root_mode(mode):
    """Root finder of the single-function entry points below (LtpLimits::root_mode): 0 exact QR, 1 fast + rest-start rule."""
    lib().emul_set_root_mode(int(mode))


def ltp_opt_braking(ts, lim, v0, a0):
    lim, t, q, d = _a(lim), np.zeros(7), np.zeros(1), np.zeros(1)
    lib().emul_ltp_opt_braking(ts, _ptr(lim), v0, a0, _ptr(q), _ptr(t), _ptr(d))
    return True, q[0], t, d[0]


def ltp_opt_switch_times(ts, lim, q_goal, q0, v0, a0, v_drive):
    lim, t, d, m = _a(lim), np.zeros(7), np.zeros(1), np.zeros(1, np.int32)
    ok = lib().emul_ltp_opt_switch_times(ts, _ptr(lim), q_goal, q0, v0, a0, v_drive, _ptr(t), _ptr(d), _ptr(m))
    return bool(ok), t, d[0], int(m[0])


def ltp_time_scaling(ts, lim, q_goal, q0, v0, a0, direction, t_required):
    lim, t, vd, m = _a(lim), np.zeros(7), np.zeros(1), np.zeros(1, np.int32)
    ok = lib().emul_ltp_time_scaling(ts, _ptr(lim), q_goal, q0, v0, a0, direction, t_required, _ptr(t), _ptr(vd), _ptr(m))
    return bool(ok), t, vd[0], int(m[0])


def ltt_interpolate(traj, sample_time, sdd, v_max, a_max, j_max=None, starting_index=0):
    traj = _a(traj)
    L, _, J = traj.shape
    sdd = _a(sdd).reshape(-1, 4)
    out = np.zeros((sdd.shape[0], 4, J))
    lib().emul_ltt_interpolate(J, L, _ptr(traj), sample_time, starting_index, sdd.shape[0], _ptr(sdd), _ptr(_a(v_max)),
                               _ptr(_a(a_max)), _ptr(out))
    return out


def human_capsules(cfg, kp, kv, t_reach):
    H = cfg.n_ext + 2 * cfg.n_bodies
    out = np.zeros((H, 7))
    lib().emul_human_capsules(C.byref(cfg), _ptr(_a(kp)), _ptr(_a(kv)), t_reach, _ptr(out))
    return out


def human_keypoints(tree, rows, info, place):
    """[n,K,3] site positions from frame rows [n,7+3K], clip info [n,7] and placements [n,7] (kernels_human.cuh on the host)."""
    K = len(tree["names"])
    rows, info, place = _a(rows), _a(info), _a(place)
    n = rows.shape[0]
    out = np.zeros((n, K, 3))
    parent, anchor = _a(tree["parent"], np.int32), _a(tree["anchor"])
    rc = lib().emul_human_keypoints(K, _ptr(parent), _ptr(anchor), n, _ptr(rows), _ptr(info), _ptr(place), _ptr(out))
    if rc:
        raise ValueError(f"human_build_tree failed with {rc}")
    return out


def capsule_markers(caps):
    """PlotCapsule pos / size / mat for caps [n,7] (geom_math.cuh capsule_marker on the host)."""
    caps = _a(caps)
    n = caps.shape[0]
    pos, size, mat = np.zeros((n, 3)), np.zeros((n, 3)), np.zeros((n, 3, 3))
    lib().emul_capsule_markers(n, _ptr(caps), _ptr(pos), _ptr(size), _ptr(mat))
    return pos, size, mat
