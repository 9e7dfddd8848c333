"""ctypes binding of include/hrg_shield.h (libhrg_shield.so).

The library is the product: there is no CPU fallback. Loading fails loudly when the CUDA extension has not been built
(`python -c "import __graft_entry__ as g; g.build()"`), and `hrg_shield_create` fails with HRG_ENODEVICE when no CUDA
device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .config import ShieldConfig

MAX_JOINTS, MAX_KEYPOINTS, MAX_BODIES, MAX_EXT = 8, 64, 24, 8
OK, EINVAL, ENODEVICE, ECUDA, ENOMEM = 0, -1, -2, -3, -4
SHIELD_OFF, SHIELD_SSM, SHIELD_PFL = 0, 1, 2
ST_REPLANNED, ST_REPLAN_FAILED, ST_NOT_PLANNABLE, ST_JOINT_LIMIT, ST_LTT_TRUNCATED, ST_NEAR_THRESHOLD = 1, 2, 4, 8, 16, 32

PROF_SLOTS = ("k_measure", "k_pre", "k_plan", "k_replan_plan", "_spare", "k_replan_deep", "k_replan_sample",
              "k_pfl_ltt", "k_replan_tail", "k_reach_verify_p0", "k_verify_stage_vel", "k_verify_stage_acc",
              "k_reach_verify_p1", "k_select")  # HRG_PROF_* of include/hrg_shield.h

_D = C.c_double
_I = C.c_int32


class HrgConfig(C.Structure):
    """Mirror of `struct HrgConfig` (include/hrg_shield.h)."""
    _fields_ = [
        ("nb_joints", _I), ("n_keypoints", _I), ("n_bodies", _I), ("n_ext", _I), ("ltt_capacity", _I), ("root_mode", _I),
        ("sample_time", _D), ("max_s_stop", _D), ("v_safe", _D), ("alpha_i_max", _D), ("secure_radius", _D),
        ("meas_err_pos", _D), ("meas_err_vel", _D), ("delay", _D),
        ("q_min", _D * MAX_JOINTS), ("q_max", _D * MAX_JOINTS), ("v_max", _D * MAX_JOINTS), ("a_max", _D * MAX_JOINTS),
        ("j_max", _D * MAX_JOINTS), ("a_max_ltt", _D * MAX_JOINTS), ("j_max_ltt", _D * MAX_JOINTS),
        ("link_T", (_D * 16) * MAX_JOINTS), ("link_caps", (_D * 7) * MAX_JOINTS),
        ("body_prox", _I * MAX_BODIES), ("body_dist", _I * MAX_BODIES), ("body_thickness", _D * MAX_BODIES),
        ("body_vmax", (_D * 2) * MAX_BODIES), ("body_amax", (_D * 2) * MAX_BODIES),
        ("ext_base", _I * MAX_EXT), ("reserved1", _I * MAX_EXT),
        ("ext_thickness", _D * MAX_EXT), ("ext_length", _D * MAX_EXT), ("ext_vmax", _D * MAX_EXT),
    ]


class HrgControllerConfig(C.Structure):
    _fields_ = [("nb_joints", C.c_int32), ("has_position_limits", C.c_int32)] + [
        (n, _D * MAX_JOINTS) for n in ("input_min", "input_max", "output_min", "output_max", "position_min", "position_max",
                                       "kp", "kd", "torque_min", "torque_max")]


class HrgStats(C.Structure):
    _fields_ = [("n_steps", C.c_int64), ("n_unsafe", C.c_int64), ("n_branch", C.c_int64 * 3), ("n_replan", C.c_int64),
                ("n_replan_failed", C.c_int64), ("n_near_threshold", C.c_int64), ("ltt_len_sum", C.c_int64),
                ("min_dist", _D), ("n_ltt_truncated", C.c_int64), ("n_pfl_rejected", C.c_int64)]


STATS_WORDS = C.sizeof(HrgStats) // 8  # the struct as int64 words (min_dist = word 9 as a bit pattern)


def make_cfg(sc: ShieldConfig, ltt_capacity: int = 1024) -> HrgConfig:
    """Flatten the three parsed YAML configs into the constant tables the kernels read."""
    r, t, h = sc.robot, sc.traj, sc.human
    J = r.nb_joints
    if J > MAX_JOINTS:
        raise ValueError(f"nb_joints {J} > {MAX_JOINTS}")
    if h.n_keypoints > MAX_KEYPOINTS or h.n_bodies > MAX_BODIES or h.n_ext > MAX_EXT:
        raise ValueError("human model exceeds the compiled table sizes")
    for name in ("q_min_allowed", "q_max_allowed", "v_max_allowed", "a_max_allowed", "j_max_allowed", "a_max_ltt",
                 "j_max_ltt"):
        if len(getattr(t, name)) < J:
            raise ValueError(f"trajectory config: {name} shorter than nb_joints")
    c = HrgConfig()
    c.nb_joints, c.n_keypoints, c.n_bodies, c.n_ext, c.ltt_capacity = J, h.n_keypoints, h.n_bodies, h.n_ext, int(ltt_capacity)
    c.root_mode = int(os.environ.get("HRG_ROOT_MODE", DEFAULT_ROOT_MODE))
    c.sample_time, c.max_s_stop, c.v_safe, c.alpha_i_max = sc.sample_time, t.max_s_stop, t.v_safe, t.alpha_i_max
    c.secure_radius = r.secure_radius
    c.meas_err_pos, c.meas_err_vel, c.delay = h.measurement_error_pos, h.measurement_error_vel, h.delay
    for j in range(J):
        c.q_min[j], c.q_max[j] = t.q_min_allowed[j], t.q_max_allowed[j]
        c.v_max[j], c.a_max[j], c.j_max[j] = t.v_max_allowed[j], t.a_max_allowed[j], t.j_max_allowed[j]
        c.a_max_ltt[j], c.j_max_ltt[j] = t.a_max_ltt[j], t.j_max_ltt[j]
        for k in range(16):
            c.link_T[j][k] = r.transformation_matrices[j, k]
        for k in range(7):
            c.link_caps[j][k] = r.enclosures[j, k]
    for b in range(h.n_bodies):
        p, d = int(h.body_prox[b]), int(h.body_dist[b])
        c.body_prox[b], c.body_dist[b], c.body_thickness[b] = p, d, h.body_thickness[b]
        c.body_vmax[b][0], c.body_vmax[b][1] = h.joint_v_max[p], h.joint_v_max[d]
        c.body_amax[b][0], c.body_amax[b][1] = h.joint_a_max[p], h.joint_a_max[d]
    for x in range(h.n_ext):
        c.ext_base[x] = int(h.ext_base[x])
        c.ext_thickness[x], c.ext_length[x], c.ext_vmax[x] = h.ext_thickness[x], h.ext_length[x], h.ext_v_max[x]
    return c


_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HRG_LIB_PATH") or os.path.join(_HERE, "libhrg_shield.so")  # HRG_LIB_PATH: profiling builds (profiles/build_trace.sh)
_lib = None

_vp, _dp, _u8p, _i32p, _u32p = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p  # device pointers as ints

# every symbol include/hrg_shield.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "hrg_shield_create": (C.c_int, [C.POINTER(HrgConfig), C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.POINTER(C.c_void_p)]),
    "hrg_shield_destroy": (C.c_int, [C.c_void_p]),
    "hrg_ltt_capacity_bound": (C.c_int, [C.POINTER(HrgConfig)]),
    "hrg_shield_enable_pfl": (C.c_int, [C.c_void_p]),
    "hrg_shield_reset": (C.c_int, [C.c_void_p, _u8p, _dp, _dp, _dp, _i32p, _vp]),
    "hrg_shield_human_measurement": (C.c_int, [C.c_void_p, _u8p, _dp, _dp, _vp]),
    "hrg_shield_new_goal": (C.c_int, [C.c_void_p, _u8p, _dp, _dp, _vp]),
    "hrg_shield_set_ltt": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_int, C.c_void_p,
                                     C.c_void_p]),
    "hrg_shield_set_ltt_batch": (C.c_int, [C.c_void_p, C.c_int, _i32p, _dp, C.c_int, _i32p, C.c_double, _i32p, _dp, _i32p,
                                           _vp]),
    "hrg_shield_store_capsules": (C.c_int, [C.c_void_p, C.c_int]),
    "hrg_shield_step": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _dp, _dp, _u8p, _u8p, _u32p, _vp]),
    "hrg_shield_tick_device": (C.c_int, [C.c_void_p, _dp, _dp, _u8p, _dp, _dp, _dp, _dp, _dp, _u8p, _u8p, _u32p, _vp]),
    "hrg_shield_tick_host": (C.c_int, [C.c_void_p] + [C.c_void_p] * 8 + [_vp]),
    "hrg_shield_upload_inputs": (C.c_int, [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [_vp]),
    "hrg_shield_upload_inputs_sparse": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                                  C.c_void_p, _vp]),
    "hrg_shield_tick_staged": (C.c_int, [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [_vp]),
    "hrg_shield_get_capsules": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _vp]),
    "hrg_capsule_markers": (C.c_int, [C.c_int, _dp, _dp, _dp, _dp, _vp]),
    "hrg_shield_ltt_max_window": (C.c_int, [C.c_void_p, _i32p, C.c_int, _dp, _dp, _vp]),
    "hrg_shield_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, _vp]),
    "hrg_shield_profile": (C.c_int, [C.c_void_p, C.c_int]),
    "hrg_shield_profile_read": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "hrg_shield_profile_read_kernels": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "hrg_shield_num_envs": (C.c_int, [C.c_void_p]),
    "hrg_shield_num_human_capsules": (C.c_int, [C.c_void_p]),
    "hrg_shield_launch_count": (C.c_int64, [C.c_void_p]),
    "hrg_last_error": (C.c_char_p, []),
    "hrg_controller_set_goal": (C.c_int, [C.c_void_p, C.POINTER(HrgControllerConfig), _u8p, _dp, _dp, _dp, _dp, _vp]),
    "hrg_controller_torques": (C.c_int, [C.c_void_p, C.POINTER(HrgControllerConfig)] + [_dp] * 8 + [_vp]),
    "hrg_human_create": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_double, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "hrg_human_destroy": (C.c_int, [C.c_void_p]),
    "hrg_human_reset": (C.c_int, [C.c_void_p, _u8p, _dp, _i32p, _vp]),
    "hrg_human_keypoints": (C.c_int, [C.c_void_p, C.c_void_p, _dp, _i32p, _i32p, _vp]),
    "hrg_shield_tick_replay_host": (C.c_int, [C.c_void_p, C.c_void_p] + [C.c_void_p] * 8 + [_vp]),
    "hrg_shield_tick_replay_piped": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 8 + [_vp, _vp]),
    "hrg_human_launch_count": (C.c_int64, [C.c_void_p]),
    "hrg_shield_state_bytes": (C.c_int64, [C.c_void_p]),
    "hrg_shield_state_save": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "hrg_shield_state_load": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "hrg_microbench_fp64": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "hrg_microbench_fp64_latency": (C.c_int, [C.c_int, C.c_void_p]),
    "hrg_stage_capsule_dist": (C.c_int, [C.c_int, _dp, _dp, _dp, _vp]),
    "hrg_stage_plan_safety_shield": (C.c_int, [C.c_double, C.c_int, _dp, _u8p, _dp, _dp, _vp]),
    "hrg_stage_ltp_plan": (C.c_int, [C.POINTER(HrgConfig), C.c_int, _dp, _dp, _dp, _dp, _dp, _i32p, C.c_int, _vp]),
    "hrg_stage_reach_verify": (C.c_int, [C.POINTER(HrgConfig), C.c_int] + [_dp] * 8 + [_dp, _dp, _dp, _u8p, _vp]),
    "hrg_stage_reach_verify_noalloc": (C.c_int, [C.POINTER(HrgConfig), C.c_int] + [_dp] * 8 + [_dp, _dp, _dp, _u8p, _vp]),
}


# planner root finder (HrgConfig.root_mode): 1 = derivative-chain finder with the exact QR iteration as the arbiter of close
# calls, 0 = the exact QR iteration for every polynomial. HRG_ROOT_MODE overrides.
DEFAULT_ROOT_MODE = 1


class ShieldLibraryError(RuntimeError):
    pass


def load_library():
    """dlopen libhrg_shield.so and type every entry point. Raises if the CUDA extension is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ShieldLibraryError(
            f"{LIB_PATH} not found: the CUDA extension is not built (run __graft_entry__.build()). "
            "There is no CPU fallback for the shield hot path.")
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(L, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def ltt_capacity_bound(sc: ShieldConfig) -> int:
    """Samples per LTT slot that can never be exceeded by a plan inside the joint limits (hrg_ltt_capacity_bound): with at
    least this `ltt_capacity` the library never rejects a plan the reference would accept."""
    cfg = make_cfg(sc, 2)
    n = load_library().hrg_ltt_capacity_bound(C.byref(cfg))
    if n < 0:
        check(n, "hrg_ltt_capacity_bound")
    return int(n)


def microbench_fp64(device: int = 0, iters: int = 4096) -> dict:
    """Measured register-resident FP64 peaks of `device`: {"dfma"|"dmul_dadd"|"dadd": {"gflops", "ginst", "ms"}}."""
    out = {}
    for mode, name in enumerate(("dfma", "dmul_dadd", "dadd")):
        g, i, ms = C.c_double(0), C.c_double(0), C.c_double(0)
        check(load_library().hrg_microbench_fp64(int(device), mode, int(iters), C.addressof(g), C.addressof(i),
                                                 C.addressof(ms)), "hrg_microbench_fp64")
        out[name] = {"gflops": g.value, "ginst": i.value, "ms": ms.value}
    return out


def microbench_fp64_latency(device: int = 0) -> dict:
    """Dependent-chain latency (SM cycles per operation, one warp alone) of the FP64 operations the planner is made of."""
    out = (C.c_double * 6)()
    check(load_library().hrg_microbench_fp64_latency(int(device), C.addressof(out)), "hrg_microbench_fp64_latency")
    return dict(zip(("dadd", "dmul", "dfma", "ddiv_plus_dadd", "dsqrt_plus_dadd", "dadd_cmp_select"), [round(x, 2) for x in out]))


def check(rc: int, what: str = ""):
    if rc != OK:
        msg = load_library().hrg_last_error()
        raise ShieldLibraryError(f"{what or 'hrg call'} failed with code {rc}: {msg.decode() if msg else ''}")
