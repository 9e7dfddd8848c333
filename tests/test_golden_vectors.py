"""CPU: pins (a) the oracle (oracle/_ref = the reference C++ built against the Eigen shim) and (b) the device arithmetic
(human_robot_gym_b200/csrc/*.cuh compiled for the host by tests/host_emul) against the known-answer tests the reference
ships for this path. Every vector below is quoted from the cited reference test; nothing here needs a GPU.

SS = human_robot_gym/controllers/failsafe_controller/sara-shield/safety_shield, LTP = SS/external/LongTermPlanner.
"""
import os

import numpy as np
import pytest

from oracle import ref
from tests import parity_util as pu
from tests.host_emul import emul

needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference at build time)")

IMPLS = [pytest.param(ref, marks=needs_ref, id="oracle_ref"), pytest.param(emul, id="device_math_on_host")]


# ---- LTP/tests/src/roots_tests.cc:9-32 ------------------------------------------------------------------------------
@needs_ref
def test_roots_degree6_known_answer():
    coef = [144, -1008, 2448, 3024.01920000000, -15768.1344000000, 0, 22752.4032012800]
    smallest, re, im = ref.poly_roots(coef)
    want = [(-1.67276, 0.0), (-1.35687, 0.0), (2.00001, 0.0), (2.09261, 0.0), (2.9685, 2.79663), (2.9685, -2.79663)]
    # the reference test runs roots<float> and also pins the ORDER of the eigenvalues; in double the two nearly equal real
    # roots 2.00001 / 2.09261 are resolved to 1e-4
    got = sorted(zip(re, im), key=lambda z: (round(z[0], 3), -z[1]))
    for (gr, gi), (wr, wi) in zip(got, sorted(want, key=lambda z: (round(z[0], 3), -z[1]))):
        assert abs(gr - wr) < 2e-4 and abs(gi - wi) < 2e-4
    assert sum(1 for x in im if x == 0.0) == 4          # imag() == 0 exactly for the real roots (roots.h:46)
    assert abs(smallest - 2.00001) < 2e-4               # getSmallestPositiveNonComplexRoot
    assert abs(emul.poly_root(coef) - smallest) < 1e-9  # device root finder, same polynomial


# ---- LTP/tests/src/long_term_planner_tests.cc:12-50 -----------------------------------------------------------------
@pytest.mark.parametrize("impl", IMPLS)
def test_opt_braking(impl):
    v_0 = [0, -1.875, -1.875, -0.875, -0.875, 0.5]
    a_0 = [0, 1, -1, 1, -1, -2]
    a_max = [2, 2, 2, 4, 4, 4]
    j_max = [4, 4, 4, 4, 4, 2]
    q_goal = [0, -1.0104, -1.9896, -0.2604, -0.7396, -0.4167]
    t_rel = [[0, 0, 0], [0.25, 0.5, 0.5], [0.75, 0.5, 0.5], [0.25, 0, 0.5], [0.75, 0, 0.5], [1.5, 0, 0.5]]
    for i in range(6):
        lim = [-3.1, 3.1, 10, a_max[i], j_max[i]]
        for sgn in ([1] if i == 0 else [1, -1]):
            ok, q, t, _ = impl.ltp_opt_braking(0.001, lim, sgn * v_0[i], sgn * a_0[i])
            assert ok
            assert np.abs(t[:3] - t_rel[i]).max() < 0.01
            assert abs(q - sgn * q_goal[i]) < 0.01


# ---- LTP/tests/src/long_term_planner_tests.cc:52-110 ----------------------------------------------------------------
OST_T = [[0.0] * 7,
         [0.25, 0.5, 1.0, 1.5, 2.0, 2.5, 3.0],
         [0.5, 0.5, 0.75, 1.25, 1.75, 2.25, 2.75],
         [0.25, 0.75, 1.25, 1.75, 2.25, 2.25, 2.75],
         [0.5, 0.5, 0.75, 1.25, 1.75, 1.75, 2.25],
         [0.25, 0.5, 1.0, 1.0, 1.5, 2.0, 2.5],
         [0.5, 0.5, 0.5, 0.5, 1.25, 1.75, 2.25],
         [0.25, 0.75, 1.25, 1.25, 1.75, 1.75, 2.25],
         [0.5, 0.5, 0.5, 0.5, 1.25, 1.25, 1.75]]


@pytest.mark.parametrize("impl", IMPLS)
def test_opt_switch_times(impl):
    v_max = [2, 2, 2, 1, 1, 8, 8, 8, 8]
    q_goal = [-1.0, 2.927, 2.8854, 0.2396, 0.6354, 1.927, 1.8854, -0.2604, 0.1354]
    v_0 = [0.0, 0.625, 1.875, -0.875, 0.875, 0.625, 1.875, -0.875, 0.875]
    a_0 = [0.0, 1.0, -1.0, 1.0, -1.0, 1.0, -1.0, 1.0, -1.0]
    for i in range(9):
        lim = [-3.1, 3.1, v_max[i], 2, 4]
        for sgn in ([1] if i == 0 else [1, -1]):
            ok, t, _, _ = impl.ltp_opt_switch_times(0.001, lim, sgn * q_goal[i], sgn * -1.0, sgn * v_0[i], sgn * a_0[i],
                                                    v_max[i])
            assert ok
            assert np.abs(t - OST_T[i]).max() < 1e-3, (i, sgn, t)


# ---- LTP/tests/src/long_term_planner_tests.cc:196-255 (this one goes through the eigenvalue solver) -------------------
@pytest.mark.parametrize("impl", IMPLS)
def test_time_scaling(impl):
    a_max = [2] * 9 + [4, 4, 4]
    j_max = [4] * 11 + [2]
    q_goal = [-1.0, 2.927, 2.8854, 0.2396, 0.6354, -7.0104, -8.9896, -3.896, -7.9433, -5.1746, -6.6538, -8.4167]
    v_0 = [0.0, 0.625, 1.875, -0.875, 0.875, -3.875, -3.875, -1.875, -1.875, -2.875, -2.875, -1.5]
    a_0 = [0.0, 1, -1, 1, -1, 1, -1, 1, -2, 1, -1, -2]
    dirs = [1.0, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1]
    t = [[0, 0, 0, 0, 0, 0, 0],
         [0.25, 0.5, 1, 1.5, 2, 2.5, 3],
         [0.5, 0.5, 0.75, 1.25, 1.75, 2.25, 2.75],
         [0.25, 0.75, 1.25, 1.75, 2.25, 2.25, 2.75],
         [0.5, 0.5, 0.75, 1.25, 1.75, 1.75, 2.25],
         [0.25, 0.75, 1.25, 1.75, 2.25, 2.75, 3.25],
         [0.75, 1.25, 1.75, 2.25, 2.75, 3.25, 3.75],
         [0.25, 0.5, 1, 4.2928, 4.6464, 4.6464, 5],
         [0.75, 0.75, 1, 2.4372, 2.9372, 3.4997, 3.9997],
         [0.25, 0.25, 0.75, 1.25, 1.9571, 1.9571, 2.6642],
         [0.75, 0.75, 1.25, 1.75, 2.4571, 2.4571, 3.1642],
         [1.5, 1.5, 2, 2.5, 3.5, 3.5, 4.5]]
    t_required = [0, 3, 2.75, 2.75, 2.25, 3.25, 3.75, 5, 3.9997, 2.6642, 3.1642, 4.5]
    for i in range(12):
        lim = [-3.1, 3.1, 4, a_max[i], j_max[i]]
        for sgn in ([1] if i == 0 else [1, -1]):
            ok, tt, _, _ = impl.ltp_time_scaling(0.001, lim, sgn * q_goal[i], sgn * -1.0, sgn * v_0[i], sgn * a_0[i],
                                                 sgn * dirs[i], t_required[i])
            assert ok, (i, sgn)
            assert np.abs(tt[:3] - t[i][:3]).max() < 0.1, (i, sgn, tt)   # the reference checks the first three times
            assert abs(tt[6] - t_required[i]) < 0.1


@needs_ref
def test_device_planner_equals_oracle_on_the_reference_grid():
    """GridTimeScalingTest-style sweep (LTP tests :330-407), coarser: both sides must agree to 1e-9 on every case."""
    lim = [-6, 7, 1.0, 2.0, 15.0]
    n = 0
    for q_goal in np.arange(-2.0, 3.01, 0.5):
        for v_0 in (-0.8, -0.3, 1e-6, 0.4, 0.9):
            for a_0 in (-1.5, -0.5, 0.0, 0.7):
                if abs(v_0 + 0.5 * a_0 * abs(a_0) / lim[4]) > lim[2]:
                    continue
                ok_r, t_r, d_r, m_r = ref.ltp_opt_switch_times(0.004, lim, q_goal, 0.5, v_0, a_0, lim[2])
                ok_e, t_e, d_e, m_e = emul.ltp_opt_switch_times(0.004, lim, q_goal, 0.5, v_0, a_0, lim[2])
                assert ok_r == ok_e and m_r == m_e and d_r == d_e
                assert np.abs(t_r - t_e).max() < 1e-9
                if not ok_r or t_r[6] < 0.02:
                    continue
                for inc in (0.05, 0.2, 1.0):
                    a = ref.ltp_time_scaling(0.004, lim, q_goal, 0.5, v_0, a_0, d_r, t_r[6] + inc)
                    b = emul.ltp_time_scaling(0.004, lim, q_goal, 0.5, v_0, a_0, d_r, t_r[6] + inc)
                    assert a[0] == b[0] and a[3] == b[3]
                    assert np.abs(a[1] - b[1]).max() < 1e-9 and (a[2] == b[2] or abs(a[2] - b[2]) < 1e-9)
                    n += 1
    assert n > 300


# ---- SS/tests/src/verify_iso_tests.cc:17-47 -------------------------------------------------------------------------
@pytest.mark.parametrize("impl", IMPLS)
def test_capsule_collision_cases(impl):
    def cap(p1, p2, r):
        return list(p1) + list(p2) + [r]
    c1, c2, c3 = cap((0, 0, 0), (0, 0, 1), 1.0), cap((1, 0, 0), (1, 0, 1), 1.0), cap((2.01, 0, 0), (2.01, 0, 1), 1.0)
    d = impl.capsule_dist([c1, c1], [c2, c3])
    assert d[0] < 0.0 and not (d[1] < 0.0)                       # CapsuleCollisionTest
    assert abs(d[0] - (-1.0)) < 1e-12 and abs(d[1] - 0.01) < 1e-12
    r_caps = [cap((0, 0, 0), (0, 0, 1), 0.1), cap((0, 0, 1), (0, 0, 2), 0.1)]
    h1 = [cap((1, 0, 0), (1, 0, 1), 0.5), cap((1, 0, 1), (1, 0, 2), 0.5)]
    h2 = [cap((1, 0, 0), (0, 0, 1), 0.5), cap((1, 0, 1), (0, 0, 2), 0.5)]

    def collides(hs):
        pairs = [(r, h) for r in r_caps for h in hs]
        return bool((impl.capsule_dist([p[0] for p in pairs], [p[1] for p in pairs]) < 0.0).any())
    assert not collides(h1)                                      # RobotHumanCollisionTest
    assert collides(h2)
    assert (not collides(h1)) or (not collides(h2))              # VerifyHumanReachTest: safe iff ANY model is free


# ---- SS/tests/src/long_term_traj_tests.cc:196-228 -------------------------------------------------------------------
@pytest.mark.parametrize("impl", IMPLS)
def test_ltt_interpolate_known_answers(impl):
    traj = np.zeros((2, 4, 2))
    traj[0, :, 0] = [0.0, 1.0, 1.0, -1.0]
    traj[1, :, 0] = [0.104833333333333, 1.095, 0.9, 0.0]
    out = impl.ltt_interpolate(traj, 0.1, [[0.01, 1.0, 0.0, 0.0], [0.01, 0.5, 0.0, 0.0], [0.01, 0.5, 1.0, 0.2]],
                               [10.0, 10.0], [10.0, 10.0], [100.0, 100.0])[:, :, 0]
    want = [[0.010049833333333, 1.00995, 0.99, -1.0], [0.010049833333333, 0.504975, 0.2475, -0.1250],
            [0.010049833333333, 0.504975, 1.25745, 1.56199]]
    assert np.abs(out - np.array(want)).max() < 1e-5


# ---- SS/tests/src/human_reach_tests.cc:69-205 ------------------------------------------------------------------------
def _human_cfg(mocap):
    from human_robot_gym_b200.cabi import make_cfg
    from human_robot_gym_b200.config import ShieldConfig
    files = pu.default_files("schunk", mocap)
    return files, make_cfg(ShieldConfig.from_files(0.004, *files), 16)


def _human_caps(impl, mocap, frames, t_command, t_brake):
    """frames: [(keypoints [K,3], t_meas)], measured in order; returns capsules ordered Pos, Vel, Accel."""
    files, cfg = _human_cfg(mocap)
    if impl is ref:
        rb = ref.RefBatch(1, 0.004, *files, np.zeros(6), np.zeros(6))
        for kp, t in frames:
            rb.human_measurement(np.asarray(kp, float)[None], t)
        return rb.human_reach(0, t_command, t_brake)
    pos = np.asarray(frames[-1][0], float)
    vel = np.zeros_like(pos)
    if len(frames) > 1:                                           # HumanReach::measurement (human_reach.cc:52-74)
        vel = (pos - np.asarray(frames[-2][0], float)) * (1 / (frames[-1][1] - frames[-2][1]))
    return emul.human_capsules(cfg, pos.ravel(), vel.ravel(), t_command - frames[-1][1] + t_brake)


@pytest.mark.parametrize("impl", IMPLS)
def test_human_reach_single_joint(impl):
    cfgs = pu.configs()
    m = cfgs["human_reach_test_single_joint"]
    caps = _human_caps(impl, "human_reach_test_single_joint", [([[1, 2, 3]], 1.0)], 1.01, 0.1)
    n_ext = len(m.get("extremities") or [])
    vel = caps[n_ext]
    assert np.abs(vel[:6] - [1, 2, 3, 1, 2, 3]).max() == 0.0
    assert abs(vel[6] - 0.16) < 1e-12                             # HumanReachAnalysisVelTest
    caps = _human_caps(impl, "human_reach_test_single_joint", [([[1, 2, 3]], 1.0), ([[2, 3, 3]], 2.0)], 2.01, 0.1)
    t = (2.01 - 2.0) + 0.1 + m["delay"]
    r = 0.1 / 2.0 + m["measurement_error_pos"] + m["measurement_error_vel"] * t + 0.5 * 10 * t ** 2
    nxt = np.array([2, 3, 3]) + (np.array([2, 3, 3]) - np.array([1, 2, 3])) * t
    acc = caps[n_ext + len(m["bodies"])]
    assert np.abs(acc[:3] - nxt).max() < 1e-12 and np.abs(acc[3:6] - nxt).max() < 1e-12
    assert abs(acc[6] - r) < 1e-12                                # HumanReachAnalysisAccTest


@pytest.mark.parametrize("impl", IMPLS)
def test_human_reach_with_measurement_error(impl):
    m = pu.configs()["human_reach_test_single_joint_error"]
    n_ext = len(m.get("extremities") or [])
    caps = _human_caps(impl, "human_reach_test_single_joint_error", [([[1, 2, 3]], 1.0)], 1.01, 0.1)
    t = 0.01 + 0.1 + m["delay"]
    assert abs(caps[n_ext][6] - (0.05 + m["measurement_error_pos"] + t * 1.0)) < 1e-12   # ...VelTestError
    caps = _human_caps(impl, "human_reach_test_single_joint_error", [([[1, 2, 3]], 1.0), ([[2, 3, 3]], 2.0)], 2.01, 0.1)
    r = 0.05 + m["measurement_error_pos"] + m["measurement_error_vel"] * t + 0.5 * 10 * t ** 2
    assert abs(caps[n_ext + len(m["bodies"])][6] - r) < 1e-12                            # ...AccTestError


@pytest.mark.parametrize("impl", IMPLS)
def test_human_reach_arm_pos(impl):
    m = pu.configs()["human_reach_test_arm_pos"]
    caps = _human_caps(impl, "human_reach_test_arm_pos", [([[0, 0, 0], [0.2, 0.2, 0.2], [0.4, 0.4, 0.4]], 1.0)], 1.01, 0.1)
    t = 0.01 + 0.1 + m["delay"]
    assert np.abs(caps[0][:6]).max() == 0.0
    assert abs(caps[0][6] - (0.725 + 0.0 + m["measurement_error_pos"] + 2 * t)) < 1e-12  # HumanReachAnalysisPos


# ---- SS/tests/bindings/test_safety_shield.py:151-290: golden robot capsules after one step ----------------------------
GOLDEN_ROBOT_CAPS = {
    "modrob1": (0.01, [
        [4.419824882251160e-04, -0.046901920258750, 0.316102435183982, 5.633175125972268e-04, -0.039612679720870, 0.309257564827811, 0.128755148876288],
        [-0.001475548520496, 0.069582202005415, 0.376756288468218, 0.001008748287424, 0.041671798076183, 0.695527111508655, 0.135477093612307],
        [0.022974546488243, -0.098358202251421, 0.769133504587120, 0.090252153491381, -0.112369997514008, 0.810089495473077, 0.123147121411940],
        [0.113983176898700, -0.074843993014393, 0.970834416653274, 0.114016223039576, -0.035085806401094, 1.136119883362162, 0.132674969625049],
        [0.068707349990918, -0.131206299703527, 1.193086500002462, 0.068707349990918, -0.131206299703527, 1.193086500002462, 0.127657916394952],
        [None, -0.101349999585660, 1.223845648629919, 0.324549986083302, -0.101349999554536, 1.223754351487657, 0.046806563054797]]),
    "schunk": (0.02, [
        [0, -0.015, 0.1553, 0, -0.015, 0.1553, 0.09],
        [2.295154367360221e-04, 0.114880476266314, 0.505203783848448, -6.617154367385000e-04, 0.114044823733687, 0.155205916151551, 0.074020642029512],
        [-1.594314036332950e-15, 1.110223024625157e-15, 0.4903, -1.594314036332950e-15, 1.110223024625157e-15, 0.4903, 0.09],
        [-4.693429213306548e-05, -0.005883360941466, 0.606104849281791, -2.365707868427609e-06, 0.061878260941472, 0.804871950718208, 0.07],
        [-1.545847699099969e-15, 3.205768983605140e-15, 0.8065, -1.545847699099969e-15, 3.205768983605140e-15, 0.8065, 0.07],
        [4.932409872672667e-04, 0.003349675805265, 0.905303127433903, -4.926409872713459e-04, 0.003181224194743, 0.9713, 0.072548874802886]]),
}


def _one_step_capsules(impl, robot):
    files = pu.default_files(robot, "cmu_mocap_no_hand")
    kp = np.zeros((1, 21, 3))
    if impl is ref:
        b = ref.RefBatch(1, 0.004, *files, np.zeros(6), np.zeros(6))
        b.new_goal(np.zeros((1, 6)))
        b.human_measurement(kp, 0.0)
        out = b.step(0.004)
        return b.capsules(42)[0][0], bool(out["is_safe"][0])
    from human_robot_gym_b200.cabi import make_cfg
    from human_robot_gym_b200.config import ShieldConfig
    b = emul.EmulBatch(make_cfg(ShieldConfig.from_files(0.004, *files), 64), 1, np.zeros(6), np.zeros(6))
    b.new_goal(np.zeros((1, 6)))
    b.human_measurement(kp, 0.0)
    out = b.step(0.004)
    return b.capsules()[0][0], bool(out["is_safe"][0])


@pytest.mark.parametrize("robot", ["modrob1", "schunk"])
@pytest.mark.parametrize("impl", IMPLS)
def test_robot_reach_capsules_golden(impl, robot):
    secure, want = GOLDEN_ROBOT_CAPS[robot]
    caps, is_safe = _one_step_capsules(impl, robot)
    assert not is_safe                                            # test_getSafety: human at the origin -> unsafe
    for link, row in enumerate(want):
        for k, w in enumerate(row):
            if w is None:                                         # value the reference test skips ("manually set")
                continue
            assert abs(caps[link][k] - (w + (secure if k == 6 else 0.0))) <= 1e-5, (link, k)


# ---- SS/tests/src/safety_shield_tests.cc:29-53: planSafetyShield on a coarsened copy of the reference's grid -----------
@needs_ref
def test_plan_safety_shield_grid_device_math_equals_oracle():
    vel = np.arange(-10, 11, 2) / 10.0
    acc = np.arange(-40, 40, 3) / 5.0
    ve = np.arange(0, 101, 4) / 100.0
    V, A, E = np.meshgrid(vel, acc, ve, indexing="ij")
    cases = np.stack([np.zeros(V.size), V.ravel(), A.ravel(), E.ravel(), np.full(V.size, 10.0), np.full(V.size, 100.0)], 1)
    files = pu.default_files("schunk", "cmu_mocap_no_hand")
    rb = ref.RefBatch(1, 0.001, *files, np.zeros(6), np.zeros(6))
    ok_r, ph_r, fin_r = rb.plan_safety_shield(cases)
    ok_e, ph_e, fin_e = emul.plan_safety_shield(0.001, cases)
    assert ok_r.all() and np.array_equal(ok_r, ok_e)
    assert np.abs(fin_r[:, 1] - cases[:, 3]).max() < 1e-6 and np.abs(fin_r[:, 2]).max() < 1e-6   # the reference's own checks
    assert np.array_equal(ph_r, ph_e) and np.array_equal(fin_r, fin_e)                           # bit-identical


# ---- closed loop: the tick logic of the kernels (host build) against the oracle, decision by decision ------------------
def _closed_loop(files, n, ticks, seed, shield_type=1):
    from human_robot_gym_b200.cabi import make_cfg
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario
    sc = ShieldConfig.from_files(0.004, *files)
    scen = SyntheticScenario(n, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, seed)
    pose = np.zeros((n, 6))
    chk = ref.RefBatch(n, 0.004, *files, pose, scen.init_q, shield_type=shield_type, threads=4)
    em = emul.EmulBatch(make_cfg(sc, 1024), n, pose, scen.init_q, shield_type)
    phase = np.arange(n) % 25
    alive = np.ones(n, bool)
    mis = near = 0
    worst = 0.0
    hist = np.zeros(3, int)
    for k in range(ticks):
        t = k * 0.004
        kp = scen.keypoints(t)
        chk.human_measurement(kp, t)
        em.human_measurement(kp, t)
        g = phase == (k % 25)
        if g.any():
            goal = scen.next_goal(g)
            chk.new_goal(goal, mask=g)
            em.new_goal(goal, mask=g)
        a, b = chk.step(t), em.step(t)
        bad = ((a["is_safe"] != b["is_safe"]) | (a["branch"] != b["branch"])) & alive
        if bad.any():
            md = chk.capsules(em.H)[2]
            nr = ((b["status"] & 32) != 0) | (np.abs(np.where(np.isfinite(md), md, 1.0)).min(axis=1) < 1e-6)
            near += int((bad & nr).sum())
            mis += int((bad & ~nr).sum())
            alive &= ~bad
        for key in ("q", "dq", "ddq", "dddq"):
            worst = max(worst, float(np.abs(a[key][alive] - b[key][alive]).max()))
        hist += np.bincount(a["branch"][alive], minlength=3)[:3]
    return mis, near, worst, hist


@needs_ref
@pytest.mark.parametrize("robot,mocap,traj,n,ticks,seed,shield_type", [
    ("schunk", "mujoco_mocap", None, 48, 200, 1001, 1),
    ("panda", "mujoco_mocap", None, 32, 150, 1002, 1),
    ("schunk", "mujoco_mocap_full_body", "trajectory_parameters_schunk_long", 24, 120, 1004, 1),
    ("schunk", "mujoco_mocap", None, 32, 200, 1004, 2)])   # PFL (robot_human_handover_cart.yaml:24)
def test_tick_logic_matches_oracle_closed_loop(robot, mocap, traj, n, ticks, seed, shield_type):
    mis, near, worst, hist = _closed_loop(pu.default_files(robot, mocap, traj), n, ticks, seed, shield_type)
    assert mis == 0, (mis, near)
    assert near <= 1
    assert worst < 1e-9            # north_star tolerance is 1e-5; same IEEE operations give ~1e-13
    assert hist[0] > 0 and hist[2] > 0


@needs_ref
@pytest.mark.parametrize("robot", ["iiwa", "jaco", "kinova3", "sawyer", "ur5e", "modrob0", "modrob1"])
@pytest.mark.parametrize("shield_type", [1, 2])
def test_tick_logic_matches_oracle_on_every_shipped_robot(robot, shield_type):
    """The other robots the reference ships configs for (SS/config/robot_parameters_*.yaml, trajectory_parameters_*.yaml:
    3 / 6 / 7 joints, their own kinematic chains and capsules, ur5e with its own velocity limits), SSM and PFL."""
    mis, near, worst, hist = _closed_loop(pu.default_files(robot, "mujoco_mocap", None), 48, 160, 4242, shield_type)
    assert mis == 0 and near <= 1, (mis, near)
    assert worst < 1e-7            # north_star: 1e-5; root_mode 1 moves a switch time by up to ~1e-9 relative (jaco: 3e-9 here)
    assert hist[0] > 0 and hist.sum() >= 48 * 160 - 48


# ---- sliding-window maxima of a LongTermTraj ----------------------------------------------------------
def _window_golden():
    """long_term_traj_tests.cc:66-193 (MaxAccWindowTest / MaxJerkWindowTest): 7 samples, 2 joints, k = 3."""
    a = np.zeros((7, 4, 2))
    a[:, 2, 0] = [12, 1, 78, 90, 57, 89, 56]
    a[:, 2, 1] = [13, 2, 79, 91, 58, 90, 57]
    a[:, 3] = a[:, 2]  # the jerk test uses the same numbers
    want = np.array([[78, 79], [90, 91], [90, 91], [90, 91], [89, 90], [89, 90], [56, 57]], dtype=float)
    return a, want


@needs_ref
def test_max_window_reference_reproduces_its_own_test():
    a, want = _window_golden()
    acc, jerk = ref.ltt_max_window(a, 0.001, 3)
    assert (acc == want).all() and (jerk == want).all()


def test_max_window_host_class_matches_golden_and_reference():
    from human_robot_gym_b200.safety_shield_py import LongTermTraj, Motion
    a, want = _window_golden()
    ms = [Motion(i * 0.001, a[i, 0], a[i, 1], a[i, 2], a[i, 3]) for i in range(7)]
    ltt = LongTermTraj(ms, 0.001, sliding_window_k=3)
    for i in range(7):
        assert ltt.getMaxAccelerationWindow(i) == want[i].tolist() and ltt.getMaxJerkWindow(i) == want[i].tolist()
    if ref.available():  # random trajectories, windows shorter and longer than the trajectory, signed values
        rng = np.random.default_rng(3)
        for L, k in ((1, 5), (9, 1), (40, 50), (137, 50)):
            tr = rng.normal(size=(L, 4, 3))
            ltt = LongTermTraj([Motion(i * 0.004, *tr[i]) for i in range(L)], 0.004, sliding_window_k=k)
            acc, jerk = ref.ltt_max_window(tr, 0.004, k)
            assert (np.array([ltt.getMaxAccelerationWindow(i) for i in range(L)]) == acc).all()
            assert (np.array([ltt.getMaxJerkWindow(i) for i in range(L)]) == jerk).all()


# ---- capsule -> MuJoCo marker (PlotCapsule), golden outputs of the imported reference (tests/golden/make_plot_capsule_fixture.py)
def test_capsule_markers_match_reference_plot_capsule():
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "plot_capsules.npz"))
    pos, size, mat = emul.capsule_markers(z["caps"])
    assert np.array_equal(pos, z["pos"]) and np.array_equal(size, z["size"])
    assert np.abs(mat - z["mat"]).max() < 1e-15
    assert (z["size"][200:210, 2] == 0.00001).all() and np.array_equal(z["mat"][225], np.eye(3))  # sphere / -z branches hit


# ---- the derivative-chain root finder (csrc/ltp_roots_fast.cuh) against the exact restatement of roots.h:22-50 ---------
# HrgConfig.root_mode 1 (the product default) answers "smallest real root > 1e-7" without the companion-matrix QR iteration
# and hands the close calls (margin < kRootMargin = 1e4 units of the QR iteration's noise) to the exact path. Pinned here:
# wherever the margin is trusted the verdict root / no root is the QR iteration's and the value agrees to 1e-9 relative on
# the planner's own polynomials (1e-7 = kRootValueTol on adversarial ones).
ROOT_MARGIN = 1e4


def _check_fast_roots(coefs, what, tol=1e-9):
    fast, margin, exact = emul.poly_root_fast_batch(coefs)
    assert not (margin == -1).any(), f"{what}: mixed-degree and per-degree form differ"
    trusted = margin >= ROOT_MARGIN
    fin_f, fin_e = np.isfinite(fast), np.isfinite(exact)
    bad_class = trusted & (fin_f != fin_e)
    assert not bad_class.any(), f"{what}: verdict differs on {coefs[bad_class][:3]} (margin {margin[bad_class][:3]})"
    both = trusted & fin_f & fin_e
    rel = np.abs(fast[both] - exact[both]) / np.abs(exact[both])
    assert rel.size == 0 or rel.max() < tol, f"{what}: root differs by {rel.max()}"
    return trusted.mean(), (rel.max() if rel.size else 0.0)


def test_fast_roots_match_qr_on_planner_polynomials():
    """Polynomials harvested from the bench workload (profiles/harvest_polys.py: every optSwitchTimes quartic, also the ones
    k_replan_plan solves ahead of need, and every timeScaling candidate of degree 4, 5, 6; Schunk and Panda)."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "planner_polys.npz"))
    for deg in (4, 5, 6):
        frac, worst = _check_fast_roots(z[f"deg{deg}"], f"degree {deg}")
        print(f"degree {deg}: {len(z[f'deg{deg}'])} polynomials, trusted {frac:.4f}, worst relative difference {worst:.2e}")
        assert frac > (0.95 if deg == 4 else 0.5)     # degree 5 / 6 only occur deep in the ladder, often from rest


@pytest.mark.parametrize("deg", [4, 5, 6])
def test_fast_roots_near_double_roots(deg):
    """(x - r)(x - r (1 + d)) q(x) and (x - r)^2 q(x) + e: pairs of real roots that merge / turn complex as d, e -> 0 - the
    classification the QR iteration decides by rounding. The margin has to hand exactly those to the exact path."""
    rng = np.random.default_rng(40 + deg)
    rows = []
    for _ in range(4000):
        r = 10 ** rng.uniform(-3, 1.5)
        rest = rng.uniform(-5, 5, deg - 2) * rng.choice([1, 1, 0.01], deg - 2)   # the other roots, some tiny
        d = rng.choice([0.0, 1e-15, 1e-13, 1e-11, 1e-9, 1e-7, 1e-5, 1e-3, 1e-1])
        p = np.poly(np.concatenate([[r, r * (1 + d)], rest])) * rng.uniform(0.5, 200)
        e = rng.choice([0.0, 0.0, 1e-14, -1e-14, 1e-10, -1e-10, 1e-6, -1e-6]) * np.abs(p).max()
        p[-1] += e
        rows.append(p)
    # kRootValueTol: a trusted root is one the QR iteration's noise cannot move by more than 1e-7 relative
    frac, worst = _check_fast_roots(np.array(rows), f"near-double degree {deg}", tol=1e-7)
    print(f"degree {deg}: trusted {frac:.3f}, worst relative difference {worst:.2e}")
    assert 0.2 < frac < 0.95                          # the merged pairs are NOT trusted, the separated ones are


def test_fast_roots_from_rest_structure():
    """x^k q(x) with exact trailing zeros (a start from rest): a k-fold zero eigenvalue the QR iteration returns as noise -
    never trusted for k >= 2."""
    rng = np.random.default_rng(7)
    for deg, k in ((4, 2), (5, 4), (6, 4), (6, 2)):
        c = np.zeros((500, deg + 1))
        c[:, :deg + 1 - k] = rng.uniform(-200, 200, (500, deg + 1 - k))
        fast, margin, exact = emul.poly_root_fast_batch(c)
        assert (margin < ROOT_MARGIN).all()


def test_exact_roots_against_lapack():
    """Independent cross-check of the eigenvalue restatement: the smallest real root > 1e-7 of the exact QR
    path (the device's restatement of roots.h + Eigen's real Schur form, bit-identical to oracle/eigen_shim) against
    numpy.roots (LAPACK's Hessenberg QR on the same companion matrix) on the planner's own polynomials. Compared where the
    verdict is not a close call (margin >= 1e4): same verdict, value within 1e-8."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "planner_polys.npz"))
    rng = np.random.default_rng(3)
    for deg in (4, 5, 6):
        c = z[f"deg{deg}"]
        c = c[rng.permutation(len(c))[:3000]]
        _, margin, exact = emul.poly_root_fast_batch(c)
        n_cmp = bad = 0
        worst = 0.0
        for p, m, r in zip(c, margin, exact):
            if m < ROOT_MARGIN:
                continue
            lam = np.roots(p)
            real = lam[(np.abs(lam.imag) <= 1e-9 * np.maximum(np.abs(lam.real), 1e-300)) & (lam.real > 1e-7)].real
            want = real.min() if real.size else np.inf
            n_cmp += 1
            if np.isfinite(want) != np.isfinite(r):
                bad += 1
            elif np.isfinite(r):
                worst = max(worst, abs(r - want) / abs(want))
        print(f"degree {deg}: {n_cmp} compared with LAPACK, verdict mismatches {bad}, worst relative difference {worst:.2e}")
        assert n_cmp > 1000 and bad == 0 and worst < 1e-8


def test_time_scaling_ladder_same_in_both_root_modes():
    """timeScaling (LTP/src/long_term_planner.cc:358-645) with the derivative-chain roots + the rest-start rejection rule
    (root_mode 1) against the exact eigen-solve for every polynomial (root_mode 0): same verdict, same profile flag, same
    switch times - for moving starts and, above all, for starts from exact rest, whose candidate polynomials x^k q(x) the QR
    iteration answers with noise roots (csrc/ltp_scaling.cuh: ltp_cand_rejected_from_rest)."""
    rng = np.random.default_rng(2024)
    lims = [[-2.9, 2.9, 1.0, 2.0, 15.0],                    # Schunk / Panda: v_max_allowed, a_max_ltt, j_max_ltt
            [-6.28318530718, 6.28318530718, 3.14, 2.0, 15.0], [-6.28318530718, 6.28318530718, 6.28, 2.0, 15.0],   # ur5e
            [-2.9, 2.9, 2.6, 10.0, 400.0], [-3.1, 3.1, 1.0, 4.0, 40.0]]
    n_rest = n_acc = n = 0
    try:
        for trial in range(10000):
            lim = lims[trial % len(lims)]
            rest = (trial // len(lims)) % 2 == 0
            q0 = rng.uniform(0.7 * lim[0], 0.7 * lim[1])
            v0 = 0.0 if rest else rng.uniform(-0.8, 0.8) * lim[2]
            a0 = 0.0 if rest else rng.uniform(-0.8, 0.8) * lim[3]
            q_goal = float(np.clip(q0 + rng.choice([0.002, 0.05, 0.5, 2.0, 8.0]) * rng.uniform(-1, 1), lim[0], lim[1]))
            emul.set_root_mode(0)
            ok, t_opt, direction, _ = emul.ltp_opt_switch_times(0.004, lim, q_goal, q0, v0, a0, lim[2])
            if not ok:
                continue
            t_req = t_opt[6] * rng.uniform(1.02, 4.0) + rng.choice([0.0, 0.05, 1.0])
            res = []
            for mode in (0, 1):
                emul.set_root_mode(mode)
                res.append(emul.ltp_time_scaling(0.004, lim, q_goal, q0, v0, a0, direction, t_req))
            (ok0, T0, vd0, m0), (ok1, T1, vd1, m1) = res
            assert ok0 == ok1 and m0 == m1, (trial, lim, q_goal, q0, v0, a0, t_req, res)
            assert np.abs(T0 - T1).max() <= 1e-9 * max(1.0, np.abs(T0).max()), (trial, T0, T1)
            assert vd0 == vd1 or abs(vd0 - vd1) <= 1e-9 * max(1.0, abs(vd0)), (trial, vd0, vd1)
            n += 1
            n_rest += rest
            n_acc += ok0
    finally:
        emul.set_root_mode(0)
    print(f"{n} ladders ({n_rest} from rest), {n_acc} with an accepted candidate")
    assert n > 7000 and n_rest > 2500 and 0.2 * n < n_acc < n
