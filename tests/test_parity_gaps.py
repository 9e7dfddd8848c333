"""Inputs the first-round parity tests never fed ("Parity gaps"): non-zero base poses, mid-episode resets,
ShieldType::OFF and mixed-type batches, branch 1 of determineNextMotion, the reach + verify stage with a base pose, the
BASELINE config sizes, and the trajectory-capacity contract. Every scenario runs twice:

  * `-m "not gpu"`: the tick logic of csrc/*.cuh compiled for the host (tests/host_emul) against the compiled reference
    (oracle/_ref) - small batches, so the GPU-less container already knows the logic is right;
  * `-m gpu`: the CUDA path through the C ABI (BatchedSafetyShield) against the compiled reference - the parity tests proper.
"""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import ref
from tests import parity_util as pu

needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built")
gpu = pytest.mark.gpu
THREADS = max(1, min(32, os.cpu_count() or 1))

# (id, robot, kwargs)   -- kwargs of closed_loop_compare shared by the CPU (small) and the GPU (full) form
SCENARIOS = {
    # (a) per-env base pose (SS/src/robot_reach.cc:5-17,41-52; failsafe_controller.py:160-189 passes the robot's real base)
    "base_pose_schunk": ("schunk", dict(base_pose="random", seed=2101)),
    "base_pose_panda": ("panda", dict(base_pose="random", seed=2102)),
    # (b) masked mid-episode reset of a random third every 40 ticks with new base pose / init_qpos / shield type
    #     (SS/src/safety_shield.cc:164-201, SS/src/human_reach.cc:44-50: the keypoint velocity survives the reset)
    "reset_mid_episode": ("schunk", dict(base_pose="random", reset_period=40, seed=2103)),
    "reset_new_types": ("schunk", dict(base_pose="random", reset_period=40, reset_types=[0, 1, 2], seed=2104)),
    # (c) ShieldType::OFF (SS/src/safety_shield.cc:617, :41,154,174) alone and mixed with SSM / PFL in one batch
    "shield_off": ("schunk", dict(shield_type=0, seed=2105)),
    "mixed_types": ("schunk", dict(shield_type="mixed", seed=2106)),
    "mixed_types_panda_reset": ("panda", dict(shield_type="mixed", base_pose="random", reset_period=50,
                                              reset_types=[0, 1, 2], seed=2107)),
    # (h) configuration values no shipped YAML carries (`cfg`: overrides of the mocap / trajectory file, pu.variant_files):
    #     measurement delay and errors (SS/src/human_reach.cc:9-24 hands them to reach_lib::System: reach time + delay, radii +
    #     errors; every shipped mocap file has zeros), PFL parameters (v_safe, alpha_i_max), limits that differ per joint
    "meas_delay_errors": ("schunk", dict(cfg=dict(mocap=dict(delay=0.05, measurement_error_pos=0.02,
                                                             measurement_error_vel=0.1)), seed=5401)),
    "meas_errors_panda_pfl": ("panda", dict(cfg=dict(mocap=dict(delay=0.012, measurement_error_pos=0.005,
                                                                 measurement_error_vel=0.3)), shield_type=2, seed=5402)),
    "pfl_parameters": ("schunk", dict(cfg=dict(traj=dict(v_safe=0.25, alpha_i_max=1.0, max_s_stop=0.35)), shield_type=2,
                                      seed=5403)),
    # (i) measurements that do not arrive on every tick and carry an older timestamp (masked humanMeasurement: the reach
    #     time of SS/src/human_reach.cc:76-80 then differs from env to env and grows between measurements)
    "irregular_measurements": ("schunk", dict(meas_every=3, meas_lag=0.002, seed=5501)),
    "irregular_measurements_panda_pfl": ("panda", dict(meas_every=4, meas_lag=0.0013, shield_type=2, seed=5502)),
    "per_joint_limits": ("schunk", dict(cfg=dict(mocap=dict(delay=0.02, measurement_error_pos=0.01),
                                                 traj=dict(a_max_allowed=[6.0, 10.0, 8.0, 10.0, 5.0, 10.0],
                                                           j_max_allowed=[150.0, 400.0, 300.0, 400.0, 200.0, 400.0],
                                                           a_max_ltt=[1.0, 2.0, 1.5, 2.0, 1.0, 2.0],
                                                           j_max_ltt=[8.0, 15.0, 10.0, 15.0, 8.0, 15.0],
                                                           v_max_allowed=[0.7, 1.3, 1.0, 0.5, 2.0, 1.0])), seed=5404)),
}


def _run(name, n_envs, ticks, backend, **extra):
    robot, kw = SCENARIOS[name]
    kw = dict(kw, **extra)
    if isinstance(kw.get("shield_type"), str):
        kw["shield_type"] = np.arange(n_envs) % 3          # OFF, SSM, PFL interleaved
    cfg = kw.pop("cfg", None)
    files = pu.variant_files(robot, **cfg) if cfg else pu.default_files(robot)
    res = pu.closed_loop_compare(files, n_envs=n_envs, ticks=ticks, backend=backend, **kw)
    print(name, backend, res)
    pu.assert_parity(res)
    if kw.get("reset_period"):
        assert res["n_reset"] > n_envs // 2, res
    if np.all(np.asarray(kw.get("shield_type", 1)) == 0) and not kw.get("reset_types"):
        assert res["unsafe_frac"] == 0.0 and res["branch_hist"][2] == 0, res     # OFF: always safe, never verified
    else:
        assert res["n_verified"] > 0, res
    return res


@needs_ref
@pytest.mark.parametrize("name", list(SCENARIOS))
def test_scenario_host_logic_matches_reference(name):
    _run(name, 60, 130, "host_emul")


@gpu
@pytest.mark.parametrize("name", list(SCENARIOS))
def test_scenario_cuda_matches_reference(name):
    _run(name, 768, 150, "cuda")


@gpu
def test_scenario_cuda_cascade_with_resets():
    """The product configuration (no stored capsules -> verification cascade) under resets, base poses and mixed types."""
    _run("mixed_types_panda_reset", 1000, 150, "cuda", store_capsules=False)


# (d) branch 1 of determineNextMotion: all three "selected-trajectory index" values of north_star
@needs_ref
def test_branch1_host_logic_matches_reference():
    res = pu.closed_loop_compare(pu.default_files("schunk"), n_envs=128, ticks=200, seed=5, backend="host_emul",
                                 scenario=pu.AdaptivePulseScenario)
    print(res)
    pu.assert_parity(res)
    assert min(res["branch_hist"]) > 0, res


@gpu
@pytest.mark.parametrize("robot,n_envs,store", [("schunk", 2048, True), ("panda", 1024, False)])
def test_all_three_branches_cuda_match_reference(robot, n_envs, store):
    res = pu.closed_loop_compare(pu.default_files(robot), n_envs=n_envs, ticks=220, seed=5,
                                 scenario=pu.AdaptivePulseScenario, store_capsules=store)
    print(res)
    pu.assert_parity(res)
    assert min(res["branch_hist"]) > 100, res          # recovery, potential AND failsafe branch, each compared exactly


# (e) the reach + verify stage with a base pose, on explicit inputs
@gpu
@pytest.mark.parametrize("robot", ["schunk", "panda"])
def test_stage_reach_verify_with_base_pose_matches_reference(robot):
    import torch

    from human_robot_gym_b200 import cabi
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import rest_pose
    L = cabi.load_library()
    files = pu.default_files(robot)
    sc = ShieldConfig.from_files(0.004, *files)
    cfg = cabi.make_cfg(sc, 8)
    J, K, H = cfg.nb_joints, cfg.n_keypoints, cfg.n_ext + 2 * cfg.n_bodies
    rng = np.random.default_rng(77)
    n = 3000
    pose = pu.random_base_pose(rng, n)
    pose[:50] = 0.0                                                  # identity base too
    q_lo, q_hi = np.asarray(sc.traj.q_min_allowed), np.asarray(sc.traj.q_max_allowed)
    q1 = rng.uniform(q_lo, q_hi, (n, J))
    q2 = np.clip(q1 + rng.uniform(-0.3, 0.3, (n, J)) * (rng.random((n, 1)) < 0.9), q_lo, q_hi)   # 10 % resting links
    s_diff, t_reach = rng.uniform(0, 0.25, n), rng.uniform(0.004, 0.6, n)
    alpha = rng.uniform(0, 6, (n, J))
    # human next to the robot's (moved) base so that all three models decide
    kp = rest_pose(pu.mocap_joint_names(files[2]))[None] + pose[:, None, :3] + rng.uniform(-1.2, 1.2, (n, 1, 3)) * [1, 1, 0.3]
    kv = rng.uniform(-1.5, 1.5, (n, K, 3)) * (rng.random((n, 1, 1)) < 0.8)
    chk = ref.RefBatch(1, 0.004, *files, np.zeros(6), np.zeros(J))
    r_c, h_c, md_c, safe_c = chk.stage_reach_verify(pose, q1, q2, s_diff, t_reach, alpha, kp, kv, H)
    dev = lambda a: torch.as_tensor(np.ascontiguousarray(a), device="cuda:0")  # noqa: E731
    ins = [dev(x) for x in (pose, q1, q2, s_diff, t_reach, alpha, kp, kv)]
    robot_t = torch.zeros(n, J, 7, dtype=torch.float64, device="cuda:0")
    human_t = torch.zeros(n, H, 7, dtype=torch.float64, device="cuda:0")
    md_t = torch.zeros(n, 3, dtype=torch.float64, device="cuda:0")
    safe_t = torch.zeros(n, dtype=torch.uint8, device="cuda:0")
    cabi.check(L.hrg_stage_reach_verify(C.byref(cfg), n, *[x.data_ptr() for x in ins], robot_t.data_ptr(), human_t.data_ptr(),
                                        md_t.data_ptr(), safe_t.data_ptr(), None), "hrg_stage_reach_verify")
    r_g, h_g, md_g, safe_g = (x.cpu().numpy() for x in (robot_t, human_t, md_t, safe_t))
    assert np.abs(r_g - r_c).max() < 1e-12 and np.abs(h_g - h_c).max() < 1e-12     # north_star: 1e-5
    assert np.abs(md_g - md_c).max() < 1e-12
    near = np.abs(md_c).min(axis=1) < 1e-6
    assert np.array_equal(safe_g[~near], safe_c[~near]) and near.sum() <= 2
    assert 0.1 < safe_c.mean() < 0.9                                               # both verdicts


# (f) the BASELINE configuration sizes (config 2: Panda 4 096; config 4: 16 384 full-body, long horizon, SSM and PFL)
@gpu
@pytest.mark.parametrize("robot,mocap,traj,n_envs,ticks,shield_type,seed", [
    ("panda", "mujoco_mocap", None, 4096, 100, 1, 1002),
    ("schunk", "mujoco_mocap_full_body", "trajectory_parameters_schunk_long", 16384, 100, 1, 1004),
    ("schunk", "mujoco_mocap_full_body", "trajectory_parameters_schunk_long", 16384, 100, 2, 1004)])
def test_config_sizes_match_reference(robot, mocap, traj, n_envs, ticks, shield_type, seed):
    chk0 = pu.make_checker
    pu.make_checker = lambda *a, **k: chk0(*a, **dict(k, threads=THREADS))       # all host cores for the reference side
    try:
        res = pu.closed_loop_compare(pu.default_files(robot, mocap, traj), n_envs=n_envs, ticks=ticks, seed=seed,
                                     shield_type=shield_type, store_capsules=False, ltt_capacity=512)
    finally:
        pu.make_checker = chk0
    print(res)
    pu.assert_parity(res)
    assert res["branch_hist"][0] > 0 and res["branch_hist"][2] > 0 and res["gpu_stats"]["n_replan"] > n_envs


# (g) the one documented divergence: a plan longer than ltt_capacity
def test_ltt_capacity_bound_covers_every_plan():
    """hrg_ltt_capacity_bound against the reference planner on adversarial problems: goals at the opposite joint limit,
    full speed / acceleration the wrong way at the start. No plan may need more samples than the bound."""
    from human_robot_gym_b200 import cabi
    from human_robot_gym_b200.config import ShieldConfig
    if not ref.available():
        pytest.skip("needs oracle/_ref")
    for robot in ("schunk", "panda"):
        sc = ShieldConfig.from_files(0.004, *pu.default_files(robot))
        t = sc.traj
        bound = cabi.ltt_capacity_bound(sc)
        J = sc.robot.nb_joints
        lo, hi = np.asarray(t.q_min_allowed[:J]), np.asarray(t.q_max_allowed[:J])
        v, a = np.asarray(t.v_max_allowed[:J]), np.asarray(t.a_max_ltt[:J])
        rng = np.random.default_rng(3)
        longest, n_ok = 0, 0
        for i in range(300):
            sgn = rng.choice([-1.0, 1.0], J)
            wrong_way = i % 2                       # start moving AWAY from the goal, just far enough from the limit to brake
            q0 = np.where(sgn > 0, lo, hi) + sgn * (0.01 + 0.62 * wrong_way + rng.uniform(0, 0.3, J) * (i % 3 == 0))
            goal = np.where(sgn > 0, hi, lo) - sgn * 0.01   # (the planner rejects end points that round outside the limits)
            v0 = -sgn * v * rng.uniform(0.5, 1.0, J) * wrong_way
            a0 = -sgn * a * rng.uniform(0.0, 0.5, J) * (i % 4 == 1)
            Lr, _ = ref.ltp_plan(0.004, t.q_min_allowed, t.q_max_allowed, t.v_max_allowed, t.a_max_ltt, t.j_max_ltt, goal,
                                 q0, v0, a0, 8)
            if Lr > 0:
                n_ok += 1
                longest = max(longest, Lr)
        assert n_ok > 100 and 0.5 * bound < longest <= bound, (robot, longest, bound, n_ok)


@gpu
def test_small_ltt_capacity_is_reported_not_silent():
    """With ltt_capacity below the plan length the replan is rejected (the reference would accept it): the env keeps its
    old trajectory, HRG_ST_LTT_TRUNCATED is set, HrgStats.n_ltt_truncated counts it and check_health() raises. With the
    default capacity (hrg_ltt_capacity_bound) the same goals are planned and followed."""
    import torch

    from human_robot_gym_b200 import cabi
    from human_robot_gym_b200.batched import BatchedSafetyShield
    files = pu.default_files("schunk")
    n = 64
    kp = torch.from_numpy(np.tile(np.array([5.0, 5.0, 1.0]), (n, 23, 1)))
    goal = torch.from_numpy(np.tile(np.array([1.5, -1.0, 1.2, 0.8, -0.9, 1.0]), (n, 1)))     # ~2.5 s = ~640 samples away
    small = BatchedSafetyShield(n, 0.004, *files, device="cuda:0", ltt_capacity=128)
    full = BatchedSafetyShield(n, 0.004, *files, device="cuda:0")
    assert full.ltt_capacity == cabi.ltt_capacity_bound(full.config) >= 2000
    for s in (small, full):
        for k in range(30):
            s.humanMeasurement(kp, k * 0.004)
            if k == 2:
                s.newLongTermTrajectory(goal)
            s.step(k * 0.004)
    st = small.stats()
    assert st["n_ltt_truncated"] > 0 and st["n_replan_failed"] >= st["n_ltt_truncated"]
    assert (small.status.cpu().numpy() & cabi.ST_LTT_TRUNCATED).all()      # retried and rejected every tick, like a failed plan
    assert float(small.q.abs().max()) == 0.0                                # stays on the old trajectory
    with pytest.raises(cabi.ShieldLibraryError, match="ltt_capacity"):
        small.check_health()
    assert full.check_health()["n_ltt_truncated"] == 0 and float(full.q.abs().max()) > 1e-3
    small.close()
    full.close()


# (j) edge inputs, on the host build of the tick logic (CPU only): what a caller can hand over without breaking any rule of the
# reference's API. Same lockstep comparison as everywhere else; ltt_capacity 2048 because moves from limit to limit are long.
EDGE_MODES = ["static_human", "collapsed_human", "dup_timestamps", "time_backwards", "human_far", "human_inside_robot",
              "goals_outside_limits", "goals_at_limits", "init_at_limits", "init_beyond_limits"]


@needs_ref
@pytest.mark.parametrize("robot,shield_type", [("schunk", 1), ("panda", 2)])
@pytest.mark.parametrize("mode", EDGE_MODES)
def test_edge_inputs_host_logic_matches_reference(mode, robot, shield_type):
    """static_human: the same keypoints every tick (zero velocities). collapsed_human: all keypoints in one point (every human
    capsule a sphere: the exact-equality sphere dispatch of RL/include/intersections.hpp:105-126). dup_timestamps: every second
    measurement repeats the previous time (the dt < 1e-7 branch of SS/src/human_reach.cc:52-62 keeps the old velocity).
    time_backwards: some measurements carry an older time than the one before (negative dt, sign-flipped velocities).
    human_far / human_inside_robot: never / always in collision. goals outside / exactly at the joint limits
    (LongTermPlanner::checkInputs, LTP/src/long_term_planner.cc:68-77). init_at_limits / init_beyond_limits: the robot starts
    on / 2 % outside its limits (checkMotionForJointLimits, SS/src/safety_shield.cc:502-509)."""
    from human_robot_gym_b200.cabi import make_cfg
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario
    from tests.host_emul import emul
    n, ticks, seed = 48, 130, 77
    files = pu.default_files(robot)
    sc = ShieldConfig.from_files(0.004, *files)
    scen = SyntheticScenario(n, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, seed)
    qlo, qhi = np.asarray(sc.traj.q_min_allowed), np.asarray(sc.traj.q_max_allowed)
    rng = np.random.default_rng(seed)
    init_q = scen.init_q.copy()
    if mode in ("init_at_limits", "init_beyond_limits"):
        init_q = np.where(rng.random(init_q.shape) < 0.5, qlo, qhi)
        if mode == "init_beyond_limits":
            init_q = init_q * np.where(rng.random(init_q.shape) < 0.3, 1.02, 1.0)
    pose = np.zeros((n, 6))
    chk = ref.RefBatch(n, 0.004, *files, pose, init_q, shield_type=shield_type, threads=4)
    em = emul.EmulBatch(make_cfg(sc, 2048), n, pose, init_q, shield_type)
    phase = np.arange(n) % 25
    alive = np.ones(n, bool)
    mis = near = 0
    worst = 0.0
    hist = np.zeros(3, int)
    kp0 = scen.keypoints(0.0)
    for k in range(ticks):
        t = tm = k * 0.004
        kp = scen.keypoints(t)
        if mode == "static_human":
            kp = kp0
        elif mode == "collapsed_human":
            kp = np.broadcast_to(kp0[:, :1, :], kp0.shape) + 0.001 * k * np.array([1.0, 0.0, 0.0])
        elif mode == "dup_timestamps":
            tm = (k // 2) * 0.008
        elif mode == "time_backwards":
            tm = t - (0.02 if k % 7 == 3 else 0.0)
        elif mode == "human_far":
            kp = kp + np.array([50.0, 0.0, 0.0])
        elif mode == "human_inside_robot":
            kp = kp * 0.05
        chk.human_measurement(kp, tm)
        em.human_measurement(kp, np.full(n, tm))
        g = phase == (k % 25)
        if g.any():
            goal = scen.next_goal(g)
            if mode == "goals_outside_limits":
                goal = goal + np.where(rng.random(goal.shape) < 0.3, rng.choice([-1.0, 1.0], goal.shape) * (qhi - qlo) * 0.6, 0.0)
            elif mode == "goals_at_limits":
                goal = np.where(rng.random(goal.shape) < 0.5, qlo, qhi)
            chk.new_goal(goal, mask=g)
            em.new_goal(goal, mask=g)
        a, b = chk.step(t), em.step(np.full(n, t))
        assert not (b["status"] & 16).any()                       # HRG_ST_LTT_TRUNCATED: the capacity contract, not this test
        bad = ((a["is_safe"] != b["is_safe"]) | (a["branch"] != b["branch"])) & alive
        if bad.any():
            md = chk.capsules(em.H)[2]
            nr = ((b["status"] & 32) != 0) | (np.abs(np.where(np.isfinite(md), md, 1.0)).min(axis=1) < 1e-6)
            near += int((bad & nr).sum())
            mis += int((bad & ~nr).sum())
            alive &= ~bad
        for key in ("q", "dq", "ddq", "dddq"):
            assert np.isfinite(b[key]).all() and np.isfinite(a[key]).all()
            worst = max(worst, float(np.abs(a[key][alive] - b[key][alive]).max()))
        hist += np.bincount(a["branch"][alive], minlength=3)[:3]
    print(mode, robot, mis, near, worst, hist.tolist())
    assert mis == 0 and near <= 1, (mis, near)
    assert worst < 1e-7, worst                                    # north_star: 1e-5
    if mode == "human_far":
        assert hist[2] == 0
    if mode == "human_inside_robot":
        assert hist[2] > hist[0]
