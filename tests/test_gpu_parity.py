"""-m gpu: the CUDA path, called through the C ABI, against the CPU checker on identical seeded inputs."""
import ctypes as C

import numpy as np
import pytest

from tests import parity_util as pu

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _lib():
    from human_robot_gym_b200 import cabi
    return cabi, cabi.load_library()


def _dev(a, dtype=None):
    return torch.as_tensor(np.ascontiguousarray(a), device="cuda:0") if dtype is None else torch.as_tensor(
        np.ascontiguousarray(a), dtype=dtype, device="cuda:0")


def random_capsules(rng, n):
    c = np.zeros((n, 7))
    c[:, :3] = rng.uniform(-1, 1, (n, 3))
    c[:, 3:6] = c[:, :3] + rng.uniform(-0.6, 0.6, (n, 3))
    c[:, 6] = rng.uniform(0.02, 0.3, n)
    return c


def capsule_cases(seed=7, n=200000):
    rng = np.random.default_rng(seed)
    c1, c2 = random_capsules(rng, n), random_capsules(rng, n)
    k = n // 10
    c1[:k, 3:6] = c1[:k, :3]                      # sphere vs capsule
    c2[k:2 * k, 3:6] = c2[k:2 * k, :3]            # capsule vs sphere
    c1[2 * k:3 * k, 3:6] = c1[2 * k:3 * k, :3]    # sphere vs sphere
    c2[2 * k:3 * k, 3:6] = c2[2 * k:3 * k, :3]
    # exactly parallel, axis aligned (denom == 0 branch incl. the chained comparisons)
    s = slice(3 * k, 4 * k)
    ax = rng.integers(0, 3, k)
    for c in (c1, c2):
        p = np.round(rng.uniform(-1, 1, (k, 3)), 2)
        c[s, :3] = p
        c[s, 3:6] = p
        c[s, 3 + 0] += np.where(ax == 0, np.round(rng.uniform(0.1, 1, k), 2), 0)
        c[s, 3 + 1] += np.where(ax == 1, np.round(rng.uniform(0.1, 1, k), 2), 0)
        c[s, 3 + 2] += np.where(ax == 2, np.round(rng.uniform(0.1, 1, k), 2), 0)
    # touching: distance within ~1e-7 of the threshold
    s = slice(4 * k, 5 * k)
    c2[s, 6] = 0.0
    return c1, c2


def test_capsule_distance_matches_reference():
    from oracle import ref
    cabi, L = _lib()
    c1, c2 = capsule_cases()
    want = ref.capsule_dist(c1, c2)
    d1, d2 = _dev(c1), _dev(c2)
    out = torch.zeros(c1.shape[0], dtype=torch.float64, device="cuda:0")
    cabi.check(L.hrg_stage_capsule_dist(c1.shape[0], d1.data_ptr(), d2.data_ptr(), out.data_ptr(), None))
    got = out.cpu().numpy()
    both_nan = np.isnan(want) & np.isnan(got)
    assert np.array_equal(np.isnan(want), np.isnan(got))
    diff = np.abs(np.where(both_nan, 0, want - got))
    assert diff.max() <= 1e-12, diff.max()  # tolerance of north_star is 1e-5; IEEE-identical arithmetic gives ~0
    flips = ((want < 0) != (got < 0)) & ~both_nan
    assert flips.sum() == 0


def test_plan_safety_shield_grid():
    """The reference's own 21 x 80 x 101 grid (SS/tests/src/safety_shield_tests.cc:29-53): a_max 10, j_max 100."""
    cabi, L = _lib()
    vel = np.arange(-10, 11) / 10.0
    acc = np.arange(-40, 40) / 5.0
    ve = np.arange(101) / 100.0
    V, A, E = np.meshgrid(vel, acc, ve, indexing="ij")
    cases = np.stack([np.zeros(V.size), V.ravel(), A.ravel(), E.ravel(), np.full(V.size, 10.0), np.full(V.size, 100.0)], 1)
    n = cases.shape[0]
    ok = torch.zeros(n, dtype=torch.uint8, device="cuda:0")
    ph = torch.zeros(n, 6, dtype=torch.float64, device="cuda:0")
    fin = torch.zeros(n, 3, dtype=torch.float64, device="cuda:0")
    d_cases = _dev(cases)
    cabi.check(L.hrg_stage_plan_safety_shield(0.001, n, d_cases.data_ptr(), ok.data_ptr(), ph.data_ptr(),
                                              fin.data_ptr(), None))
    ok, fin = ok.cpu().numpy(), fin.cpu().numpy()
    assert ok.all()
    assert np.abs(fin[:, 1] - cases[:, 3]).max() < 1e-6   # final_vel ~ ve
    assert np.abs(fin[:, 2]).max() < 1e-6                  # final_acc ~ 0
    from oracle import ref
    if ref.available():
        files = pu.default_files("schunk", "cmu_mocap_no_hand")
        rb = ref.RefBatch(1, 0.001, *files, np.zeros(6), np.zeros(6))
        ok_r, ph_r, fin_r = rb.plan_safety_shield(cases)
        assert np.array_equal(ok_r, ok)
        assert np.abs(ph_r - ph.cpu().numpy()).max() < 1e-12
        assert np.abs(fin_r - fin).max() < 1e-12


def test_long_term_planner_matches_reference():
    from oracle import ref
    if not ref.available():
        pytest.skip("needs oracle/_ref")
    from human_robot_gym_b200.cabi import make_cfg
    from human_robot_gym_b200.config import ShieldConfig
    cabi, L = _lib()
    files = pu.default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    cap = 1024
    cfg = make_cfg(sc, cap)
    J = cfg.nb_joints
    rng = np.random.default_rng(11)
    n = 512
    q0 = rng.uniform(-1.5, 1.5, (n, J))
    v0 = rng.uniform(-0.6, 0.6, (n, J)) * (rng.random((n, 1)) < 0.7)
    a0 = rng.uniform(-1.5, 1.5, (n, J)) * (rng.random((n, 1)) < 0.7)
    goal = np.clip(q0 + rng.uniform(-1.0, 1.0, (n, J)) * rng.choice([0.05, 0.3, 1.0], (n, 1)), sc.traj.q_min_allowed,
                   sc.traj.q_max_allowed)
    traj = torch.zeros(n, cap, 4, J, dtype=torch.float64, device="cuda:0")
    length = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    d_goal, d_q0, d_v0, d_a0 = _dev(goal), _dev(q0), _dev(v0), _dev(a0)  # keep the device copies alive over the call
    cabi.check(L.hrg_stage_ltp_plan(C.byref(cfg), n, d_goal.data_ptr(), d_q0.data_ptr(), d_v0.data_ptr(),
                                    d_a0.data_ptr(), traj.data_ptr(), length.data_ptr(), cap, None))
    length, traj = length.cpu().numpy(), traj.cpu().numpy()
    t = sc.traj
    n_ok, worst = 0, 0.0
    for i in range(n):
        Lr, tr = ref.ltp_plan(0.004, t.q_min_allowed, t.q_max_allowed, t.v_max_allowed, t.a_max_ltt, t.j_max_ltt, goal[i],
                              q0[i], v0[i], a0[i], cap)
        assert (Lr if Lr <= cap else -Lr) == length[i], (i, Lr, length[i])
        if Lr > 0 and Lr <= cap:
            n_ok += 1
            worst = max(worst, float(np.abs(tr - traj[i, :Lr]).max()))
    assert n_ok > n // 4
    assert worst < 1e-5, worst


@pytest.mark.parametrize("robot,n_envs,ticks,seed", [("schunk", 512, 150, 1001), ("panda", 512, 150, 1002)])
def test_closed_loop_matches_checker(robot, n_envs, ticks, seed):
    res = pu.closed_loop_compare(pu.default_files(robot), n_envs=n_envs, ticks=ticks, seed=seed)
    print(res)
    pu.assert_parity(res)                             # decisions exact; setpoints, distances, capsules < 1e-5 (north_star)
    assert res["branch_hist"][0] > 0 and res["branch_hist"][2] > 0   # branch 1: tests/test_parity_gaps.py


@pytest.mark.parametrize("robot,n_envs,ticks,seed", [("schunk", 1024, 150, 1011), ("panda", 512, 150, 1012)])
def test_closed_loop_exact_root_mode_matches_checker(robot, n_envs, ticks, seed, monkeypatch):
    """HrgConfig.root_mode 0: every planner polynomial through the exact restatement of the reference's QR iteration (the
    arbiter the default mode 1 falls back to), product configuration."""
    monkeypatch.setenv("HRG_ROOT_MODE", "0")
    res = pu.closed_loop_compare(pu.default_files(robot), n_envs=n_envs, ticks=ticks, seed=seed, store_capsules=False)
    print(res)
    pu.assert_parity(res)


@pytest.mark.parametrize("robot,n_envs,ticks,seed,shield_type", [("schunk", 1024, 150, 3001, 1), ("panda", 517, 150, 3002, 1),
                                                                  ("schunk", 384, 150, 3004, 2)])
def test_verification_cascade_matches_checker(robot, n_envs, ticks, seed, shield_type):
    """Product configuration (no stored capsules): pass 0 of reach + verify runs as the Pos -> Vel -> Accel cascade over
    compacted survivor lists. Decisions and setpoints against the reference; distances of the evaluated models."""
    res = pu.closed_loop_compare(pu.default_files(robot), n_envs=n_envs, ticks=ticks, seed=seed, shield_type=shield_type,
                                 store_capsules=False)
    print(res)
    pu.assert_parity(res)
    assert 0.02 < res["unsafe_frac"] < 0.6, res  # both verdicts, i.e. all three cascade outcomes occur


@pytest.mark.parametrize("n_envs", [1, 37, 95])
def test_ragged_batches_match_checker(n_envs):
    """Batch sizes that are not a multiple of the 32-env tile / 4-group warp: a one-env batch, partial last tiles, partial
    replan groups."""
    res = pu.closed_loop_compare(pu.default_files("schunk"), n_envs=n_envs, ticks=120, seed=2000 + n_envs)
    print(res)
    assert res["mismatch_excl_near"] == 0, res
    assert res["max_setpoint_diff"] < 1e-5 and res["max_min_dist_diff"] < 1e-5 and res["max_capsule_diff"] < 1e-5, res
    assert res["gpu_stats"]["n_steps"] == n_envs * 120
    assert res["gpu_stats"]["n_replan"] > 0


def test_full_size_batch_position_invariance():
    """BASELINE size (65 536 shields on one GPU) through a size-independent property: every env is an exact replica of
    one of 128 source envs, scattered over the batch by a fixed permutation. All replicas of a source env must produce
    bit-identical setpoints and decisions on every tick (tiles, replan groups, work counters and the order in which the
    replan list is compacted must not leak into results), and the 128 source envs must agree bit for bit with a 128-env
    batch of their own - which the closed-loop tests above tie to the reference."""
    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario
    files = pu.default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n_src, n, ticks = 128, 65536, 60
    scen = SyntheticScenario(n_src, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 4242)
    rng = np.random.default_rng(7)
    src = rng.permutation(np.arange(n) % n_src)          # source env of every batch slot
    src_t = torch.from_numpy(src).cuda()
    small = BatchedSafetyShield(n_src, 0.004, *files, init_qpos=scen.init_q, device="cuda:0", ltt_capacity=512)
    big = BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q[src], device="cuda:0", ltt_capacity=512)
    phase = np.arange(n_src) % 25
    n_unsafe = 0
    for k in range(ticks):
        t = k * 0.004
        kp = torch.from_numpy(scen.keypoints(t)).cuda()
        small.humanMeasurement(kp, t)
        big.humanMeasurement(kp[src_t].contiguous(), t)
        gm = phase == (k % 25)
        if gm.any():
            g = torch.from_numpy(scen.next_goal(gm)).cuda()
            gm_t = torch.from_numpy(gm).cuda()
            small.newLongTermTrajectory(g, env_mask=gm_t)
            big.newLongTermTrajectory(g[src_t].contiguous(), env_mask=gm_t[src_t].contiguous())
        small.step(t)
        big.step(t)
        for name in ("q", "dq", "ddq", "dddq", "is_safe", "branch", "status"):
            a, b = getattr(small, name), getattr(big, name)
            assert torch.equal(a[src_t], b), (k, name)
        n_unsafe += int((small.is_safe == 0).sum())
    st = big.stats()
    assert st["n_steps"] == n * ticks and st["n_replan"] > 0
    assert 0 < n_unsafe < n_src * ticks  # both verdicts occur


def test_pfl_shield_matches_checker():
    """ShieldType PFL (BASELINE config 4b, robot_human_handover_cart.yaml:24): per-sample Cartesian velocities of every
    planned LTT, alpha_i from their differences, v_limit failsafe, is_safe || is_under_v_limit."""
    res = pu.closed_loop_compare(pu.default_files("schunk"), n_envs=384, ticks=150, seed=1004, shield_type=2)
    print(res)
    assert res["mismatch_excl_near"] == 0, res
    assert res["max_setpoint_diff"] < 1e-5 and res["max_min_dist_diff"] < 1e-5 and res["max_capsule_diff"] < 1e-5, res
    assert res["gpu_stats"]["n_replan"] > 0


def test_full_body_long_horizon():
    files = pu.default_files("schunk", "mujoco_mocap_full_body", "trajectory_parameters_schunk_long")
    res = pu.closed_loop_compare(files, n_envs=256, ticks=100, seed=1004)
    print(res)
    assert res["mismatch_excl_near"] == 0, res
    assert res["max_setpoint_diff"] < 1e-5 and res["max_min_dist_diff"] < 1e-5, res


def test_host_buffer_ticks_agree():
    """hrg_shield_tick_host, the double-buffered upload/tick_staged pair and the device-tensor verbs give bit-identical
    setpoints and flags (they run the same kernels; only the staging differs)."""
    from human_robot_gym_b200.batched import BatchedSafetyShield, PipelinedHostShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario
    files = pu.default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n, ticks = 300, 80
    scen = SyntheticScenario(n, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 77)
    mk = lambda: BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0", ltt_capacity=1024)
    a, b, c = mk(), mk(), mk()
    d = PipelinedHostShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0", chunks=3, ltt_capacity=1024)
    pin = lambda *shape, dt=torch.float64: torch.zeros(*shape, dtype=dt).pin_memory()
    outs = {k: (pin(n, 6), pin(n, 6), pin(n, 6), pin(n, dt=torch.uint8)) for k in "bcd"}
    kp = [torch.from_numpy(scen.keypoints(k * 0.004)).pin_memory() for k in range(ticks + 1)]
    tt = [torch.full((n,), k * 0.004, dtype=torch.float64).pin_memory() for k in range(ticks + 1)]
    phase = np.arange(n) % 25
    masks = [torch.from_numpy((phase == (k % 25)).astype(np.uint8)).pin_memory() for k in range(ticks + 1)]
    goals = [torch.from_numpy(scen.next_goal()).pin_memory() for _ in range(ticks + 1)]
    cs = torch.cuda.Stream()
    c.upload_inputs(0, kp[0], tt[0], masks[0], goals[0], cs)
    for k in range(ticks):
        a.humanMeasurement(kp[k], k * 0.004)
        a.newLongTermTrajectory(goals[k], env_mask=masks[k])
        a.step(k * 0.004)
        b.tick_host(kp[k], tt[k], masks[k], goals[k], *outs["b"])
        c.upload_inputs((k + 1) & 1, kp[k + 1], tt[k + 1], masks[k + 1], goals[k + 1], cs)
        c.tick_staged(k & 1, *outs["c"])
        d.tick_host(kp[k], tt[k], masks[k], goals[k], *outs["d"])
        torch.cuda.synchronize()
        qa, dqa, ddqa = a.q.cpu(), a.dq.cpu(), a.ddq.cpu()
        fa = (a.is_safe | (a.branch << 1)).cpu()
        for key in "bcd":
            q, dq, ddq, fl = outs[key]
            assert torch.equal(q, qa) and torch.equal(dq, dqa) and torch.equal(ddq, ddqa), (key, k)
            assert torch.equal(fl, fa), (key, k)
    assert a.stats()["n_replan"] > 0 and a.stats() == b.stats() == c.stats()
    for x in (a, b, c, d):
        x.close()


def test_ltt_max_window_matches_reference():
    """getMaxAccelerationWindow / getMaxJerkWindow of the live LTT of every env: at the shield's own
    index, and at explicit indices (before the start, inside, past the end). A maximum adds no rounding of its own, so the
    values agree as well as the planned trajectories themselves do (tolerance 1e-5, north_star; observed ~1e-12)."""
    from oracle import ref
    if not ref.available():
        pytest.skip("needs oracle/_ref (the window tables are the reference's own)")
    seen = {"n": 0}
    rng = np.random.default_rng(5)
    holder = {}

    def collect(k, a, b):
        if k % 20 != 19:
            return
        gpu, chk = holder["gpu"], holder["chk"]
        for idx in (None, rng.integers(-5, 400, gpu.n_envs).astype(np.int32)):
            acc, jerk = gpu.ltt_max_window(index=None if idx is None else torch.from_numpy(idx))
            ra, rj = chk.live_max_window(idx)
            d = max(float(np.abs(acc.cpu().numpy() - ra).max()), float(np.abs(jerk.cpu().numpy() - rj).max()))
            seen["max_diff"] = max(seen.get("max_diff", 0.0), d)
            seen["nonzero"] = seen.get("nonzero", 0) + int((ra != 0).sum())
        seen["n"] += 1

    # closed_loop_compare owns both shields; reach them through the factory hooks
    import human_robot_gym_b200.batched as batched
    orig_gpu, orig_chk = batched.BatchedSafetyShield, pu.make_checker

    class Spy(orig_gpu):
        def __init__(self, *a, **kw):
            super().__init__(*a, **kw)
            holder["gpu"] = self

    def spy_checker(*a, **kw):
        c, kind = orig_chk(*a, **kw)
        holder["chk"] = c
        return c, kind

    batched.BatchedSafetyShield, pu.make_checker = Spy, spy_checker
    try:
        res = pu.closed_loop_compare(pu.default_files("schunk"), n_envs=203, ticks=120, seed=77, collect=collect)
    finally:
        batched.BatchedSafetyShield, pu.make_checker = orig_gpu, orig_chk
    print("ltt_max_window:", seen)
    assert res["mismatch_excl_near"] == 0 and seen["n"] == 6 and seen["nonzero"] > 1000, (res, seen)
    assert seen["max_diff"] < 1e-5, seen


def test_capsule_markers_match_reference_plot_capsule():
    """hrg_capsule_markers against PlotCapsule outputs generated by importing the reference (tests/golden)."""
    import os
    cabi, lib = _lib()
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "plot_capsules.npz"))
    n = z["caps"].shape[0]
    caps = _dev(z["caps"])
    pos, size = torch.zeros(n, 3, dtype=torch.float64, device="cuda:0"), torch.zeros(n, 3, dtype=torch.float64, device="cuda:0")
    mat = torch.zeros(n, 9, dtype=torch.float64, device="cuda:0")
    cabi.check(lib.hrg_capsule_markers(n, caps.data_ptr(), pos.data_ptr(), size.data_ptr(), mat.data_ptr(), None))
    torch.cuda.synchronize()
    assert np.array_equal(pos.cpu().numpy(), z["pos"]) and np.array_equal(size.cpu().numpy(), z["size"])
    assert np.abs(mat.cpu().numpy().reshape(n, 3, 3) - z["mat"]).max() < 1e-14


def test_set_ltt_batch_equals_one_by_one():
    """hrg_shield_set_ltt_batch: verdicts and the following closed loop identical to installing the same
    user trajectories env by env with hrg_shield_set_ltt; the verdict cases are those of test_setLTT
    (SS/tests/bindings/test_safety_shield.py:115-131) and safety_shield.cc:699-730."""
    from human_robot_gym_b200.batched import BatchedSafetyShield
    files = pu.default_files("schunk")
    n, J = 40, 6
    a = BatchedSafetyShield(n, 0.004, *files, device="cuda:0")
    b = BatchedSafetyShield(n, 0.004, *files, device="cuda:0")
    z = np.zeros(J)
    good = np.array([[z, z, z, z], [z, z, np.r_[1.0, z[1:]], z], [np.r_[0.5, z[1:]], np.r_[1.0, z[1:]], np.r_[-1.0, z[1:]], z],
                     [np.r_[1.0, z[1:]], z, z, z]])                      # [4 samples, 4, J]
    bad_start = good.copy(); bad_start[0, 0, 0] = 0.5                    # start position differs -> 2
    bad_end = good.copy(); bad_end[3] = good[2]                          # does not end stopped -> 4
    kinds = [good, bad_start, bad_end]
    envs = np.array([3, 7, 8, 21, 39, 0], dtype=np.int32)
    which = [0, 1, 0, 2, 0, 0]
    lengths = np.array([4, 4, 4, 4, 4, 3], dtype=np.int32)               # the last one is truncated: moving at its end -> 4
    traj = np.zeros((len(envs), 5, 4, J))                                 # Lmax 5 > every length
    for m, k in enumerate(which):
        traj[m, :4] = kinds[k]
    want = [a.set_ltt(int(e), traj[m, :lengths[m]], 1.0, 0, 1.0) for m, e in enumerate(envs)]
    got = b.set_ltt_batch(envs, traj, lengths, 1.0, alpha_i=1.0).cpu().numpy().tolist()
    assert got == want == [0, 2, 0, 4, 0, 4]
    bad = b.set_ltt_batch(np.array([n + 5], dtype=np.int32), traj[:1], lengths[:1], 1.0).cpu().numpy().tolist()
    assert bad == [-1]
    kp = np.tile(np.array([3.0, 3.0, 1.0]), (n, a.K, 1))
    for k in range(40):
        t = k * 0.004
        for s in (a, b):
            s.humanMeasurement(torch.from_numpy(kp), t)
            s.step(t)
        assert torch.equal(a.q, b.q) and torch.equal(a.ddq, b.ddq) and torch.equal(a.is_safe, b.is_safe)
    assert float(a.q[3].abs().max()) > 0 and float(a.q[1].abs().max()) == 0  # env 3 follows its user trajectory, env 1 rests
    a.close()
    b.close()


@pytest.mark.gpu
def test_state_dict_round_trip_is_bit_identical():
    """state_dict / load_state_dict (hrg_shield_state_*): after a restore the closed loop repeats itself bit for bit, replans
    and pending trajectories included; a blob of another shape is refused."""
    from human_robot_gym_b200 import cabi
    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario
    files = pu.default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n = 300
    scen = SyntheticScenario(n, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 4242)
    sh = BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0", ltt_capacity=512, config=sc)
    goals = [torch.from_numpy(scen.next_goal()).cuda() for _ in range(4)]
    mask = torch.from_numpy((np.arange(n) % 25)[None] == np.arange(25)[:, None]).to(torch.uint8).cuda()

    def run(k0, k1):
        out = []
        for k in range(k0, k1):
            t = torch.full((n,), k * 0.004, dtype=torch.float64, device="cuda:0")
            kp = torch.from_numpy(scen.keypoints(k * 0.004)).cuda()
            sh.tick_device(kp, t, mask[k % 25], goals[(k // 25) % 4])
            out.append(torch.cat([sh.q.flatten(), sh.dq.flatten(), sh.ddq.flatten(), sh.is_safe.double(), sh.branch.double()]).clone())
        return torch.stack(out)

    run(0, 40)
    state = sh.state_dict()
    a = run(40, 90)
    sh.load_state_dict(state)
    b = run(40, 90)
    assert torch.equal(a, b)
    other = BatchedSafetyShield(n + 1, 0.004, *files, init_qpos=np.zeros((n + 1, 6)), device="cuda:0", ltt_capacity=512, config=sc)
    with pytest.raises(cabi.ShieldLibraryError):
        other.load_state_dict(state)


@pytest.mark.gpu
def test_sparse_goal_upload_equals_dense():
    """hrg_shield_upload_inputs_sparse ((env, goal) pairs) against hrg_shield_upload_inputs (mask + goal of every env): the
    same ticks bit for bit, including ticks without goals and an out-of-range index that must be dropped."""
    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario
    files = pu.default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n, J = 257, 6
    scen = SyntheticScenario(n, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 99)
    mk = lambda: BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0", ltt_capacity=512, config=sc)  # noqa: E731
    a, b = mk(), mk()
    pin = dict(pin_memory=True)
    outs = [[torch.zeros(n, J, dtype=torch.float64, **pin) for _ in range(3)] + [torch.zeros(n, dtype=torch.uint8, **pin)]
            for _ in range(2)]
    goals = [scen.next_goal() for _ in range(3)]
    for k in range(70):
        kp = torch.from_numpy(scen.keypoints(k * 0.004)).pin_memory()
        t = torch.full((n,), k * 0.004, dtype=torch.float64).pin_memory()
        sel = np.flatnonzero(np.arange(n) % 25 == k % 25) if k % 3 else np.zeros(0, dtype=np.int64)   # some ticks: no goals
        g = goals[(k // 25) % 3]
        mask = np.zeros(n, dtype=np.uint8)
        mask[sel] = 1
        if len(sel):
            a.upload_inputs(0, kp, t, torch.from_numpy(mask).pin_memory(), torch.from_numpy(g).pin_memory())
        else:
            a.upload_inputs(0, kp, t)
        idx = np.concatenate([sel, [n + 5]]).astype(np.int32) if len(sel) else None      # + one index out of range
        val = np.concatenate([g[sel], np.zeros((1, J))]) if len(sel) else None
        b.upload_inputs_sparse(0, kp, t, None if idx is None else torch.from_numpy(idx).pin_memory(),
                               None if val is None else torch.from_numpy(np.ascontiguousarray(val)).pin_memory())
        a.tick_staged(0, *outs[0])
        b.tick_staged(0, *outs[1])
        torch.cuda.synchronize()
        for x, y in zip(outs[0], outs[1]):
            assert torch.equal(x, y), k
    assert a.stats()["n_replan"] == b.stats()["n_replan"] > n


@pytest.mark.parametrize("robot,shield_type,store", [("modrob0", 1, False), ("modrob0", 2, True), ("ur5e", 1, False),
                                                     ("iiwa", 1, True), ("jaco", 2, False), ("sawyer", 1, False)])
def test_closed_loop_other_shipped_robots(robot, shield_type, store):
    """Robots the reference ships configs for besides Schunk and Panda: 3 joints (modrob0: the generic-J forms of the sampler
    and of the PFL constructor), 6 with their own velocity limits (ur5e), 7 (iiwa, jaco, sawyer); SSM and PFL, with and
    without stored capsules (plain reach + verify / the model cascade)."""
    res = pu.closed_loop_compare(pu.default_files(robot), n_envs=256, ticks=150, seed=5100, shield_type=shield_type,
                                 store_capsules=store)
    print(robot, res)
    pu.assert_parity(res)
    assert res["gpu_stats"]["n_replan"] > 0 and res["branch_hist"][0] > 0


@pytest.mark.parametrize("robot,sample_time,shield_type", [("schunk", 0.001, 1), ("schunk", 0.002, 2), ("panda", 0.01, 1),
                                                           ("schunk", 0.008, 1)])
def test_closed_loop_other_sample_times(robot, sample_time, shield_type):
    """The shield's sample time is a constructor argument (bindings :46-62; human-robot-gym runs 0.004, the reference's
    tests 0.001): LTT sampling and the planner's rounding of switch times to whole samples; at 8-10 ms the reference starts to
    reject plans (`t_rel[i] < -eps`, LTP/src/long_term_planner.cc:341-343), and so must the device, env for env."""
    res = pu.closed_loop_compare(pu.default_files(robot), n_envs=128, ticks=200, seed=5300, shield_type=shield_type,
                                 sample_time=sample_time, store_capsules=False, ltt_capacity=4096)
    print(robot, sample_time, res)
    pu.assert_parity(res)
    assert res["gpu_stats"]["n_replan"] > 0 and res["branch_hist"][0] > 0 and res["branch_hist"][2] > 0
