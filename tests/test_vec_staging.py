"""Host staging adapter between a vectorised env and the batched shield."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from human_robot_gym_b200.vec_staging import VecEnvShieldAdapter  # noqa: E402


class _FakeShield:
    """Records what one tick_host call receives and answers with values derived from its inputs."""
    n_envs, J, K = 5, 3, 4
    device = "cpu"

    def __init__(self):
        self.calls = []
        self.synced = 0

    def tick_host(self, kp, t, goal_mask, goal_q, q, dq, ddq, flags):
        self.calls.append((kp.clone(), t.clone(), None if goal_mask is None else goal_mask.clone(),
                           None if goal_q is None else goal_q.clone()))
        q[:] = kp.sum(dim=(1, 2))[:, None] + torch.arange(self.J, dtype=torch.float64)
        dq[:] = t[:, None].expand(-1, self.J)
        ddq[:] = 0 if goal_q is None else goal_q * goal_mask[:, None]
        flags[:] = torch.tensor([1, 4, 5, 0, 1], dtype=torch.uint8)

    def synchronize(self):
        self.synced += 1


def test_gather_tick_scatter():
    sh = _FakeShield()
    ad = VecEnvShieldAdapter(sh, pin_memory=False)
    rng = np.random.default_rng(0)
    kp = rng.normal(size=(5, 4, 3))
    for i in range(5):
        ad.set_human_measurement(i, kp[i].tolist(), 0.004 * i)
    ad.set_goal(3, [0.1, 0.2, 0.3])
    ad.step()
    got_kp, got_t, gm, gq = sh.calls[0]
    assert np.array_equal(got_kp.numpy(), kp) and np.allclose(got_t.numpy(), 0.004 * np.arange(5))
    assert gm.tolist() == [0, 0, 0, 1, 0] and gq[3].tolist() == [0.1, 0.2, 0.3] and sh.synced == 1
    q, dq, ddq = ad.desired_motion(3)
    assert np.allclose(q, kp[3].sum() + np.arange(3)) and np.allclose(dq, 0.012) and np.allclose(ddq, [0.1, 0.2, 0.3])
    assert [ad.get_safety(i) for i in range(5)] == [True, False, True, False, True]
    assert [ad.branch(i) for i in range(5)] == [0, 2, 2, 0, 0]
    # the goal mask is consumed by the tick; a tick without goals passes NULL for both goal arguments
    ad.put_rows(range(5), kp, np.zeros(5))
    ad.step()
    assert sh.calls[1][2] is None and sh.calls[1][3] is None and ad.ticks == 2


def test_missing_measurement_and_bad_shapes_fail_loudly():
    ad = VecEnvShieldAdapter(_FakeShield(), pin_memory=False)
    for i in range(4):
        ad.set_human_measurement(i, np.zeros((4, 3)), 0.0)
    with pytest.raises(RuntimeError, match="1 env"):
        ad.step()
    with pytest.raises(ValueError):
        ad.set_human_measurement(0, np.zeros((3, 3)), 0.0)
    with pytest.raises(ValueError):
        ad.set_goal(0, [0.0, 0.0])
    ad.step(require_all=False)  # explicit opt-out: stale rows are re-sent
    # rows from worker processes: measurements for some envs, goals for others
    ad.put_rows([0, 1, 2, 3, 4], np.ones((5, 4, 3)), np.arange(5.0), goal_qpos=np.full((2, 3), 7.0), goal_envs=[1, 4])
    assert ad.goal_mask.tolist() == [0, 1, 0, 0, 1] and ad.goal[4].tolist() == [7.0, 7.0, 7.0]


@pytest.mark.gpu
def test_adapter_matches_direct_calls():
    """Per-env writes + one batched host tick give the setpoints of the plain humanMeasurement / newLongTermTrajectory /
    step sequence, bit for bit."""
    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.synthetic import SyntheticScenario
    from tests.parity_util import default_files, mocap_joint_names
    files = default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n = 70
    scen = SyntheticScenario(n, mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 21)
    a = BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0")
    b = BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0")
    ad = VecEnvShieldAdapter(b)
    for k in range(80):
        t = k * 0.004
        kp = scen.keypoints(t)
        gmask = (np.arange(n) % 25) == (k % 25)
        g = scen.next_goal(gmask)
        a.humanMeasurement(torch.from_numpy(kp), t)
        if gmask.any():
            a.newLongTermTrajectory(torch.from_numpy(g), env_mask=torch.from_numpy(gmask))
        a.step(t)
        for i in range(n):
            ad.set_human_measurement(i, kp[i], t)
            if gmask[i]:
                ad.set_goal(i, g[i])
        ad.step()
        assert np.array_equal(ad.q, a.q.cpu().numpy()) and np.array_equal(ad.ddq, a.ddq.cpu().numpy())
        assert np.array_equal(ad.flags & 1, a.is_safe.cpu().numpy()) and np.array_equal(ad.flags >> 1, a.branch.cpu().numpy())
    a.close()
    b.close()
