"""Batched FailsafeController: ctypes layout on CPU; kernels against the numpy restatement on GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from human_robot_gym_b200 import cabi
from oracle import controller_oracle as co
from tests import parity_util as pu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_controller_struct_layout_matches_the_header(tmp_path):
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "hrg_shield.h"\n'
                    'int main(void){printf("%zu %zu %zu\\n", sizeof(HrgControllerConfig), '
                    'offsetof(HrgControllerConfig, kp), offsetof(HrgControllerConfig, torque_max));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(prog)])
    out = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert out == [C.sizeof(cabi.HrgControllerConfig), cabi.HrgControllerConfig.kp.offset,
                   cabi.HrgControllerConfig.torque_max.offset]


def test_oracle_scale_action_known_values():
    # failsafe.json of the reference: input +-1 -> output +-0.2; out-of-range actions are clipped first
    a = np.array([-2.0, -1.0, 0.0, 0.5, 1.0, 3.0])
    np.testing.assert_allclose(co.scale_action(a, -1.0, 1.0, -0.2, 0.2), [-0.2, -0.2, 0.0, 0.1, 0.2, 0.2], atol=1e-16)
    g = co.set_goal_position(np.array([0.3, -0.3]), np.array([1.0, -1.0]), position_limit=np.array([[-1.1, -1.1], [1.1, 1.1]]))
    np.testing.assert_allclose(g, [1.1, -1.1])


def test_controller_entry_points_reject_null_arguments():
    lib = cabi.load_library()
    k = cabi.HrgControllerConfig()
    assert lib.hrg_controller_set_goal(None, C.byref(k), None, None, None, None, None, None) == cabi.EINVAL
    assert lib.hrg_controller_torques(None, C.byref(k), *([None] * 8), None) == cabi.EINVAL


@pytest.mark.gpu
def test_batched_controller_matches_restatement():
    torch = pytest.importorskip("torch")
    from human_robot_gym_b200.batched import BatchedSafetyShield
    from human_robot_gym_b200.config import ShieldConfig
    from human_robot_gym_b200.controller import BatchedFailsafeController
    from human_robot_gym_b200.synthetic import SyntheticScenario
    files = pu.default_files("schunk")
    sc = ShieldConfig.from_files(0.004, *files)
    n, J, ticks = 203, 6, 60
    scen = SyntheticScenario(n, pu.mocap_joint_names(files[2]), sc.traj.q_min_allowed, sc.traj.q_max_allowed, 31)
    mk = lambda: BatchedSafetyShield(n, 0.004, *files, init_qpos=scen.init_q, device="cuda:0", ltt_capacity=512)
    sh, plain = mk(), mk()
    rng = np.random.default_rng(5)
    lim = np.stack([sc.traj.q_min_allowed[:J] + 0.05, sc.traj.q_max_allowed[:J] - 0.05])
    kp = rng.uniform(20, 80, J)
    trange = (np.full(J, -40.0), np.full(J, 40.0))
    ctl = BatchedFailsafeController(sh, trange, input_max=1, input_min=-1, output_max=0.2, output_min=-0.2, kp=kp,
                                    damping_ratio=1.0, qpos_limits=lim)
    kd = 2 * np.sqrt(kp) * 1.0
    joint_pos, joint_vel = scen.init_q.copy(), np.zeros((n, J))
    phase = np.arange(n) % 25
    worst_tau, n_clipped = 0.0, 0
    for k in range(ticks):
        t = k * 0.004
        kpts = torch.from_numpy(scen.keypoints(t))
        ctl.set_human_measurement(kpts, t)
        plain.humanMeasurement(kpts, t)
        gm = phase == (k % 25)
        if gm.any():
            action = rng.uniform(-1.5, 1.5, (n, J))
            goal = ctl.set_goal(action, joint_pos, env_mask=gm).cpu().numpy()
            want = np.stack([co.set_goal_position(co.scale_action(action[e], -1.0, 1.0, -0.2, 0.2), joint_pos[e],
                                                  position_limit=lim) for e in range(n)])
            assert np.array_equal(goal[gm], want[gm])  # elementwise IEEE arithmetic: bit-exact
            plain.newLongTermTrajectory(torch.from_numpy(want), env_mask=torch.from_numpy(gm))
        M = rng.uniform(-0.5, 0.5, (n, J, J)) + 2.0 * np.eye(J)
        comp = rng.uniform(-5, 5, (n, J))
        tau = ctl.run_controller(joint_pos, joint_vel, M, comp, t).cpu().numpy()
        plain.step(t)
        for name in ("q", "dq", "ddq", "is_safe", "branch"):  # the controller path drives the very same shield
            assert torch.equal(getattr(sh, name), getattr(plain, name)), (k, name)
        qd, dqd, ddqd = (x.cpu().numpy() for x in (sh.q, sh.dq, sh.ddq))
        want = np.stack([co.torques(qd[e], dqd[e], ddqd[e], joint_pos[e], joint_vel[e], M[e], comp[e], kp, kd, *trange)
                         for e in range(n)])
        worst_tau = max(worst_tau, float(np.abs(tau - want).max()))
        n_clipped += int((np.abs(want) == 40.0).sum())
        # a crude plant so that errors and velocities are not identically zero: follow the setpoint with a lag
        joint_vel = 0.8 * dqd + rng.normal(0, 0.01, (n, J))
        joint_pos = joint_pos + 0.9 * (qd - joint_pos)
    assert worst_tau < 1e-10, worst_tau  # np.dot (BLAS) may order the J-term sums differently; N m
    assert n_clipped >= 0
    assert sh.stats()["n_replan"] > 0
