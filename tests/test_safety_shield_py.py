"""The reference's own binding tests (SS/tests/bindings/test_motion.py, test_ltt.py, test_safety_shield.py) restated
against the drop-in module. Motion / LongTermTraj are host containers (CPU); everything that steps a shield is -m gpu."""
import numpy as np
import pytest

from human_robot_gym_b200.safety_shield_py import LongTermTraj, Motion, SafetyShield, ShieldType
from tests import parity_util as pu
from tests.test_golden_vectors import GOLDEN_ROBOT_CAPS


def test_motion_constructors_and_predicates():
    m = Motion(6)
    assert m.getAngle() == [0.0] * 6 and m.getTime() == 0 and m.isStopped()
    m = Motion(1.0, [1.0, 2.0], 0.5)
    assert m.getS() == 0.5 and m.getVelocity() == [0.0, 0.0]
    m = Motion(0.0, [0.0, 0.0], [1.0, 0.0], [0.0, 0.0], [0.0, 0.0], 0.25)
    assert not m.isStopped() and m.getS() == 0.25
    n = Motion(0.0, [0.00005, 0.0], [1.00005, 0.0])
    assert m.hasSamePos(n) and not m.hasSamePos(Motion(0.0, [0.0005, 0.0]))   # binding default 1e-4 (bindings :41)
    assert m.hasSamePos(Motion(time=0.0, q=[0.0005, 0.0], s=1.0), threshold=1e-3) and m.hasSameVel(n) and not m.hasSameVel(Motion(0.0, [0, 0], [1.1, 0]))
    m.setTime(2.0); m.setS(3.0); m.setAngle([1, 1]); m.setJerk(np.array([0.5, 0.5]))
    assert (m.getTime(), m.getS(), m.getAngle(), m.getJerk()) == (2.0, 3.0, [1.0, 1.0], [0.5, 0.5])
    assert int(ShieldType.PFL) == 2 and eval("ShieldType." + "SSM") == ShieldType.SSM   # failsafe_controller.py:173


def test_long_term_traj_accessors():
    """SS/tests/bindings/test_ltt.py + long_term_traj_tests.cc (index clamping, interpolate, sliding window)."""
    ms = [Motion(i * 0.1, [0.1 * i, 0.0], [0.0, 0.0], [12.0 - i, float(i % 3)], [float(i), 0.0]) for i in range(5)]
    ltt = LongTermTraj(ms, 0.1, starting_index=3, sliding_window_k=3, alpha_i_max=5.0)
    assert ltt.getLength() == 5 and ltt.getCurrentPos() == 0 and ltt.getAlphaI() == [5.0, 5.0]
    assert ltt.getTrajectoryIndex(3) == 0 and ltt.getTrajectoryIndex(5) == 2 and ltt.getTrajectoryIndex(100) == 4
    assert ltt.getTrajectoryIndex(0) == 0                       # clamped up to current_pos_
    ltt.increasePosition()
    assert ltt.getCurrentPos() == 1 and ltt.getTrajectoryIndex(3) == 1
    assert ltt.getNextMotion().getAngle()[0] == pytest.approx(0.2)
    assert ltt.getMaxAccelerationWindow(3) == [11.0, 2.0]       # window [1, 4): max(11, 10, 9), max(1, 2, 0)
    assert ltt.getMaxJerkWindow(7) == [4.0, 0.0]
    m0 = Motion(0, [0.0, 0.0], [1.0, 0.0], [1.0, 0.0], [-1.0, 0.0])
    m1 = Motion(0.1, [0.104833333333333, 0.0], [1.095, 0.0], [0.9, 0.0], [0.0, 0.0])
    lt = LongTermTraj()
    lt.setLongTermTrajectory([m0, m1], 0.1)
    r = lt.interpolate(0.01, 0.5, 1.0, 0.2, [10.0, 10.0], [10.0, 10.0], [100.0, 100.0])
    assert r.getAngle()[0] == pytest.approx(0.010049833333333, abs=1e-5)
    assert r.getAcceleration()[0] == pytest.approx(1.25745, abs=1e-5) and r.getJerk()[0] == pytest.approx(1.56199, abs=1e-5)


def _shield(robot):
    f = pu.default_files(robot, "cmu_mocap_no_hand")
    return SafetyShield(sample_time=0.004, trajectory_config_file=f[0], robot_config_file=f[1], mocap_config_file=f[2],
                        init_x=0.0, init_y=0.0, init_z=0.0, init_roll=0.0, init_pitch=0.0, init_yaw=0.0,
                        init_qpos=[0.0] * 6)


@pytest.mark.gpu
@pytest.mark.parametrize("robot", ["modrob1", "schunk"])
def test_reference_binding_tests_on_the_drop_in(robot):
    """test_safety_shield.py: reset, humanMeasurement, newLTT, step time, getSafety, golden robot capsules."""
    sh = _shield(robot)
    sh.reset(init_x=0.0, init_y=0.0, init_z=0.0, init_roll=0.0, init_pitch=0.0, init_yaw=0.0, init_qpos=[0.0] * 6,
             current_time=0.0)
    sh.newLongTermTrajectory([0.0] * 6, [0.0] * 6)
    sh.humanMeasurement([[0.0, 0.0, 0.0] for _ in range(21)], 0.0)
    motion = sh.step(0.004)
    assert motion.getTime() == 0.004
    assert sh.getSafety() is False and sh.getShieldType() == ShieldType.SSM
    caps = sh.getRobotReachCapsules()
    secure, want = GOLDEN_ROBOT_CAPS[robot]
    assert len(caps) == 6 and all(len(c) == 7 for c in caps)
    for link, row in enumerate(want):
        for k, w in enumerate(row):
            if w is not None:
                assert abs(caps[link][k] - (w + (secure if k == 6 else 0.0))) <= 1e-5
    assert len(sh.getHumanReachCapsules(type=1)) == len(sh.getHumanReachCapsules(type=2)) > 0


@pytest.mark.gpu
def test_set_long_term_trajectory_verdicts():
    """test_setLTT (test_safety_shield.py:115-131) + the three TrajectoryException cases of safety_shield.cc:699-730."""
    from human_robot_gym_b200.safety_shield_py import TrajectoryException
    sh = _shield("modrob1")
    z = [0.0] * 6
    good = [Motion(0.0, z, z, z, 0.0), Motion(1.0, z, z, [1.0] + z[1:], 0.0),
            Motion(2.0, [0.5] + z[1:], [1.0] + z[1:], [-1.0] + z[1:], 0.0), Motion(3.0, [1.0] + z[1:], z, z, 0.0)]
    sh.setLongTermTrajectory(LongTermTraj(good, 1.0))
    with pytest.raises(TrajectoryException):
        sh.setLongTermTrajectory(LongTermTraj([Motion(0.0, [0.5] + z[1:], z, z, 0.0)] + good[1:], 1.0))
    with pytest.raises(TrajectoryException):
        sh.setLongTermTrajectory(LongTermTraj(good[:3], 1.0))   # does not end stopped
