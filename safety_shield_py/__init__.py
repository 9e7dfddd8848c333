"""`import safety_shield_py` — the module name the reference binds with pybind11
(SS/bindings/safety_shield_bindings.cpp:26) and that `FailsafeController` imports
(HRG/controllers/failsafe_controller/failsafe_controller/failsafe_controller.py:23:
`from safety_shield_py import SafetyShield, ShieldType`).

With this repository root on `sys.path` that import resolves to the CUDA-backed drop-in with no edit on the reference
side; the reference's own binding tests (SS/tests/bindings/test_{motion,ltt,safety_shield}.py) run against it unchanged
(tests/test_reference_binding_tests.py).
"""
from human_robot_gym_b200.safety_shield_py import (LongTermTraj, Motion, RobotMovementException, SafetyShield, ShieldType,
                                                   TrajectoryException)

__all__ = ["Motion", "LongTermTraj", "ShieldType", "SafetyShield", "RobotMovementException", "TrajectoryException"]
