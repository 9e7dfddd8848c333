"""TEST INFRASTRUCTURE ONLY - numpy restatement of the FailsafeController arithmetic.

PARITY UNPINNED: the reference has no test or fixture for this controller, and the arithmetic it inherits lives in a
third-party dependency that is absent from /root/reference and from this image - robosuite==1.3.2 (pinned in the
reference's requirements.txt:8 / environment.yml:13). The functions below restate that published code and the reference's
own call sites; only tests/ may import this module.

  scale_action       robosuite/controllers/base_controller.py (Controller.scale_action), called at
                     HRG/controllers/failsafe_controller/failsafe_controller/failsafe_controller.py:285
  set_goal_position  robosuite/utils/control_utils.py, called at failsafe_controller.py:289-294
  torques            failsafe_controller.py:352-369 (+ Controller.clip_torques)
"""
import numpy as np


def scale_action(action, input_min, input_max, output_min, output_max):
    action_scale = abs(output_max - output_min) / abs(input_max - input_min)
    action_output_transform = (output_max + output_min) / 2.0
    action_input_transform = (input_max + input_min) / 2.0
    action = np.clip(action, input_min, input_max)
    return (action - action_input_transform) * action_scale + action_output_transform


def set_goal_position(delta, current_position, position_limit=None, set_pos=None):
    goal_position = set_pos if set_pos is not None else current_position + delta
    if position_limit is not None:
        goal_position = np.clip(goal_position, position_limit[0], position_limit[1])
    return goal_position


def torques(q_des, dq_des, ddq_des, joint_pos, joint_vel, mass_matrix, torque_compensation, kp, kd, torque_min, torque_max):
    """One env: the body of FailsafeController.run_controller after the shield step."""
    position_error = q_des - joint_pos
    vel_pos_error = dq_des - joint_vel
    feedback_torque = np.multiply(np.array(position_error), np.array(kp)) + np.multiply(vel_pos_error, kd)
    tau = np.dot(mass_matrix, feedback_torque + ddq_des) + torque_compensation
    return np.clip(tau, torque_min, torque_max)
