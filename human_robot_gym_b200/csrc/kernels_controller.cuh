// Batched FailsafeController: the two per-tick computations of the shield's immediate caller that are
// not MuJoCo's - the goal from the policy action, and the PD+ torque law. One thread per env; the tables come by value.
// Reference: HRG/controllers/failsafe_controller/failsafe_controller/failsafe_controller.py:252-300,312-373 and the
// robosuite 1.3.2 helpers it calls (Controller.scale_action, control_utils.set_goal_position, Controller.clip_torques).
#pragma once
#include "state.cuh"

namespace hrg {

// scale_action: action = clip(action, input_min, input_max);
// (action - (input_max + input_min) / 2) * (|output_max - output_min| / |input_max - input_min|) + (output_max + output_min) / 2
HRG_HD double ctrl_scale_action(const HrgControllerConfig& k, int j, double a) {
  const double scale = fabs(k.output_max[j] - k.output_min[j]) / fabs(k.input_max[j] - k.input_min[j]);
  const double out_t = (k.output_max[j] + k.output_min[j]) / 2.0, in_t = (k.input_max[j] + k.input_min[j]) / 2.0;
  a = fmin(fmax(a, k.input_min[j]), k.input_max[j]);  // np.clip = minimum(maximum(a, lo), hi)
  return (a - in_t) * scale + out_t;
}

// set_goal(): goal per env-major [N,J]; also performs SafetyShield::newLongTermTrajectory's bookkeeping (new_goal_env).
__global__ void k_ctrl_set_goal(DevState S, HrgControllerConfig k, const uint8_t* __restrict__ mask,
                                const double* __restrict__ action, const double* __restrict__ joint_pos,
                                const double* __restrict__ set_qpos, double* __restrict__ goal_out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= S.N) return;
  if (mask && !mask[e]) return;
  const int J = S.J;
  double g[HRG_MAX_JOINTS];
  for (int j = 0; j < J; ++j) {
    const size_t o = (size_t)e * J + j;
    double goal = set_qpos ? set_qpos[o] : joint_pos[o] + ctrl_scale_action(k, j, action[o]);
    if (k.has_position_limits) goal = fmin(fmax(goal, k.position_min[j]), k.position_max[j]);
    g[j] = goal;
    if (goal_out) goal_out[o] = goal;
  }
  new_goal_env(S, *S.cfg, e, g);
}

// run_controller(): feedback = e * kp + de * kd; tau = M . (feedback + ddq_des) + torque_compensation; clip.
__global__ void k_ctrl_torques(int N, int J, HrgControllerConfig k, const double* __restrict__ q_des,
                               const double* __restrict__ dq_des, const double* __restrict__ ddq_des,
                               const double* __restrict__ joint_pos, const double* __restrict__ joint_vel,
                               const double* __restrict__ mass, const double* __restrict__ comp,
                               double* __restrict__ torques) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= N) return;
  double u[HRG_MAX_JOINTS];
  for (int j = 0; j < J; ++j) {
    const size_t o = (size_t)e * J + j;
    const double fb = (q_des[o] - joint_pos[o]) * k.kp[j] + (dq_des[o] - joint_vel[o]) * k.kd[j];
    u[j] = fb + ddq_des[o];
  }
  const double* M = mass + (size_t)e * J * J;
  for (int i = 0; i < J; ++i) {
    double acc = 0.0;
    for (int j = 0; j < J; ++j) acc += M[i * J + j] * u[j];
    const double tau = acc + comp[(size_t)e * J + i];
    torques[(size_t)e * J + i] = fmin(fmax(tau, k.torque_min[i]), k.torque_max[i]);
  }
}

}  // namespace hrg
