// Path-parameter (s) arithmetic of the shield: jerk-limited 3-phase paths, failsafe/recovery planning, LTT interpolation.
// Per-env scalar code, written as host/device inline functions so that the thread-per-env kernels are thin glue.
// SS = human_robot_gym/controllers/failsafe_controller/sara-shield/safety_shield (reference tree).
#pragma once
#include <math.h>
#include <stdint.h>

#include "../../include/hrg_shield.h"

#if defined(__CUDACC__)
#define HRG_HD __host__ __device__ __forceinline__
// Large, branchy routines that are called from several places are real calls on the device: force-inlining them blew
// k_replan up to 82k SASS instructions (1.3 MB) and 36 % of its stall samples were instruction-cache misses
// (profiles/r01_ncu_summary.md).
#define HRG_HD_CALL __host__ __device__ __noinline__
// Per-joint loops that only LOAD (or only compute) run over the compile-time bound with a `j < J` guard so that they
// unroll: the thread-per-env kernels are bound by the latency of dependent global loads, and an unrolled loop has all of
// its loads in flight at once instead of one round trip per joint.
#define HRG_UNROLL _Pragma("unroll")
#else
#define HRG_HD inline
#define HRG_HD_CALL inline
#define HRG_UNROLL
#endif

namespace hrg {

// The reference's Path (SS/include/safety_shield/path.h:30-60): ph[0..2] = end times of the three phases (relative to
// "now"; they go negative once elapsed), ph[3..5] = their jerks. `is_current_` lives in the env flag word.
struct PathS {
  double ph[6];
  double pos, vel, acc, jerk;
};

HRG_HD void path_clear(PathS& p) {
  for (int i = 0; i < 6; ++i) p.ph[i] = 0.0;
  p.pos = p.vel = p.acc = p.jerk = 0.0;
}

// Path::increment (SS/src/path.cc:15-39): the active jerk is that of the first phase whose end time is still > 1e-6;
// pos only advances while vel > -1e-6.
HRG_HD void path_increment(PathS& p, double ts) {
  const double eps = 0.000001;
  double jerk = 0.0;
  if (eps < p.ph[0])
    jerk = p.ph[3];
  else if (eps < p.ph[1])
    jerk = p.ph[4];
  else if (eps < p.ph[2])
    jerk = p.ph[5];
  p.jerk = jerk;
  if (p.vel > -eps) p.pos += p.vel * ts + p.acc * ts * ts / 2 + jerk * ts * ts * ts / 6;
  p.vel += p.acc * ts + jerk * ts * ts / 2;
  p.acc += jerk * ts;
  p.ph[0] -= ts;
  p.ph[1] -= ts;
  p.ph[2] -= ts;
}

// Path::getFinalMotion (SS/src/path.cc:41-53)
HRG_HD void path_final(const PathS& p, double& fp, double& fv, double& fa) {
  fp = p.pos;
  fv = p.vel;
  fa = p.acc;
  double l_time = 0.0;
  for (int i = 0; i < 3; ++i) {
    double dt = p.ph[i] - l_time;
    double j = p.ph[i + 3];
    fp += fv * dt + fa * dt * dt / 2 + j * dt * dt * dt / 6;
    fv += fa * dt + j * dt * dt / 2;
    fa += j * dt;
    l_time = p.ph[i];
  }
}

// Path::getMaxVelocity (SS/src/path.cc:118-139), PFL only
HRG_HD double path_max_velocity(const PathS& p) {
  double max_vel = p.vel, l_time = 0.0, vel = p.vel, acc = p.acc;
  for (int i = 0; i < 3; ++i) {
    double dt = p.ph[i] - l_time, j = p.ph[i + 3];
    if (acc > 0 && j < 0 && -acc / j < dt) {
      double mid_t = -acc / j;
      double mid_vel = vel + acc * mid_t + j * mid_t * mid_t / 2;
      max_vel = fmax(max_vel, mid_vel);
    }
    vel += acc * dt + j * dt * dt / 2;
    acc += j * dt;
    max_vel = fmax(max_vel, vel);
    l_time = p.ph[i];
  }
  return max_vel;
}

HRG_HD double round_to_timestep(double t, double ts) { return ceil(t / ts) * ts; }  // safety_shield.h:355-357

// SafetyShield::planSafetyShield (SS/src/safety_shield.cc:203-341). Mirrors the partial side effects: once past the
// first guard, pos/vel/acc of `path` are overwritten (and is_current cleared, by the caller via the return of
// `touched`) even if planning later fails; the phases only change on success.
HRG_HD_CALL bool plan_safety_shield(double pos, double vel, double acc, double ve, double a_max, double j_max, double ts,
                               PathS& path, bool& touched) {
  touched = false;
  if (a_max < 0 || fabs(acc) > a_max) return false;
  touched = true;
  path.pos = pos;
  path.vel = vel;
  path.acc = acc;
  const double epsilon = 1e-6;
  if (fabs(vel - ve) < epsilon && fabs(acc) < epsilon) {
    path.ph[0] = ts; path.ph[1] = ts; path.ph[2] = ts;
    path.ph[3] = 0.0; path.ph[4] = 0.0; path.ph[5] = 0.0;
    return true;
  }
  double t_to_a_0 = round_to_timestep(fabs(acc) / j_max, ts);
  double v_at_a_0 = vel + acc * t_to_a_0 / 2;
  if (fabs(v_at_a_0 - ve) < epsilon) {
    path.ph[0] = t_to_a_0; path.ph[1] = t_to_a_0; path.ph[2] = t_to_a_0;
    path.ph[3] = -acc / t_to_a_0; path.ph[4] = 0.0; path.ph[5] = 0.0;
    return true;
  }
  // The eight cases of :235-314 differ only in the sign of the a_max^2/acc^2 term of the trapezoid test:
  //   slow down (v_at_a_0 > ve + eps): acc > 0 -> +acc^2/2, else -acc^2/2 ; speed up: acc > 0 -> -acc^2/2, else +acc^2/2
  double a_peak, t01, t12, t23, v01, v12, v23;
  const double hsq = acc * acc / 2;
  if (v_at_a_0 > ve + epsilon) {
    double lhs = (acc > 0) ? vel - (a_max * a_max + hsq) / j_max : vel - (a_max * a_max - hsq) / j_max;
    if (lhs > ve) {
      a_peak = -a_max;
      t01 = round_to_timestep((acc - a_peak) / j_max, ts);
      t23 = round_to_timestep(-a_peak / j_max, ts);
      v01 = (acc + a_peak) * t01 / 2;
      v23 = a_peak * t23 / 2;
      t12 = round_to_timestep((ve - vel - v01 - v23) / a_peak, ts);
      v12 = a_peak * t12;
    } else {
      a_peak = -sqrt(fabs((vel - ve) * j_max + hsq));
      t01 = round_to_timestep((acc - a_peak) / j_max, ts);
      t23 = round_to_timestep(-a_peak / j_max, ts);
      v01 = (acc + a_peak) * t01 / 2;
      v23 = a_peak * t23 / 2;
      t12 = 0;
      v12 = 0;
    }
  } else {
    double lhs = (acc > 0) ? vel + (a_max * a_max - hsq) / j_max : vel + (a_max * a_max + hsq) / j_max;
    if (lhs < ve) {
      a_peak = a_max;
      t01 = round_to_timestep((a_peak - acc) / j_max, ts);
      t23 = round_to_timestep(a_peak / j_max, ts);
      v01 = (acc + a_peak) * t01 / 2;
      v23 = a_peak * t23 / 2;
      t12 = round_to_timestep((ve - vel - v01 - v23) / a_peak, ts);
      v12 = a_peak * t12;
    } else {
      a_peak = sqrt(fabs((ve - vel) * j_max + hsq));
      t01 = round_to_timestep((a_peak - acc) / j_max, ts);
      t23 = round_to_timestep(a_peak / j_max, ts);
      v01 = (acc + a_peak) * t01 / 2;
      v23 = a_peak * t23 / 2;
      t12 = 0;
      v12 = 0;
    }
  }
  double correct_int = ve - vel - v01 - v12 - v23;
  a_peak += correct_int / ((t01 + t23) / 2 + t12 + epsilon);
  double j01 = (t01 >= epsilon) ? (a_peak - acc) / t01 : 0.0;
  double j23 = (t23 >= epsilon) ? -a_peak / t23 : 0.0;
  if (t01 < 0 || t12 < 0 || t23 < 0) return false;
  path.ph[0] = t01;
  path.ph[1] = t01 + t12;
  path.ph[2] = t01 + t12 + t23;
  path.ph[3] = j01;
  path.ph[4] = 0.0;
  path.ph[5] = j23;
  return true;
}

// SafetyShield::calculateMaxAccJerk (SS/src/safety_shield.cc:343-365); the 1e-9 is added for joint 0 only.
HRG_HD void max_acc_jerk(const double* prev_speed, const HrgConfig& c, double& a_man, double& j_man) {
  const int J = c.nb_joints;
  double denom = fabs(prev_speed[0]) + fabs(c.a_max_ltt[0]) * c.max_s_stop + 1E-9;
  double min_c = (c.a_max[0] - fabs(c.a_max_ltt[0])) / denom;
  for (int i = 1; i < J; ++i) {
    denom = fabs(prev_speed[i]) + fabs(c.a_max_ltt[i]) * c.max_s_stop;
    double nc = (c.a_max[i] - fabs(c.a_max_ltt[i])) / denom;
    min_c = (nc < min_c) ? nc : min_c;
  }
  a_man = (min_c < 0) ? 0 : min_c;
  denom = fabs(prev_speed[0]) + fabs(c.a_max_ltt[0]) * c.max_s_stop + 1E-9;
  double min_d = (c.j_max[0] - 3 * fabs(c.a_max_ltt[0]) * a_man - fabs(c.j_max_ltt[0])) / denom;
  for (int i = 1; i < J; ++i) {
    denom = fabs(prev_speed[i]) + fabs(c.a_max_ltt[i]) * c.max_s_stop;
    double nd = (c.j_max[i] - 3 * fabs(c.a_max_ltt[i]) * a_man - fabs(c.j_max_ltt[i])) / denom;
    min_d = (nd < min_d) ? nd : min_d;
  }
  j_man = (min_d < 0) ? 0 : min_d;
}

HRG_HD double clampd(double v, double lo, double hi) { return (v < lo) ? lo : ((hi < v) ? hi : v); }  // std::clamp

// LongTermTraj::getTrajectoryIndex (SS/include/safety_shield/long_term_traj.h:220-228)
HRG_HD int ltt_index(int index, int starting_index, int length, int current_pos) {
  int d = index - starting_index;
  if (d > length - 1) d = length - 1;
  if (d < current_pos) d = current_pos;
  return d;
}

// Sample index + sub-sample offset of LongTermTraj::interpolate (SS/src/long_term_traj.cc:41-54):
// ind = s / Ts; ind1 = trunc(ind); dt = frac(ind) * Ts (applied even when the index gets clamped).
HRG_HD void ltt_locate(double s, double ts, int starting_index, int length, int current_pos, int& sample, double& dt) {
  double ind = s / ts;
  double intpart;
  double ind_mod = modf(ind, &intpart);
  sample = ltt_index(static_cast<int>(intpart), starting_index, length, current_pos);
  dt = ind_mod * ts;
}

// One joint of LongTermTraj::interpolate (SS/src/long_term_traj.cc:63-75). The jerk is not clamped.
HRG_HD void ltt_interp_joint(double q1, double dq1, double ddq1, double dddq1, double dt, double ds, double dds,
                             double ddds, double v_max, double a_max, double& q, double& dq, double& ddq,
                             double& dddq) {
  q = q1 + dt * dq1 + 1.0 / 2 * dt * dt * ddq1 + 1.0 / 6 * dt * dt * dt * dddq1;
  double v_max_int = dq1 + dt * ddq1 + 1.0 / 2 * dt * dt * dddq1;
  double v_int = v_max_int * ds;
  dq = clampd(v_int, -v_max, v_max);
  double a_max_int = ddq1 + dt * dddq1;
  double a_int = v_max_int * dds + ds * ds * a_max_int;
  ddq = clampd(a_int, -a_max, a_max);
  dddq = dddq1 * ds * ds * ds + 3.0 * a_max_int * dds * ds + v_max_int * ddds;
}

}  // namespace hrg
