// C-ABI batch driver around the UNMODIFIED reference sara-shield (TEST INFRASTRUCTURE, not product code).
//
// Compiled by oracle/Makefile together with the reference sources where they lie under /root/reference (nothing is
// copied into this repository) into oracle/_ref/libsara_ref.so. It holds N independent reference `SafetyShield`
// objects and drives them with the loop shape of SS/src/debug.cc:38-55
// (humanMeasurement -> [newLongTermTrajectory] -> step), dumping everything the parity tests compare.
// Built with -fno-access-control so private members of the reference classes can be read without touching its sources.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "long_term_planner/long_term_planner.h"
#include "reach_lib.hpp"
#include "safety_shield/long_term_traj.h"
#include "safety_shield/path.h"
#include "safety_shield/safety_shield.h"
#include "spdlog/spdlog.h"

using safety_shield::Motion;
using safety_shield::Path;
using safety_shield::SafetyShield;
using safety_shield::ShieldType;

namespace {
struct RefBatch {
  int n = 0, J = 0, K = 0, threads = 1;
  std::vector<std::unique_ptr<SafetyShield>> env;
};

template <class F>
void par_for(const RefBatch* b, F f) {
  int T = b->threads < 1 ? 1 : b->threads;
  if (T > b->n) T = b->n;
  if (T <= 1) {
    for (int i = 0; i < b->n; ++i) f(i);
    return;
  }
  std::vector<std::thread> th;
  for (int t = 0; t < T; ++t) {
    th.emplace_back([=]() {
      int lo = static_cast<int>(static_cast<int64_t>(b->n) * t / T);
      int hi = static_cast<int>(static_cast<int64_t>(b->n) * (t + 1) / T);
      for (int i = lo; i < hi; ++i) f(i);
    });
  }
  for (auto& x : th) x.join();
}

void put_capsule(const reach_lib::Capsule& c, double* o) {
  o[0] = c.p1_.x; o[1] = c.p1_.y; o[2] = c.p1_.z;
  o[3] = c.p2_.x; o[4] = c.p2_.y; o[5] = c.p2_.z;
  o[6] = c.r_;
}
reach_lib::Capsule get_capsule(const double* c) {
  return reach_lib::Capsule(reach_lib::Point(c[0], c[1], c[2]), reach_lib::Point(c[3], c[4], c[5]), c[6]);
}
void put_path(Path& p, double* o) {  // 11 doubles: phases[6], pos, vel, acc, jerk, is_current
  for (int i = 0; i < 6; ++i) o[i] = p.phases_[i];
  o[6] = p.pos_; o[7] = p.vel_; o[8] = p.acc_; o[9] = p.jerk_; o[10] = p.is_current_ ? 1.0 : 0.0;
}
}  // namespace

extern "C" {

// ---------------------------------------------------------------------------------------------------- batch of shields
// types != NULL: one ShieldType per env (mixed OFF / SSM / PFL batches), else `shield_type` for all.
void* ref_create_types(int n, double sample_time, const char* traj_yaml, const char* robot_yaml, const char* mocap_yaml,
                       const double* base_pose /*[n,6]*/, const double* init_q /*[n,J]*/, int J, int shield_type,
                       const int32_t* types /*[n] or NULL*/, int threads) {
  spdlog::set_level(spdlog::level::off);
  RefBatch* b = new RefBatch();
  b->n = n;
  b->J = J;
  b->threads = threads;
  b->env.resize(n);
  for (int i = 0; i < n; ++i) {
    std::vector<double> q(init_q + static_cast<size_t>(i) * J, init_q + static_cast<size_t>(i + 1) * J);
    const double* p = base_pose + static_cast<size_t>(i) * 6;
    b->env[i].reset(new SafetyShield(sample_time, traj_yaml, robot_yaml, mocap_yaml, p[0], p[1], p[2], p[3], p[4], p[5], q,
                                     static_cast<ShieldType>(types ? types[i] : shield_type)));
  }
  b->K = static_cast<int>(b->env[0]->human_reach_->n_joints_meas_);
  return b;
}
void* ref_create(int n, double sample_time, const char* traj_yaml, const char* robot_yaml, const char* mocap_yaml,
                 const double* base_pose, const double* init_q, int J, int shield_type, int threads) {
  return ref_create_types(n, sample_time, traj_yaml, robot_yaml, mocap_yaml, base_pose, init_q, J, shield_type, nullptr,
                          threads);
}
void ref_destroy(void* h) { delete static_cast<RefBatch*>(h); }
void ref_set_threads(void* h, int threads) { static_cast<RefBatch*>(h)->threads = threads; }
int ref_num_keypoints(void* h) { return static_cast<RefBatch*>(h)->K; }
int ref_num_joints(void* h) { return static_cast<RefBatch*>(h)->env[0]->nb_joints_; }

void ref_reset_types(void* h, const uint8_t* mask, const double* base_pose, const double* init_q, const double* t,
                     int shield_type, const int32_t* types /*[n] or NULL*/) {
  RefBatch* b = static_cast<RefBatch*>(h);
  par_for(b, [=](int i) {
    if (mask && !mask[i]) return;
    const double* p = base_pose + static_cast<size_t>(i) * 6;
    std::vector<double> q(init_q + static_cast<size_t>(i) * b->J, init_q + static_cast<size_t>(i + 1) * b->J);
    b->env[i]->reset(p[0], p[1], p[2], p[3], p[4], p[5], q, t[i], static_cast<ShieldType>(types ? types[i] : shield_type));
  });
}
void ref_reset(void* h, const uint8_t* mask, const double* base_pose, const double* init_q, const double* t,
               int shield_type) {
  ref_reset_types(h, mask, base_pose, init_q, t, shield_type, nullptr);
}

void ref_human_measurement(void* h, const uint8_t* mask, const double* kp /*[n,K,3]*/, const double* t /*[n]*/) {
  RefBatch* b = static_cast<RefBatch*>(h);
  par_for(b, [=](int i) {
    if (mask && !mask[i]) return;
    std::vector<reach_lib::Point> pts(b->K);
    const double* p = kp + static_cast<size_t>(i) * b->K * 3;
    for (int k = 0; k < b->K; ++k) pts[k] = reach_lib::Point(p[3 * k], p[3 * k + 1], p[3 * k + 2]);
    b->env[i]->humanMeasurement(pts, t[i]);
  });
}

void ref_new_goal(void* h, const uint8_t* mask, const double* goal_q, const double* goal_dq) {
  RefBatch* b = static_cast<RefBatch*>(h);
  par_for(b, [=](int i) {
    if (mask && !mask[i]) return;
    std::vector<double> q(goal_q + static_cast<size_t>(i) * b->J, goal_q + static_cast<size_t>(i + 1) * b->J);
    std::vector<double> dq(goal_dq + static_cast<size_t>(i) * b->J, goal_dq + static_cast<size_t>(i + 1) * b->J);
    b->env[i]->newLongTermTrajectory(q, dq);
  });
}

// One `step` for every env. Any output pointer may be NULL. branch: 0 recovery / 1 potential / 2 failsafe
// (SS/src/safety_shield.cc:511-553); the paths that decide it are not modified by determineNextMotion, so the branch can
// be read back after the call.
void ref_step(void* h, const double* t, double* q, double* dq, double* ddq, double* dddq, uint8_t* is_safe,
              uint8_t* branch) {
  RefBatch* b = static_cast<RefBatch*>(h);
  const int J = b->J;
  par_for(b, [=](int i) {
    SafetyShield& s = *b->env[i];
    Motion m = s.step(t[i]);
    const size_t o = static_cast<size_t>(i) * J;
    for (int j = 0; j < J; ++j) {
      if (q) q[o + j] = m.q_[j];
      if (dq) dq[o + j] = m.dq_[j];
      if (ddq) ddq[o + j] = m.ddq_[j];
      if (dddq) dddq[o + j] = m.dddq_[j];
    }
    if (is_safe) is_safe[i] = s.is_safe_ ? 1 : 0;
    if (branch) branch[i] = s.is_safe_ ? (s.recovery_path_.pos_ >= s.failsafe_path_.pos_ ? 0 : 1) : 2;
  });
}

// Capsules of the last step + per-model minimum capsule distance (min over robot x human of capsule_capsule_dist).
void ref_get_capsules(void* h, double* robot /*[n,J,7]*/, double* human /*[n,H,7]*/, int H, double* min_dist /*[n,3]*/) {
  RefBatch* b = static_cast<RefBatch*>(h);
  const int J = b->J;
  par_for(b, [=](int i) {
    SafetyShield& s = *b->env[i];
    if (robot)
      for (size_t j = 0; j < s.robot_capsules_.size() && j < static_cast<size_t>(J); ++j)
        put_capsule(s.robot_capsules_[j], robot + (static_cast<size_t>(i) * J + j) * 7);
    int k = 0;
    for (size_t m = 0; m < s.human_capsules_.size(); ++m) {
      double md = INFINITY;
      for (auto& hc : s.human_capsules_[m]) {
        if (human && k < H) put_capsule(hc, human + (static_cast<size_t>(i) * H + k) * 7);
        ++k;
        for (auto& rc : s.robot_capsules_) {
          double d = reach_lib::intersections::capsule_capsule_dist(rc, hc);
          if (d < md) md = d;
        }
      }
      if (min_dist && m < 3) min_dist[static_cast<size_t>(i) * 3 + m] = md;
    }
  });
}

// State dump for debugging divergences: 5 paths x 11 doubles + 8 scalars.
void ref_get_state(void* h, int i, double* out /*[63]*/) {
  SafetyShield& s = *static_cast<RefBatch*>(h)->env[i];
  put_path(s.recovery_path_, out);
  put_path(s.failsafe_path_, out + 11);
  put_path(s.failsafe_path_2_, out + 22);
  put_path(s.safe_path_, out + 33);
  put_path(s.potential_path_, out + 44);
  out[55] = s.path_s_;
  out[56] = s.path_s_discrete_;
  out[57] = s.new_ltt_;
  out[58] = s.new_goal_;
  out[59] = s.new_ltt_processed_;
  out[60] = s.recovery_path_correct_;
  out[61] = s.long_term_trajectory_.getLength();
  out[62] = s.new_long_term_trajectory_.getLength();
}

// recovery / failsafe / failsafe_2 / safe / potential path of every env (11 doubles each, see put_path) - lets a test
// scenario time its human input on the reference's own braking state (tests/parity_util.py AdaptivePulseScenario).
void ref_get_paths(void* h, double* out /*[n,5,11]*/) {
  RefBatch* b = static_cast<RefBatch*>(h);
  for (int i = 0; i < b->n; ++i) {
    SafetyShield& s = *b->env[i];
    double* o = out + static_cast<size_t>(i) * 55;
    put_path(s.recovery_path_, o);
    put_path(s.failsafe_path_, o + 11);
    put_path(s.failsafe_path_2_, o + 22);
    put_path(s.safe_path_, o + 33);
    put_path(s.potential_path_, o + 44);
  }
}

// ------------------------------------------------------------------------------------------- per-stage entry points
void ref_capsule_dist(int n, const double* c1 /*[n,7]*/, const double* c2 /*[n,7]*/, double* dist /*[n]*/) {
  for (int i = 0; i < n; ++i)
    dist[i] = reach_lib::intersections::capsule_capsule_dist(get_capsule(c1 + 7 * i), get_capsule(c2 + 7 * i));
}

// planSafetyShield on shield 0 of the batch (SS/src/safety_shield.cc:203-341) + Path::getFinalMotion.
void ref_plan_safety_shield(void* h, int n, const double* in /*[n,6]: pos vel acc ve a_max j_max*/,
                            uint8_t* ok /*[n]*/, double* phases /*[n,6]*/, double* fin /*[n,3]*/) {
  SafetyShield& s = *static_cast<RefBatch*>(h)->env[0];
  for (int i = 0; i < n; ++i) {
    const double* a = in + 6 * i;
    Path p;
    ok[i] = s.planSafetyShield(a[0], a[1], a[2], a[3], a[4], a[5], p) ? 1 : 0;
    for (int k = 0; k < 6; ++k) phases[6 * i + k] = p.phases_[k];
    p.getFinalMotion(fin[3 * i], fin[3 * i + 1], fin[3 * i + 2]);
  }
}

// LongTermPlanner::planTrajectory (LTP/src/long_term_planner.cc:7-63). Returns length (0 on failure);
// traj layout [L][4][J] (q, v, a, j per sample), truncated to max_len samples.
int ref_ltp_plan(int J, double ts, const double* q_min, const double* q_max, const double* v_max, const double* a_max,
                 const double* j_max, const double* q_goal, const double* q0, const double* v0, const double* a0,
                 double* traj, int max_len) {
  auto vec = [J](const double* p) { return std::vector<double>(p, p + J); };
  long_term_planner::LongTermPlanner ltp(J, ts, vec(q_min), vec(q_max), vec(v_max), vec(a_max), vec(j_max));
  long_term_planner::Trajectory tr;
  bool ok = ltp.planTrajectory(vec(q_goal), vec(q0), vec(v0), vec(a0), tr);
  if (!ok) return 0;
  int L = tr.length < max_len ? tr.length : max_len;
  for (int i = 0; i < L; ++i)
    for (int j = 0; j < J; ++j) {
      double* o = traj + (static_cast<size_t>(i) * 4) * J;
      o[0 * J + j] = tr.q[j][i];
      o[1 * J + j] = tr.v[j][i];
      o[2 * J + j] = tr.a[j][i];
      o[3 * J + j] = tr.j[j][i];
    }
  return tr.length;
}

// Single-joint pieces of the planner, as exposed by the reference's own test fixture
// (LTP/tests/include/long_term_planner_fixture.h:24-29). lim = {q_min, q_max, v_max, a_max, j_max}.
static long_term_planner::LongTermPlanner make_ltp1(double ts, const double* lim) {
  return long_term_planner::LongTermPlanner(1, ts, {lim[0]}, {lim[1]}, {lim[2]}, {lim[3]}, {lim[4]});
}
int ref_ltp_opt_braking(double ts, const double* lim, double v0, double a0, double* q, double* t_rel /*[7]*/, double* dir) {
  auto ltp = make_ltp1(ts, lim);
  std::array<double, 7> t{};
  bool ok = ltp.optBraking(0, v0, a0, *q, t, *dir);
  for (int i = 0; i < 7; ++i) t_rel[i] = t[i];
  return ok ? 1 : 0;
}
int ref_ltp_opt_switch_times(double ts, const double* lim, double q_goal, double q0, double v0, double a0, double v_drive,
                             double* t_out /*[7]*/, double* dir, int* mod) {
  auto ltp = make_ltp1(ts, lim);
  std::array<double, 7> t{};
  char m = 0;
  bool ok = ltp.optSwitchTimes(0, q_goal, q0, v0, a0, v_drive, t, *dir, m);
  for (int i = 0; i < 7; ++i) t_out[i] = t[i];
  *mod = m;
  return ok ? 1 : 0;
}
int ref_ltp_time_scaling(double ts, const double* lim, double q_goal, double q0, double v0, double a0, double dir,
                         double t_required, double* t_out /*[7]*/, double* v_drive, int* mod) {
  auto ltp = make_ltp1(ts, lim);
  std::array<double, 7> t{};
  char m = 0;
  bool ok = ltp.timeScaling(0, q_goal, q0, v0, a0, dir, t_required, t, *v_drive, m);
  for (int i = 0; i < 7; ++i) t_out[i] = t[i];
  *mod = m;
  return ok ? 1 : 0;
}

// LongTermTraj::interpolate (SS/src/long_term_traj.cc:41-77) on a caller-supplied trajectory [L][4][J].
void ref_ltt_interpolate(int J, int L, const double* traj, double sample_time, int starting_index, int n,
                         const double* sdd /*[n,4]: s ds dds ddds*/, const double* v_max, const double* a_max,
                         const double* j_max, double* out /*[n,4,J]*/) {
  std::vector<Motion> ms;
  for (int i = 0; i < L; ++i) {
    const double* p = traj + static_cast<size_t>(i) * 4 * J;
    ms.emplace_back(i * sample_time, std::vector<double>(p, p + J), std::vector<double>(p + J, p + 2 * J),
                    std::vector<double>(p + 2 * J, p + 3 * J), std::vector<double>(p + 3 * J, p + 4 * J));
  }
  safety_shield::LongTermTraj ltt(ms, sample_time, starting_index);
  std::vector<double> vm(v_max, v_max + J), am(a_max, a_max + J), jm(j_max, j_max + J);
  for (int k = 0; k < n; ++k) {
    Motion m = ltt.interpolate(sdd[4 * k], sdd[4 * k + 1], sdd[4 * k + 2], sdd[4 * k + 3], vm, am, jm);
    double* o = out + static_cast<size_t>(k) * 4 * J;
    for (int j = 0; j < J; ++j) {
      o[j] = m.q_[j]; o[J + j] = m.dq_[j]; o[2 * J + j] = m.ddq_[j]; o[3 * J + j] = m.dddq_[j];
    }
  }
}

// LongTermTraj::calculate_max_acc_jerk_window (SS/src/long_term_traj.cc:79-159) through the constructor, on a caller-supplied
// trajectory [L][4][J]: the full tables max_acceleration_window_ / max_jerk_window_ [L][J].
void ref_ltt_max_window(int J, int L, const double* traj, double sample_time, int k, double* out_acc, double* out_jerk) {
  std::vector<Motion> ms;
  for (int i = 0; i < L; ++i) {
    const double* p = traj + static_cast<size_t>(i) * 4 * J;
    ms.emplace_back(i * sample_time, std::vector<double>(p, p + J), std::vector<double>(p + J, p + 2 * J),
                    std::vector<double>(p + 2 * J, p + 3 * J), std::vector<double>(p + 3 * J, p + 4 * J));
  }
  safety_shield::LongTermTraj ltt(ms, sample_time, 0, k);
  for (int i = 0; i < L; ++i)
    for (int j = 0; j < J; ++j) {
      out_acc[static_cast<size_t>(i) * J + j] = ltt.max_acceleration_window_[i][j];
      out_jerk[static_cast<size_t>(i) * J + j] = ltt.max_jerk_window_[i][j];
    }
}

// getMaxAccelerationWindow(index) / getMaxJerkWindow(index) of the live LTT of every env (index = path_s_discrete_).
void ref_live_max_window(void* h, const int32_t* index /*[n] or NULL*/, double* out_acc /*[n,J]*/, double* out_jerk /*[n,J]*/) {
  RefBatch* B = static_cast<RefBatch*>(h);
  for (size_t i = 0; i < B->env.size(); ++i) {
    SafetyShield& s = *B->env[i];
    const int J = s.nb_joints_;
    const int idx = index ? index[i] : s.path_s_discrete_;
    std::vector<double> a = s.long_term_trajectory_.getMaxAccelerationWindow(idx);
    std::vector<double> jk = s.long_term_trajectory_.getMaxJerkWindow(idx);
    for (int j = 0; j < J; ++j) {
      out_acc[i * J + j] = a[j];
      out_jerk[i * J + j] = jk[j];
    }
  }
}

// HumanReach::humanReachabilityAnalysis (SS/src/human_reach.cc:76-87) of env i on its stored measurement; returns the
// number of capsules written as Pos, Vel, Accel (7 doubles each).
int ref_human_reach(void* h, int i, double t_command, double t_brake, double* out, int max_caps) {
  SafetyShield& s = *static_cast<RefBatch*>(h)->env[i];
  s.human_reach_->humanReachabilityAnalysis(t_command, t_brake);
  int k = 0;
  for (auto& set : s.human_reach_->getAllCapsules())
    for (auto& c : set)
      if (k < max_caps) put_capsule(c, out + 7 * k++);
  return k;
}

// RobotReach::reach (SS/src/robot_reach.cc:61-98) + HumanReach::humanReachabilityAnalysis (human_reach.cc:76-87) +
// VerifyISO::verify_human_reach (verify_iso.cc:19-33) on explicit inputs, one case after the other on the objects of env 0
// (its base pose, measurement buffers and capsules are overwritten): the stage hrg_stage_reach_verify mirrors.
void ref_stage_reach_verify(void* h, int n, const double* base_pose /*[n,6]*/, const double* q_start /*[n,J]*/,
                            const double* q_goal /*[n,J]*/, const double* s_diff /*[n]*/, const double* t_reach /*[n]*/,
                            const double* alpha_i /*[n,J]*/, const double* kp_pos /*[n,K,3]*/,
                            const double* kp_vel /*[n,K,3]*/, double* robot /*[n,J,7]*/, double* human /*[n,H,7]*/, int H,
                            double* min_dist /*[n,3]*/, uint8_t* is_safe /*[n]*/) {
  RefBatch* b = static_cast<RefBatch*>(h);
  SafetyShield& s = *b->env[0];
  const int J = b->J, K = b->K;
  for (int i = 0; i < n; ++i) {
    const double* p = base_pose + static_cast<size_t>(i) * 6;
    s.robot_reach_->reset(p[0], p[1], p[2], p[3], p[4], p[5]);
    std::vector<double> q1(q_start + static_cast<size_t>(i) * J, q_start + static_cast<size_t>(i + 1) * J);
    std::vector<double> q2(q_goal + static_cast<size_t>(i) * J, q_goal + static_cast<size_t>(i + 1) * J);
    std::vector<double> al(alpha_i + static_cast<size_t>(i) * J, alpha_i + static_cast<size_t>(i + 1) * J);
    Motion m1(0.0, q1), m2(0.0, q2);
    std::vector<reach_lib::Capsule> rc = s.robot_reach_->reach(m1, m2, s_diff[i], al);
    for (int k = 0; k < K; ++k) {
      const double* a = kp_pos + (static_cast<size_t>(i) * K + k) * 3;
      const double* v = kp_vel + (static_cast<size_t>(i) * K + k) * 3;
      s.human_reach_->joint_pos_[k] = reach_lib::Point(a[0], a[1], a[2]);
      s.human_reach_->joint_vel_[k] = reach_lib::Point(v[0], v[1], v[2]);
    }
    s.human_reach_->last_meas_timestep_ = 0.0;
    s.human_reach_->humanReachabilityAnalysis(0.0, t_reach[i]);
    std::vector<std::vector<reach_lib::Capsule>> hc = s.human_reach_->getAllCapsules();
    if (is_safe) is_safe[i] = s.verify_->verify_human_reach(rc, hc) ? 1 : 0;
    for (int j = 0; j < J && robot; ++j) put_capsule(rc[j], robot + (static_cast<size_t>(i) * J + j) * 7);
    int k = 0;
    for (size_t m = 0; m < hc.size(); ++m) {
      double md = INFINITY;
      for (auto& c : hc[m]) {
        if (human && k < H) put_capsule(c, human + (static_cast<size_t>(i) * H + k) * 7);
        ++k;
        for (auto& r : rc) md = std::min(md, reach_lib::intersections::capsule_capsule_dist(r, c));
      }
      if (min_dist && m < 3) min_dist[static_cast<size_t>(i) * 3 + m] = md;
    }
  }
}

// roots() + getSmallestPositiveNonComplexRoot (LTP/include/long_term_planner/roots.h:22-50)
double ref_poly_roots(int ncoef, const double* coef, double* re, double* im) {
  Eigen::VectorXd pv(ncoef);
  for (int i = 0; i < ncoef; ++i) pv(i) = coef[i];
  auto r = long_term_planner::roots<double>(pv);
  for (int i = 0; i < ncoef - 1; ++i) {
    re[i] = r(i, 0).real();
    im[i] = r(i, 0).imag();
  }
  return long_term_planner::getSmallestPositiveNonComplexRoot<double>(r);
}

}  // extern "C"
