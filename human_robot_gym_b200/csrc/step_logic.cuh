// Per-env control logic of SafetyShield::step (SS/src/safety_shield.cc:555-647), split at the two points where the
// tick hands over to another kernel: (1) before LongTermPlanner replanning, (2) before reach + verify.
// Each function is the body of a thread-per-env kernel.
#pragma once
#include "state.cuh"

namespace hrg {

// newLongTermTrajectory (SS/src/safety_shield.cc:672-697)
HRG_HD void new_goal_env(const DevState& S, const HrgConfig& c, int e, const double* goal_q /*[J] of this env*/) {
  for (int j = 0; j < S.J; ++j) {
    S.goal_q[(size_t)j * S.N + e] = clampd(goal_q[j], c.q_min[j], c.q_max[j]);
    S.replan_q[(size_t)j * S.N + e] = -1234567;
  }
  uint32_t f = S.flags[e];
  f |= F_NEW_GOAL;
  f &= ~(F_NEW_LTT | F_NEW_LTT_PROCESSED);
  S.flags[e] = f;
}

// computesPotentialTrajectory(is_safe_, next_motion_.getVelocity()) (:367-456). `f` is the env's flag word (updated in
// place, caller stores it). Outputs the goal motion's angles, its path position s_d, the slot it was interpolated from
// and t_brake (= potential_path_.getPhase(3)).
HRG_HD void potential_env(const DevState& S, const HrgConfig& c, int e, uint32_t& f, double* gq, double& s_goal,
                          double& t_brake_out, int& gslot_out) {
  const int J = S.J, N = S.N;
  const double ts = c.sample_time;
  const int live = (f & F_LIVE_SLOT) ? 1 : 0, pend = 1 - live;
  const uint32_t type = (f & F_TYPE_MASK) >> F_TYPE_SHIFT;
  const bool v = (f & F_IS_SAFE) != 0;
  // :370-376 advance the LTT cursors to the travelled path position
  double path_s = S.path_s[e];
  int sd = S.path_sd[e];
  if (path_s >= (sd + 1) * ts) {
    int lc = S.ltt_cur[(size_t)live * N + e], ll = S.ltt_len[(size_t)live * N + e];
    int pc = S.ltt_cur[(size_t)pend * N + e], pl = S.ltt_len[(size_t)pend * N + e];
    while (path_s >= (sd + 1) * ts) {
      lc = (lc + 1 < ll - 1) ? lc + 1 : ll - 1;
      if (f & F_NEW_LTT) pc = (pc + 1 < pl - 1) ? pc + 1 : pl - 1;
      sd++;
    }
    S.ltt_cur[(size_t)live * N + e] = lc;
    if (f & F_NEW_LTT) S.ltt_cur[(size_t)pend * N + e] = pc;
    S.path_sd[e] = sd;
  }
  PathS rec, fs, fs2;
  path_load(S, P_REC, e, rec);
  path_load(S, P_FS, e, fs);
  path_load(S, P_FS2, e, fs2);
  // :378-388
  if (v && (f & F_REC_CORRECT)) {
    f |= F_REC_CURRENT;
    fs = fs2;
  } else {
    f &= ~F_REC_CURRENT;
    path_increment(fs, ts);
  }
  // :391-398
  double prev[HRG_MAX_JOINTS];
  HRG_UNROLL
  for (int j = 0; j < HRG_MAX_JOINTS; ++j)
    if (j < J) prev[j] = S.prev_dq[(size_t)j * N + e];
  double a_man, j_man;
  max_acc_jerk(prev, c, a_man, j_man);
  // :402-407 plan a recovery path unless already on it
  bool touched;
  if (!(f & F_REC_CURRENT)) {
    bool ok = plan_safety_shield(fs.pos, fs.vel, fs.acc, 1, a_man, j_man, ts, rec, touched);
    f = ok ? (f | F_REC_CORRECT) : (f & ~F_REC_CORRECT);
  }
  // :409-426
  if (f & F_REC_CORRECT) {
    path_increment(rec, ts);
    bool ok2;
    if (type == HRG_SHIELD_PFL) {
      // planPFLFailsafe (:458-500) needs the per-sample Cartesian velocities of the LTT
      ok2 = plan_safety_shield(rec.pos, rec.vel, rec.acc, 0.0, a_man, j_man, ts, fs2, touched);
      if (ok2) {
        double s_d, ds_d, dds_d;
        path_final(fs2, s_d, ds_d, dds_d);
        const int slot = (f & F_NEW_LTT) ? pend : live;
        // LongTermTraj::getMaxofMaximumCartesianVelocityWithS (SS/src/long_term_traj.cc:165-176)
        int i = S.ltt_cur[(size_t)slot * N + e];
        const int len = S.ltt_len[(size_t)slot * N + e];
        const double* cv = S.ltt_cart + ((size_t)e * 2 + slot) * S.L;
        const double* tm = S.ltt_time + ((size_t)e * 2 + slot) * S.L;
        double vmax = cv[i];
        while (i < len) {
          vmax = fmax(vmax, cv[i]);
          if (tm[i] > s_d) break;
          ++i;
        }
        const bool under_iso = vmax <= c.v_safe;
        double v_limit = under_iso ? rec.vel : c.v_safe / vmax;
        ok2 = plan_safety_shield(rec.pos, rec.vel, rec.acc, v_limit, a_man, j_man, ts, fs2, touched);
        bool under = under_iso ? true : (path_max_velocity(fs2) < v_limit);
        f = under ? (f | F_UNDER_V_LIMIT) : (f & ~F_UNDER_V_LIMIT);
      }
    } else {
      ok2 = plan_safety_shield(rec.pos, rec.vel, rec.acc, 0, a_man, j_man, ts, fs2, touched);
    }
    if (!ok2 || rec.pos < fs.pos) f &= ~F_REC_CORRECT;
  }
  path_store(S, P_REC, e, rec);
  path_store(S, P_FS, e, fs);
  path_store(S, P_FS2, e, fs2);
  // :428-449 goal of the potential path
  const PathS& pot = (f & F_REC_CORRECT) ? fs2 : fs;
  double s_d, ds_d, dds_d;
  path_final(pot, s_d, ds_d, dds_d);
  const double ddds_d = fs.jerk;
  double gdq[HRG_MAX_JOINTS], gddq[HRG_MAX_JOINTS], gdddq[HRG_MAX_JOINTS];
  const int gslot = (f & F_NEW_LTT) ? pend : live;
  ltt_interpolate(S, c, e, gslot, s_d, ds_d, dds_d, ddds_d, gq, gdq, gddq, gdddq);
  s_goal = s_d;
  t_brake_out = pot.ph[2];
  gslot_out = gslot;
}

// Middle part of step() (:606-637): potential trajectory, joint-limit check, hand-over to reach + verify.
// `last_meas` = HumanReach::last_meas_timestep_ (S.last_meas[e], or - inside a tick that measures - the tick's own
// measurement time, so that this part does not have to wait for k_measure).
HRG_HD void plan_env(const DevState& S, const HrgConfig& c, int e, double t, uint32_t& f, double last_meas) {
  const int J = S.J, N = S.N;
  const uint32_t type = (f & F_TYPE_MASK) >> F_TYPE_SHIFT;
  // :612-614
  if ((f & F_NEW_LTT) && !(f & F_NEW_LTT_PROCESSED)) f &= ~F_IS_SAFE;
  double gq[HRG_MAX_JOINTS], s_d, t_brake;
  int gslot;
  potential_env(S, c, e, f, gq, s_d, t_brake, gslot);
  // :617-637
  f &= ~F_VERIFY;
  if (type != HRG_SHIELD_OFF) {
    bool in_limits = true;
    for (int j = 0; j < J; ++j)
      if (gq[j] < c.q_min[j] || gq[j] > c.q_max[j]) in_limits = false;
    if (!in_limits) {
      f &= ~F_IS_SAFE;
      f |= (uint32_t)HRG_ST_JOINT_LIMIT << F_STATUS_SHIFT;
    } else {
      f |= F_VERIFY;
      double cq[HRG_MAX_JOINTS], al[HRG_MAX_JOINTS];  // all loads first: a store in between would serialise them
      HRG_UNROLL
      for (int j = 0; j < HRG_MAX_JOINTS; ++j)
        if (j < J) {
          cq[j] = cur_at(S, j, e);
          al[j] = S.ltt_alpha[((size_t)gslot * J + j) * N + e];  // getAlphaI() (:606-610)
        }
      const double cur_s = cur_at(S, 4 * J, e), last = last_meas;
      HRG_UNROLL
      for (int j = 0; j < HRG_MAX_JOINTS; ++j)
        if (j < J) {
          ri_at(S, j, e) = cq[j];
          ri_at(S, J + j, e) = gq[j];
          ri_at(S, 2 * J + j, e) = al[j];
        }
      ri_at(S, 3 * J, e) = s_d - cur_s;             // goal_motion.getS() - current_motion.getS()
      ri_at(S, 3 * J + 1, e) = t - last + t_brake;  // human_reach.cc:79
    }
  } else {
    f |= F_IS_SAFE;
  }
}

// First part of step() (:556-605). Returns true if the env must wait for the LongTermPlanner (k_replan).
HRG_HD bool pre_env(const DevState& S, const HrgConfig& c, int e, uint32_t& f) {
  const int J = S.J, N = S.N;
  f &= ~(0xFFFFu << F_STATUS_SHIFT);
  f &= ~F_PENDING_REPLAN;
  int live = (f & F_LIVE_SLOT) ? 1 : 0;
  // getCurrentMotion (:649-661): always on the OLD long-term trajectory
  PathS p;
  path_load(S, (f & F_REC_CURRENT) ? P_REC : P_FS, e, p);
  double cur[4 * HRG_MAX_JOINTS];
  ltt_interpolate(S, c, e, live, p.pos, p.vel, p.acc, p.jerk, cur, cur + J, cur + 2 * J, cur + 3 * J);
  for (int k = 0; k < 4 * J; ++k) cur_at(S, k, e) = cur[k];
  cur_at(S, 4 * J, e) = p.pos;
  // :562-568 commit the pending LTT
  if ((f & F_NEW_LTT) && (f & F_NEW_LTT_PROCESSED)) {
    bool stopped = true;  // Motion::isStopped (SS/include/safety_shield/motion.h:142-149)
    for (int j = 0; j < J; ++j)
      if (fabs(cur[J + j]) > 1e-4 || fabs(cur[2 * J + j]) > 1e-4 || fabs(cur[3 * J + j]) > 1e-4) stopped = false;
    if ((f & F_IS_SAFE) || stopped) {
      f ^= F_LIVE_SLOT;
      f &= ~(F_NEW_LTT | F_NEW_GOAL);
    }
  }
  // :570-605
  if (f & F_NEW_GOAL) {
    bool plannable = true;
    for (int j = 0; j < J; ++j)
      if (fabs(cur[2 * J + j]) > c.a_max_ltt[j]) plannable = false;
    if (plannable) {
      bool last_close = true;
      if (f & F_NEW_LTT) {
        for (int j = 0; j < J; ++j)
          if (fabs(cur[j] - S.replan_q[(size_t)j * N + e]) > 0.01) {
            last_close = false;
            break;
          }
      } else {
        last_close = false;
      }
      if (!last_close) {
        f |= F_PENDING_REPLAN;
        return true;
      }
      f |= F_NEW_LTT;
      f &= ~F_NEW_LTT_PROCESSED;
    } else {
      f &= ~F_NEW_LTT;
      f |= (uint32_t)HRG_ST_NOT_PLANNABLE << F_STATUS_SHIFT;
    }
  }
  return false;
}

// determineNextMotion (:511-553); the caller sets new_ltt_processed_ (:641). Writes the setpoints (env-major [N][J], any pointer may be NULL).
HRG_HD void select_env(const DevState& S, const HrgConfig& c, int e, uint32_t& f, double* q_out, double* dq_out,
                       double* ddq_out, double* dddq_out) {
  const int J = S.J, N = S.N;
  const double ts = c.sample_time;
  const int live = (f & F_LIVE_SLOT) ? 1 : 0;
  const uint32_t type = (f & F_TYPE_MASK) >> F_TYPE_SHIFT;
  if (type == HRG_SHIELD_PFL && (f & F_VERIFY) && (f & F_UNDER_V_LIMIT)) f |= F_IS_SAFE;  // :631-633
  double s_d, ds_d, dds_d, ddds_d;
  int slot = live, branch;
  if (f & F_IS_SAFE) {
    PathS rec, fs;
    path_load(S, P_REC, e, rec);
    path_load(S, P_FS, e, fs);
    PathS pot;
    path_load(S, (f & F_REC_CORRECT) ? P_FS2 : P_FS, e, pot);
    if (rec.pos >= fs.pos) {
      s_d = rec.pos; ds_d = rec.vel; dds_d = rec.acc; ddds_d = rec.jerk;
      branch = 0;
    } else {
      path_increment(pot, ts);
      s_d = pot.pos; ds_d = pot.vel; dds_d = pot.acc; ddds_d = pot.jerk;
      branch = 1;
    }
    if (f & F_NEW_LTT) slot = 1 - live;
    path_store(S, P_SAFE, e, pot);
  } else {
    PathS safe;
    path_load(S, P_SAFE, e, safe);
    path_increment(safe, ts);
    s_d = safe.pos; ds_d = safe.vel; dds_d = safe.acc; ddds_d = safe.jerk;
    path_store(S, P_SAFE, e, safe);
    branch = 2;
  }
  S.path_s[e] = s_d;
  double q[HRG_MAX_JOINTS], dq[HRG_MAX_JOINTS], ddq[HRG_MAX_JOINTS], dddq[HRG_MAX_JOINTS];
  ltt_interpolate(S, c, e, slot, s_d, ds_d, dds_d, ddds_d, q, dq, ddq, dddq);
  for (int j = 0; j < J; ++j) {
    S.prev_dq[(size_t)j * N + e] = dq[j];
    if (q_out) q_out[(size_t)e * J + j] = q[j];
    if (dq_out) dq_out[(size_t)e * J + j] = dq[j];
    if (ddq_out) ddq_out[(size_t)e * J + j] = ddq[j];
    if (dddq_out) dddq_out[(size_t)e * J + j] = dddq[j];
  }
  f = (f & ~F_BRANCH_MASK) | ((uint32_t)branch << F_BRANCH_SHIFT);
}

}  // namespace hrg
