// Sampling of the planned jerk profiles into a long-term trajectory (LongTermPlanner::getTrajectory,
// LTP/src/long_term_planner.cc:706-841) and the per-joint pieces of planTrajectory (:7-63).
#pragma once

namespace hrg {

struct JointPlan {
  double T[7];   // cumulative jerk switch times (t_scaled)
  double dir;    // direction of motion
  double v_drive;
  int mod;       // mod_jerk_profile
};

// checkInputs for one joint (:68-77)
HRG_HD bool ltp_check_inputs_joint(const LtpLimits& L, double q_min, double q_max, double q_0, double v_0, double a_0) {
  if (q_0 < q_min || q_0 > q_max || fabs(v_0) > L.v_max || fabs(a_0) > L.a_max) return false;
  if (fabs(v_0 + 0.5 * a_0 * fabs(a_0) / L.j_max) > L.v_max) return false;
  return true;
}

// The sampled jerk of one joint is a piecewise-constant fill plus up to 10 sub-sample corrections at fixed indices
// (:751-807). Keeping it in this closed form lets the sampler stream q,v,a,j out without a scratch jerk array.
struct JerkSampler {
  int s[7];        // switch times in samples: floor/ceil alternating (:751-757)
  double jp[7];    // jerk of each phase (:734-744)
  int nc;
  int cidx[10];
  double cval[10];
  bool phase4;
};

HRG_HD void ltp_sampler_init(const LtpLimits& L, const JointPlan& P, JerkSampler& S) {
  const double ts = L.t_sample;
  double fr[7];
  for (int k = 0; k < 7; ++k) {
    int prof = (P.mod == 1) ? ((k == 0 || k == 4) ? -1 : ((k == 2 || k == 6) ? 1 : 0))
                            : ((k == 0 || k == 6) ? 1 : ((k == 2 || k == 4) ? -1 : 0));
    S.jp[k] = P.dir * L.j_max * prof;
    fr[k] = P.T[k] - ts * floor(P.T[k] / ts);
    S.s[k] = (k & 1) ? (int)ceil(P.T[k] / ts) : (int)floor(P.T[k] / ts);
  }
  int n = 0;
#define HRG_CORR(I, V) \
  do {                 \
    S.cidx[n] = (I);   \
    S.cval[n] = (V);   \
    ++n;               \
  } while (0)
  if (S.s[2] >= S.s[1]) {  // phase 2 exists (:768-776)
    HRG_CORR(S.s[0] + 1, fr[0] / ts * S.jp[0]);
    if (S.s[1] > 0) HRG_CORR(S.s[1], (1 - fr[1] / ts) * S.jp[2]);
    HRG_CORR(S.s[2] + 1, fr[2] / ts * S.jp[2]);
  } else if (S.s[1] > 0) {  // (:777-783) two terms added one after the other
    HRG_CORR(S.s[1], fr[0] / ts * S.jp[0]);
    HRG_CORR(S.s[1], (fr[2] - fr[0]) / ts * S.jp[2]);
  }
  if (S.s[3] > 0) HRG_CORR(S.s[3], (1 - fr[3] / ts) * S.jp[4]);  // (:786-788)
  if (S.s[2] - S.s[0] > 0) {                                      // (:790-800)
    HRG_CORR(S.s[4] + 1, fr[4] / ts * S.jp[4]);
  } else if (S.s[4] > 0) {
    HRG_CORR(S.s[4], fr[4] / ts * S.jp[4]);
    HRG_CORR(S.s[4], fr[0] / ts * S.jp[0]);
    HRG_CORR(S.s[4], (fr[2] - fr[0]) / ts * S.jp[2]);
  }
  if (S.s[5] > 0) HRG_CORR(S.s[5], (1 - fr[5] / ts) * S.jp[6]);  // (:803-805)
  HRG_CORR(S.s[6] + 1, fr[6] / ts * S.jp[6]);                    // (:807)
#undef HRG_CORR
  S.nc = n;
  S.phase4 = S.s[3] - S.s[2] > 2;
}

// j_traj[i]: the std::fill sequence of :759-766 (a later phase overwrites an earlier one where the rounded switch
// samples overlap), then the corrections in program order.
HRG_HD double ltp_jerk_at(const JerkSampler& S, int i) {
  double j = 0.0;
  if (i < S.s[0]) j = S.jp[0];
  for (int k = 1; k < 7; ++k)
    if (i >= S.s[k - 1] && i < S.s[k]) j = S.jp[k];
  for (int k = 0; k < S.nc; ++k)
    if (S.cidx[k] == i) j = j + S.cval[k];
  return j;
}

// traj_len contribution of one joint (:716-719)
HRG_HD int ltp_joint_len(const LtpLimits& L, const JointPlan& P) { return (int)ceil(P.T[6] / L.t_sample) + 1; }

// Integrates one joint over `len` samples (:810-831) and streams the samples out: out[i*stride + {0,J,2J,3J}] =
// q, v, a, j. Returns the final position (for the joint-limit test of :59-61).
HRG_HD double ltp_sample_joint(const LtpLimits& L, const JointPlan& P, double q_0, double v_0, double a_0, int len,
                               double* out, int J, size_t stride) {
  JerkSampler S;
  ltp_sampler_init(L, P, S);
  const double ts = L.t_sample;
  double jj = ltp_jerk_at(S, 0);
  double a = a_0 + ts * jj;
  double v = v_0 + ts * a;
  double q = q_0 + ts * v;
  out[0] = q; out[J] = v; out[2 * J] = a; out[3 * J] = jj;
  const double v_cruise = P.v_drive * P.dir;
  for (int i = 1; i < len; ++i) {
    jj = ltp_jerk_at(S, i);
    if (i <= S.s[6])
      a = a + ts * jj;
    else
      a = 0.0;
    if (S.phase4 && i >= S.s[2] + 1 && i < S.s[3] - 1)
      v = v_cruise;
    else if (i <= S.s[6])
      v = v + ts * a;
    else
      v = 0.0;
    q = q + ts * v;
    double* o = out + (size_t)i * stride;
    o[0] = q; o[J] = v; o[2 * J] = a; o[3 * J] = jj;
  }
  return q;
}

#if defined(__CUDACC__)
// Warp-cooperative form of ltp_sample_joint for k_replan_sample: lane = joint gl + 8 * sub, every lane holds the plan
// of joint `row` (= gl, or J-1 for the shadow lanes gl >= J). The trajectory is produced in chunks of 32 samples:
//   1. all 32 lanes evaluate the piecewise-constant jerk fill (:759-766) of sample c0 + lane, joint after joint, into a
//      [8][32] shared buffer (the 7-way range test is the expensive part of ltp_jerk_at and has no serial dependence);
//   2. one lane per joint adds the sub-sample corrections (:767-807) that fall into the chunk, in program order;
//   3. every lane runs the reference's sequential recurrence a += ts*j, v += ts*a, q += ts*v over the chunk (identical
//      roundings); sub-lane s stores component s of {q, v, a, j}, so one store instruction writes a whole [4][J] sample.
// `len` must be warp-uniform. Returns the final position (joint-limit test of :59-61).
__device__ __forceinline__ double ltp_sample_joint_warp(const LtpLimits& L, const JointPlan& P, double q_0, double v_0,
                                                        double a_0, int len, double* traj /*[len][4][J]*/, int J, int gl,
                                                        int row, int sub, bool joint_lane,
                                                        JerkSampler* samp /*[8] shared*/, double* jbuf /*[8][32] shared*/) {
  const int lane = gl + 8 * sub;
  if (sub == 0) ltp_sampler_init(L, P, samp[gl]);
  __syncwarp();
  const JerkSampler& S = samp[row];
  const double ts = L.t_sample;
  const double v_cruise = P.v_drive * P.dir;
  const int s6 = S.s[6];
  // cruise window [cr_lo, cr_hi) as one unsigned compare; empty when phase 4 is short (:818)
  const int cr_lo = S.s[2] + 1;
  const unsigned cr_span = (S.phase4 && S.s[3] - 1 > cr_lo) ? (unsigned)(S.s[3] - 1 - cr_lo) : 0u;
  const bool p1 = sub == 1, p2 = sub == 2, p3 = sub == 3;
  // shadow lanes (gl >= J) carry joint J-1 and store the same values to the same address as its real lane
  unsigned off = (unsigned)(sub * J + row);
  const unsigned stride = 4u * J;
  const double* jb = jbuf + row * 32;
  double a = a_0, v = v_0, q = q_0;
  for (int c0 = 0; c0 < len; c0 += 32) {
    const int i_mine = c0 + lane;
    for (int jj = 0; jj < J; ++jj) {
      const JerkSampler& Q = samp[jj];
      double j = 0.0;
      if (i_mine < Q.s[0]) j = Q.jp[0];
#pragma unroll
      for (int k = 1; k < 7; ++k)
        if (i_mine >= Q.s[k - 1] && i_mine < Q.s[k]) j = Q.jp[k];
      jbuf[jj * 32 + lane] = j;
    }
    __syncwarp();
    if (sub == 0 && joint_lane)
      for (int k = 0; k < S.nc; ++k) {
        const int d = S.cidx[k] - c0;
        if (d >= 0 && d < 32) jbuf[gl * 32 + d] = jbuf[gl * 32 + d] + S.cval[k];
      }
    __syncwarp();
    const int n = (len - c0) < 32 ? (len - c0) : 32;
    int u = 0;
    if (c0 == 0) {  // the first sample integrates unconditionally (:810-813)
      const double jj = jb[0];
      a = a_0 + ts * jj;
      v = v_0 + ts * a;
      q = q_0 + ts * v;
      double val = q;
      val = p1 ? v : val;
      val = p2 ? a : val;
      val = p3 ? jj : val;
      traj[off] = val;
      off += stride;
      u = 1;
    }
#pragma unroll 4
    for (; u < n; ++u) {
      const int i = c0 + u;
      const double jj = jb[u];
      const bool live = i <= s6;
      const bool cruise = (unsigned)(i - cr_lo) < cr_span;
      const double an = a + ts * jj;
      a = live ? an : 0.0;
      double vn = v + ts * a;
      vn = live ? vn : 0.0;
      v = cruise ? v_cruise : vn;
      q = q + ts * v;
      double val = q;
      val = p1 ? v : val;
      val = p2 ? a : val;
      val = p3 ? jj : val;
      traj[off] = val;
      off += stride;
    }
    __syncwarp();
  }
  return q;
}
#endif

// Sequential planTrajectory for one env (:7-63) — used where one thread owns a whole env (host tests, stage entry);
// the replan kernel runs the same per-joint calls with one lane per joint. traj layout [len][4][J].
// Returns the trajectory length, 0 on failure, -len if len > cap (nothing is written then).
// `mat`: 36 doubles of root-finder scratch (shared memory on the device, optional on the host).
HRG_HD int ltp_plan_env_seq(const HrgConfig& c, const double* q_goal, const double* q_0, const double* v_0,
                            const double* a_0, double* traj, int cap, double* mat = nullptr) {
  const int J = c.nb_joints;
  JointPlan opt[HRG_MAX_JOINTS], sc[HRG_MAX_JOINTS];
  LtpLimits L[HRG_MAX_JOINTS];
  OptRootCache rc[HRG_MAX_JOINTS];  // the quartic roots of optSwitchTimes, shared by all its calls for a joint
  for (int j = 0; j < J; ++j) {
    L[j].v_max = c.v_max[j]; L[j].a_max = c.a_max_ltt[j]; L[j].j_max = c.j_max_ltt[j]; L[j].t_sample = c.sample_time;
    L[j].mat = mat;
    L[j].root_mode = c.root_mode;
    if (!ltp_check_inputs_joint(L[j], c.q_min[j], c.q_max[j], q_0[j], v_0[j], a_0[j])) return 0;
  }
#if defined(HRG_TRACE_PREFILL) && !defined(__CUDA_ARCH__)
  for (int j = 0; j < J; ++j)  // host-only instrumentation: what k_replan_plan's unconditional prefill solves
    for (int w = 0; w < 2; ++w) {
      OptRootCache tmp;
      opt_cache_clear(tmp);
      opt_prefill_root(L[j], q_goal[j], q_0[j], v_0[j], a_0[j], w, tmp);
    }
#endif
  for (int j = 0; j < J; ++j) {
    ltp_zero7(opt[j].T);
    opt_cache_clear(rc[j]);
    if (!ltp_opt_switch_times(L[j], q_goal[j], q_0[j], v_0[j], a_0[j], L[j].v_max, opt[j].T, opt[j].dir, opt[j].mod, &rc[j]))
      return 0;
  }
  double t_required = -1;
  int slowest = -1;
  for (int j = 0; j < J; ++j)
    if (opt[j].T[6] > t_required) {
      t_required = opt[j].T[6];
      slowest = j;
    }
  if (slowest == -1) return 0;
  int len = 0;
  for (int j = 0; j < J; ++j) {
    sc[j] = opt[j];
    sc[j].v_drive = L[j].v_max;
    ltp_zero7(sc[j].T);
    if (j != slowest)
      ltp_time_scaling(L[j], q_goal[j], q_0[j], v_0[j], a_0[j], opt[j].dir, t_required, sc[j].T, sc[j].v_drive,
                       sc[j].mod, &rc[j]);
    double mx = sc[j].T[0];
    for (int k = 1; k < 7; ++k) mx = (mx < sc[j].T[k]) ? sc[j].T[k] : mx;
    if (mx <= 0.0)
      for (int k = 0; k < 7; ++k) sc[j].T[k] = opt[j].T[k];
    int l = ltp_joint_len(L[j], sc[j]);
    len = l > len ? l : len;
  }
  if (len > cap) return -len;
  bool ok = true;
  for (int j = 0; j < J; ++j) {
    double qf = ltp_sample_joint(L[j], sc[j], q_0[j], v_0[j], a_0[j], len, traj + j, J, (size_t)4 * J);
    if (qf < c.q_min[j] || qf > c.q_max[j]) ok = false;
  }
  return ok ? len : 0;
}

}  // namespace hrg
