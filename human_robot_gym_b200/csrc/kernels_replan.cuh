// calculateLongTermTrajectory (SS/src/safety_shield.cc:732-760) -> LongTermPlanner::planTrajectory
// (LTP/src/long_term_planner.cc:7-63) for the envs on the replan list: the replan chain of a tick.
//
// One PROBLEM PER LANE. The planner is branchy FP64 code, 85 % of it companion-matrix eigen-solves; the first design gave
// every (env, joint) problem an 8-lane group that shared each solve through a matrix in shared memory (four divergent
// groups per warp, ~10 of 32 lanes active, 0.34 ms per tick for ~2 100 replans). Three changes replace it:
//   * the solver keeps the matrix in registers (ltp_roots_reg.cuh), so a lane needs no partner and a warp runs 32 solves in
//     one instruction stream;
//   * optSwitchTimes' two quartic roots do not depend on v_drive (OptRootCache, ltp_profiles.cuh): they are solved once
//     per (env, joint) in k_replan_plan and handed to every later call, which leaves ONE eigen-solve per timeScaling
//     candidate instead of up to three;
//   * kernels follow the data flow of one env instead of one kernel per planner routine:
//       k_replan_plan    lane = (env, joint), floor(32 / J) envs per warp: checkInputs + optSwitchTimes (:20-37), slowest
//                        joint through shuffles over the env's lanes (:39-45), the two closed-form timeScaling candidates
//                        (:378-436); joints that accept neither go on the `deep` list
//       k_replan_deep    lane = (deep joint, polynomial candidate 2..7), tasks ordered by candidate so that a warp solves
//                        polynomials of one degree (:449-629); all six are evaluated, the first accepted one in the
//                        reference's order is taken later - what its sequential ladder returns
//       k_replan_sample  lane = (env, joint): getTrajectory (:706-841) with the reference's sequential recurrence per
//                        joint, joint-limit test (:59-61)
//       [k_pfl_ltt       PFL handles: per-sample Cartesian velocities (pfl_math.cuh)]
//       k_replan_tail    thread = env: install the trajectory and run the rest of the env's step up to the reach / verify
//                        hand-over (plan_env)
// Work is fetched from device counters (S.replan_ctr); grids only bound residency.
#pragma once
#include "ltp_math.cuh"
#include "step_logic.cuh"
#include "pfl_math.cuh"

namespace hrg {

enum { CTR_PLAN = 0, CTR_SAMPLE = 1, CTR_DEEP = 2, CTR_DEEP_COUNT = 3, N_CTR = 4 };
constexpr int kReplanBlock = 32;  // one warp per block: a tick has a few hundred busy warps; this spreads them over all SM sub-partitions

__device__ __forceinline__ LtpLimits joint_limits(const HrgConfig& c, int j) {
  LtpLimits L;
  L.v_max = c.v_max[j]; L.a_max = c.a_max_ltt[j]; L.j_max = c.j_max_ltt[j]; L.t_sample = c.sample_time;
  L.mat = nullptr;
  return L;
}

// The `mx <= 0` fallback of planTrajectory (:47-50) and the joint's sample count.
__device__ __forceinline__ void finish_scaled(const LtpLimits& L, const PlanRec& opt, const double* T, double v_drive,
                                              int mod, PlanRec& sc) {
  sc = opt;
  double mx = T[0];
  for (int k = 1; k < 7; ++k) mx = (mx < T[k]) ? T[k] : mx;
  if (!(mx <= 0.0))
    for (int k = 0; k < 7; ++k) sc.T[k] = T[k];
  sc.v_drive = v_drive;
  sc.mod = mod;
  JointPlan P;
  for (int k = 0; k < 7; ++k) P.T[k] = sc.T[k];
  sc.len = ltp_joint_len(L, P);
}

__device__ __forceinline__ void cache_from_rec(const PlanRec& r, double q_goal, double q_0, double v_0, double a_0,
                                               OptRootCache& rc) {
  rc.q_goal = q_goal; rc.q_0 = q_0; rc.v_0 = v_0; rc.a_0 = a_0;
  rc.root[0] = r.root[0]; rc.root[1] = r.root[1];
  rc.have = (r.have & 3) | 4;
}

// Fetches the next `per_warp` list entries for the warp; returns the first one, or -1 when the list is exhausted.
__device__ __forceinline__ int fetch_warp_tasks(int32_t* ctr, int per_warp, int ntask, int lane) {
  int base = 0;
  if (lane == 0) base = atomicAdd(ctr, per_warp);
  base = __shfl_sync(0xffffffffu, base, 0);
  return base < ntask ? base : -1;
}

__global__ void __launch_bounds__(kReplanBlock) k_replan_plan(DevState S) {
  const HrgConfig& c = *S.cfg;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int J = S.J, N = S.N;
  const int E = 32 / J;                          // envs per warp
  const int sub = lane / J;                      // env of this lane within the warp (>= E: spare lane)
  const int j = lane - sub * J;
  const int seg0 = (sub < E ? sub : 0) * J;      // first lane of this lane's env
  const int count = *S.replan_count;
  for (;;) {
    const int base = fetch_warp_tasks(S.replan_ctr + CTR_PLAN, E, count, lane);
    if (base < 0) break;
    const int idx = base + sub;
    const bool act = sub < E && idx < count;
    const LtpLimits L = joint_limits(c, act ? j : 0);
    PlanRec opt;
    ltp_zero7(opt.T);
    opt.dir = 0.0; opt.v_drive = L.v_max; opt.mod = 0; opt.len = 0;
    opt.root[0] = opt.root[1] = 0.0; opt.have = 0; opt.pad = 0;
    OptRootCache rc;
    opt_cache_clear(rc);
    double q_0 = 0.0, v_0 = 0.0, a_0 = 0.0, q_goal = 0.0;
    bool ok_j = false;
    int e = 0;
    if (act) {
      e = S.replan_list[idx];
      q_0 = cur_at(S, j, e); v_0 = cur_at(S, J + j, e); a_0 = cur_at(S, 2 * J + j, e);
      q_goal = S.goal_q[(size_t)j * N + e];
      ok_j = ltp_check_inputs_joint(L, c.q_min[j], c.q_max[j], q_0, v_0, a_0);
      if (ok_j) ok_j = ltp_opt_switch_times(L, q_goal, q_0, v_0, a_0, L.v_max, opt.T, opt.dir, opt.mod, &rc);
    }
    // slowest joint of the env (:39-45) in the reference's loop order; it stops at the first joint whose checkInputs /
    // optSwitchTimes fails: ok = all joints ok
    PlanHead h;
    h.t_required = -1;
    h.slowest = -1;
    bool ok = true;
    for (int k = 0; k < J; ++k) {
      const double tk = __shfl_sync(full, opt.T[6], seg0 + k);
      const int okk = __shfl_sync(full, ok_j ? 1 : 0, seg0 + k);
      ok = ok && okk != 0;
      if (tk > h.t_required) {
        h.t_required = tk;
        h.slowest = k;
      }
    }
    if (h.slowest == -1) ok = false;
    h.ok = ok ? 1 : 0;
    if (!act) continue;  // (spare lanes wait in the next fetch's shuffle)
    opt.len = h.ok;
    PlanRec sc = opt;
    if (ok) {
      double T[7];
      ltp_zero7(T);
      double v_drive = L.v_max;
      int mod = opt.mod;
      bool deep = false;
      if (j != h.slowest) {
        double v_m = v_0, a_m = a_0;
        if (opt.dir < 0) {
          v_m = -v_m;
          a_m = -a_m;
        }
        const double v_orig = opt.dir * v_m, a_orig = opt.dir * a_m;
        bool acc = false;
        for (int cand = 0; cand < 2 && !acc; ++cand) {
          double A[7];
          int ncoef;
          v_drive = ltp_scaling_setup(L, cand, q_goal, q_0, v_m, a_m, opt.dir, h.t_required, A, ncoef);
          acc = ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, v_drive, h.t_required, T, mod, &rc);
        }
        deep = !acc;
      }
      if (deep) {
        const int slot = atomicAdd(S.replan_ctr + CTR_DEEP_COUNT, 1);
        S.deep_list[slot] = idx * J + j;
        sc.len = -(slot + 1);
      } else {
        finish_scaled(L, opt, T, v_drive, mod, sc);
      }
    }
    opt.root[0] = rc.root[0]; opt.root[1] = rc.root[1]; opt.have = rc.have & 3;  // whatever has been solved so far
    S.plan_opt[(size_t)idx * HRG_MAX_JOINTS + j] = opt;
    S.plan_sc[(size_t)idx * HRG_MAX_JOINTS + j] = sc;
    if (j == 0) S.plan_head[idx] = h;
  }
}

// One (deep-list slot, polynomial candidate) task per lane. Tasks are ordered by candidate - the degree-6 one first, it
// has the longest solve - so that the 32 lanes of a warp run the same solver.
__global__ void __launch_bounds__(kReplanBlock) k_replan_deep(DevState S) {
  const HrgConfig& c = *S.cfg;
  const int lane = threadIdx.x & 31;
  const int J = S.J, N = S.N;
  const int n_deep = S.replan_ctr[CTR_DEEP_COUNT];
  const int ntask = n_deep * 6;
  for (;;) {
    const int base = fetch_warp_tasks(S.replan_ctr + CTR_DEEP, 32, ntask, lane);
    if (base < 0) break;
    const int dt = base + lane;
    if (dt >= ntask) continue;
    const int cls = dt / n_deep, slot = dt - cls * n_deep;
    const int cand = cls == 0 ? 7 : (cls == 1 ? 4 : (cls < 4 ? cls : cls + 1));  // 7, 4, 2, 3, 5, 6
    const int task = S.deep_list[slot];
    const int idx = task / J, j = task - idx * J;
    const int e = S.replan_list[idx];
    const PlanHead h = S.plan_head[idx];
    const PlanRec opt = S.plan_opt[(size_t)idx * HRG_MAX_JOINTS + j];
    const LtpLimits L = joint_limits(c, j);
    const double q_0 = cur_at(S, j, e), v_raw = cur_at(S, J + j, e), a_raw = cur_at(S, 2 * J + j, e);
    const double q_goal = S.goal_q[(size_t)j * N + e];
    OptRootCache rc;
    cache_from_rec(opt, q_goal, q_0, v_raw, a_raw, rc);
    double v_0 = v_raw, a_0 = a_raw;
    if (opt.dir < 0) {
      v_0 = -v_0;
      a_0 = -a_0;
    }
    const double v_orig = opt.dir * v_0, a_orig = opt.dir * a_0;
    PlanRec r;
    ltp_zero7(r.T);
    r.dir = opt.dir;
    r.mod = opt.mod;
    r.root[0] = r.root[1] = 0.0; r.have = 0; r.pad = 0;
    double A[7];
    int ncoef;
    ltp_scaling_setup(L, cand, q_goal, q_0, v_0, a_0, opt.dir, h.t_required, A, ncoef);
    r.v_drive = ltp_scaling_vdrive(L, cand, smallest_positive_real_root(L, A, ncoef), v_0, a_0);
    r.len = ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, r.v_drive, h.t_required, r.T, r.mod, &rc) ? 1 : 0;
    S.deep_res[(size_t)slot * 6 + (cand - 2)] = r;
  }
}

// ---- getTrajectory with one lane per (env, joint) -------------------------------------------------------------------
// ltp_sampler_init / ltp_jerk_at (ltp_traj.cuh) with every table index a compile-time constant, so the sampler lives in
// registers: the up to ten sub-sample corrections get twelve fixed slots in program order (index -1 = not applied).
struct JerkSamplerReg {
  int s[7];
  double jp[7];
  int cidx[12];
  double cval[12];
};
__device__ __forceinline__ void sampler_reg_init(const LtpLimits& L, const JointPlan& P, JerkSamplerReg& S) {
  const double ts = L.t_sample;
  double fr[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    const int prof = (P.mod == 1) ? ((k == 0 || k == 4) ? -1 : ((k == 2 || k == 6) ? 1 : 0))
                                  : ((k == 0 || k == 6) ? 1 : ((k == 2 || k == 4) ? -1 : 0));
    S.jp[k] = P.dir * L.j_max * prof;
    fr[k] = P.T[k] - ts * floor(P.T[k] / ts);
    S.s[k] = (k & 1) ? (int)ceil(P.T[k] / ts) : (int)floor(P.T[k] / ts);
  }
#pragma unroll
  for (int k = 0; k < 12; ++k) {
    S.cidx[k] = -1;
    S.cval[k] = 0.0;
  }
#define HRG_CORR(SLOT, I, V) \
  do {                       \
    S.cidx[SLOT] = (I);      \
    S.cval[SLOT] = (V);      \
  } while (0)
  if (S.s[2] >= S.s[1]) {  // phase 2 exists (:768-776)
    HRG_CORR(0, S.s[0] + 1, fr[0] / ts * S.jp[0]);
    if (S.s[1] > 0) HRG_CORR(1, S.s[1], (1 - fr[1] / ts) * S.jp[2]);
    HRG_CORR(2, S.s[2] + 1, fr[2] / ts * S.jp[2]);
  } else if (S.s[1] > 0) {  // (:777-783) two terms added one after the other
    HRG_CORR(3, S.s[1], fr[0] / ts * S.jp[0]);
    HRG_CORR(4, S.s[1], (fr[2] - fr[0]) / ts * S.jp[2]);
  }
  if (S.s[3] > 0) HRG_CORR(5, S.s[3], (1 - fr[3] / ts) * S.jp[4]);  // (:786-788)
  if (S.s[2] - S.s[0] > 0) {                                         // (:790-800)
    HRG_CORR(6, S.s[4] + 1, fr[4] / ts * S.jp[4]);
  } else if (S.s[4] > 0) {
    HRG_CORR(7, S.s[4], fr[4] / ts * S.jp[4]);
    HRG_CORR(8, S.s[4], fr[0] / ts * S.jp[0]);
    HRG_CORR(9, S.s[4], (fr[2] - fr[0]) / ts * S.jp[2]);
  }
  if (S.s[5] > 0) HRG_CORR(10, S.s[5], (1 - fr[5] / ts) * S.jp[6]);  // (:803-805)
  HRG_CORR(11, S.s[6] + 1, fr[6] / ts * S.jp[6]);                    // (:807)
#undef HRG_CORR
}
__device__ __forceinline__ double jerk_reg_at(const JerkSamplerReg& S, int i) {
  double j = 0.0;
  if (i < S.s[0]) j = S.jp[0];
#pragma unroll
  for (int k = 1; k < 7; ++k)
    if (i >= S.s[k - 1] && i < S.s[k]) j = S.jp[k];
#pragma unroll
  for (int k = 0; k < 12; ++k)
    if (S.cidx[k] == i) j = j + S.cval[k];
  return j;
}

// WITH_PFL only decides who owns alpha_i afterwards (k_pfl_ltt for PFL envs); the sampling itself is the same.
__global__ void __launch_bounds__(kReplanBlock) k_replan_sample(DevState S) {
  const HrgConfig& c = *S.cfg;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int J = S.J;
  const int E = 32 / J;
  const int sub = lane / J;
  const int j = lane - sub * J;
  const int seg0 = (sub < E ? sub : 0) * J;
  const int count = *S.replan_count;
  for (;;) {
    const int base = fetch_warp_tasks(S.replan_ctr + CTR_SAMPLE, E, count, lane);
    if (base < 0) break;
    const int idx = base + sub;
    const bool act = sub < E && idx < count;
    const LtpLimits L = joint_limits(c, act ? j : 0);
    int e = 0, pend = 0, len_j = 0;
    bool ok = false;
    JointPlan sc;
    ltp_zero7(sc.T);
    sc.dir = 0.0; sc.v_drive = 0.0; sc.mod = 0;
    double q_0 = 0.0, v_0 = 0.0, a_0 = 0.0;
    if (act) {
      e = S.replan_list[idx];
      pend = (S.flags[e] & F_LIVE_SLOT) ? 0 : 1;
      ok = S.plan_head[idx].ok != 0;
      q_0 = cur_at(S, j, e); v_0 = cur_at(S, J + j, e); a_0 = cur_at(S, 2 * J + j, e);
      if (ok) {
        PlanRec rec = S.plan_sc[(size_t)idx * HRG_MAX_JOINTS + j];
        if (rec.len < 0) {
          // a joint from the deep list: first accepted polynomial candidate in the reference's order, else "no valid
          // solution" (:640-644: zero times, v_drive = v_max, standard profile)
          const PlanRec* res = S.deep_res + (size_t)(-rec.len - 1) * 6;
          const PlanRec opt = S.plan_opt[(size_t)idx * HRG_MAX_JOINTS + j];
          double T[7];
          ltp_zero7(T);
          double v_drive = L.v_max;
          int mod = 0;
          for (int k = 0; k < 6; ++k)
            if (res[k].len) {
              for (int i = 0; i < 7; ++i) T[i] = res[k].T[i];
              v_drive = res[k].v_drive;
              mod = res[k].mod;
              break;
            }
          finish_scaled(L, opt, T, v_drive, mod, rec);
        }
        for (int k = 0; k < 7; ++k) sc.T[k] = rec.T[k];
        sc.dir = rec.dir; sc.v_drive = rec.v_drive; sc.mod = rec.mod;
        len_j = rec.len;
      }
    }
    // trajectory length of the env = max over its joints (:716-719)
    int len = 0;
    for (int k = 0; k < J; ++k) {
      const int other = __shfl_sync(full, len_j, seg0 + k);
      len = other > len ? other : len;
    }
    uint32_t trunc = 0;
    if (ok) {
      if (len > S.L) {
        trunc = 2;
        ok = false;
      } else if (len < 1) {
        ok = false;
      }
    }
    const int my_len = (act && ok) ? len : 0;
    int wlen = my_len;  // samples this warp integrates
    for (int o = 16; o > 0; o >>= 1) {
      const int other = __shfl_xor_sync(full, wlen, o);
      wlen = other > wlen ? other : wlen;
    }
    bool in = true;
    if (wlen > 0) {
      JerkSamplerReg Q;
      sampler_reg_init(L, sc, Q);
      const double ts = L.t_sample;
      const double v_cruise = sc.v_drive * sc.dir;
      const int s6 = Q.s[6];
      const int cr_lo = Q.s[2] + 1;
      const bool phase4 = Q.s[3] - Q.s[2] > 2;
      const unsigned cr_span = (phase4 && Q.s[3] - 1 > cr_lo) ? (unsigned)(Q.s[3] - 1 - cr_lo) : 0u;
      double* out = ltt_sample_w(S, e, pend, 0) + j;  // sample i: out[i * 4J + {0, J, 2J, 3J}] = q, v, a, j
      const size_t stride = (size_t)4 * J;
      double a = a_0, v = v_0, q = q_0;
      for (int i = 0; i < wlen; ++i) {
        const double jj = jerk_reg_at(Q, i);
        if (i < my_len) {
          if (i == 0) {  // the first sample integrates unconditionally (:810-813)
            a = a_0 + ts * jj;
            v = v_0 + ts * a;
            q = q_0 + ts * v;
          } else {
            const bool live = i <= s6;
            const bool cruise = (unsigned)(i - cr_lo) < cr_span;
            const double an = a + ts * jj;
            a = live ? an : 0.0;
            double vn = v + ts * a;
            vn = live ? vn : 0.0;
            v = cruise ? v_cruise : vn;
            q = q + ts * v;
          }
          out[0] = q;
          out[J] = v;
          out[2 * J] = a;
          out[3 * J] = jj;
          out += stride;
        }
      }
      if (my_len > 0) in = !(q < c.q_min[j] || q > c.q_max[j]);  // (:59-61)
    }
    // every joint of the env must end inside its limits
    bool all_in = true;
    for (int k = 0; k < J; ++k) {
      const int in_k = __shfl_sync(full, in ? 1 : 0, seg0 + k);  // (never behind a short-circuit: every lane must shuffle)
      all_in = all_in && in_k != 0;
    }
    if (act && j == 0) {
      // verdict for k_pfl_ltt / k_replan_tail: bit 0 = trajectory written and valid, bit 1 = longer than the slot;
      // slowest = its length
      PlanHead o = S.plan_head[idx];
      o.ok = ((ok && all_in) ? 1 : 0) | (int)trunc;
      o.slowest = len;
      S.plan_head[idx] = o;
    }
  }
}

// Thread per replanning env: install the new trajectory (safety_shield.cc:593-604, long_term_traj.h:108-116) and run the
// rest of the env's step up to the reach / verify hand-over. The envs that do not replan got this from k_plan while the
// chain was running.
__global__ void __launch_bounds__(128) k_replan_tail(DevState S, const double* __restrict__ t) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= *S.replan_count) return;
  const HrgConfig& c = *S.cfg;
  const int J = S.J, N = S.N;
  const int e = S.replan_list[idx];
  uint32_t f = S.flags[e];
  const int pend = (f & F_LIVE_SLOT) ? 0 : 1;
  const PlanHead h = S.plan_head[idx];
  const bool ok = (h.ok & 1) != 0;
  uint32_t status = HRG_ST_REPLANNED;
  if (h.ok & 2) status |= HRG_ST_LTT_TRUNCATED;
  const bool pfl = ((f & F_TYPE_MASK) >> F_TYPE_SHIFT) == HRG_SHIELD_PFL;
  if (ok) {
    for (int j = 0; j < J; ++j) {
      S.replan_q[(size_t)j * N + e] = cur_at(S, j, e);  // last_replan_start_motion_ = current_motion (:593)
      // SSM / OFF: alpha_i_ = alpha_i_max (long_term_traj.h:108-116); PFL: written by k_pfl_ltt
      if (!pfl) S.ltt_alpha[((size_t)pend * J + j) * N + e] = c.alpha_i_max;
    }
    S.ltt_len[(size_t)pend * N + e] = h.slowest;
    S.ltt_cur[(size_t)pend * N + e] = 0;
    S.ltt_start[(size_t)pend * N + e] = S.path_sd[e];
    S.ltt_t0[(size_t)pend * N + e] = S.path_s[e];
    S.ltt_ts[(size_t)pend * N + e] = c.sample_time;
    f |= F_NEW_LTT;
    f &= ~F_NEW_LTT_PROCESSED;
  } else {
    f &= ~F_NEW_LTT;
    status |= HRG_ST_REPLAN_FAILED;
  }
  f &= ~F_PENDING_REPLAN;
  f |= status << F_STATUS_SHIFT;
  plan_env(S, c, e, t[e], f);
  S.flags[e] = f;
}

}  // namespace hrg
