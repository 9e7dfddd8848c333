// calculateLongTermTrajectory (SS/src/safety_shield.cc:732-760) -> LongTermPlanner::planTrajectory
// (LTP/src/long_term_planner.cc:7-63) for the envs on the replan list: the replan chain of a tick.
//
// ~2 100 of 65 536 envs replan per tick and the tick cannot end before the slowest of them has its trajectory: the chain is
// a LATENCY problem (the GPU has ten times more lanes than there are problems), so everything here is shaped to shorten the
// dependent chain of one warp (history and measurements: DESIGN.md section 3, profiles/r02_summary.md):
//   * the planner's polynomial roots come from the derivative-chain root finder (ltp_roots_fast.cuh); the exact restatement of
//     the reference's eigen-solve (ltp_roots_reg.cuh, one problem per lane, matrix in registers) only decides the close calls;
//   * optSwitchTimes' two quartic roots do not depend on v_drive (OptRootCache, ltp_profiles.cuh): they are found once per
//     (env, joint) in k_replan_plan and handed to every later call;
//   * kernels follow the data flow of one env instead of one kernel per planner routine:
//       k_replan_plan    lane = (env, joint, half), 2 envs per warp: checkInputs + the two quartic roots side by side,
//                        optSwitchTimes (:20-37), slowest joint through shuffles over the env's lanes (:39-45), the two
//                        closed-form timeScaling candidates side by side (:378-436); joints that accept neither go on the
//                        `deep` list
//       k_replan_deep    8-lane group = deep joint, lane = polynomial candidate 3..8 (:449-629) in one instruction stream
//                        for all degrees; the first accepted candidate in the reference's order wins (a ballot); exact QR
//                        only for close calls in front of it; starts from rest settled by ltp_cand_rejected_from_rest
//       k_replan_sample  2 envs per warp, lane = (env, joint): getTrajectory (:706-841) - jerk filled phase by phase into
//                        shared memory, the reference's sequential recurrence as three skewed chains, joint-limit test
//                        (:59-61)
//       [k_pfl_ltt       PFL handles: per-sample Cartesian velocities (pfl_math.cuh)]
//       k_replan_tail    4 envs per warp: install the trajectory and run the rest of the env's step up to the reach / verify
//                        hand-over (plan_env)
// Work is fetched from device counters (S.replan_ctr); grids only bound residency, and registers are capped so that every
// task of a tick has a resident warp.
#pragma once
#include "ltp_math.cuh"
#include "step_logic.cuh"
#include "pfl_math.cuh"

namespace hrg {

enum { CTR_PLAN = 0, CTR_SAMPLE = 1, CTR_DEEP = 2, CTR_DEEP_COUNT = 3, CTR_SELECT_DONE = 4, N_CTR = 5 };
// Residency: a tick has up to N / 25 replanning envs (2 621 at 65 536) and every task wants a warp of its own AT ONCE - a
// second round doubles the kernel. Registers per thread are capped accordingly (148 SMs x minBlocks warps).
#ifndef HRG_PLAN_MINBLOCKS
#define HRG_PLAN_MINBLOCKS 12
#endif
#ifndef HRG_SAMPLE_MINBLOCKS
#define HRG_SAMPLE_MINBLOCKS 20
#endif
constexpr int kReplanBlock = 32;  // one warp per block: a tick has a few hundred busy warps; this spreads them over all SM sub-partitions

__device__ __forceinline__ LtpLimits joint_limits(const HrgConfig& c, int j) {
  LtpLimits L;
  L.v_max = c.v_max[j]; L.a_max = c.a_max_ltt[j]; L.j_max = c.j_max_ltt[j]; L.t_sample = c.sample_time;
  L.mat = nullptr;
  L.root_mode = c.root_mode;
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

// Lane = (env, joint, half): two lanes per joint, floor(16 / J) envs per warp. A tick has far fewer problems than the GPU has
// lanes, so lanes are spent on shortening each warp's dependent chain - the kernel takes as long as its slowest warp:
//   1. half h solves quartic h of optSwitchTimes ahead of time (opt_prefill_root): every lane of the warp is in the
//      eigen-solver at the same program point, and the two solves of a joint run side by side;
//   2. half 0 runs checkInputs + optSwitchTimes with v_max (closed forms only now); the env's slowest joint comes from
//      shuffles over its lanes;
//   3. half h evaluates closed-form timeScaling candidate h (:378-436) - the reference tries (1) first and (2) only if
//      (1) is rejected, so half 0's result wins when both are accepted.
__global__ void __launch_bounds__(kReplanBlock, HRG_PLAN_MINBLOCKS) k_replan_plan(DevState S, int envs_per_warp) {
  HRG_TRACE_SPAN(0);
  hrg_grid_dep_launch();
  const HrgConfig& c = *S.cfg;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int J = S.J, N = S.N;
  const int E = envs_per_warp;                   // envs per warp, <= 16 / J
  const int half = lane & 1, pairi = lane >> 1;
  const int sub = pairi / J;                     // env of this lane within the warp (>= E: spare lane)
  const int j = pairi - sub * J;
  const int seg0 = (sub < E ? sub : 0) * 2 * J;  // first lane of this lane's env
  const int count = *S.replan_count;
  for (;;) {
    const int base = fetch_warp_tasks(S.replan_ctr + CTR_PLAN, E, count, lane);
    if (base < 0) break;
    HRG_TRACE_T0();
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
    if (act) {
      const int e = S.replan_list[idx];
      q_0 = cur_at(S, j, e); v_0 = cur_at(S, J + j, e); a_0 = cur_at(S, 2 * J + j, e);
      q_goal = S.goal_q[(size_t)j * N + e];
      ok_j = ltp_check_inputs_joint(L, c.q_min[j], c.q_max[j], q_0, v_0, a_0);
      if (ok_j) opt_prefill_root(L, q_goal, q_0, v_0, a_0, half, rc);  // 1.
    }
#if defined(HRG_TASK_TRACE)
    const unsigned long long tr_p1 = trace_now();
#endif
    {  // the partner's root
      const double other = __shfl_xor_sync(full, rc.root[half], 1);
      rc.root[half ^ 1] = other;
      if (ok_j) rc.have |= 3;
    }
    if (act && ok_j && half == 0)  // 2.
      ok_j = ltp_opt_switch_times(L, q_goal, q_0, v_0, a_0, L.v_max, opt.T, opt.dir, opt.mod, &rc);
    // half 1 gets its joint's result; then the slowest joint of the env (:39-45) in the reference's loop order; the
    // reference stops at the first joint whose checkInputs / optSwitchTimes fails: ok = all joints ok
    opt.dir = __shfl_sync(full, opt.dir, lane & ~1);
    opt.mod = __shfl_sync(full, opt.mod, lane & ~1);
    PlanHead h;
    h.t_required = -1;
    h.slowest = -1;
    bool ok = true;
    for (int k = 0; k < J; ++k) {
      const double tk = __shfl_sync(full, opt.T[6], seg0 + 2 * k);
      const int okk = __shfl_sync(full, ok_j ? 1 : 0, seg0 + 2 * k);
      ok = ok && okk != 0;
      if (tk > h.t_required) {
        h.t_required = tk;
        h.slowest = k;
      }
    }
    if (h.slowest == -1) ok = false;
    h.ok = ok ? 1 : 0;
#if defined(HRG_TASK_TRACE)
    const unsigned long long tr_p2 = trace_now();
#endif
    // 3. candidate `half`
    double T[7];
    ltp_zero7(T);
    double v_drive = L.v_max;
    int mod = opt.mod;
    bool acc = false;
    const bool scale = act && ok && j != h.slowest;
    if (scale) {
      double v_m = v_0, a_m = a_0;
      if (opt.dir < 0) {
        v_m = -v_m;
        a_m = -a_m;
      }
      const double v_orig = opt.dir * v_m, a_orig = opt.dir * a_m;
      double A[7];
      int ncoef;
      v_drive = ltp_scaling_setup(L, half, q_goal, q_0, v_m, a_m, opt.dir, h.t_required, A, ncoef);
      acc = ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, v_drive, h.t_required, T, mod, &rc);
    }
    // half 0 keeps its own candidate if accepted, else takes half 1's
    const int acc1 = __shfl_xor_sync(full, acc ? 1 : 0, 1);
    const bool take1 = half == 0 && !acc && acc1 != 0;
    {
      const double vd1 = __shfl_xor_sync(full, v_drive, 1);
      const int mod1 = __shfl_xor_sync(full, mod, 1);
      if (take1) {
        v_drive = vd1;
        mod = mod1;
      }
#pragma unroll
      for (int k = 0; k < 7; ++k) {
        const double t1 = __shfl_xor_sync(full, T[k], 1);
        if (take1) T[k] = t1;
      }
    }
    HRG_TRACE_TASK(0, base / E, 0);
#if defined(HRG_TASK_TRACE)
    if (lane == 0 && base / E < kTraceTasks) {
      unsigned long long* r3 = g_task_trace[3][base / E];
      r3[0] = tr_p1 - trace_t0;        // checkInputs + the two quartic roots
      r3[1] = tr_p2 - tr_p1;           // optSwitchTimes + slowest joint
      r3[2] = trace_now() - tr_p2;     // closed-form candidates
      r3[3] = 1;
    }
#endif
    if (!act || half != 0) continue;  // (the other lanes wait in the next fetch's shuffle)
    opt.len = h.ok;
    PlanRec sc = opt;
    if (ok) {
      if (scale && !acc && !take1) {
        const int slot = atomicAdd(S.replan_ctr + CTR_DEEP_COUNT, 1);
        S.deep_list[slot] = idx * J + j;
        sc.len = -(slot + 1);
      } else {
        if (!scale) {  // the slowest joint keeps its time-optimal plan (:41-45): zero times -> the `mx <= 0` fallback
          ltp_zero7(T);
          v_drive = L.v_max;
          mod = opt.mod;
        }
        finish_scaled(L, opt, T, v_drive, mod, sc);
      }
    }
    opt.root[0] = rc.root[0]; opt.root[1] = rc.root[1]; opt.have = rc.have & 3;
    S.plan_opt[(size_t)idx * HRG_MAX_JOINTS + j] = opt;
    S.plan_sc[(size_t)idx * HRG_MAX_JOINTS + j] = sc;
    if (j == 0) S.plan_head[idx] = h;
  }
}

// One deep-list joint per 8-lane group, lane c of the group = polynomial candidate c + 2 in the reference's order (:449-629,
// degrees 4, 4, 5, 4, 4, 6); `joints_per_warp` <= 4 groups are used. The reference walks the six candidates one after the
// other and takes the first one it accepts; here all six are evaluated side by side:
//   A. every lane finds its root with the derivative-chain finder (ltp_roots_fast.cuh) and, where that verdict is
//      unambiguous, runs the optSwitchTimes trial;
//   B. a candidate whose root is a close call (margin below kRootMargin: the QR iteration's rounding decides, typically a
//      start from rest, whose polynomials have a multiple root at 0) is settled by the exact QR restatement - but only if it
//      comes BEFORE the first candidate already accepted in A, since the reference never gets to the later ones.
// With root_mode 0 every candidate is a close call by definition and B is the whole kernel.
__global__ void __launch_bounds__(kReplanBlock) k_replan_deep(DevState S, int joints_per_warp) {
  HRG_TRACE_SPAN(1);
  hrg_grid_dep_wait();
  hrg_grid_dep_launch();
  const HrgConfig& c = *S.cfg;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int J = S.J, N = S.N;
  const int n_deep = S.replan_ctr[CTR_DEEP_COUNT];
  const int grp = lane >> 3, ci = lane & 7;
  for (;;) {
    const int base = fetch_warp_tasks(S.replan_ctr + CTR_DEEP, joints_per_warp, n_deep, lane);
    if (base < 0) break;
    HRG_TRACE_T0();
    const int slot = base + grp;
    const bool act = grp < joints_per_warp && slot < n_deep && ci < 6;
    const int cand = ci + 2;
    PlanRec r;
    ltp_zero7(r.T);
    r.dir = 0.0; r.v_drive = 0.0; r.mod = 0; r.len = 0;
    r.root[0] = r.root[1] = 0.0; r.have = 0; r.pad = 0;
    LtpLimits L = joint_limits(c, 0);
    OptRootCache rc;
    opt_cache_clear(rc);
    double A[7];
    int ncoef = 5;
    double q_0 = 0.0, q_goal = 0.0, v_0 = 0.0, a_0 = 0.0, v_orig = 0.0, a_orig = 0.0, t_required = 0.0;
    bool need_qr = false;
    if (act) {
      const int task = S.deep_list[slot];
      const int idx = task / J, j = task - idx * J;
      const int e = S.replan_list[idx];
      const PlanHead h = S.plan_head[idx];
      const PlanRec opt = S.plan_opt[(size_t)idx * HRG_MAX_JOINTS + j];
      L = joint_limits(c, j);
      q_0 = cur_at(S, j, e);
      const double v_raw = cur_at(S, J + j, e), a_raw = cur_at(S, 2 * J + j, e);
      q_goal = S.goal_q[(size_t)j * N + e];
      cache_from_rec(opt, q_goal, q_0, v_raw, a_raw, rc);
      v_0 = v_raw;
      a_0 = a_raw;
      if (opt.dir < 0) {
        v_0 = -v_0;
        a_0 = -a_0;
      }
      v_orig = opt.dir * v_0;
      a_orig = opt.dir * a_0;
      t_required = h.t_required;
      r.dir = opt.dir;
      r.mod = opt.mod;
      ltp_scaling_setup(L, cand, q_goal, q_0, v_0, a_0, opt.dir, t_required, A, ncoef);
      // A.
      need_qr = true;
      if (L.root_mode == 1) {
        double margin;
        const double root = fr_root_mixed(A, ncoef, margin);
        if (margin >= kRootMargin) {
          need_qr = false;
          r.v_drive = ltp_scaling_vdrive(L, cand, root, v_0, a_0);
          r.len = ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, r.v_drive, t_required, r.T, r.mod, &rc) ? 1 : 0;
        } else if (ltp_cand_rejected_from_rest(L, cand, A, ncoef, q_goal, q_0, v_0, a_0, v_orig, a_orig, t_required, &rc)) {
          HRG_TRACE_COUNT(2);
          need_qr = false;  // a start from rest: rejected whichever way the QR iteration's rounding falls (ltp_scaling.cuh)
          ltp_zero7(r.T);
          r.mod = opt.mod;
        }
      }
    }
    // B.
    const unsigned acc = (__ballot_sync(full, act && !need_qr && r.len != 0) >> (grp * 8)) & 0x3fu;
    const int first_acc = acc ? (__ffs(acc) - 1) : 6;
    if (act && need_qr && ci < first_acc) {
      HRG_TRACE_COUNT(0);
      r.v_drive = ltp_scaling_vdrive(L, cand, rr_root(A, ncoef), v_0, a_0);
      r.len = ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, r.v_drive, t_required, r.T, r.mod, &rc) ? 1 : 0;
    }
    if (act) S.deep_res[(size_t)slot * 6 + ci] = r;
    __syncwarp(full);
    HRG_TRACE_TASK(1, base / joints_per_warp, n_deep);
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

// Install the new trajectory of one replanning env (safety_shield.cc:593-604, long_term_traj.h:108-116) and run the rest of
// the env's step up to the reach / verify hand-over. The envs that do not replan got this from k_plan while the chain was
// running. `h` = the sampler's verdict (ok bit 0: trajectory valid, bit 1: longer than the slot; slowest = its length).
__device__ __forceinline__ void replan_tail_env(const DevState& S, const HrgConfig& c, int e, const PlanHead& h, double t,
                                             double last_meas) {
  const int J = S.J, N = S.N;
  uint32_t f = S.flags[e];
  const int pend = (f & F_LIVE_SLOT) ? 0 : 1;
  const bool ok = (h.ok & 1) != 0;
  uint32_t status = HRG_ST_REPLANNED;
  if (h.ok & 2) status |= HRG_ST_LTT_TRUNCATED;
  const bool pfl = ((f & F_TYPE_MASK) >> F_TYPE_SHIFT) == HRG_SHIELD_PFL;
  if (ok) {
    double cq[HRG_MAX_JOINTS];  // all loads first: a store in between would serialise them
#pragma unroll
    for (int j = 0; j < HRG_MAX_JOINTS; ++j)
      if (j < J) cq[j] = cur_at(S, j, e);
    const int sd = S.path_sd[e];
    const double ps = S.path_s[e];
#pragma unroll
    for (int j = 0; j < HRG_MAX_JOINTS; ++j)
      if (j < J) {
        S.replan_q[(size_t)j * N + e] = cq[j];  // last_replan_start_motion_ = current_motion (:593)
        // SSM / OFF: alpha_i_ = alpha_i_max (long_term_traj.h:108-116); PFL: written by k_pfl_ltt
        if (!pfl) S.ltt_alpha[((size_t)pend * J + j) * N + e] = c.alpha_i_max;
      }
    S.ltt_len[(size_t)pend * N + e] = h.slowest;
    S.ltt_cur[(size_t)pend * N + e] = 0;
    S.ltt_start[(size_t)pend * N + e] = sd;
    S.ltt_t0[(size_t)pend * N + e] = ps;
    S.ltt_ts[(size_t)pend * N + e] = c.sample_time;
    f |= F_NEW_LTT;
    f &= ~F_NEW_LTT_PROCESSED;
  } else {
    f &= ~F_NEW_LTT;
    status |= HRG_ST_REPLAN_FAILED;
  }
  f &= ~F_PENDING_REPLAN;
  f |= status << F_STATUS_SHIFT;
  plan_env(S, c, e, t, f, last_meas);
  S.flags[e] = f;
}

// Thread per replanning env. A kernel of its own: inside the sampler its ~250 registers would cap the residency of the
// sampling loop (fused and uncapped: 9 warps / SM, a second round above 1 332 replans; capped to 128: the loop spills and
// runs at half speed - both measured, profiles/r02_summary.md). PFL handles run k_pfl_ltt between the two.
// What replan_tail_env / plan_env will read of env e, asked for ahead of time by all lanes of the env's warp (field-major
// state: every field is its own cache line, and the tail is one lane chasing ~100 of them one after the other).
template <bool L1>
__device__ __forceinline__ void prefetch_tail_inputs(const DevState& S, int e, int lane, int nl = 32) {
  const int J = S.J;
  const size_t N = (size_t)S.N;
  auto pf = [](const void* p) {
    if (L1)
      asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
    else
      asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
  };
  for (int k = lane; k < N_PATHS * PATH_FIELDS; k += nl) pf(S.paths + (size_t)k * N + e);
  for (int k = lane; k < 4 * J + 1; k += nl) pf(S.cur + (size_t)k * N + e);
  for (int k = lane; k < J; k += nl) pf(S.prev_dq + (size_t)k * N + e);
  for (int k = lane; k < 2 * J; k += nl) pf(S.ltt_alpha + (size_t)k * N + e);
  for (int k = lane; k < 2; k += nl) {
    pf(S.ltt_len + (size_t)k * N + e);
    pf(S.ltt_cur + (size_t)k * N + e);
    pf(S.ltt_start + (size_t)k * N + e);
    pf(S.ltt_t0 + (size_t)k * N + e);
    pf(S.ltt_ts + (size_t)k * N + e);
  }
  if (lane == 2 % nl) pf(S.path_s + e);
  if (lane == 3 % nl) pf(S.path_sd + e);
}

constexpr int kTailBlock = 32;
__global__ void __launch_bounds__(kTailBlock) k_replan_tail(DevState S, const double* __restrict__ t,
                                                            const double* __restrict__ meas_t) {
  hrg_grid_dep_wait();
  hrg_grid_dep_launch();
  // four envs per warp: the 8 lanes of a slot first pull their env's state into L1 (the tail is one lane chasing ~100
  // cache lines one after the other: out of L1 that is ~40 cycles each instead of ~300 out of L2), then lane 0 of the slot
  // runs the tail
  const int lane = threadIdx.x & 31, g = lane >> 3, l8 = lane & 7;
  const int count = *S.replan_count;
  for (int base = blockIdx.x * 4; base < count; base += gridDim.x * 4) {
    const int idx = base + g;
    const bool valid = idx < count;
    const int e = valid ? S.replan_list[idx] : 0;
    if (valid) {
      prefetch_tail_inputs<true>(S, e, l8, 8);
      // the head of the trajectory just planned (the goal motion is interpolated there) and the live sample in use
      const int pend = (S.flags[e] & F_LIVE_SLOT) ? 0 : 1;
      const char* p0 = reinterpret_cast<const char*>(ltt_sample(S, e, pend, 0));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(p0 + 128 * l8));
    }
    __syncwarp();
    if (valid && l8 == 0) replan_tail_env(S, *S.cfg, e, S.plan_head[idx], t[e], meas_t ? meas_t[e] : S.last_meas[e]);
    __syncwarp();
  }
}

// One env per warp (one warp per block, so the blocks of a tick are all resident at once and spread over every
// scheduler). getTrajectory has two parts of very different shape: the jerk of sample i of joint r - piecewise constant over
// the seven phases plus up to twelve one-sample corrections - and the reference's sequential recurrence a += ts j,
// v += ts a, q += ts v, which cannot be reordered without changing roundings. Per super-chunk of kSampleChunk samples the
// warp first FILLS the jerk of every joint into shared memory phase by phase (all 32 lanes, a handful of instructions per
// phase that touches the chunk - evaluating jerk_reg_at per sample, ~100 instructions each, made the earlier producer /
// consumer forms throughput bound: 13 M warp instructions per tick), applies the corrections in the reference's order, and
// then lanes 0..J-1 integrate the chunk and stream the samples out.
constexpr int kSampleBlock = 32;
constexpr int kSampleChunk = 128;
// Envs per warp: env g of a warp owns lanes [8g, 8g + J). The recurrence is one instruction stream for all of them, so two
// envs per warp halve the warp instructions of the kernel (it is issue bound: 2 621 warps of ~12 k instructions on 592
// schedulers) at the price of a loop as long as the longer of the two trajectories.
#ifndef HRG_SAMPLE_ENVS
#define HRG_SAMPLE_ENVS 2
#endif
constexpr int kSampleEnvs = HRG_SAMPLE_ENVS;
static_assert(kSampleEnvs == 1 || kSampleEnvs == 2 || kSampleEnvs == 4, "envs per sampler warp");
// JT: the joint count as a compile-time constant (6 Schunk, 7 Panda; 0 = read it from the state): with it the sample stride
// 4J is an immediate and the four stores of a sample need no 64-bit pointer arithmetic (a third of the recurrence loop).
template <int JT>
__global__ void __launch_bounds__(kSampleBlock, HRG_SAMPLE_MINBLOCKS) k_replan_sample(DevState S) {
  HRG_TRACE_SPAN(2);
  hrg_grid_dep_wait();
  hrg_grid_dep_launch();
  __shared__ JerkSamplerReg q_s[kSampleEnvs][HRG_MAX_JOINTS];
  __shared__ double jbuf_s[kSampleEnvs][HRG_MAX_JOINTS][kSampleChunk + 1];  // + 1: the owner lanes read one column of it
  const HrgConfig& c = *S.cfg;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int J = JT > 0 ? JT : S.J;
  const int g = lane >> 3, jl8 = lane & 7;          // env slot of the lane, joint of the lane
  const bool slot = g < kSampleEnvs;
  const int gs = slot ? g : 0;
  const int j = jl8 < J ? jl8 : 0;
  const int count = *S.replan_count;
  // the lane's share of the jerk fill: the kSampleEnvs * J joints of the warp get 32 / (kSampleEnvs * J) lanes each
  const int TJ = kSampleEnvs * J, lpj = 32 / TJ;
  const int fr = lane / lpj, fl = lane - fr * lpj;
  const bool filler = fr < TJ;
  const int fg = filler ? fr / J : 0, fj = filler ? fr - (fr / J) * J : 0;
  for (;;) {
    int base = 0;
    if (lane == 0) base = atomicAdd(S.replan_ctr + CTR_SAMPLE, kSampleEnvs);
    base = __shfl_sync(full, base, 0);
    if (base >= count) break;
    HRG_TRACE_T0();
    const int idx = base + gs;
    const bool valid = slot && idx < count;
    const bool own = valid && jl8 < J;  // the lane that owns joint j of env slot g
    const int e = valid ? S.replan_list[idx] : -1;
#pragma unroll
    for (int k = 0; k < kSampleEnvs; ++k) {  // for k_replan_tail, which follows: one lane per env chasing ~100 cache lines
      const int ek = __shfl_sync(full, e, 8 * k);
      if (ek >= 0) prefetch_tail_inputs<false>(S, ek, lane);
    }
    const int pend = (valid && (S.flags[e] & F_LIVE_SLOT)) ? 0 : 1;
    PlanHead head;
    head.ok = 0; head.slowest = -1; head.t_required = 0.0;
    if (valid) head = S.plan_head[idx];
    bool ok = valid && head.ok != 0;
    const LtpLimits L = joint_limits(c, j);
    JointPlan sc;
    ltp_zero7(sc.T);
    sc.dir = 0.0; sc.v_drive = 0.0; sc.mod = 0;
    double q_0 = 0.0, v_0 = 0.0, a_0 = 0.0;
    int len_j = 0;
    if (own) {
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
        sampler_reg_init(L, sc, q_s[gs][j]);
      }
    }
    __syncwarp(full);
    int len = len_j;  // trajectory length of the env = max over its joints (:716-719): over the 8 lanes of its slot
    for (int o = 4; o >= 1; o >>= 1) {
      const int other = __shfl_xor_sync(full, len, o);
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
    const int len_run = ok ? len : 0;  // samples this slot integrates
    int len_warp = len_run;            // ... and the longest of the warp
    for (int o = 16; o >= 8; o >>= 1) {
      const int other = __shfl_xor_sync(full, len_warp, o);
      len_warp = other > len_warp ? other : len_warp;
    }
    const int flen = __shfl_sync(full, len_run, 8 * fg);  // length of the env whose joint this lane helps to fill
    bool in = true;
    {
      const double ts = L.t_sample;
      const double v_cruise = sc.v_drive * sc.dir;
      const JerkSamplerReg& Q = q_s[gs][j];
      const bool run = own && ok;
      const int s6 = run ? Q.s[6] : 0;
      const int cr_lo = run ? Q.s[2] + 1 : 1;
      const bool phase4 = run && Q.s[3] - Q.s[2] > 2;
      const unsigned cr_span = (phase4 && Q.s[3] - 1 > cr_lo) ? (unsigned)(Q.s[3] - 1 - cr_lo) : 0u;
      double* out_q = run ? ltt_sample_w(S, e, pend, 0) + j : nullptr;  // sample i: [i * 4J + {0, J, 2J, 3J}] = q, v, a, j
      double *out_v = out_q + J, *out_a = out_q + 2 * J;
      const size_t stride = (size_t)4 * J;
      double a = a_0, v = v_0, q = q_0;
      int fs[7];
      double fjp[7];
      if (flen > 0) {
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          fs[k] = q_s[fg][fj].s[k];
          fjp[k] = q_s[fg][fj].jp[k];
        }
      } else {
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          fs[k] = 0;
          fjp[k] = 0.0;
        }
      }
      for (int c0 = 0; c0 < len_warp; c0 += kSampleChunk) {
        // fill: every joint of the warp gets 32 / (envs * J) lanes, each evaluates the phase of its samples from the
        // joint's switch indices held in registers (jerk_reg_at without the corrections: the LAST matching range wins) -
        // branch free, all joints at once
        if (filler) {
          double* jb = jbuf_s[fg][fj];
          const int fn = (flen - c0 < kSampleChunk) ? flen - c0 : kSampleChunk;
          for (int u = fl; u < fn; u += lpj) {
            const int i = c0 + u;
            double jv = (i < fs[0]) ? fjp[0] : 0.0;
#pragma unroll
            for (int k = 1; k < 7; ++k) jv = (i >= fs[k - 1] && i < fs[k]) ? fjp[k] : jv;
            jb[u] = jv;
          }
        }
        __syncwarp(full);
        if (run && c0 < len) {
          const int c1 = (c0 + kSampleChunk < len) ? c0 + kSampleChunk : len;
          double* jb = jbuf_s[gs][j];
          // the one-sample corrections of the lane's own joint, in slot order (two of them can hit one sample)
#pragma unroll
          for (int k = 0; k < 12; ++k) {
            const int ci = Q.cidx[k];
            if (ci >= c0 && ci < c1) jb[ci - c0] = jb[ci - c0] + Q.cval[k];
          }
          // The recurrence, skewed by one sample per stage so that the three chains of an iteration are independent
          // (a single warp issues in order: un-skewed, every sample was one ~60-cycle dependent chain a -> v -> q):
          // iteration i computes a_i, v_(i-1) and q_(i-2), each from the values the PREVIOUS iteration left - the
          // operations on every value, and their order, are the reference's (:810-831; its unconditional first sample is
          // the general formula with live = true, cruise = false, which is what i = 0 gives: s6 >= 0, cr_lo >= 1).
          auto stage_q = [&]() {  // q_(i-2) = q_(i-3) + ts v_(i-2)
            q = q + ts * v;
            out_q[0] = q;
            out_q += stride;
          };
          auto stage_v = [&](int iv) {  // v_(i-1) from v_(i-2) and a_(i-1)
            const bool live = iv <= s6;
            const bool cruise = (unsigned)(iv - cr_lo) < cr_span;
            double vn = v + ts * a;
            vn = live ? vn : 0.0;
            v = cruise ? v_cruise : vn;
            out_v[0] = v;
            out_v += stride;
          };
          auto stage_a = [&](int i) {  // a_i from a_(i-1) and the jerk of sample i
            const double jj = jb[i - c0];
            const bool live = i <= s6;
            const double an = a + ts * jj;
            a = live ? an : 0.0;
            out_a[0] = a;
            out_a[J] = jj;
            out_a += stride;
          };
          int i = c0;
          for (; i < 2 && i < c1; ++i) {  // pipeline fill (first chunk only)
            if (i >= 1) stage_v(i - 1);
            stage_a(i);
          }
#pragma unroll 4
          for (; i < c1; ++i) {  // steady state: three independent chains per iteration, no guards
            stage_q();
            stage_v(i - 1);
            stage_a(i);
          }
          if (c1 == len) {  // pipeline drain
            if (len >= 2) stage_q();
            stage_v(len - 1);
            stage_q();
          }
        }
        __syncwarp(full);
      }
      if (run) in = !(q < c.q_min[j] || q > c.q_max[j]);  // (:59-61)
    }
    const unsigned inb = __ballot_sync(full, in);
    const bool all_in = ((inb >> (8 * gs)) & 0xffu) == 0xffu;  // every joint of the env must end inside its limits
    if (valid && jl8 == 0) {
      // verdict for k_pfl_ltt / k_replan_tail: bit 0 = trajectory written and valid, bit 1 = longer than the slot;
      // slowest = its length
      PlanHead o = head;
      o.ok = ((ok && all_in) ? 1 : 0) | (int)trunc;
      o.slowest = len;
      S.plan_head[idx] = o;
    }
    HRG_TRACE_TASK(2, base / kSampleEnvs, len_warp);
    __syncwarp(full);  // q_s, jbuf_s are re-used by the next envs
  }
}

}  // namespace hrg
