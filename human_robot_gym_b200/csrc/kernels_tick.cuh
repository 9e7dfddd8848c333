// Thread-per-env kernels of one shield tick (scalar path logic; coalesced field-major state) and the replan kernel
// (one warp per replanning env).
#pragma once
#include "ltp_math.cuh"
#include "step_logic.cuh"
#include "pfl_math.cuh"

namespace hrg {

constexpr int kTickBlock = 128;

__global__ void k_new_goal(DevState S, const uint8_t* __restrict__ mask, const double* __restrict__ goal_q) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= S.N) return;
  if (mask && !mask[e]) return;
  new_goal_env(S, *S.cfg, e, goal_q + (size_t)e * S.J);
}

// First part of step() for every env: current motion, LTT commit, replan trigger -> compacted replan list. Kept apart
// from k_plan because the replan chain - the critical path of the tick - only waits for this much.
// goal_mask / goal_q != NULL: newLongTermTrajectory for the masked envs first (the tick entry points fold k_new_goal in).
__global__ void __launch_bounds__(kTickBlock) k_pre(DevState S, const uint8_t* __restrict__ goal_mask,
                                                    const double* __restrict__ goal_q) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= S.N) return;
  const HrgConfig& c = *S.cfg;
  if (goal_mask && goal_mask[e]) new_goal_env(S, c, e, goal_q + (size_t)e * S.J);
  uint32_t f = S.flags[e];
  const bool replan = pre_env(S, c, e, f);
  if (replan) {
    const int slot = atomicAdd(S.replan_count, 1);
    S.replan_list[slot] = e;
  }
  S.replan_mark[e] = replan ? 1 : 0;
  S.flags[e] = f;
}

// Middle part of step() for the envs that do NOT replan this tick (the others get it from k_replan_sample once their
// new trajectory exists). Runs next to the replan chain.
__global__ void __launch_bounds__(kTickBlock) k_plan(DevState S, const double* __restrict__ t) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= S.N || S.replan_mark[e]) return;
  const HrgConfig& c = *S.cfg;
  uint32_t f = S.flags[e];
  plan_env(S, c, e, t[e], f);
  S.flags[e] = f;
}

constexpr int kReplanBlock = 64;
#ifndef HRG_REPLAN_MINBLOCKS
#define HRG_REPLAN_MINBLOCKS 12  // resident 64-thread blocks per SM the single-lane replan kernels are compiled for
#endif

// ---- calculateLongTermTrajectory (SS/src/safety_shield.cc:732-760) -> LongTermPlanner::planTrajectory
// (LTP/src/long_term_planner.cc:7-63) for the envs on the replan list, as four kernels. The planner is branchy FP64
// code whose cost differs by two orders of magnitude between joints (up to eight profile shapes, six of them behind a
// companion-matrix eigenvalue solve, each followed by an optSwitchTimes trial that may solve two more quartics). The
// whole batch only holds a few thousand such problems per tick, so the kernel time is the LONGEST dependent chain, not
// the instruction total: lanes that diverge serialise, and a warp that walks the whole ladder is the tail everybody
// waits for. Hence the unit of scheduling is as small as the reference's data flow allows, and every problem gets an
// 8-lane group whose lanes split the eigen-solves (see kGroupLanes):
//   k_replan_opt     1 group per (env, joint)            checkInputs + optSwitchTimes (:20-37)
//   k_replan_scale   1 group per (env, joint)            slowest joint -> t_required (:39-45); the two closed-form
//                                                        timeScaling candidates (:378-436); joints that need the
//                                                        polynomial candidates go on the `deep` list
//   k_replan_deep    1 group per (deep joint, candidate) the six polynomial candidates (:449-629), all evaluated; the
//                                                        first accepted one in the reference's order is taken later,
//                                                        which is what its sequential ladder returns
//   k_replan_sample  1 warp per env                      getTrajectory (:706-841), joint-limit test (:59-61), LTT
//                                                        install, and the rest of the env's step up to reach / verify.
// Work is fetched from device counters (S.replan_ctr); grids only bound residency.
enum { CTR_OPT = 0, CTR_SCALE = 1, CTR_DEEP = 2, CTR_DEEP_COUNT = 3, N_CTR = 4 };

// Task packing of the planning kernels: one (env, joint[, candidate]) problem per 8-lane group, four groups per warp.
// The 8 lanes of a group run the problem's scalar code redundantly (identical inputs, identical results) and share each
// of its eigen-solves through the cooperative root finder (ltp_math.cuh). Earlier packings - one problem per lane with
// 1..32 problems per warp - are in profiles/r01_replan_packing.md: one problem per warp had the shortest chains but was
// instruction-fetch bound, many per warp serialised on divergent QR iteration counts.
constexpr int kGroupLanes = 8, kGroupsPerWarp = 4;
// Groups of one warp that hold different problems diverge and are issued one after the other, so a warp's chain is up to
// the SUM of its groups' chains. k_replan_opt / _scale have ~12 600 problems per tick and need 4 groups per warp to be
// resident in one wave; k_replan_deep has ~3 500 and is pure tail latency at 8 % occupancy, so it spreads them thinner.
#ifndef HRG_DEEP_GROUPS
#define HRG_DEEP_GROUPS 2
#endif

__device__ __forceinline__ LtpLimits joint_limits(const HrgConfig& c, int j, double* mat) {
  LtpLimits L;
  L.v_max = c.v_max[j]; L.a_max = c.a_max_ltt[j]; L.j_max = c.j_max_ltt[j]; L.t_sample = c.sample_time;
  L.mat = mat;
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

// Fetches the next four tasks of counter `ctr` for the warp: group g gets base + g. Returns -1 for the whole warp when
// the list is exhausted, -2 for a group beyond its end (whole 8-lane groups leave together).
template <int G = kGroupsPerWarp>
__device__ __forceinline__ int fetch_group_task(int32_t* ctr, int ntask, int lane) {
  int base = 0;
  if (lane == 0) base = atomicAdd(ctr, G);
  base = __shfl_sync(0xffffffffu, base, 0);
  if (base >= ntask) return -1;
  const int g = lane >> 3, t = base + g;
  return (g < G && t < ntask) ? t : -2;
}
__device__ __forceinline__ LtpLimits group_limits(const HrgConfig& c, int j, double* mat, int lane) {
  LtpLimits L = joint_limits(c, j, mat);
  L.coop_mask = 0xFFu << (lane & 24);
  L.coop_lane = lane & 7;
  return L;
}

// checkInputs + optSwitchTimes (planTrajectory :20-37) of one (replanning env, joint) per group.
__global__ void __launch_bounds__(kReplanBlock, HRG_REPLAN_MINBLOCKS) k_replan_opt(DevState S) {
  __shared__ double mat_s[kReplanBlock / 32][kGroupsPerWarp][37];
  const HrgConfig& c = *S.cfg;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int gl = lane & 7, grp = lane >> 3;
  const int J = S.J, N = S.N;
  const int ntask = *S.replan_count * J;
  for (;;) {
    const int task = fetch_group_task(S.replan_ctr + CTR_OPT, ntask, lane);
    if (task == -1) break;
    if (task < 0) continue;
    const int idx = task / J, j = task - idx * J;
    const int e = S.replan_list[idx];
    const LtpLimits L = group_limits(c, j, mat_s[wib][grp], lane);
    const double q_0 = cur_at(S, j, e), v_0 = cur_at(S, J + j, e), a_0 = cur_at(S, 2 * J + j, e);
    const double q_goal = S.goal_q[(size_t)j * N + e];
    PlanRec opt;
    ltp_zero7(opt.T);
    opt.dir = 0.0; opt.v_drive = L.v_max; opt.mod = 0; opt.len = 0;
    bool ok = ltp_check_inputs_joint(L, c.q_min[j], c.q_max[j], q_0, v_0, a_0);
    if (ok) ok = ltp_opt_switch_times(L, q_goal, q_0, v_0, a_0, L.v_max, opt.T, opt.dir, opt.mod);
    opt.len = ok ? 1 : 0;
    if (gl == 0) S.plan_opt[(size_t)idx * HRG_MAX_JOINTS + j] = opt;
  }
}

// Slowest joint of the env (:39-45, every group of the env derives it from the J records of k_replan_opt) and the two
// closed-form timeScaling candidates of this group's joint (:378-436); a joint that accepts neither goes on the deep
// list. The reference stops at the first joint whose checkInputs / optSwitchTimes fails: ok = all joints ok.
__global__ void __launch_bounds__(kReplanBlock, HRG_REPLAN_MINBLOCKS) k_replan_scale(DevState S) {
  __shared__ double mat_s[kReplanBlock / 32][kGroupsPerWarp][37];
  const HrgConfig& c = *S.cfg;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int gl = lane & 7, grp = lane >> 3;
  const int J = S.J, N = S.N;
  const int ntask = *S.replan_count * J;
  for (;;) {
    const int task = fetch_group_task(S.replan_ctr + CTR_SCALE, ntask, lane);
    if (task == -1) break;
    if (task < 0) continue;
    const int idx = task / J, j = task - idx * J;
    const int e = S.replan_list[idx];
    const LtpLimits L = group_limits(c, j, mat_s[wib][grp], lane);
    const PlanRec* rec = S.plan_opt + (size_t)idx * HRG_MAX_JOINTS;
    PlanRec opt = rec[j];
    PlanHead h;
    h.t_required = -1;
    h.slowest = -1;
    bool ok = true;
    for (int k = 0; k < J; ++k) {
      const double tk = rec[k].T[6];
      ok = ok && rec[k].len != 0;
      if (tk > h.t_required) {
        h.t_required = tk;
        h.slowest = k;
      }
    }
    if (h.slowest == -1) ok = false;
    h.ok = ok ? 1 : 0;
    opt.len = h.ok;
    PlanRec sc = opt;
    if (ok) {
      double q_0 = cur_at(S, j, e), v_0 = cur_at(S, J + j, e), a_0 = cur_at(S, 2 * J + j, e);
      const double q_goal = S.goal_q[(size_t)j * N + e];
      double T[7];
      ltp_zero7(T);
      double v_drive = L.v_max;
      int mod = opt.mod;
      bool deep = false;
      if (j != h.slowest) {
        if (opt.dir < 0) {
          v_0 = -v_0;
          a_0 = -a_0;
        }
        const double v_orig = opt.dir * v_0, a_orig = opt.dir * a_0;
        bool acc = false;
        for (int cand = 0; cand < 2 && !acc; ++cand) {
          double A[7];
          int ncoef;
          v_drive = ltp_scaling_setup(L, cand, q_goal, q_0, v_0, a_0, opt.dir, h.t_required, A, ncoef);
          acc = ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, v_drive, h.t_required, T, mod);
        }
        deep = !acc;
      }
      if (deep) {
        if (gl == 0) {
          const int slot = atomicAdd(S.replan_ctr + CTR_DEEP_COUNT, 1);
          S.deep_list[slot] = idx * J + j;
          sc.len = -(slot + 1);
        }
      } else {
        finish_scaled(L, opt, T, v_drive, mod, sc);
      }
    }
    if (gl == 0) {
      S.plan_sc[(size_t)idx * HRG_MAX_JOINTS + j] = sc;
      if (j == 0) S.plan_head[idx] = h;
    }
  }
}

// One (deep-list slot, polynomial candidate 2..7) task per 8-lane group, HRG_DEEP_GROUPS groups per warp. The 8 lanes run the task's
// scalar code redundantly (identical inputs, identical results) and share every eigen-solve of it - the candidate's own
// polynomial and those inside optSwitchTimes - through the cooperative root finder (ltp_math.cuh).
__global__ void __launch_bounds__(kReplanBlock, HRG_REPLAN_MINBLOCKS) k_replan_deep(DevState S) {
  __shared__ double mat_s[kReplanBlock / 32][HRG_DEEP_GROUPS][37];
  const HrgConfig& c = *S.cfg;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int gl = lane & 7, grp = lane >> 3;
  const int J = S.J, N = S.N;
  const int ntask = S.replan_ctr[CTR_DEEP_COUNT] * 6;
  for (;;) {
    const int dt = fetch_group_task<HRG_DEEP_GROUPS>(S.replan_ctr + CTR_DEEP, ntask, lane);
    if (dt == -1) break;
    if (dt < 0) continue;  // whole 8-lane groups leave together
    const int slot = dt / 6, cand = 2 + (dt - slot * 6);
    const int task = S.deep_list[slot];
    const int idx = task / J, j = task - idx * J;
    const int e = S.replan_list[idx];
    const PlanHead h = S.plan_head[idx];
    const PlanRec opt = S.plan_opt[(size_t)idx * HRG_MAX_JOINTS + j];
    LtpLimits L = joint_limits(c, j, mat_s[wib][grp]);
    L.coop_mask = 0xFFu << (lane & 24);
    L.coop_lane = gl;
    double q_0 = cur_at(S, j, e), v_0 = cur_at(S, J + j, e), a_0 = cur_at(S, 2 * J + j, e);
    const double q_goal = S.goal_q[(size_t)j * N + e];
    if (opt.dir < 0) {
      v_0 = -v_0;
      a_0 = -a_0;
    }
    const double v_orig = opt.dir * v_0, a_orig = opt.dir * a_0;
    PlanRec r;
    ltp_zero7(r.T);
    r.dir = opt.dir;
    r.mod = opt.mod;
    double A[7];
    int ncoef;
    ltp_scaling_setup(L, cand, q_goal, q_0, v_0, a_0, opt.dir, h.t_required, A, ncoef);
    r.v_drive = ltp_scaling_vdrive(L, cand, smallest_positive_real_root(L, A, ncoef), v_0, a_0);
    r.len = ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, r.v_drive, h.t_required, r.T, r.mod) ? 1 : 0;
    if (gl == 0) S.deep_res[(size_t)slot * 6 + (cand - 2)] = r;
  }
}

// One warp per env: lane = joint gl + 8 * sub. Each sub-lane evaluates the jerk of one of four consecutive samples and
// stores one of the four components; the integration itself stays the reference's sequential recurrence (bit-identical
// sums). Lane 0 then finishes the env's step up to the reach / verify hand-over.
// WITH_PFL = false is the build every SSM/OFF-only handle launches: the PFL branch is compiled out of it (its call would
// otherwise cost the sampler registers even when never taken).
// PFL handles run it twice around k_pfl_ltt: phase 1 = sampling only (verdict and length parked in plan_head), phase 2 = the
// env tail only; phase 0 (SSM / OFF handles) does both in one launch.
template <bool WITH_PFL>
__global__ void __launch_bounds__(128) k_replan_sample(DevState S, const double* __restrict__ t, int phase) {
  __shared__ JerkSampler samp_s[4][8];
  __shared__ double jbuf_s[4][8 * 32];
  const HrgConfig& c = *S.cfg;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int gl = lane & 7, sub = lane >> 3;
  const int idx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (idx >= *S.replan_count) return;  // whole warps leave together
  const int J = S.J, N = S.N;
  const int j = gl < J ? gl : J - 1;
  const bool joint_lane = gl < J;
  const int e = S.replan_list[idx];
  uint32_t f = S.flags[e];
  const int pend = (f & F_LIVE_SLOT) ? 0 : 1;
  const PlanHead h = S.plan_head[idx];
  uint32_t status = HRG_ST_REPLANNED;
  bool ok = h.ok != 0;
  int len = 0;
  LtpLimits L;
  L.v_max = c.v_max[j]; L.a_max = c.a_max_ltt[j]; L.j_max = c.j_max_ltt[j]; L.t_sample = c.sample_time;
  const double q_0 = cur_at(S, j, e), v_0 = cur_at(S, J + j, e), a_0 = cur_at(S, 2 * J + j, e);
  if (WITH_PFL && phase == 2) {  // phase 1 left: bit 0 = trajectory installed, bit 1 = longer than the slot; slowest = its length
    ok = (h.ok & 1) != 0;
    if (h.ok & 2) status |= HRG_ST_LTT_TRUNCATED;
    len = h.slowest;
  } else if (ok) {
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
    JointPlan sc;
    for (int k = 0; k < 7; ++k) sc.T[k] = rec.T[k];
    sc.dir = rec.dir; sc.v_drive = rec.v_drive; sc.mod = rec.mod;
    len = rec.len;
    for (int o = 4; o > 0; o >>= 1) {
      const int other = __shfl_xor_sync(full, len, o);
      len = other > len ? other : len;
    }
    if (len > S.L) {
      status |= HRG_ST_LTT_TRUNCATED;
      ok = false;
    } else if (len < 1) {
      ok = false;
    }
    if (ok) {
      const double qf = ltp_sample_joint_warp(L, sc, q_0, v_0, a_0, len, ltt_sample_w(S, e, pend, 0), J, gl, j, sub,
                                              joint_lane, samp_s[wib], jbuf_s[wib]);
      const bool in = !(qf < c.q_min[j] || qf > c.q_max[j]);
      ok = __all_sync(full, in);
    }
  }
  if (WITH_PFL && phase == 1) {
    if (lane == 0) {
      PlanHead o = h;
      o.ok = (ok ? 1 : 0) | ((status & HRG_ST_LTT_TRUNCATED) ? 2 : 0);
      o.slowest = len;
      S.plan_head[idx] = o;
    }
    return;  // k_pfl_ltt builds the PFL part of the trajectory (long_term_traj.cc:5-39), phase 2 finishes the env
  }
  const bool pfl = WITH_PFL && ((f & F_TYPE_MASK) >> F_TYPE_SHIFT) == HRG_SHIELD_PFL;
  if (ok && joint_lane && sub == 0) {
    S.replan_q[(size_t)j * N + e] = q_0;  // last_replan_start_motion_ = current_motion (:593)
    // SSM/OFF: alpha_i_ = alpha_i_max (long_term_traj.h:108-116)
    if (!pfl) S.ltt_alpha[((size_t)pend * J + j) * N + e] = c.alpha_i_max;
  }
  __syncwarp(full);
  if (lane == 0) {
    if (ok) {
      S.ltt_len[(size_t)pend * N + e] = len;
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
}

// determineNextMotion + setpoint write-out + statistics. Also re-arms the replan list for the next tick.
__global__ void __launch_bounds__(kTickBlock)
    k_select(DevState S, double* q, double* dq, double* ddq, double* dddq, uint8_t* is_safe, uint8_t* branch,
             uint32_t* status) {
  constexpr int NC = 9;
  __shared__ unsigned long long s_cnt[NC];
  __shared__ double s_min;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (threadIdx.x < NC) s_cnt[threadIdx.x] = 0ull;
  if (threadIdx.x == 0) s_min = INFINITY;
  __syncthreads();
  if (e == 0) {
    *S.replan_count = 0;
    for (int k = 0; k < N_CTR; ++k) S.replan_ctr[k] = 0;
    S.verify_cnt[0] = S.verify_cnt[1] = 0;
  }
  unsigned long long my[NC] = {0, 0, 0, 0, 0, 0, 0, 0, 0};  // unsafe, b0, b1, b2, replan, failed, near, len_sum, truncated
  double my_min = INFINITY;
  if (e < S.N) {
    const HrgConfig& c = *S.cfg;
    uint32_t f = S.flags[e];
    select_env(S, c, e, f, q, dq, ddq, dddq);
    f |= F_NEW_LTT_PROCESSED | F_STEPPED;  // safety_shield.cc:641
    S.flags[e] = f;
    const uint32_t st = f >> F_STATUS_SHIFT;
    const int br = (f & F_BRANCH_MASK) >> F_BRANCH_SHIFT;
    if (is_safe) is_safe[e] = (f & F_IS_SAFE) ? 1 : 0;
    if (branch) branch[e] = (uint8_t)br;
    if (status) status[e] = st;
    if (S.stats) {
      my[0] = (f & F_IS_SAFE) ? 0 : 1;
      my[1 + br] = 1;
      my[4] = (st & HRG_ST_REPLANNED) ? 1 : 0;
      my[5] = (st & HRG_ST_REPLAN_FAILED) ? 1 : 0;
      my[6] = (st & HRG_ST_NEAR_THRESHOLD) ? 1 : 0;
      my[8] = (st & HRG_ST_LTT_TRUNCATED) ? 1 : 0;
      if ((st & HRG_ST_REPLANNED) && !(st & HRG_ST_REPLAN_FAILED))
        my[7] = (unsigned long long)S.ltt_len[(size_t)((f & F_LIVE_SLOT) ? 0 : 1) * S.N + e];
      if (f & F_VERIFY) {
        my_min = fmin(S.min_dist[e], fmin(S.min_dist[(size_t)S.N + e], S.min_dist[(size_t)2 * S.N + e]));
      }
    }
  }
  if (S.stats) {
    for (int k = 0; k < NC; ++k) {
      unsigned long long v = my[k];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_cnt[k], v);
    }
    for (int o = 16; o > 0; o >>= 1) my_min = fmin(my_min, __shfl_xor_sync(0xffffffffu, my_min, o));
    if ((threadIdx.x & 31) == 0 && my_min < INFINITY) {
      // monotone CAS min on the bit pattern (values may be negative => compare as doubles)
      unsigned long long* a = reinterpret_cast<unsigned long long*>(&s_min);
      unsigned long long old = *a;
      while (__longlong_as_double((long long)old) > my_min) {
        unsigned long long prev = atomicCAS(a, old, (unsigned long long)__double_as_longlong(my_min));
        if (prev == old) break;
        old = prev;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int n_here = S.N - blockIdx.x * blockDim.x;
      if (n_here > (int)blockDim.x) n_here = blockDim.x;
      unsigned long long* g = reinterpret_cast<unsigned long long*>(S.stats);
      atomicAdd(&g[0], (unsigned long long)n_here);
      for (int k = 0; k < 8; ++k)
        if (s_cnt[k]) atomicAdd(&g[1 + k], s_cnt[k]);
      if (s_cnt[8]) atomicAdd(reinterpret_cast<unsigned long long*>(&S.stats->n_ltt_truncated), s_cnt[8]);
      if (s_min < INFINITY) {
        unsigned long long* a = reinterpret_cast<unsigned long long*>(&S.stats->min_dist);
        unsigned long long old = *a;
        while (__longlong_as_double((long long)old) > s_min) {
          unsigned long long prev = atomicCAS(a, old, (unsigned long long)__double_as_longlong(s_min));
          if (prev == old) break;
          old = prev;
        }
      }
    }
  }
}

// SafetyShield ctor tail / reset (SS/src/safety_shield.cc:147-160,164-201): clear flags and paths, seat the base pose,
// install the 1-sample LTT at init_qpos, prime the paths with computesPotentialTrajectory(is_safe_, 0) +
// determineNextMotion(is_safe_). The human keypoint buffers are deliberately left alone: HumanReach::reset only sets
// last_meas_timestep_ = -1 (human_reach.cc:44-50), so the old velocity survives until two new measurements arrived.
__global__ void k_reset(DevState S, const uint8_t* __restrict__ mask, const double* __restrict__ base_pose,
                        const double* __restrict__ init_q, const int32_t* __restrict__ shield_type) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= S.N) return;
  if (mask && !mask[e]) return;
  const HrgConfig& c = *S.cfg;
  const int J = S.J, N = S.N;
  uint32_t type = shield_type ? (uint32_t)shield_type[e] : (uint32_t)HRG_SHIELD_SSM;
  if (type > HRG_SHIELD_PFL) type = HRG_SHIELD_SSM;
  if (type == HRG_SHIELD_PFL && !S.ltt_cart) {  // not a PFL-capable handle (hrg_shield_enable_pfl): keep SSM, count it
    type = HRG_SHIELD_SSM;
    if (S.stats) atomicAdd(reinterpret_cast<unsigned long long*>(&S.stats->n_pfl_rejected), 1ull);
  }
  uint32_t f = S.flags[e] & F_UNDER_V_LIMIT;  // is_under_v_limit_ is not touched by reset()
  f |= (type << F_TYPE_SHIFT) & F_TYPE_MASK;
  if (type == HRG_SHIELD_OFF) f |= F_IS_SAFE;
  const double* p = base_pose + (size_t)e * 6;
  double B[12];
  base_pose_matrix(p[0], p[1], p[2], p[3], p[4], p[5], B);
  for (int k = 0; k < 12; ++k) S.base_T[(size_t)k * N + e] = B[k];
  S.last_meas[e] = -1;
  PathS z;
  path_clear(z);
  for (int w = 0; w < N_PATHS; ++w) path_store(S, w, e, z);
  S.path_s[e] = 0;
  S.path_sd[e] = 0;
  // live slot 0 = {Motion(0.0, init_qpos)}
  double* m = ltt_sample_w(S, e, 0, 0);
  for (int j = 0; j < J; ++j) {
    m[j] = init_q[(size_t)e * J + j];
    m[J + j] = 0.0; m[2 * J + j] = 0.0; m[3 * J + j] = 0.0;
    S.prev_dq[(size_t)j * N + e] = 0.0;
    S.ltt_alpha[((size_t)0 * J + j) * N + e] = (type == HRG_SHIELD_PFL) ? 0.0 : c.alpha_i_max;
  }
  if (S.ltt_cart) {  // PFL ctor on the one-sample LTT: zero joint velocity -> zero Cartesian velocity, alpha_i = 0
    S.ltt_cart[((size_t)e * 2 + 0) * S.L] = 0.0;
    S.ltt_time[((size_t)e * 2 + 0) * S.L] = 0.0;
  }
  for (int s = 0; s < 2; ++s) {
    S.ltt_len[(size_t)s * N + e] = 1;
    S.ltt_cur[(size_t)s * N + e] = 0;
    S.ltt_start[(size_t)s * N + e] = 0;
    S.ltt_t0[(size_t)s * N + e] = 0.0;
    S.ltt_ts[(size_t)s * N + e] = c.sample_time;
  }
  double gq[HRG_MAX_JOINTS], s_d, t_brake;
  int gslot;
  potential_env(S, c, e, f, gq, s_d, t_brake, gslot);
  select_env(S, c, e, f, nullptr, nullptr, nullptr, nullptr);
  S.flags[e] = f;
  for (int k = 0; k < 3; ++k) S.min_dist[(size_t)k * N + e] = INFINITY;
}

// ---- stage entry points ----------------------------------------------------------------------------------------
__global__ void k_stage_capsule_dist(int n, const double* __restrict__ c1, const double* __restrict__ c2,
                                     double* __restrict__ dist) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  dist[i] = capsule_capsule_dist(cap_load(c1 + (size_t)i * 7), cap_load(c2 + (size_t)i * 7));
}

__global__ void k_stage_plan_safety_shield(double ts, int n, const double* __restrict__ in, uint8_t* ok,
                                           double* phases, double* fin) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* a = in + (size_t)i * 6;
  PathS p;
  path_clear(p);
  bool touched;
  ok[i] = plan_safety_shield(a[0], a[1], a[2], a[3], a[4], a[5], ts, p, touched) ? 1 : 0;
  for (int k = 0; k < 6; ++k) phases[(size_t)i * 6 + k] = p.ph[k];
  path_final(p, fin[(size_t)i * 3], fin[(size_t)i * 3 + 1], fin[(size_t)i * 3 + 2]);
}

__global__ void k_stage_ltp_plan(const HrgConfig* __restrict__ cfg, int n, const double* goal, const double* q0,
                                 const double* v0, const double* a0, double* traj, int32_t* length, int cap) {
  __shared__ double mat_s[64][37];  // launched with 64 threads per block
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int J = cfg->nb_joints;
  length[i] = ltp_plan_env_seq(*cfg, goal + (size_t)i * J, q0 + (size_t)i * J, v0 + (size_t)i * J, a0 + (size_t)i * J,
                               traj + (size_t)i * cap * 4 * J, cap, mat_s[threadIdx.x]);
}

}  // namespace hrg
