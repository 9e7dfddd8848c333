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

// Lane = (env, joint, half): two lanes per joint, floor(16 / J) envs per warp. A tick has far fewer problems than the GPU has
// lanes, so lanes are spent on shortening each warp's dependent chain - the kernel takes as long as its slowest warp:
//   1. half h solves quartic h of optSwitchTimes ahead of time (opt_prefill_root): every lane of the warp is in the
//      eigen-solver at the same program point, and the two solves of a joint run side by side;
//   2. half 0 runs checkInputs + optSwitchTimes with v_max (closed forms only now); the env's slowest joint comes from
//      shuffles over its lanes;
//   3. half h evaluates closed-form timeScaling candidate h (:378-436) - the reference tries (1) first and (2) only if
//      (1) is rejected, so half 0's result wins when both are accepted.
__global__ void __launch_bounds__(kReplanBlock) k_replan_plan(DevState S, int envs_per_warp) {
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

// One (deep-list slot, polynomial candidate) task per lane. Tasks are ordered by candidate - the degree-6 one first, it
// has the longest solve - so that the 32 lanes of a warp run the same solver.
// `per_warp` <= 32 tasks per warp: the lanes of a warp run the solver's three-way case split and its unrolled row / column
// positions one after the other wherever they disagree, so fewer problems per warp = a shorter chain per warp, and the GPU
// has warps to spare.
__global__ void __launch_bounds__(kReplanBlock) k_replan_deep(DevState S, int per_warp) {
  const HrgConfig& c = *S.cfg;
  const int lane = threadIdx.x & 31;
  const int J = S.J, N = S.N;
  const int n_deep = S.replan_ctr[CTR_DEEP_COUNT];
  const int ntask = n_deep * 6;
  for (;;) {
    const int base = fetch_warp_tasks(S.replan_ctr + CTR_DEEP, per_warp, ntask, lane);
    if (base < 0) break;
    const int dt = base + lane;
    if (lane >= per_warp || dt >= ntask) continue;
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

// One env per block of two warps. getTrajectory has two parts of very different shape: the jerk of sample i of joint r
// (a 7-way range test plus up to twelve sub-sample corrections, ~90 instructions, independent for every (r, i)), and the
// reference's sequential recurrence a += ts j, v += ts a, q += ts v (a dozen instructions that cannot be reordered without
// changing roundings). The kernel takes as long as the LONGEST trajectory of the tick, so only the recurrence may sit on the
// dependent chain: warp 1 produces the jerk of chunk c + 1 (lane = sample, one joint after the other) into one half of a
// double buffer while lanes 0..J-1 of warp 0 integrate chunk c out of the other half and stream the samples out.
constexpr int kSampleBlock = 64;
__global__ void __launch_bounds__(kSampleBlock) k_replan_sample(DevState S) {
  __shared__ JerkSamplerReg q_s[HRG_MAX_JOINTS];
  __shared__ double jbuf_s[2][HRG_MAX_JOINTS][33];
  __shared__ int len_s[HRG_MAX_JOINTS];
  __shared__ int idx_s;
  const HrgConfig& c = *S.cfg;
  const unsigned full = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int J = S.J;
  const int count = *S.replan_count;
  for (;;) {
    if (tid == 0) idx_s = atomicAdd(S.replan_ctr + CTR_SAMPLE, 1);
    __syncthreads();
    const int idx = idx_s;
    if (idx >= count) break;
    const int e = S.replan_list[idx];
    const int pend = (S.flags[e] & F_LIVE_SLOT) ? 0 : 1;
    const PlanHead head = S.plan_head[idx];
    bool ok = head.ok != 0;
    const bool jl = warp == 0 && lane < J;  // the lane that owns joint `lane`
    const int j = lane < J ? lane : 0;
    const LtpLimits L = joint_limits(c, j);
    JointPlan sc;
    ltp_zero7(sc.T);
    sc.dir = 0.0; sc.v_drive = 0.0; sc.mod = 0;
    double q_0 = 0.0, v_0 = 0.0, a_0 = 0.0;
    if (jl) {
      q_0 = cur_at(S, j, e); v_0 = cur_at(S, J + j, e); a_0 = cur_at(S, 2 * J + j, e);
      int len_j = 0;
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
        sampler_reg_init(L, sc, q_s[j]);
      }
      len_s[j] = len_j;
    }
    __syncthreads();
    int len = 0;  // trajectory length of the env = max over its joints (:716-719)
    for (int k = 0; k < J; ++k) len = len_s[k] > len ? len_s[k] : len;
    uint32_t trunc = 0;
    if (ok) {
      if (len > S.L) {
        trunc = 2;
        ok = false;
      } else if (len < 1) {
        ok = false;
      }
    }
    bool in = true;
    if (ok) {
      const int nchunk = (len + 31) >> 5;
      auto produce = [&](int ch) {  // warp 1: jerk of samples ch * 32 + lane of every joint
        const int i = (ch << 5) + lane;
        for (int r = 0; r < J; ++r) jbuf_s[ch & 1][r][lane] = jerk_reg_at(q_s[r], i);
      };
      if (warp == 1) produce(0);
      __syncthreads();
      const double ts = L.t_sample;
      const double v_cruise = sc.v_drive * sc.dir;
      const JerkSamplerReg& Q = q_s[j];
      const int s6 = Q.s[6];
      const int cr_lo = Q.s[2] + 1;
      const bool phase4 = Q.s[3] - Q.s[2] > 2;
      const unsigned cr_span = (phase4 && Q.s[3] - 1 > cr_lo) ? (unsigned)(Q.s[3] - 1 - cr_lo) : 0u;
      double* out = ltt_sample_w(S, e, pend, 0) + j;  // sample i: out[i * 4J + {0, J, 2J, 3J}] = q, v, a, j
      const size_t stride = (size_t)4 * J;
      double a = a_0, v = v_0, q = q_0;
      for (int ch = 0; ch < nchunk; ++ch) {
        if (warp == 1) {
          if (ch + 1 < nchunk) produce(ch + 1);
        } else if (jl) {
          const double* jb = jbuf_s[ch & 1][j];
          const int i0 = ch << 5;
          const int n = (len - i0) < 32 ? (len - i0) : 32;
          int u = 0;
          if (ch == 0) {  // the first sample integrates unconditionally (:810-813)
            const double jj = jb[0];
            a = a_0 + ts * jj;
            v = v_0 + ts * a;
            q = q_0 + ts * v;
            out[0] = q; out[J] = v; out[2 * J] = a; out[3 * J] = jj;
            out += stride;
            u = 1;
          }
#pragma unroll 4
          for (; u < n; ++u) {
            const int i = i0 + u;
            const double jj = jb[u];
            const bool live = i <= s6;
            const bool cruise = (unsigned)(i - cr_lo) < cr_span;
            const double an = a + ts * jj;
            a = live ? an : 0.0;
            double vn = v + ts * a;
            vn = live ? vn : 0.0;
            v = cruise ? v_cruise : vn;
            q = q + ts * v;
            out[0] = q; out[J] = v; out[2 * J] = a; out[3 * J] = jj;
            out += stride;
          }
        }
        __syncthreads();
      }
      if (jl) in = !(q < c.q_min[j] || q > c.q_max[j]);  // (:59-61)
    }
    if (warp == 0) {
      const bool all_in = __all_sync(full, in);  // every joint of the env must end inside its limits
      if (lane == 0) {
        // verdict for k_pfl_ltt / k_replan_tail: bit 0 = trajectory written and valid, bit 1 = longer than the slot;
        // slowest = its length
        PlanHead o = head;
        o.ok = ((ok && all_in) ? 1 : 0) | (int)trunc;
        o.slowest = len;
        S.plan_head[idx] = o;
      }
    }
    __syncthreads();  // idx_s, q_s, len_s are re-used by the next env
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
