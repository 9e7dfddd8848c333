// Thread-per-env kernels of one shield tick (scalar path logic; coalesced field-major state). The replan chain is in
// kernels_replan.cuh.
#pragma once
#include "kernels_replan.cuh"

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
  HRG_TRACE_SPAN(5);
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

// Middle part of step() for the envs that do NOT replan this tick (the others get it from k_replan_tail once their
// new trajectory exists). Runs next to the replan chain.
// meas_t: the measurement time of a tick that measures (k_measure then runs NEXT to this kernel and writes the same value to
// last_meas), NULL = the stored last_meas.
__global__ void __launch_bounds__(kTickBlock) k_plan(DevState S, const double* __restrict__ t,
                                                     const double* __restrict__ meas_t) {
  HRG_TRACE_SPAN(6);
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= S.N || S.replan_mark[e]) return;
  const HrgConfig& c = *S.cfg;
  uint32_t f = S.flags[e];
  plan_env(S, c, e, t[e], f, meas_t ? meas_t[e] : S.last_meas[e]);
  S.flags[e] = f;
}

// determineNextMotion + setpoint write-out + statistics. Also re-arms the replan list for the next tick.
__global__ void __launch_bounds__(kTickBlock)
    k_select(DevState S, double* q, double* dq, double* ddq, double* dddq, uint8_t* is_safe, uint8_t* branch,
             uint32_t* status, int which) {
  // which: -1 = every env; 0 = the envs that did not replan this tick (runs while the replan chain is still busy: they are
  // 96 % of the batch and nothing of theirs depends on it); 1 = the replanned envs, and the tick's last launch (re-arms the
  // work counters).
  HRG_TRACE_SPAN(which == 0 ? 10 : 12);
  hrg_grid_dep_wait();
  hrg_grid_dep_launch();
  constexpr int NC = 10;
  __shared__ unsigned long long s_cnt[NC];
  __shared__ double s_min;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int e = tid;
  if (threadIdx.x < NC) s_cnt[threadIdx.x] = 0ull;
  if (threadIdx.x == 0) s_min = INFINITY;
  __syncthreads();
  // re-arm the work lists for the next tick: each launch resets what only kernels BEFORE it on its own stream read
  // (which 0 runs behind the cascade stages on the main stream, which 1 behind the replan chain and pass 1 on the side
  // stream; the two may overlap). (Walking the replan list instead of testing every env's mark was measured: the few dozen
  // warps it concentrates the scattered loads of ~2 000 envs in take 18 us against 9 us for the marks.)
  if (tid == 0 && which != 0) {
    *S.replan_count = 0;
    for (int k = 0; k < N_CTR; ++k) S.replan_ctr[k] = 0;
  }
  if (tid == 0 && which != 1) S.verify_cnt[0] = S.verify_cnt[1] = 0;
  unsigned long long my[NC] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // unsafe, b0, b1, b2, replan, failed, near, len_sum, truncated, steps
  double my_min = INFINITY;
  if (e < S.N && (which < 0 || (int)S.replan_mark[e] == which)) {
    my[9] = 1;
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
      unsigned long long* g = reinterpret_cast<unsigned long long*>(S.stats);
      if (s_cnt[9]) atomicAdd(&g[0], s_cnt[9]);
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
