// Device-resident state of a batch of N shields: structure-of-arrays FP64 in HBM.
//
// Every kernel of the tick maps consecutive LANES to consecutive ENVS (thread-per-env path logic; lane-per-env tiles of
// 32 envs in reach + verify), so all per-env state is stored field-major [field][N]: a warp reads or writes 32
// consecutive doubles = one fully used 256 B request per field. The only env-major records are the long-term
// trajectories (one sample = 4J contiguous doubles gathered per env per interpolation).
// The reference holds the same data as one heap-allocated SafetyShield object per env
// (SS/include/safety_shield/safety_shield.h:70-293).
#pragma once
#include "geom_math.cuh"
#include "task_trace.cuh"

// Programmatic dependent launch (sm_90+): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may be
// scheduled while its predecessor on the stream is still running; it must not touch the predecessor's results before
// hrg_grid_dep_wait() (returns once the predecessor has completed and its writes are visible). hrg_grid_dep_launch() lets the
// NEXT kernel of the stream be staged. Both are no-ops in a kernel launched the ordinary way.
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void hrg_grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void hrg_grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#else
inline void hrg_grid_dep_wait() {}
inline void hrg_grid_dep_launch() {}
#endif

namespace hrg {

// flag word (one per env)
enum : uint32_t {
  F_IS_SAFE = 1u << 0,        // is_safe_
  F_NEW_LTT = 1u << 1,        // new_ltt_
  F_NEW_GOAL = 1u << 2,       // new_goal_
  F_NEW_LTT_PROCESSED = 1u << 3,
  F_REC_CORRECT = 1u << 4,    // recovery_path_correct_
  F_UNDER_V_LIMIT = 1u << 5,  // is_under_v_limit_ (PFL)
  F_REC_CURRENT = 1u << 6,    // recovery_path_.isCurrent()
  F_LIVE_SLOT = 1u << 7,      // which LTT slot is long_term_trajectory_ (the other one is new_long_term_trajectory_)
  F_VERIFY = 1u << 8,         // this step runs reach + verify
  F_PENDING_REPLAN = 1u << 9, // this step waits for k_replan
  F_TYPE_SHIFT = 10,          // 2 bits: ShieldType
  F_TYPE_MASK = 3u << 10,
  F_STEPPED = 1u << 12,       // at least one step() since reset (capsule getters)
  F_BRANCH_SHIFT = 13,        // 2 bits: branch of the last determineNextMotion
  F_BRANCH_MASK = 3u << 13,
  F_STATUS_SHIFT = 16,        // HRG_ST_* bits of the last step live in the high half
};

enum { P_REC = 0, P_FS = 1, P_FS2 = 2, P_SAFE = 3, N_PATHS = 4, PATH_FIELDS = 10 };

// Hand-over records between the three replan kernels, indexed by position on the replan list.
struct PlanRec {   // one joint: cumulative jerk switch times, direction, cruise velocity, profile flag, samples needed
  double T[7];
  double dir, v_drive;
  double root[2];   // optSwitchTimes' two quartic roots of this joint, once solved (OptRootCache, ltp_profiles.cuh)
  int32_t mod, len;
  int32_t have, pad;  // bit 0 / 1: root[0] / root[1] valid
};
struct PlanHead {  // one env: time of the slowest joint, its index, and whether every joint passed checkInputs/optSwitchTimes
  double t_required;
  int32_t slowest, ok;
};

struct DevState {
  int N, J, K, H, L;  // envs, joints, keypoints, human capsules, LTT capacity (samples per slot)
  const HrgConfig* cfg;
  // ---- field-major [..][N]
  double* paths;      // [N_PATHS][PATH_FIELDS][N]
  double* path_s;     // [N]
  int32_t* path_sd;   // [N] path_s_discrete_
  uint32_t* flags;    // [N]
  double* prev_dq;    // [J][N]  next_motion_.getVelocity()
  double* goal_q;     // [J][N]  new_goal_motion_
  double* replan_q;   // [J][N]  last_replan_start_motion_
  int32_t* ltt_len;   // [2][N]
  int32_t* ltt_cur;   // [2][N]  current_pos_
  int32_t* ltt_start; // [2][N]  starting_index_
  double* ltt_t0;     // [2][N]  time of sample 0 (= path_s_ at planning; PFL scan)
  double* ltt_ts;     // [2][N]  sample_time_ of the LTT (differs from the shield's only for user-supplied LTTs)
  double* ltt_alpha;  // [2][J][N] alpha_i_ of the LTT
  double* base_T;     // [12][N] rows 0..2 of the base pose matrix
  double* kp_pos;     // [K*3][N]
  double* kp_vel;     // [K*3][N]
  double* last_meas;  // [N]
  // ---- env-major [N][..]
  double* ltt;        // [N][2][L][4][J]  q, dq, ddq, dddq per sample
  double* ltt_cart;   // [N][2][L] max Cartesian velocity per sample (PFL; NULL otherwise)
  double* ltt_time;   // [N][2][L] Motion time of each sample (PFL scan of getMaxofMaximumCartesianVelocityWithS)
  // ---- per-step hand-over between the kernels of one tick, field-major
  double* cur;        // [4J+1][N]: q, dq, ddq, dddq of getCurrentMotion(), then its s
  double* reach_in;   // [3J+2][N]: q_start[J], q_goal[J], alpha_i[J], s_diff, t_reach
  double* min_dist;   // [3][N]
  int32_t* replan_list;   // [N]
  int32_t* replan_count;  // [1]
  int32_t* replan_ctr;    // [N_CTR] work counters of the replan kernels + length of the deep list (kernels_replan.cuh CTR_*)
  PlanRec* plan_opt;      // [N][HRG_MAX_JOINTS] time-optimal plan per joint (k_replan_plan); len = this joint ok
  PlanRec* plan_sc;       // [N][HRG_MAX_JOINTS] plan scaled to the slowest joint (k_replan_plan / k_replan_sample); len < 0 = -(deep slot + 1)
  PlanHead* plan_head;    // [N]
  int32_t* deep_list;     // [N * J] (env, joint) tasks whose closed-form timeScaling candidates were rejected
  PlanRec* deep_res;      // [N * J][6] result of each polynomial candidate; len = accepted
  uint8_t* replan_mark;   // [N] 1 = this tick's reach + verify of the env runs after k_replan (second pass)
  // verification cascade (kernels_reach.cuh): envs whose Pos / Vel model collides go on to the next model
  int32_t* verify_list;   // [2][N] envs that still need the Vel (0) / Accel (1) model this tick
  int32_t* verify_cnt;    // [2]
  double* link_caps_buf;  // [7J][N] this tick's swept link capsules of the envs on verify_list, field-major
  // optional per-step capsule dump (drop-in getters); NULL = not stored
  double* robot_caps;  // [N][J][7]
  double* human_caps;  // [N][H][7]
  HrgStats* stats;
};

HRG_HD int cur_fields(int J) { return 4 * J + 1; }
HRG_HD int reach_fields(int J) { return 3 * J + 2; }
HRG_HD double& cur_at(const DevState& S, int field, int e) { return S.cur[(size_t)field * S.N + e]; }
HRG_HD double& ri_at(const DevState& S, int field, int e) { return S.reach_in[(size_t)field * S.N + e]; }

HRG_HD void path_load(const DevState& S, int which, int e, PathS& p) {
  const double* b = S.paths + (size_t)which * PATH_FIELDS * S.N + e;
  for (int i = 0; i < 6; ++i) p.ph[i] = b[(size_t)i * S.N];
  p.pos = b[(size_t)6 * S.N];
  p.vel = b[(size_t)7 * S.N];
  p.acc = b[(size_t)8 * S.N];
  p.jerk = b[(size_t)9 * S.N];
}
HRG_HD void path_store(const DevState& S, int which, int e, const PathS& p) {
  double* b = S.paths + (size_t)which * PATH_FIELDS * S.N + e;
  for (int i = 0; i < 6; ++i) b[(size_t)i * S.N] = p.ph[i];
  b[(size_t)6 * S.N] = p.pos;
  b[(size_t)7 * S.N] = p.vel;
  b[(size_t)8 * S.N] = p.acc;
  b[(size_t)9 * S.N] = p.jerk;
}
HRG_HD const double* ltt_sample(const DevState& S, int e, int slot, int i) {
  return S.ltt + (((size_t)e * 2 + slot) * S.L + i) * 4 * S.J;
}
HRG_HD double* ltt_sample_w(const DevState& S, int e, int slot, int i) {
  return S.ltt + (((size_t)e * 2 + slot) * S.L + i) * 4 * S.J;
}

// LongTermTraj::interpolate (SS/src/long_term_traj.cc:41-77) of slot `slot` of env e at (s, ds, dds, ddds).
// out: q, dq, ddq, dddq with stride `os` between joints (so it can write env-major or registers).
HRG_HD void ltt_interpolate(const DevState& S, const HrgConfig& c, int e, int slot, double s, double ds, double dds,
                            double ddds, double* q, double* dq, double* ddq, double* dddq) {
  int sample;
  double dt;
  ltt_locate(s, S.ltt_ts[(size_t)slot * S.N + e], S.ltt_start[(size_t)slot * S.N + e], S.ltt_len[(size_t)slot * S.N + e],
             S.ltt_cur[(size_t)slot * S.N + e], sample, dt);
  const double* m = ltt_sample(S, e, slot, sample);
  const int J = S.J;
  double m0[HRG_MAX_JOINTS], m1[HRG_MAX_JOINTS], m2[HRG_MAX_JOINTS], m3[HRG_MAX_JOINTS];
  HRG_UNROLL
  for (int j = 0; j < HRG_MAX_JOINTS; ++j)
    if (j < J) {
      m0[j] = m[j]; m1[j] = m[J + j]; m2[j] = m[2 * J + j]; m3[j] = m[3 * J + j];
    }
  HRG_UNROLL
  for (int j = 0; j < HRG_MAX_JOINTS; ++j)
    if (j < J)
      ltt_interp_joint(m0[j], m1[j], m2[j], m3[j], dt, ds, dds, ddds, c.v_max[j], c.a_max[j], q[j], dq[j], ddq[j],
                       dddq[j]);
}

}  // namespace hrg
