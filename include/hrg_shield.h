/* hrg_shield.h — C ABI of the batched B200 safety shield (libhrg_shield.so).
 *
 * Drop-in boundary for the sara-shield hot path of human-robot-gym. Each entry point replaces one method of the
 * reference's pybind module `safety_shield_py` (SS/bindings/safety_shield_bindings.cpp:82-115, where
 * SS = human_robot_gym/controllers/failsafe_controller/sara-shield/safety_shield), now over a batch of N independent
 * shields held as structure-of-arrays FP64 device memory owned by an opaque handle.
 *
 * Conventions: plain C types only; every array argument is a DEVICE pointer to caller-owned memory unless the
 * function name ends in `_host` (then it is a HOST pointer, pinned for best speed, and the call performs the
 * host<->device copies on `stream`); nothing synchronises implicitly; every function returns 0 on success or a
 * negative HRG_E* code (no exceptions cross the ABI). `stream` is a cudaStream_t passed as void*.
 * There is no CPU fallback: without a CUDA device hrg_shield_create fails with HRG_ENODEVICE.
 */
#ifndef HRG_SHIELD_H
#define HRG_SHIELD_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HRG_MAX_JOINTS 8     /* Schunk 6, Panda 7 */
#define HRG_MAX_KEYPOINTS 64 /* mujoco_mocap.yaml 23, cmu_mocap*.yaml 21-31 */
#define HRG_MAX_BODIES 24    /* shipped models 8; synthetic full body 16 */
#define HRG_MAX_EXT 8        /* shipped models 2; synthetic full body 4 */

#define HRG_OK 0
#define HRG_EINVAL (-1)
#define HRG_ENODEVICE (-2)
#define HRG_ECUDA (-3)
#define HRG_ENOMEM (-4)

/* ShieldType (SS/include/safety_shield/safety_shield.h:55) */
#define HRG_SHIELD_OFF 0
#define HRG_SHIELD_SSM 1
#define HRG_SHIELD_PFL 2

/* Per-env status bits reported by hrg_shield_step (replace the reference's logged-and-swallowed exceptions). */
#define HRG_ST_REPLANNED 0x1u        /* LongTermPlanner ran this tick (safety_shield.cc:589) */
#define HRG_ST_REPLAN_FAILED 0x2u    /* planTrajectory returned false (safety_shield.cc:600) */
#define HRG_ST_NOT_PLANNABLE 0x4u    /* |ddq| > a_max_ltt (safety_shield.cc:572,603) */
#define HRG_ST_JOINT_LIMIT 0x8u      /* goal motion outside [q_min,q_max] (safety_shield.cc:619) */
#define HRG_ST_LTT_TRUNCATED 0x10u   /* planned trajectory longer than ltt_capacity: replan rejected */
#define HRG_ST_NEAR_THRESHOLD 0x20u  /* some capsule distance within 1e-6 m of 0 (excluded from flag parity) */

/* Constant tables: the content of the reference's three YAML files (safety_shield.cc:61-131), already ordered the
 * way the reference EVALUATES them (bodies/extremities in std::map order). Filled by the host layer. */
typedef struct HrgConfig {
  int32_t nb_joints;    /* robot_parameters_*.yaml nb_joints */
  int32_t n_keypoints;  /* len(joint_names) of the mocap yaml */
  int32_t n_bodies;     /* Vel and Accel capsules each */
  int32_t n_ext;        /* Pos spheres */
  int32_t ltt_capacity; /* samples per long-term-trajectory slot (two slots per env) */
  int32_t root_mode;    /* planner root finder: 0 = every polynomial through the restatement of the reference's QR
                         * iteration (roots.h:22-50); 1 = derivative-chain root finder, QR only where its margin to a
                         * different verdict is small (csrc/ltp_roots_fast.cuh) */
  double sample_time;
  double max_s_stop, v_safe, alpha_i_max; /* trajectory_parameters_*.yaml */
  double secure_radius;                   /* robot_parameters_*.yaml */
  double meas_err_pos, meas_err_vel, delay; /* mocap yaml */
  double q_min[HRG_MAX_JOINTS], q_max[HRG_MAX_JOINTS];
  double v_max[HRG_MAX_JOINTS], a_max[HRG_MAX_JOINTS], j_max[HRG_MAX_JOINTS]; /* *_max_allowed */
  double a_max_ltt[HRG_MAX_JOINTS], j_max_ltt[HRG_MAX_JOINTS];
  double link_T[HRG_MAX_JOINTS][16];  /* transformation_matrices, row-major 4x4 per joint */
  double link_caps[HRG_MAX_JOINTS][7]; /* enclosures: p1 p2 r in link frame */
  int32_t body_prox[HRG_MAX_BODIES], body_dist[HRG_MAX_BODIES];
  double body_thickness[HRG_MAX_BODIES];
  double body_vmax[HRG_MAX_BODIES][2], body_amax[HRG_MAX_BODIES][2]; /* joint_v_max/joint_a_max of prox, dist */
  int32_t ext_base[HRG_MAX_EXT];
  int32_t reserved1[HRG_MAX_EXT];
  double ext_thickness[HRG_MAX_EXT], ext_length[HRG_MAX_EXT], ext_vmax[HRG_MAX_EXT];
} HrgConfig;

/* Device-reduced per-step statistics (the only cross-GPU exchange: one tiny all_reduce of this struct). */
typedef struct HrgStats {
  int64_t n_steps;          /* env-steps accumulated */
  int64_t n_unsafe;         /* steps with is_safe == false */
  int64_t n_branch[3];      /* determineNextMotion branch histogram: recovery / potential / failsafe */
  int64_t n_replan;         /* LongTermPlanner invocations */
  int64_t n_replan_failed;
  int64_t n_near_threshold; /* steps with a capsule distance within 1e-6 of 0 */
  int64_t ltt_len_sum;      /* sum of planned LTT lengths (mean L = ltt_len_sum / (n_replan - n_replan_failed)) */
  double min_dist;          /* smallest capsule distance seen */
  int64_t n_ltt_truncated;  /* replans rejected because the plan needs more than ltt_capacity samples: the ONE place where
                             * this library can diverge from the reference (its LTT is an unbounded std::vector); must be 0
                             * for parity - size ltt_capacity with hrg_ltt_capacity_bound() */
  int64_t n_pfl_rejected;   /* envs reset to ShieldType PFL on a handle without PFL buffers (kept as SSM); must be 0 */
} HrgStats;

/* Constant tables of the batched FailsafeController: robosuite 1.3.2 Controller.scale_action ranges,
 * position limits of set_goal_position, PD gains and actuator range of JointPositionController, all per joint. */
typedef struct HrgControllerConfig {
  int32_t nb_joints;
  int32_t has_position_limits; /* 0: set_goal_position(..., position_limit=None) */
  double input_min[HRG_MAX_JOINTS], input_max[HRG_MAX_JOINTS];
  double output_min[HRG_MAX_JOINTS], output_max[HRG_MAX_JOINTS];
  double position_min[HRG_MAX_JOINTS], position_max[HRG_MAX_JOINTS];
  double kp[HRG_MAX_JOINTS], kd[HRG_MAX_JOINTS];
  double torque_min[HRG_MAX_JOINTS], torque_max[HRG_MAX_JOINTS]; /* actuator_range */
} HrgControllerConfig;

typedef struct HrgShield HrgShield;

/* SafetyShield(sample_time, trajectory_config_file, robot_config_file, mocap_config_file, init_x..init_yaw, init_qpos,
 * shield_type) — safety_shield_bindings.cpp:84-96 / safety_shield.cc:55-162 — for n_envs shields on CUDA device `device`.
 * base_pose [N,6] = (x,y,z,roll,pitch,yaw), init_q [N,J], shield_type [N] (or NULL: all SSM); HOST pointers here
 * (construction is not on the hot path). The shields are primed exactly like the reference ctor. */
int hrg_shield_create(const HrgConfig* cfg, int n_envs, int device, const double* base_pose_host,
                      const double* init_q_host, const int32_t* shield_type_host, HrgShield** out);
int hrg_shield_destroy(HrgShield* s);

/* Upper bound on the number of samples LongTermPlanner::getTrajectory (LTP/src/long_term_planner.cc:706-841) can emit for
 * ANY reachable start state and goal inside the joint limits of cfg: per joint, the time to brake from (v_max, a_max_ltt),
 * cross the whole range [q_min, q_max] with a v_max / a_max_ltt / j_max_ltt profile and settle, over sample_time (+ slack).
 * A handle whose ltt_capacity is at least this never truncates, i.e. never diverges from the reference's unbounded LTT. */
int hrg_ltt_capacity_bound(const HrgConfig* cfg);

/* Make the handle PFL-capable (per-sample Cartesian velocity + time buffers of both LTT slots, 2 x N x L doubles each).
 * Implied by a PFL entry in hrg_shield_create's shield_type; needed before hrg_shield_reset may switch envs to PFL.
 * Allocates both buffers or neither. Synchronous (cudaMalloc); not for the tick path. */
int hrg_shield_enable_pfl(HrgShield* s);

/* SafetyShield.reset(init_x.., init_qpos, current_time, shield_type) — bindings :97-106 / safety_shield.cc:164-201.
 * env_mask [N] u8 or NULL (all). shield_type [N] i32 or NULL (all SSM); PFL entries need a PFL-capable handle (see
 * hrg_shield_enable_pfl) - on a handle without, such envs are kept as SSM and counted in HrgStats.n_pfl_rejected.
 * Never allocates. */
int hrg_shield_reset(HrgShield* s, const uint8_t* env_mask, const double* base_pose, const double* init_q,
                     const double* t, const int32_t* shield_type, void* stream);

/* SafetyShield.humanMeasurement(human_measurement, time) — bindings :110 / safety_shield.h:521-539 /
 * human_reach.cc:52-74. keypoints [N,K,3], t [N]. env_mask [N] u8 or NULL. */
int hrg_shield_human_measurement(HrgShield* s, const uint8_t* env_mask, const double* keypoints, const double* t,
                                 void* stream);

/* SafetyShield.newLongTermTrajectory(goal_position, goal_velocity) — bindings :108 / safety_shield.cc:672-697.
 * goal_q, goal_dq [N,J]; env_mask [N] u8 or NULL. */
int hrg_shield_new_goal(HrgShield* s, const uint8_t* env_mask, const double* goal_q, const double* goal_dq,
                        void* stream);

/* SafetyShield.setLongTermTrajectory(traj) — bindings :109 / safety_shield.cc:699-730 — for ONE env.
 * traj [L,4,J] (q,dq,ddq,dddq per sample) HOST pointer; sample_time/starting_index/alpha_i [J] are the LongTermTraj's
 * own (long_term_traj.h:108-116). Synchronous. Returns in *verdict:
 * 0 installed, 1 RobotMovementException, 2 start position differs, 3 start velocity differs, 4 does not end stopped. */
int hrg_shield_set_ltt(HrgShield* s, int env, const double* traj_host, int length, double sample_time,
                       int starting_index, const double* alpha_i_host, int32_t* verdict_host);

/* setLongTermTrajectory for MANY envs at once, DEVICE pointers, asynchronous on `stream`: trajectory m
 * (traj [n_traj, max_len, 4, J], its first length[m] samples) goes to env env_index[m] (distinct envs); starting_index [n_traj]
 * or NULL (0), alpha_i [n_traj, J]; verdict [n_traj] as hrg_shield_set_ltt, -1 = env or length out of range. */
int hrg_shield_set_ltt_batch(HrgShield* s, int n_traj, const int32_t* env_index, const double* traj, int max_len,
                             const int32_t* length, double sample_time, const int32_t* starting_index,
                             const double* alpha_i, int32_t* verdict, void* stream);

/* Keep the robot/human capsules of every step for hrg_shield_get_capsules (off by default: it adds 56(J+H) bytes of
 * writes per env-step that only the visualisation getters need). Takes effect from the next step. */
int hrg_shield_store_capsules(HrgShield* s, int on);

/* SafetyShield.step(cycle_begin_time) — bindings :107 / safety_shield.cc:555-647 — for every env.
 * t [N]; outputs (any may be NULL): q,dq,ddq,dddq [N,J]; is_safe [N] u8 (getSafety, bindings :113);
 * branch [N] u8 0 recovery / 1 potential / 2 failsafe (safety_shield.cc:511-553); status [N] HRG_ST_* bits. */
int hrg_shield_step(HrgShield* s, const double* t, double* q, double* dq, double* ddq, double* dddq,
                    uint8_t* is_safe, uint8_t* branch, uint32_t* status, void* stream);

/* One control tick on DEVICE-resident inputs: [newLongTermTrajectory(goal_q) for the envs of goal_mask] +
 * humanMeasurement(keypoints, t) + step(t) — failsafe_controller.py:300,310,329. Same results as the three separate calls;
 * the measurement kernel is scheduled next to the start of the replan chain instead of in front of it (only reach + verify
 * read the keypoints). goal_mask / goal_q may be NULL (no new goals this tick). Outputs as hrg_shield_step. */
int hrg_shield_tick_device(HrgShield* s, const double* keypoints, const double* t, const uint8_t* goal_mask,
                           const double* goal_q, double* q, double* dq, double* ddq, double* dddq, uint8_t* is_safe,
                           uint8_t* branch, uint32_t* status, void* stream);

/* humanMeasurement + [newLongTermTrajectory for masked envs] + step with HOST buffers: one H2D of the inputs, the
 * kernels, one D2H of the setpoints, all on `stream`; the caller synchronises the stream. goal_mask/goal_q may be NULL.
 * out_host [N, 3J+1] doubles per env: q[J], dq[J], ddq[J], then is_safe + 2*branch packed as a double is NOT used —
 * flags_host [N] u8 = is_safe | branch<<1. */
int hrg_shield_tick_host(HrgShield* s, const double* keypoints_host, const double* t_host, const uint8_t* goal_mask_host,
                         const double* goal_q_host, double* q_host, double* dq_host, double* ddq_host,
                         uint8_t* flags_host, void* stream);

/* Double-buffered form of hrg_shield_tick_host for callers whose next inputs exist before the current tick has finished
 * (the human of human-robot-gym is an animation replay, human_env.py:1710-1767; goals only change every 25th tick):
 * hrg_shield_upload_inputs copies one tick's HOST inputs into staging slot 0/1 on `copy_stream`;
 * hrg_shield_tick_staged runs humanMeasurement + [newLongTermTrajectory] + step on that slot on `stream` (it waits for the
 * upload) and copies the setpoints back. Uploading tick k+1 while tick k computes hides the host->device copy. A slot is
 * not overwritten before the tick that consumed it has finished (events inside the handle). */
int hrg_shield_upload_inputs(HrgShield* s, int slot, const double* keypoints_host, const double* t_host,
                             const uint8_t* goal_mask_host, const double* goal_q_host, void* copy_stream);
int hrg_shield_tick_staged(HrgShield* s, int slot, double* q_host, double* dq_host, double* ddq_host,
                           uint8_t* flags_host, void* stream);
/* hrg_shield_upload_inputs with the goals as (env, goal) pairs: goal_env_host [n_goals] i32 (distinct envs, out-of-range
 * entries are dropped), goal_q_host [n_goals, J]; n_goals = 0: no new goals this tick. newLongTermTrajectory is called for
 * one env in 25 per tick (failsafe_controller.py:252-300 once per policy step), so this saves the mask and the 8J bytes of
 * goal per env that the dense form sends for all of them. */
int hrg_shield_upload_inputs_sparse(HrgShield* s, int slot, const double* keypoints_host, const double* t_host, int n_goals,
                                    const int32_t* goal_env_host, const double* goal_q_host, void* copy_stream);

/* getRobotReachCapsules / getHumanReachCapsules — bindings :111-112 / safety_shield.h:549-574 — of the last step.
 * robot [N,J,7], human [N,H,7] with H = n_ext + 2*n_bodies ordered Pos, Vel, Accel; min_dist [N,3] = per human model
 * the minimum capsule_capsule_dist over robot x human capsules (+inf where the step skipped verification).
 * Any pointer may be NULL. */
int hrg_shield_get_capsules(HrgShield* s, double* robot, double* human, double* min_dist, void* stream);

/* PlotCapsule(p1, p2, r).pos / .size / .mat - HRG/controllers/failsafe_controller/failsafe_controller/plot_capsule.py:30-86,
 * what FailsafeController.get_robot_capsules / get_human_capsules (failsafe_controller.py:383-427) hand to the MuJoCo
 * viewer - for n capsules: caps [n,7] (e.g. the output of hrg_shield_get_capsules) -> pos [n,3], size [n,3] = (r, r, half
 * length), mat [n,9] row-major rotation of the z axis onto the capsule axis. A headless device buffer: no viewer here. */
int hrg_capsule_markers(int n, const double* caps, double* pos, double* size, double* mat, void* stream);

/* LongTermTraj.getMaxAccelerationWindow(index) / getMaxJerkWindow(index) - SS/include/safety_shield/long_term_traj.h:243-259,
 * SS/src/long_term_traj.cc:79-159 - of every env's live long-term trajectory: per joint the maximum of the signed
 * acceleration / jerk over samples [t, min(t + k, L)), t = getTrajectoryIndex(index). index [N] i32 or NULL (each env's
 * path_s_discrete_); window_k <= 0: the shield's sliding_window_k_ = floor(max_s_stop / sample_time). max_acc, max_jerk [N,J]
 * (either may be NULL). Evaluated on demand (the step path never reads the windows, safety_shield.cc:392). */
int hrg_shield_ltt_max_window(HrgShield* s, const int32_t* index, int window_k, double* max_acc, double* max_jerk,
                              void* stream);

/* Accumulated statistics since the last clear; `out` is a DEVICE pointer to one HrgStats. */
int hrg_shield_stats(HrgShield* s, HrgStats* out, int clear, void* stream);

/* Checkpoint of every env's shield state (paths, flags, both long-term-trajectory slots, human measurement history, base
 * poses, statistics) - SURVEY.md 5 "state_dict / load_state_dict"; the reference has no counterpart (its state lives in the
 * C++ object). HOST buffers, synchronous, between ticks. A blob only loads into a handle of the same shape (envs, joints,
 * keypoints, ltt_capacity); a PFL blob makes the handle PFL-capable. After a load the next tick continues bit-identically. */
int64_t hrg_shield_state_bytes(const HrgShield* s);
int hrg_shield_state_save(HrgShield* s, void* dst_host, int64_t bytes);
int hrg_shield_state_load(HrgShield* s, const void* src_host, int64_t bytes);

/* Per-kernel device timing of the tick for the next `max_steps` steps (0 = off): a CUDA event is recorded on the step's
 * stream after EVERY kernel launch of a step (the replan chain is not overlapped while profiling, so each interval is one
 * kernel). hrg_shield_profile_read_kernels synchronises on the last event and returns the summed milliseconds per launch
 * slot (ms_host[HRG_PROF_SLOTS], indices below) and the number of steps they cover; hrg_shield_profile_read returns the same
 * folded into four groups (ms_host[4]: measure+pre+plan, the replan chain, reach+verify, select). */
#define HRG_PROF_MEASURE 0    /* k_measure */
#define HRG_PROF_PRE 1        /* k_pre */
#define HRG_PROF_PLAN 2       /* k_plan */
#define HRG_PROF_PLAN_LTT 3   /* k_replan_plan */
#define HRG_PROF_SPARE 4      /* (unused) */
#define HRG_PROF_DEEP 5       /* k_replan_deep */
#define HRG_PROF_SAMPLE 6     /* k_replan_sample */
#define HRG_PROF_PFL 7        /* k_pfl_ltt (PFL handles) */
#define HRG_PROF_TAIL 8       /* k_replan_tail */
#define HRG_PROF_VERIFY_P0 9  /* k_reach_verify pass 0 (Pos model of the cascade, or the full kernel) */
#define HRG_PROF_STAGE_VEL 10 /* k_verify_stage<1> */
#define HRG_PROF_STAGE_ACC 11 /* k_verify_stage<2> */
#define HRG_PROF_VERIFY_P1 12 /* k_reach_verify pass 1 (replanned envs) */
#define HRG_PROF_SELECT 13    /* k_select */
#define HRG_PROF_SLOTS 14
int hrg_shield_profile(HrgShield* s, int max_steps);
int hrg_shield_profile_read_kernels(HrgShield* s, double* ms_host, int* steps_host);
int hrg_shield_profile_read(HrgShield* s, double* ms_host, int* steps_host);

int hrg_shield_num_envs(const HrgShield* s);
int hrg_shield_num_human_capsules(const HrgShield* s);
/* Number of kernels this library has launched on behalf of `s` since creation (bench.py's gpu_launches claim). */
int64_t hrg_shield_launch_count(const HrgShield* s);
const char* hrg_last_error(void);

/* ---- batched FailsafeController (the shield's immediate caller); all DEVICE pointers, cfg HOST ------- */
/* FailsafeController.set_goal(action, set_qpos) — failsafe_controller.py:252-300: scaled = scale_action(action)
 * (robosuite/controllers/base_controller.py: clip to the input range, affine map onto the output range),
 * goal = set_goal_position(scaled, joint_pos, position_limit, set_pos) (robosuite/utils/control_utils.py: joint_pos +
 * scaled, or set_pos, clipped to the limits), then newLongTermTrajectory(goal, 0) for the envs of env_mask.
 * action, joint_pos [N,J]; set_qpos [N,J] or NULL; goal_out [N,J] or NULL. */
int hrg_controller_set_goal(HrgShield* s, const HrgControllerConfig* cfg, const uint8_t* env_mask, const double* action,
                            const double* joint_pos, const double* set_qpos, double* goal_out, void* stream);
/* FailsafeController.run_controller torque law — failsafe_controller.py:352-369: tau = M (kp e + kd de + ddq_des) +
 * torque_compensation, clipped to the actuator range (clip_torques). q_des, dq_des, ddq_des are hrg_shield_step's
 * outputs; joint_pos, joint_vel, torque_comp, torques [N,J]; mass_matrix [N,J,J] row-major. */
int hrg_controller_torques(HrgShield* s, const HrgControllerConfig* cfg, const double* q_des, const double* dq_des,
                           const double* ddq_des, const double* joint_pos, const double* joint_vel,
                           const double* mass_matrix, const double* torque_comp, double* torques, void* stream);

/* ---- animation -> keypoints: the human of human-robot-gym is a replay of recorded clips ---------------
 * Replaces, for the shield's input, `_control_human` + `sim.forward()` + `sim.data.get_site_xpos("Human_" + element)`
 * (HRG/environments/manipulation/human_env.py:1710-1767, :459-463): the kinematic tree of
 * HRG/models/assets/human/human.xml (mocap Pelvis, then per body three hinges z, y, x about the body's site) evaluated on
 * the device for N envs, each replaying its own list of clips with its own placement.
 * Only the placement depends on the env, so hrg_human_create walks the tree ONCE per frame of the clip bank (k_human_bake)
 * and hrg_human_keypoints is a gather + rigid transform per env per tick (csrc/kernels_human.cuh). */
typedef struct HrgHuman HrgHuman;

/* Tree: n_bodies bodies below the Pelvis in parent-before-child order (the order of HumanObject.joint_elements,
 * HRG/models/objects/human/human.py:58-82, = the keypoint order of mujoco_mocap.yaml); parent[b] = -1 for children of the
 * Pelvis; anchor [n_bodies,3] = joint/site position in the (common, untransformed) body frame. All HOST pointers.
 * Clips: frames [sum(clip_len), 7 + 3 n_bodies] = Pelvis_pos xyz, Pelvis_quat xyzw, then angle x,y,z per body (radians);
 * clip_info [n_clips,7] = position_offset xyz + orientation_quat xyzw of the clip's *_info.json.
 * step_length = simulation_step_freq / human_animation_freq (human_env.py:1462-1465); list_len = clips per env list
 * (`_n_animations_to_sample_at_resets`). */
int hrg_human_create(int n_bodies, const int32_t* parent_host, const double* anchor_host, int n_clips,
                     const int32_t* clip_len_host, const double* frames_host, const double* clip_info_host,
                     double step_length, int n_envs, int list_len, int device, HrgHuman** out);
int hrg_human_destroy(HrgHuman* h);
/* Env reset (human_env.py:1640-1657): placement [N,7] = human_pos_offset xyz + human_rot_offset quat wxyz; clip_list
 * [N,list_len] = the env's `_human_animation_ids`; index, animation_start_time := 0. DEVICE pointers, env_mask [N] u8 or
 * NULL (all). Envs replay clip 0 at the origin until their first reset; clip ids outside [0, n_clips) are clamped. */
int hrg_human_reset(HrgHuman* h, const uint8_t* env_mask, const double* placement, const int32_t* clip_list,
                    void* stream);
/* `_control_human(force_update=True)` at low_level_time [N] (i64, DEVICE) for every env, then the site positions:
 * keypoints [N,n_bodies,3] (DEVICE) in the layout hrg_shield_human_measurement / hrg_shield_tick_device consume.
 * clip_out, frame_out [N] i32 or NULL: which clip / frame each env showed. */
int hrg_human_keypoints(HrgHuman* h, const int64_t* low_level_time, double* keypoints, int32_t* clip_out,
                        int32_t* frame_out, void* stream);
/* One control tick of envs whose human is a clip replay, with HOST buffers: `_control_human` + `human_measurement` +
 * [newLongTermTrajectory] + humanMeasurement + step (human_env.py:503-524) - the clock low_level_time_host [N] i64, t_host [N]
 * and the goals go up, the keypoints are produced and consumed on the device, the setpoints come back (as
 * hrg_shield_tick_host). `h` and `s` must hold the same number of envs and keypoints on the same device. */
int hrg_shield_tick_replay_host(HrgShield* s, HrgHuman* h, const int64_t* low_level_time_host, const double* t_host,
                                const uint8_t* goal_mask_host, const double* goal_q_host, double* q_host, double* dq_host,
                                double* ddq_host, uint8_t* flags_host, void* stream);
/* hrg_shield_tick_replay_host with the setpoints of the tick copied to the host on `copy_stream` out of output slot 0/1, so
 * that the copy overlaps the NEXT tick's kernels (the shield follows its own trajectory: a tick never needs the host to have
 * seen the previous setpoints - SS/src/safety_shield.cc:555-647 reads no measured robot state). The host buffers of a slot are
 * valid once `copy_stream` has passed the call (synchronise it, or call again on the same slot); alternate the slots. */
int hrg_shield_tick_replay_piped(HrgShield* s, HrgHuman* h, int slot, const int64_t* low_level_time_host, const double* t_host,
                                 const uint8_t* goal_mask_host, const double* goal_q_host, double* q_host, double* dq_host,
                                 double* ddq_host, uint8_t* flags_host, void* stream, void* copy_stream);
int64_t hrg_human_launch_count(const HrgHuman* h);

/* Register-resident FP64 throughput of CUDA device `device` (the FP-pipe roofline denominator, not in
 * MEASURED_PEAKS.json): mode 0 = DFMA chains, 1 = DMUL + DADD pairs (what this library's -fmad=false arithmetic issues),
 * 2 = DADD only; 8 independent chains per thread, 2048 threads per SM, `iters` x 32 chain steps per thread. Synchronous.
 * gflops_host: achieved GFLOP/s (FMA = 2 flop); ginst_host (may be NULL): 1e9 thread-level FP64 instructions per second -
 * the unit of ncu's smsp__sass_thread_inst_executed_op_d{add,mul,fma}_pred_on; ms_host (may be NULL): best launch time. */
int hrg_microbench_fp64(int device, int mode, int iters, double* gflops_host, double* ginst_host, double* ms_host);
/* Dependent-issue latency in SM cycles of one FP64 operation in a chain, one warp alone on its SM: cycles_host[6] = DADD, DMUL,
 * DFMA, IEEE division (+ one DADD), sqrt (+ one DADD), DADD + compare + select. The replan chain of a tick is a few hundred
 * warps running dependent FP64 code, so these - not the throughput peaks - bound it. */
int hrg_microbench_fp64_latency(int device, double* cycles_host);

/* ---- per-stage entry points (parity tests + ncu); all DEVICE pointers --------------------------------------- */
/* capsule_capsule_dist (RL/include/intersections.hpp:105-124) for n independent pairs: c1,c2 [n,7] -> dist [n]. */
int hrg_stage_capsule_dist(int n, const double* c1, const double* c2, double* dist, void* stream);
/* planSafetyShield + Path::getFinalMotion (safety_shield.cc:203-341, path.cc:41-53) for n cases:
 * in [n,6] = pos vel acc ve a_max j_max -> ok [n] u8, phases [n,6], fin [n,3]. */
int hrg_stage_plan_safety_shield(double sample_time, int n, const double* in, uint8_t* ok, double* phases, double* fin,
                                 void* stream);
/* LongTermPlanner::planTrajectory (LTP/src/long_term_planner.cc:7-63) for n independent problems with the limits of
 * cfg: goal,q0,v0,a0 [n,J] -> traj [n,cap,4,J], length [n] i32 (0 = failed, -L = longer than cap). */
int hrg_stage_ltp_plan(const HrgConfig* cfg, int n, const double* goal, const double* q0, const double* v0,
                       const double* a0, double* traj, int32_t* length, int cap, void* stream);
/* reach + verify only (robot_reach.cc:61-98, human_reach.cc:76-87, verify_iso.cc:19-33) on explicit inputs:
 * base_pose [n,6], q_start,q_goal [n,J], s_diff,t_reach [n], alpha_i [n,J], kp_pos,kp_vel [n,K,3] ->
 * robot [n,J,7], human [n,H,7], min_dist [n,3], is_safe [n]. */
int hrg_stage_reach_verify(const HrgConfig* cfg, int n, const double* base_pose, const double* q_start,
                           const double* q_goal, const double* s_diff, const double* t_reach, const double* alpha_i,
                           const double* kp_pos, const double* kp_vel, double* robot, double* human, double* min_dist,
                           uint8_t* is_safe, void* stream);

/* hrg_stage_reach_verify without any allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). Same arguments and results as hrg_stage_reach_verify (bit-identical); the packed scratch state lives in a
 * static __device__ array of HRG_NOALLOC_SCRATCH_BYTES inside the library, which every GPU that loads the library holds
 * until exit. Limit: sizeof(HrgConfig) + n * (8 * (17 + 3J + 6K) + 4) bytes must fit, i.e. n <= 6 041 for Schunk (J 6) /
 * 5 938 for Panda (J 7) with the 23-keypoint mujoco_mocap human, 2 463 at HRG_MAX_JOINTS / HRG_MAX_KEYPOINTS; a larger n is
 * rejected with HRG_EINVAL and a message naming the limit (hrg_last_error). Calls no cudaMalloc, cudaFree or allocating
 * library routine; launches four kernels on `stream` on the CURRENT device and, unlike hrg_stage_reach_verify, does not
 * synchronise. All calls on a device share the one array: calls that overlap on the GPU (different streams, or the
 * same stream from two threads without ordering) race - serialise them; the result fields are cleared at the start of
 * every call, the input fields are fully rewritten. */
#define HRG_NOALLOC_SCRATCH_BYTES (8u << 20)
int hrg_stage_reach_verify_noalloc(const HrgConfig* cfg, int n, const double* base_pose, const double* q_start,
                                   const double* q_goal, const double* s_diff, const double* t_reach,
                                   const double* alpha_i, const double* kp_pos, const double* kp_vel, double* robot,
                                   double* human, double* min_dist, uint8_t* is_safe, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HRG_SHIELD_H */
