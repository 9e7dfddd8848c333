// libhrg_shield.so — C ABI of the batched B200 safety shield (see include/hrg_shield.h).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo (see __graft_entry__.build()).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "kernels_reach.cuh"
#include "kernels_tick.cuh"
#include "kernels_controller.cuh"
#include "microbench.cuh"

using namespace hrg;

namespace {
thread_local std::string g_err;

int fail(int code, const char* what, cudaError_t ce = cudaSuccess) {
  g_err = what;
  if (ce != cudaSuccess) {
    g_err += ": ";
    g_err += cudaGetErrorString(ce);
  }
  return code;
}
#define CK(call)                                        \
  do {                                                  \
    cudaError_t _e = (call);                            \
    if (_e != cudaSuccess) return fail(HRG_ECUDA, #call, _e); \
  } while (0)

int cfg_check(const HrgConfig* c) {
  if (!c) return fail(HRG_EINVAL, "cfg is NULL");
  if (c->nb_joints < 1 || c->nb_joints > HRG_MAX_JOINTS) return fail(HRG_EINVAL, "nb_joints out of range");
  if (c->n_keypoints < 1 || c->n_keypoints > HRG_MAX_KEYPOINTS) return fail(HRG_EINVAL, "n_keypoints out of range");
  if (c->n_bodies < 0 || c->n_bodies > HRG_MAX_BODIES) return fail(HRG_EINVAL, "n_bodies out of range");
  if (c->n_ext < 0 || c->n_ext > HRG_MAX_EXT) return fail(HRG_EINVAL, "n_ext out of range");
  if (!(c->sample_time > 0)) return fail(HRG_EINVAL, "sample_time must be > 0");
  for (int b = 0; b < c->n_bodies; ++b)
    if (c->body_prox[b] < 0 || c->body_prox[b] >= c->n_keypoints || c->body_dist[b] < 0 ||
        c->body_dist[b] >= c->n_keypoints)
      return fail(HRG_EINVAL, "body keypoint index out of range");
  for (int x = 0; x < c->n_ext; ++x)
    if (c->ext_base[x] < 0 || c->ext_base[x] >= c->n_keypoints) return fail(HRG_EINVAL, "extremity index out of range");
  return HRG_OK;
}
}  // namespace

struct HrgShield {
  DevState S;
  HrgConfig cfg;
  int device = 0;
  int sm_count = 148;
  bool pfl = false;
  long long launches = 0;
  std::vector<void*> allocs;
  std::vector<std::pair<void*, size_t>> persistent;  // what a checkpoint holds (hrg_shield_state_*)
  // staging for the *_host tick
  double* d_kp = nullptr;
  double* d_t = nullptr;
  double* d_goal = nullptr;
  uint8_t* d_goal_mask = nullptr;
  // second staging slot + events of the double-buffered host tick (hrg_shield_upload_inputs / hrg_shield_tick_staged)
  double* d_kp2[2] = {nullptr, nullptr};
  double* d_t2[2] = {nullptr, nullptr};
  double* d_goal2[2] = {nullptr, nullptr};
  uint8_t* d_goal_mask2[2] = {nullptr, nullptr};
  int32_t* d_goal_idx2[2] = {nullptr, nullptr};  // sparse goals of hrg_shield_upload_inputs_sparse: env indices ...
  double* d_goal_sp2[2] = {nullptr, nullptr};    // ... and their goals [n_goals, J]
  bool slot_has_goal[2] = {false, false};
  bool slot_used[2] = {false, false};
  cudaEvent_t ev_up[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
  double* d_q = nullptr;
  double* d_dq = nullptr;
  double* d_ddq = nullptr;
  uint8_t* d_safe = nullptr;
  uint8_t* d_branch = nullptr;
  uint8_t* d_flags8 = nullptr;
  // second set of output buffers + events of the pipelined replay tick (hrg_shield_tick_replay_piped)
  double* d_out2[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
  uint8_t* d_safe2[2] = {nullptr, nullptr};
  uint8_t* d_branch2[2] = {nullptr, nullptr};
  uint8_t* d_flags82[2] = {nullptr, nullptr};
  cudaEvent_t ev_out[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
  bool out_used[2] = {false, false};
  size_t reach_smem = 0, reach_smem_p0 = 0;
  size_t measure_smem = 0;
  // k_replan runs on its own stream, concurrently with the first reach + verify pass (envs that do not replan)
  cudaStream_t side = nullptr, meas = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_meas = nullptr;
  bool overlap = true;
  bool pdl_now = false;  // (pdl && this tick runs forked: verify_pass0 reads it)
  int pdl = 0;  // programmatic dependent launch (HRG_PDL): 0 off, 1 on the replan chain's stream, 2 on both branches
  bool debug_sync = false;
  int deep_per_warp = 2;  // deep-list joints per warp of k_replan_deep, <= 4 (HRG_DEEP_PER_WARP)
  int plan_envs_per_warp = 2;  // envs per warp of k_replan_plan, <= 16 / J (HRG_PLAN_ENVS_PER_WARP)
  bool cascade = true;  // verification cascade for pass 0 (HRG_NO_CASCADE=1 or stored capsules: full kernel)
  // optional per-kernel timing of the tick (bench.py's roofline leg): one event after every launch of a step
  bool prof = false;
  std::vector<cudaEvent_t> ev;
  std::vector<int> ev_id;  // HRG_PROF_* slot of the kernel that ENDS at this event; -1 = start of a step
  size_t ev_used = 0;

  template <class T>
  int alloc(T** p, size_t n) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, n * sizeof(T) + 16);
    if (e != cudaSuccess) return fail(HRG_ENOMEM, "cudaMalloc", e);
    e = cudaMemset(q, 0, n * sizeof(T) + 16);
    if (e != cudaSuccess) return fail(HRG_ECUDA, "cudaMemset", e);
    allocs.push_back(q);
    *p = static_cast<T*>(q);
    return HRG_OK;
  }
};

namespace {
// Launch with programmatic stream serialisation (pdl): the kernel may be staged while its predecessor on the stream still
// runs and synchronises on it with hrg_grid_dep_wait() - takes the ~3.5 us launch gap off a kernel boundary.
template <class... KArgs, class... Args>
inline cudaError_t launch_k(void (*k)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, k, static_cast<KArgs>(args)...);
}
inline int grid_for(int n, int block) { return (n + block - 1) / block; }
inline int tile_grid(int n) { return (n + kTileEnvs - 1) / kTileEnvs; }

// (env index, goal) pairs -> the mask / goal arrays the tick reads; indices out of range are dropped
__global__ void k_scatter_goals(int n, int J, int N, const int32_t* __restrict__ idx, const double* __restrict__ val,
                                uint8_t* __restrict__ mask, double* __restrict__ goal) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int e = idx[i];
  if (e < 0 || e >= N) return;
  mask[e] = 1;
  for (int j = 0; j < J; ++j) goal[(size_t)e * J + j] = val[(size_t)i * J + j];
}

__global__ void k_pack_flags(int n, const uint8_t* __restrict__ safe, const uint8_t* __restrict__ branch,
                             uint8_t* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint8_t)(safe[i] | (branch[i] << 1));
}
}  // namespace

extern "C" {

const char* hrg_last_error(void) { return g_err.c_str(); }

int hrg_shield_create(const HrgConfig* cfg, int n_envs, int device, const double* base_pose_host,
                      const double* init_q_host, const int32_t* shield_type_host, HrgShield** out) {
  if (!out) return fail(HRG_EINVAL, "out is NULL");
  *out = nullptr;
  int rc = cfg_check(cfg);
  if (rc) return rc;
  if (n_envs < 1) return fail(HRG_EINVAL, "n_envs must be >= 1");
  if (cfg->ltt_capacity < 2) return fail(HRG_EINVAL, "ltt_capacity must be >= 2");
  if (!base_pose_host || !init_q_host) return fail(HRG_EINVAL, "base_pose/init_q is NULL");
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev < 1) return fail(HRG_ENODEVICE, "no CUDA device (this library has no CPU fallback)", ce);
  if (device < 0 || device >= ndev) return fail(HRG_EINVAL, "device index out of range");
  CK(cudaSetDevice(device));
  HrgShield* h = new (std::nothrow) HrgShield();
  if (!h) return fail(HRG_ENOMEM, "host allocation failed");
  h->cfg = *cfg;
  h->device = device;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  h->sm_count = prop.multiProcessorCount;
  DevState& S = h->S;
  const size_t N = n_envs;
  const int J = cfg->nb_joints, K = cfg->n_keypoints, H = cfg->n_ext + 2 * cfg->n_bodies, L = cfg->ltt_capacity;
  S.N = n_envs; S.J = J; S.K = K; S.H = H; S.L = L;
  h->pfl = false;
  if (shield_type_host)
    for (size_t i = 0; i < N; ++i) h->pfl |= (shield_type_host[i] == HRG_SHIELD_PFL);
  HrgConfig* dcfg = nullptr;
#define AL(ptr, count)                         \
  do {                                         \
    rc = h->alloc(&(ptr), (size_t)(count));    \
    if (rc) {                                  \
      hrg_shield_destroy(h);                   \
      return rc;                               \
    }                                          \
  } while (0)
  AL(dcfg, 1);
  S.cfg = dcfg;
  AL(S.paths, (size_t)N_PATHS * PATH_FIELDS * N);
  AL(S.path_s, N);
  AL(S.path_sd, N);
  AL(S.flags, N);
  AL(S.prev_dq, (size_t)J * N);
  AL(S.goal_q, (size_t)J * N);
  AL(S.replan_q, (size_t)J * N);
  AL(S.ltt_len, 2 * N);
  AL(S.ltt_cur, 2 * N);
  AL(S.ltt_start, 2 * N);
  AL(S.ltt_t0, 2 * N);
  AL(S.ltt_ts, 2 * N);
  AL(S.ltt_alpha, (size_t)2 * J * N);
  AL(S.base_T, 12 * N);
  AL(S.kp_pos, (size_t)K * 3 * N);
  AL(S.kp_vel, (size_t)K * 3 * N);
  AL(S.last_meas, N);
  AL(S.ltt, (size_t)N * 2 * L * 4 * J);
  AL(S.cur, (size_t)N * cur_fields(J));
  AL(S.reach_in, (size_t)N * reach_fields(J));
  AL(S.min_dist, 3 * N);
  AL(S.replan_list, N);
  AL(S.replan_count, 1);
  AL(S.replan_ctr, 8);
  AL(S.deep_list, (size_t)N * J);
  AL(S.deep_res, (size_t)N * J * 6);
  AL(S.replan_mark, N);
  AL(S.verify_list, (size_t)2 * N);
  AL(S.verify_cnt, 2);
  AL(S.link_caps_buf, (size_t)7 * J * N);
  AL(S.plan_opt, (size_t)N * HRG_MAX_JOINTS);
  AL(S.plan_sc, (size_t)N * HRG_MAX_JOINTS);
  AL(S.plan_head, N);
  AL(S.stats, 1);
  // the state that survives a tick (everything else is scratch the next tick rewrites): SafetyShield members as SoA
#define PERSIST(ptr, count) h->persistent.emplace_back((void*)(ptr), (size_t)(count) * sizeof(*(ptr)))
  PERSIST(S.paths, (size_t)N_PATHS * PATH_FIELDS * N);
  PERSIST(S.path_s, N); PERSIST(S.path_sd, N); PERSIST(S.flags, N);
  PERSIST(S.prev_dq, (size_t)J * N); PERSIST(S.goal_q, (size_t)J * N); PERSIST(S.replan_q, (size_t)J * N);
  PERSIST(S.ltt_len, 2 * N); PERSIST(S.ltt_cur, 2 * N); PERSIST(S.ltt_start, 2 * N);
  PERSIST(S.ltt_t0, 2 * N); PERSIST(S.ltt_ts, 2 * N); PERSIST(S.ltt_alpha, (size_t)2 * J * N);
  PERSIST(S.base_T, 12 * N); PERSIST(S.kp_pos, (size_t)K * 3 * N); PERSIST(S.kp_vel, (size_t)K * 3 * N);
  PERSIST(S.last_meas, N); PERSIST(S.ltt, (size_t)N * 2 * L * 4 * J); PERSIST(S.stats, 1);
  S.ltt_cart = nullptr;
  S.robot_caps = nullptr;
  S.human_caps = nullptr;
  S.ltt_time = nullptr;
  if (h->pfl) {
    h->pfl = false;
    rc = hrg_shield_enable_pfl(h);
    if (rc) {
      hrg_shield_destroy(h);
      return rc;
    }
  }
  // staging buffers of hrg_shield_tick_host
  AL(h->d_kp, (size_t)K * 3 * N);
  AL(h->d_t, N);
  AL(h->d_goal, (size_t)J * N);
  AL(h->d_goal_mask, N);
  h->d_kp2[0] = h->d_kp; h->d_t2[0] = h->d_t; h->d_goal2[0] = h->d_goal; h->d_goal_mask2[0] = h->d_goal_mask;
  AL(h->d_kp2[1], (size_t)K * 3 * N);
  AL(h->d_t2[1], N);
  AL(h->d_goal2[1], (size_t)J * N);
  AL(h->d_goal_mask2[1], N);
  for (int k = 0; k < 2; ++k) {
    AL(h->d_goal_idx2[k], N);
    AL(h->d_goal_sp2[k], (size_t)J * N);
  }
  AL(h->d_q, (size_t)J * N);
  AL(h->d_dq, (size_t)J * N);
  AL(h->d_ddq, (size_t)J * N);
  AL(h->d_safe, N);
  AL(h->d_branch, N);
  AL(h->d_flags8, N);
  for (int k = 0; k < 2; ++k) {
    for (int c = 0; c < 3; ++c) AL(h->d_out2[k][c], (size_t)J * N);
    AL(h->d_safe2[k], N);
    AL(h->d_branch2[k], N);
    AL(h->d_flags82[k], N);
  }
#undef AL
  h->reach_smem = (size_t)reach_smem_doubles(J, H) * sizeof(double);
  h->measure_smem = (size_t)kTileEnvs * ((K * 3) | 1) * sizeof(double);
  h->reach_smem_p0 = h->reach_smem;
  if (const char* v = getenv("HRG_REACH_P0_SMEM_KB"))  // residency experiment: pad pass 0's shared memory (fewer blocks / SM)
    if ((size_t)atoi(v) * 1024 > h->reach_smem_p0) h->reach_smem_p0 = (size_t)atoi(v) * 1024;
  cudaFuncSetAttribute(k_reach_verify<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->reach_smem_p0);
  cudaFuncSetAttribute(k_reach_verify<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->reach_smem);
  cudaFuncSetAttribute(k_reach_verify<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->reach_smem_p0);
  cudaFuncSetAttribute(k_verify_stage<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->reach_smem);
  cudaFuncSetAttribute(k_verify_stage<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->reach_smem);
  cudaFuncSetAttribute(k_measure, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->measure_smem);
  {
    int lo = 0, hi = 0;  // the replan chain is the critical path of the tick: highest priority for its blocks
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    ce = cudaStreamCreateWithPriority(&h->side, cudaStreamNonBlocking, hi);
    if (const char* v = getenv("HRG_NO_OVERLAP")) h->overlap = !(atoi(v) > 0);
    if (const char* v = getenv("HRG_PDL")) h->pdl = atoi(v);
    if (const char* v = getenv("HRG_NO_CASCADE")) h->cascade = !(atoi(v) > 0);
    if (const char* v = getenv("HRG_DEBUG_SYNC")) h->debug_sync = atoi(v) > 0;
    h->plan_envs_per_warp = 16 / J > 2 ? 2 : 16 / J;
    if (const char* v = getenv("HRG_PLAN_ENVS_PER_WARP"))
      if (atoi(v) >= 1 && atoi(v) <= 16 / J) h->plan_envs_per_warp = atoi(v);
    if (const char* v = getenv("HRG_DEEP_PER_WARP"))
      if (atoi(v) >= 1 && atoi(v) <= 4) h->deep_per_warp = atoi(v);
  }
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&h->meas, cudaStreamNonBlocking);
  for (int k = 0; k < 2 && ce == cudaSuccess; ++k) {
    ce = cudaEventCreateWithFlags(&h->ev_out[k], cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&h->ev_copied[k], cudaEventDisableTiming);
  }
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
  if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&h->ev_meas, cudaEventDisableTiming);
  for (int k = 0; k < 2 && ce == cudaSuccess; ++k) {
    ce = cudaEventCreateWithFlags(&h->ev_up[k], cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&h->ev_done[k], cudaEventDisableTiming);
  }
  if (ce == cudaSuccess) ce = cudaMemcpy(dcfg, cfg, sizeof(HrgConfig), cudaMemcpyHostToDevice);
  HrgStats st0;
  memset(&st0, 0, sizeof(st0));
  st0.min_dist = INFINITY;
  if (ce == cudaSuccess) ce = cudaMemcpy(S.stats, &st0, sizeof(st0), cudaMemcpyHostToDevice);
  // prime every shield like the reference constructor
  double *d_pose = nullptr, *d_q0 = nullptr;
  int32_t* d_type = nullptr;
  if (ce == cudaSuccess) ce = cudaMalloc(&d_pose, 6 * N * sizeof(double));
  if (ce == cudaSuccess) ce = cudaMalloc(&d_q0, (size_t)J * N * sizeof(double));
  if (ce == cudaSuccess) ce = cudaMemcpy(d_pose, base_pose_host, 6 * N * sizeof(double), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(d_q0, init_q_host, (size_t)J * N * sizeof(double), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess && shield_type_host) {
    ce = cudaMalloc(&d_type, N * sizeof(int32_t));
    if (ce == cudaSuccess) ce = cudaMemcpy(d_type, shield_type_host, N * sizeof(int32_t), cudaMemcpyHostToDevice);
  }
  if (ce == cudaSuccess) {
    k_reset<<<grid_for(n_envs, 128), 128>>>(S, nullptr, d_pose, d_q0, d_type);
    h->launches++;
    ce = cudaDeviceSynchronize();
  }
  cudaFree(d_pose);
  cudaFree(d_q0);
  cudaFree(d_type);
  if (ce != cudaSuccess) {
    hrg_shield_destroy(h);
    return fail(HRG_ECUDA, "shield initialisation", ce);
  }
  *out = h;
  return HRG_OK;
}

int hrg_shield_destroy(HrgShield* s) {
  if (!s) return HRG_OK;
  cudaSetDevice(s->device);
  for (void* p : s->allocs) cudaFree(p);
  for (cudaEvent_t e : s->ev) cudaEventDestroy(e);
  for (int k = 0; k < 2; ++k) {
    if (s->ev_out[k]) cudaEventDestroy(s->ev_out[k]);
    if (s->ev_copied[k]) cudaEventDestroy(s->ev_copied[k]);
  }
  if (s->ev_fork) cudaEventDestroy(s->ev_fork);
  if (s->meas) cudaStreamDestroy(s->meas);
  if (s->ev_join) cudaEventDestroy(s->ev_join);
  if (s->ev_meas) cudaEventDestroy(s->ev_meas);
  for (int k = 0; k < 2; ++k) {
    if (s->ev_up[k]) cudaEventDestroy(s->ev_up[k]);
    if (s->ev_done[k]) cudaEventDestroy(s->ev_done[k]);
  }
  if (s->side) cudaStreamDestroy(s->side);
  delete s;
  return HRG_OK;
}

int hrg_shield_num_envs(const HrgShield* s) { return s ? s->S.N : 0; }
int hrg_shield_num_human_capsules(const HrgShield* s) { return s ? s->S.H : 0; }
int64_t hrg_shield_launch_count(const HrgShield* s) { return s ? s->launches : 0; }

int hrg_ltt_capacity_bound(const HrgConfig* c) {
  if (cfg_check(c)) return HRG_EINVAL;
  double worst = 0.0;
  for (int j = 0; j < c->nb_joints; ++j) {
    const double v = c->v_max[j], a = c->a_max_ltt[j], jk = c->j_max_ltt[j];
    if (!(v > 0 && a > 0 && jk > 0)) return fail(HRG_EINVAL, "hrg_ltt_capacity_bound: limits must be > 0");
    const double range = fabs(c->q_max[j] - c->q_min[j]);
    const double ramp = v / a + a / jk;  // 0 -> v_max under (a, j), or back
    // brake from full speed the wrong way, overshoot by at most v * ramp, cross the range plus that overshoot, settle
    const double T = ramp + (range + v * ramp) / v + 2.0 * ramp;
    if (T > worst) worst = T;
  }
  const double n = ceil(worst / c->sample_time) + 8.0;
  return n > 2.0e9 ? 2000000000 : (int)n;
}

int hrg_shield_enable_pfl(HrgShield* s) {
  if (!s) return fail(HRG_EINVAL, "hrg_shield_enable_pfl: NULL handle");
  if (s->S.ltt_cart && s->S.ltt_time) return HRG_OK;
  CK(cudaSetDevice(s->device));
  CK(cudaDeviceSynchronize());  // kernels in flight hold DevState copies with the NULL pointers
  double *cart = nullptr, *tm = nullptr;
  const size_t nb = (size_t)s->S.N * 2 * s->S.L * sizeof(double);
  cudaError_t e = cudaMalloc(&cart, nb);
  if (e == cudaSuccess) e = cudaMalloc(&tm, nb);
  if (e == cudaSuccess) e = cudaMemset(cart, 0, nb);
  if (e == cudaSuccess) e = cudaMemset(tm, 0, nb);
  if (e != cudaSuccess) {  // both or neither
    cudaFree(cart);
    cudaFree(tm);
    return fail(HRG_ENOMEM, "hrg_shield_enable_pfl: allocation of the PFL buffers failed", e);
  }
  s->allocs.push_back(cart);
  s->allocs.push_back(tm);
  s->S.ltt_cart = cart;
  s->S.ltt_time = tm;
  s->pfl = true;
  s->persistent.emplace_back((void*)cart, nb);
  s->persistent.emplace_back((void*)tm, nb);
  return HRG_OK;
}

int hrg_shield_reset(HrgShield* s, const uint8_t* env_mask, const double* base_pose, const double* init_q,
                     const double* t, const int32_t* shield_type, void* stream) {
  if (!s || !base_pose || !init_q) return fail(HRG_EINVAL, "hrg_shield_reset: NULL argument");
  CK(cudaSetDevice(s->device));
  (void)t;  // cycle_begin_time_ is only stored by the reference (safety_shield.cc:185)
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  k_reset<<<grid_for(s->S.N, 128), 128, 0, st>>>(s->S, env_mask, base_pose, init_q, shield_type);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_shield_human_measurement(HrgShield* s, const uint8_t* env_mask, const double* keypoints, const double* t,
                                 void* stream) {
  if (!s || !keypoints || !t) return fail(HRG_EINVAL, "hrg_shield_human_measurement: NULL argument");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  k_measure<<<tile_grid(s->S.N), kMeasureWarps * 32, s->measure_smem, st>>>(s->S, env_mask, keypoints, t);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_shield_new_goal(HrgShield* s, const uint8_t* env_mask, const double* goal_q, const double* goal_dq,
                        void* stream) {
  if (!s || !goal_q) return fail(HRG_EINVAL, "hrg_shield_new_goal: NULL argument");
  CK(cudaSetDevice(s->device));
  (void)goal_dq;  // clamped and stored by the reference, never read on the step path (safety_shield.cc:682-684)
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  k_new_goal<<<grid_for(s->S.N, 128), 128, 0, st>>>(s->S, env_mask, goal_q);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

// reach + verify of the envs that do not replan this tick: the Pos-model tile kernel plus the two cascade stages over its
// survivors, or the full kernel when capsules / all distances are stored for the getters.
extern "C++" {
template <class Mark>
static void verify_pass0(HrgShield* s, cudaStream_t st, size_t smem, Mark&& mark) {
  const int N = s->S.N, J = s->S.J;
  if (s->cascade && !s->S.human_caps) {
    size_t stage_smem = (size_t)kHumDoubles * s->cfg.n_bodies * kTileEnvs * sizeof(double);
    if (stage_smem < (size_t)J * kTileEnvs * sizeof(double)) stage_smem = (size_t)J * kTileEnvs * sizeof(double);
    k_reach_verify<1, true><<<tile_grid(N), 32 * J, smem, st>>>(s->S, 0);
    mark(HRG_PROF_VERIFY_P0);
    launch_k(k_verify_stage<1>, tile_grid(N), 32 * J, stage_smem, st, s->pdl_now, s->S);
    mark(HRG_PROF_STAGE_VEL);
    launch_k(k_verify_stage<2>, tile_grid(N), 32 * J, stage_smem, st, s->pdl_now, s->S);
    mark(HRG_PROF_STAGE_ACC);
    s->launches += 2;
  } else {
    k_reach_verify<1><<<tile_grid(N), 32 * J, smem, st>>>(s->S, 0);
    mark(HRG_PROF_VERIFY_P0);
  }
}
}  // extern "C++"

// One tick's launches. `kp` != NULL folds humanMeasurement(kp, t) into the tick: the keypoints are only read by reach +
// verify (and last_meas by the path planning), so k_measure runs on the main stream next to the start of the replan chain
// instead of in front of it.
static int step_impl(HrgShield* s, const double* kp, const uint8_t* goal_mask, const double* goal_q, const double* t,
                     double* q, double* dq, double* ddq, double* dddq,
                     uint8_t* is_safe, uint8_t* branch, uint32_t* status, void* stream) {
  CK(cudaSetDevice(s->device));  // a handle on cuda:1 may be driven while another device is current
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int N = s->S.N;
  const bool prof = s->prof && s->ev_used + HRG_PROF_SLOTS + 1 <= s->ev.size();
  // profiling: everything on `st`, one event after each launch (the chain is then serialised, so the events are clean)
  auto mark = [&](int id) {
    if (s->debug_sync) {  // HRG_DEBUG_SYNC=1: wait for every kernel and name it on stderr (finds a kernel that hangs or faults)
      const cudaError_t e1 = cudaStreamSynchronize(s->side), e2 = cudaStreamSynchronize(st);
      fprintf(stderr, "[hrg] after launch slot %d: %s / %s\n", id, cudaGetErrorString(e1), cudaGetErrorString(e2));
      fflush(stderr);
    }
    if (!prof) return;
    s->ev_id[s->ev_used] = id;
    cudaEventRecord(s->ev[s->ev_used++], st);
  };
  mark(-1);
  const bool fork = s->overlap && !prof;
  const bool pdl = fork && s->pdl >= 1;  // (the profile pass records an event between the kernels: ordinary launches)
  s->pdl_now = fork && s->pdl >= 2;
  // In a tick that measures, the measurement time IS t: the path planning takes it from there instead of waiting for
  // k_measure to store it (k_measure writes the same value to last_meas)
  const double* meas_t = kp ? t : nullptr;
  cudaError_t err = cudaSuccess;
  auto keep = [&](cudaError_t e) {
    if (err == cudaSuccess && e != cudaSuccess) err = e;
  };
  if (kp && !fork) {
    k_measure<<<tile_grid(N), kMeasureWarps * 32, s->measure_smem, st>>>(s->S, nullptr, kp, t);
    mark(HRG_PROF_MEASURE);
    s->launches++;
  }
  k_pre<<<grid_for(N, kTickBlock), kTickBlock, 0, st>>>(s->S, goal_mask, goal_q);
  mark(HRG_PROF_PRE);
  if (!fork) {
    k_plan<<<grid_for(N, kTickBlock), kTickBlock, 0, st>>>(s->S, t, meas_t);
    mark(HRG_PROF_PLAN);
  }
  cudaStream_t rs = fork ? s->side : st;
  if (fork) {
    // main stream: path planning + reach / verify + determineNextMotion of the envs that do not replan;
    // side stream (high priority): the replan chain, then reach / verify + determineNextMotion of the replanned envs
    keep(cudaEventRecord(s->ev_fork, st));
    keep(cudaStreamWaitEvent(s->side, s->ev_fork, 0));
  }
  // (from the first fork to the join an early return would leave a stream un-joined for the next tick: the first error is
  // collected instead and the streams are joined in any case)
  {
    // the replan list length is only known on the device: the kernels fetch their work from device counters, their grids
    // only bound residency (one warp per block, so that the few hundred busy warps spread over all schedulers)
    const int J = s->S.J;
    const int cap = s->sm_count * 16;
    int g_plan = grid_for(N, s->plan_envs_per_warp), g_samp = grid_for(N, kSampleEnvs), g_deep = grid_for(N * J, s->deep_per_warp);
    if (g_plan > cap) g_plan = cap;
    if (g_samp > s->sm_count * HRG_SAMPLE_MINBLOCKS) g_samp = s->sm_count * HRG_SAMPLE_MINBLOCKS;
    if (g_deep > cap) g_deep = cap;
    k_replan_plan<<<g_plan, kReplanBlock, 0, rs>>>(s->S, s->plan_envs_per_warp);
    mark(HRG_PROF_PLAN_LTT);
    launch_k(k_replan_deep, g_deep, kReplanBlock, 0, rs, pdl, s->S, s->deep_per_warp);
    mark(HRG_PROF_DEEP);
    if (J == 6)
      launch_k(k_replan_sample<6>, g_samp, kSampleBlock, 0, rs, pdl, s->S);
    else if (J == 7)
      launch_k(k_replan_sample<7>, g_samp, kSampleBlock, 0, rs, pdl, s->S);
    else
      launch_k(k_replan_sample<0>, g_samp, kSampleBlock, 0, rs, pdl, s->S);
    mark(HRG_PROF_SAMPLE);
    if (s->pfl) {
      int pgrid = s->sm_count * 8;
      if (pgrid > N) pgrid = N;
      k_pfl_ltt<<<pgrid, kPflBlock, 0, rs>>>(s->S);
      mark(HRG_PROF_PFL);
      s->launches++;
    }
    {
      int g_tail = grid_for(N, 4);
      if (g_tail > s->sm_count * 16) g_tail = s->sm_count * 16;
      launch_k(k_replan_tail, g_tail, kTailBlock, 0, rs, pdl && !s->pfl, s->S, t, meas_t);
    }
    mark(HRG_PROF_TAIL);
  }
  if (fork) {
    k_plan<<<grid_for(N, kTickBlock), kTickBlock, 0, st>>>(s->S, t, meas_t);
    if (kp) {
      // k_measure: only reach + verify read the keypoints, and the path planning takes the measurement time from t, so it
      // gets a stream of its own NEXT to k_plan - launched after k_pre, the head of the chain and k_plan, which its 2 048
      // blocks would otherwise keep waiting for SM slots
      keep(cudaStreamWaitEvent(s->meas, s->ev_fork, 0));
      k_measure<<<tile_grid(N), kMeasureWarps * 32, s->measure_smem, s->meas>>>(s->S, nullptr, kp, t);
      keep(cudaEventRecord(s->ev_meas, s->meas));
      s->launches++;
    }
    if (kp) {  // reach + verify read the keypoints
      keep(cudaStreamWaitEvent(st, s->ev_meas, 0));
      keep(cudaStreamWaitEvent(s->side, s->ev_meas, 0));
    }
    verify_pass0(s, st, s->reach_smem_p0, mark);
    // determineNextMotion of the envs that did not replan
    launch_k(k_select, grid_for(N, kTickBlock), kTickBlock, 0, st, s->pdl_now && s->cascade && !s->S.human_caps, s->S, q, dq, ddq, dddq,
             is_safe, branch, status, 0);
    s->launches++;
  } else {
    verify_pass0(s, st, s->reach_smem, mark);
  }
  // the replanned envs: reach + verify pass 1 and determineNextMotion, behind the chain (fork: on its stream)
  k_reach_verify<2><<<tile_grid(N), 32 * s->S.J * 2, s->reach_smem, rs>>>(s->S, 1);
  mark(HRG_PROF_VERIFY_P1);
  launch_k(k_select, grid_for(N, kTickBlock), kTickBlock, 0, rs, pdl, s->S, q, dq, ddq, dddq, is_safe, branch, status,
           fork ? 1 : -1);
  mark(HRG_PROF_SELECT);
  if (fork) {
    keep(cudaEventRecord(s->ev_join, s->side));
    keep(cudaStreamWaitEvent(st, s->ev_join, 0));
  }
  s->launches += 9;
  keep(cudaGetLastError());
  if (err != cudaSuccess) return fail(HRG_ECUDA, "hrg_shield_step: launch failed", err);
  return HRG_OK;
}

int hrg_shield_step(HrgShield* s, const double* t, double* q, double* dq, double* ddq, double* dddq, uint8_t* is_safe,
                    uint8_t* branch, uint32_t* status, void* stream) {
  if (!s || !t) return fail(HRG_EINVAL, "hrg_shield_step: NULL argument");
  return step_impl(s, nullptr, nullptr, nullptr, t, q, dq, ddq, dddq, is_safe, branch, status, stream);
}

int hrg_shield_tick_device(HrgShield* s, const double* keypoints, const double* t, const uint8_t* goal_mask,
                           const double* goal_q, double* q, double* dq, double* ddq, double* dddq, uint8_t* is_safe,
                           uint8_t* branch, uint32_t* status, void* stream) {
  if (!s || !keypoints || !t) return fail(HRG_EINVAL, "hrg_shield_tick_device: NULL argument");
  const bool goal = goal_mask && goal_q;
  return step_impl(s, keypoints, goal ? goal_mask : nullptr, goal ? goal_q : nullptr, t, q, dq, ddq, dddq, is_safe, branch,
                   status, stream);
}

int hrg_shield_tick_host(HrgShield* s, const double* keypoints_host, const double* t_host,
                         const uint8_t* goal_mask_host, const double* goal_q_host, double* q_host, double* dq_host,
                         double* ddq_host, uint8_t* flags_host, void* stream) {
  if (!s || !keypoints_host || !t_host) return fail(HRG_EINVAL, "hrg_shield_tick_host: NULL argument");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t N = s->S.N, J = s->S.J, K3 = (size_t)s->S.K * 3;
  CK(cudaMemcpyAsync(s->d_kp, keypoints_host, N * K3 * sizeof(double), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->d_t, t_host, N * sizeof(double), cudaMemcpyHostToDevice, st));
  const bool goal = goal_mask_host && goal_q_host;
  if (goal) {
    CK(cudaMemcpyAsync(s->d_goal_mask, goal_mask_host, N, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(s->d_goal, goal_q_host, N * J * sizeof(double), cudaMemcpyHostToDevice, st));
  }
  const int rc = hrg_shield_tick_device(s, s->d_kp, s->d_t, goal ? s->d_goal_mask : nullptr, goal ? s->d_goal : nullptr,
                                        s->d_q, s->d_dq, s->d_ddq, nullptr, s->d_safe, s->d_branch, nullptr, stream);
  if (rc) return rc;
  if (q_host) CK(cudaMemcpyAsync(q_host, s->d_q, N * J * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (dq_host) CK(cudaMemcpyAsync(dq_host, s->d_dq, N * J * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (ddq_host) CK(cudaMemcpyAsync(ddq_host, s->d_ddq, N * J * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (flags_host) {
    k_pack_flags<<<grid_for((int)N, 256), 256, 0, st>>>((int)N, s->d_safe, s->d_branch, s->d_flags8);
    s->launches++;
    CK(cudaMemcpyAsync(flags_host, s->d_flags8, N, cudaMemcpyDeviceToHost, st));
  }
  return HRG_OK;
}

int hrg_shield_upload_inputs(HrgShield* s, int slot, const double* keypoints_host, const double* t_host,
                             const uint8_t* goal_mask_host, const double* goal_q_host, void* copy_stream) {
  if (!s || !keypoints_host || !t_host || slot < 0 || slot > 1)
    return fail(HRG_EINVAL, "hrg_shield_upload_inputs: bad argument");
  CK(cudaSetDevice(s->device));
  cudaStream_t cs = static_cast<cudaStream_t>(copy_stream);
  const size_t N = s->S.N, J = s->S.J, K3 = (size_t)s->S.K * 3;
  // do not overwrite the slot before the tick that consumed it last has finished with it
  if (s->slot_used[slot]) CK(cudaStreamWaitEvent(cs, s->ev_done[slot], 0));
  CK(cudaMemcpyAsync(s->d_kp2[slot], keypoints_host, N * K3 * sizeof(double), cudaMemcpyHostToDevice, cs));
  CK(cudaMemcpyAsync(s->d_t2[slot], t_host, N * sizeof(double), cudaMemcpyHostToDevice, cs));
  s->slot_has_goal[slot] = goal_mask_host && goal_q_host;
  if (s->slot_has_goal[slot]) {
    CK(cudaMemcpyAsync(s->d_goal_mask2[slot], goal_mask_host, N, cudaMemcpyHostToDevice, cs));
    CK(cudaMemcpyAsync(s->d_goal2[slot], goal_q_host, N * J * sizeof(double), cudaMemcpyHostToDevice, cs));
  }
  CK(cudaEventRecord(s->ev_up[slot], cs));
  return HRG_OK;
}

// Sparse form: only the envs whose policy period starts on this tick carry a goal (1 in 25 in human-robot-gym), so the goals
// travel as (env index, goal) pairs instead of a mask and a goal for every env - 7 % fewer bytes per tick on a link that the
// end-to-end tick saturates. goal_env_host [n_goals] i32 (distinct envs), goal_q_host [n_goals, J]; n_goals may be 0.
int hrg_shield_upload_inputs_sparse(HrgShield* s, int slot, const double* keypoints_host, const double* t_host, int n_goals,
                                    const int32_t* goal_env_host, const double* goal_q_host, void* copy_stream) {
  if (!s || !keypoints_host || !t_host || slot < 0 || slot > 1 || n_goals < 0 || n_goals > s->S.N ||
      (n_goals > 0 && (!goal_env_host || !goal_q_host)))
    return fail(HRG_EINVAL, "hrg_shield_upload_inputs_sparse: bad argument");
  CK(cudaSetDevice(s->device));
  cudaStream_t cs = static_cast<cudaStream_t>(copy_stream);
  const size_t N = s->S.N, J = s->S.J, K3 = (size_t)s->S.K * 3;
  if (s->slot_used[slot]) CK(cudaStreamWaitEvent(cs, s->ev_done[slot], 0));
  CK(cudaMemcpyAsync(s->d_kp2[slot], keypoints_host, N * K3 * sizeof(double), cudaMemcpyHostToDevice, cs));
  CK(cudaMemcpyAsync(s->d_t2[slot], t_host, N * sizeof(double), cudaMemcpyHostToDevice, cs));
  s->slot_has_goal[slot] = n_goals > 0;
  if (n_goals > 0) {
    CK(cudaMemcpyAsync(s->d_goal_idx2[slot], goal_env_host, (size_t)n_goals * sizeof(int32_t), cudaMemcpyHostToDevice, cs));
    CK(cudaMemcpyAsync(s->d_goal_sp2[slot], goal_q_host, (size_t)n_goals * J * sizeof(double), cudaMemcpyHostToDevice, cs));
    CK(cudaMemsetAsync(s->d_goal_mask2[slot], 0, N, cs));
    k_scatter_goals<<<grid_for(n_goals, 128), 128, 0, cs>>>(n_goals, (int)J, (int)N, s->d_goal_idx2[slot], s->d_goal_sp2[slot],
                                                            s->d_goal_mask2[slot], s->d_goal2[slot]);
    s->launches++;
    CK(cudaGetLastError());
  }
  CK(cudaEventRecord(s->ev_up[slot], cs));
  return HRG_OK;
}

int hrg_shield_tick_staged(HrgShield* s, int slot, double* q_host, double* dq_host, double* ddq_host,
                           uint8_t* flags_host, void* stream) {
  if (!s || slot < 0 || slot > 1) return fail(HRG_EINVAL, "hrg_shield_tick_staged: bad argument");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t N = s->S.N, J = s->S.J;
  CK(cudaStreamWaitEvent(st, s->ev_up[slot], 0));
  const bool goal = s->slot_has_goal[slot];
  const int rc = hrg_shield_tick_device(s, s->d_kp2[slot], s->d_t2[slot], goal ? s->d_goal_mask2[slot] : nullptr,
                                        goal ? s->d_goal2[slot] : nullptr, s->d_q, s->d_dq, s->d_ddq, nullptr, s->d_safe,
                                        s->d_branch, nullptr, stream);
  if (rc) return rc;
  CK(cudaEventRecord(s->ev_done[slot], st));
  s->slot_used[slot] = true;
  if (q_host) CK(cudaMemcpyAsync(q_host, s->d_q, N * J * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (dq_host) CK(cudaMemcpyAsync(dq_host, s->d_dq, N * J * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (ddq_host) CK(cudaMemcpyAsync(ddq_host, s->d_ddq, N * J * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (flags_host) {
    k_pack_flags<<<grid_for((int)N, 256), 256, 0, st>>>((int)N, s->d_safe, s->d_branch, s->d_flags8);
    s->launches++;
    CK(cudaMemcpyAsync(flags_host, s->d_flags8, N, cudaMemcpyDeviceToHost, st));
  }
  return HRG_OK;
}

}  // extern "C"
#include "hrg_shield_extra.cuh"
#include "hrg_human.cuh"
#include "hrg_noalloc.cuh"
#if defined(HRG_TASK_TRACE)
// profiling builds only (profiles/build_trace.sh): copies the task trace [4][kTraceTasks][4] u64 to the host and clears it
extern "C" int hrg_trace_read(unsigned long long* dst_host) {
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(dst_host, hrg::g_task_trace, sizeof(hrg::g_task_trace)));
  void* p = nullptr;
  CK(cudaGetSymbolAddress(&p, hrg::g_task_trace));
  CK(cudaMemset(p, 0, sizeof(hrg::g_task_trace)));
  return HRG_OK;
}
// kernel spans [16][2] u64 (earliest block start, latest block end); re-armed (start = ~0, end = 0)
extern "C" int hrg_trace_read_spans(unsigned long long* dst_host) {
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(dst_host, hrg::g_kernel_span, sizeof(hrg::g_kernel_span)));
  unsigned long long init[16][2];
  for (auto& r : init) {
    r[0] = ~0ull;
    r[1] = 0ull;
  }
  CK(cudaMemcpyToSymbol(hrg::g_kernel_span, init, sizeof(init)));
  return HRG_OK;
}
extern "C" int hrg_trace_read_counts(unsigned long long* dst_host /*[4]*/) {
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(dst_host, hrg::g_qr_count, sizeof(hrg::g_qr_count)));
  return HRG_OK;
}
#endif
