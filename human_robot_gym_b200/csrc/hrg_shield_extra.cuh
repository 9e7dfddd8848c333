// Second half of the C ABI (included at the end of hrg_shield.cu): getters, statistics, user-supplied LTT and the
// per-stage entry points used by the parity tests and by ncu.
#pragma once

namespace {

__global__ void k_copy_stats(HrgStats* dst, HrgStats* src, int clear) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    *dst = *src;
    if (clear) {
      HrgStats z;
      z.n_steps = 0; z.n_unsafe = 0; z.n_branch[0] = z.n_branch[1] = z.n_branch[2] = 0;
      z.n_replan = 0; z.n_replan_failed = 0; z.n_near_threshold = 0; z.ltt_len_sum = 0;
      z.min_dist = INFINITY;
      z.n_ltt_truncated = 0; z.n_pfl_rejected = 0;
      *src = z;
    }
  }
}

__global__ void k_capsule_markers(int n, const double* __restrict__ caps, double* __restrict__ pos,
                                  double* __restrict__ size, double* __restrict__ mat) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double c[7], p[3], sz[3], m[9];
  for (int k = 0; k < 7; ++k) c[k] = caps[(size_t)i * 7 + k];
  capsule_marker(c, p, sz, m);
  for (int k = 0; k < 3; ++k) {
    pos[(size_t)i * 3 + k] = p[k];
    size[(size_t)i * 3 + k] = sz[k];
  }
  for (int k = 0; k < 9; ++k) mat[(size_t)i * 9 + k] = m[k];
}

// LongTermTraj::getMaxAccelerationWindow / getMaxJerkWindow (SS/include/safety_shield/long_term_traj.h:243-259) of every env's
// live LTT. The reference fills max_acceleration_window_ / max_jerk_window_ for ALL samples whenever an LTT is built
// (calculate_max_acc_jerk_window, SS/src/long_term_traj.cc:79-159: entry t = max over samples [t, min(t + k, L)) of the
// signed ddq / dddq) although nothing on the step path reads them (safety_shield.cc:392); here the window that is asked for
// is evaluated on demand - a maximum is exact, so the values are bit-identical. Warp per env: lanes = (sample group,
// column of the contiguous [ddq | dddq] half of a sample), groups merged by shuffle.
__global__ void k_ltt_max_window(DevState S, const int32_t* __restrict__ index, int window_k, double* __restrict__ max_acc,
                                 double* __restrict__ max_jerk) {
  const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (e >= S.N) return;
  const int J = S.J, W = 2 * J, G = 32 / W;
  const int live = (S.flags[e] & F_LIVE_SLOT) ? 1 : 0;
  const size_t m = (size_t)live * S.N + e;
  const int len = S.ltt_len[m];
  const int k = min(len, window_k);  // long_term_traj.cc:82
  const int t = ltt_index(index ? index[e] : S.path_sd[e], S.ltt_start[m], len, S.ltt_cur[m]);
  const int g = lane / W, c = lane - g * W;
  double best = -INFINITY;
  if (g < G)
    for (int i = t + g; i < min(t + k, len); i += G) best = fmax(best, ltt_sample(S, e, live, i)[W + c]);
  for (int h = G >> 1; h >= 1; h >>= 1) {  // G is 2 for J = 6..8, 4 for J = 4, ...: fold the groups onto group 0
    const double o = __shfl_down_sync(0xffffffffu, best, h * W);
    if (g < h) best = fmax(best, o);
  }
  if (G > 1 && (G & (G - 1))) {  // G not a power of two (J = 5: G = 3): fold the last group explicitly
    const double o = __shfl_down_sync(0xffffffffu, best, (G - 1) * W);
    if (g == 0) best = fmax(best, o);
  }
  if (g == 0) {
    if (c < J) {
      if (max_acc) max_acc[(size_t)e * J + c] = best;
    } else if (max_jerk) {
      max_jerk[(size_t)e * J + (c - J)] = best;
    }
  }
}

// SafetyShield::setLongTermTrajectory (SS/src/safety_shield.cc:699-730) for one env; traj already on the device.
__device__ __forceinline__ void set_ltt_block(const DevState& S, int e, const double* __restrict__ traj, int length,
                                              double sample_time, int starting_index, const double* __restrict__ alpha,
                                              int32_t* verdict) {
  __shared__ int s_verdict;
  const HrgConfig& c = *S.cfg;
  const int J = S.J, N = S.N;
  if (threadIdx.x == 0) {
    uint32_t f = S.flags[e];
    const int live = (f & F_LIVE_SLOT) ? 1 : 0;
    PathS p;
    path_load(S, (f & F_REC_CURRENT) ? P_REC : P_FS, e, p);
    double q[HRG_MAX_JOINTS], dq[HRG_MAX_JOINTS], ddq[HRG_MAX_JOINTS], dddq[HRG_MAX_JOINTS];
    ltt_interpolate(S, c, e, live, p.pos, p.vel, p.acc, p.jerk, q, dq, ddq, dddq);
    int v = 0;
    for (int j = 0; j < J && v == 0; ++j)
      if (fabs(dq[j]) > 1e-4 || fabs(ddq[j]) > 1e-4 || fabs(dddq[j]) > 1e-4) v = 1;  // RobotMovementException
    const double* first = traj + (size_t)ltt_index(0, starting_index, length, 0) * 4 * J;
    const double* last = traj + (size_t)ltt_index(length - 1, starting_index, length, 0) * 4 * J;
    for (int j = 0; j < J && v == 0; ++j)
      if (fabs(q[j] - first[j]) > 1e-3) v = 2;  // hasSamePos (motion.h:151-160)
    for (int j = 0; j < J && v == 0; ++j)
      if (fabs(dq[j] - first[J + j]) > 1e-4) v = 3;  // hasSameVel
    for (int j = 0; j < J && v == 0; ++j)
      if (fabs(last[J + j]) > 1e-4 || fabs(last[2 * J + j]) > 1e-4 || fabs(last[3 * J + j]) > 1e-4) v = 4;
    if (v == 0) {
      const int pend = 1 - live;
      S.ltt_len[(size_t)pend * N + e] = length;
      S.ltt_cur[(size_t)pend * N + e] = 0;
      S.ltt_start[(size_t)pend * N + e] = starting_index;
      S.ltt_t0[(size_t)pend * N + e] = 0.0;
      S.ltt_ts[(size_t)pend * N + e] = sample_time;
      for (int j = 0; j < J; ++j) S.ltt_alpha[((size_t)pend * J + j) * N + e] = alpha[j];
      f |= F_NEW_LTT | F_NEW_LTT_PROCESSED;
      S.flags[e] = f;
    }
    s_verdict = v;
    *verdict = v;
  }
  __syncthreads();
  if (s_verdict == 0) {
    const int pend = (S.flags[e] & F_LIVE_SLOT) ? 0 : 1;
    double* dst = ltt_sample_w(S, e, pend, 0);
    for (int i = threadIdx.x; i < length * 4 * J; i += blockDim.x) dst[i] = traj[i];
    if (S.ltt_cart)  // a user-built LongTermTraj comes from the SSM constructor: no Cartesian velocities (long_term_traj.h:108-116)
      for (int i = threadIdx.x; i < length; i += blockDim.x) {
        S.ltt_cart[((size_t)e * 2 + pend) * S.L + i] = 0.0;
        S.ltt_time[((size_t)e * 2 + pend) * S.L + i] = i * sample_time;
      }
  }
}

__global__ void k_set_ltt(DevState S, int e, const double* __restrict__ traj, int length, double sample_time,
                          int starting_index, const double* __restrict__ alpha, int32_t* verdict) {
  set_ltt_block(S, e, traj, length, sample_time, starting_index, alpha, verdict);
}

// One block per user trajectory: block m installs traj[m] into env env_index[m] (the envs must be distinct).
__global__ void k_set_ltt_batch(DevState S, int max_len, const int32_t* __restrict__ env_index,
                                const double* __restrict__ traj, const int32_t* __restrict__ length, double sample_time,
                                const int32_t* __restrict__ starting_index, const double* __restrict__ alpha,
                                int32_t* __restrict__ verdict) {
  const int m = blockIdx.x;
  const int e = env_index[m], len = length[m];
  if (e < 0 || e >= S.N || len < 1 || len > S.L || len > max_len) {  // block-uniform
    if (threadIdx.x == 0) verdict[m] = -1;
    return;
  }
  set_ltt_block(S, e, traj + (size_t)m * max_len * 4 * S.J, len, sample_time, starting_index ? starting_index[m] : 0,
                alpha + (size_t)m * S.J, verdict + m);
}
}  // namespace

extern "C" {

int hrg_shield_store_capsules(HrgShield* s, int on) {
  if (!s) return fail(HRG_EINVAL, "hrg_shield_store_capsules: NULL handle");
  CK(cudaSetDevice(s->device));
  if (on && !s->S.robot_caps) {
    int rc = s->alloc(&s->S.robot_caps, (size_t)s->S.N * s->S.J * 7);
    if (rc) return rc;
    rc = s->alloc(&s->S.human_caps, (size_t)s->S.N * s->S.H * 7);
    if (rc) return rc;
  }
  return HRG_OK;
}

int hrg_shield_get_capsules(HrgShield* s, double* robot, double* human, double* min_dist, void* stream) {
  if (!s) return fail(HRG_EINVAL, "hrg_shield_get_capsules: NULL handle");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if ((robot || human) && !s->S.robot_caps)
    return fail(HRG_EINVAL, "capsules are not stored: call hrg_shield_store_capsules(s, 1) before stepping");
  const size_t N = s->S.N;
  if (robot) CK(cudaMemcpyAsync(robot, s->S.robot_caps, N * s->S.J * 7 * sizeof(double), cudaMemcpyDeviceToDevice, st));
  if (human) CK(cudaMemcpyAsync(human, s->S.human_caps, N * s->S.H * 7 * sizeof(double), cudaMemcpyDeviceToDevice, st));
  if (min_dist) {
    k_min_dist_out<<<grid_for((int)N, 256), 256, 0, st>>>(s->S, min_dist);
    s->launches++;
    CK(cudaGetLastError());
  }
  return HRG_OK;
}

int hrg_shield_stats(HrgShield* s, HrgStats* out, int clear, void* stream) {
  if (!s || !out) return fail(HRG_EINVAL, "hrg_shield_stats: NULL argument");
  CK(cudaSetDevice(s->device));
  k_copy_stats<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(out, s->S.stats, clear);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_shield_set_ltt(HrgShield* s, int env, const double* traj_host, int length, double sample_time,
                       int starting_index, const double* alpha_i_host, int32_t* verdict_host) {
  if (!s || !traj_host || !alpha_i_host || !verdict_host) return fail(HRG_EINVAL, "hrg_shield_set_ltt: NULL argument");
  if (env < 0 || env >= s->S.N) return fail(HRG_EINVAL, "hrg_shield_set_ltt: env out of range");
  if (length < 1 || length > s->S.L) return fail(HRG_EINVAL, "hrg_shield_set_ltt: length exceeds ltt_capacity");
  if (!(sample_time > 0)) return fail(HRG_EINVAL, "hrg_shield_set_ltt: sample_time must be > 0");
  CK(cudaSetDevice(s->device));
  const int J = s->S.J;
  double* d = nullptr;
  const size_t nb = ((size_t)length * 4 * J + J) * sizeof(double);
  CK(cudaMalloc(&d, nb + sizeof(int32_t) * 2));
  cudaError_t ce = cudaMemcpy(d, traj_host, (size_t)length * 4 * J * sizeof(double), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess)
    ce = cudaMemcpy(d + (size_t)length * 4 * J, alpha_i_host, J * sizeof(double), cudaMemcpyHostToDevice);
  int32_t* dv = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(d) + nb);
  if (ce == cudaSuccess) {
    k_set_ltt<<<1, 128>>>(s->S, env, d, length, sample_time, starting_index, d + (size_t)length * 4 * J, dv);
    s->launches++;
    ce = cudaMemcpy(verdict_host, dv, sizeof(int32_t), cudaMemcpyDeviceToHost);
  }
  cudaFree(d);
  if (ce != cudaSuccess) return fail(HRG_ECUDA, "hrg_shield_set_ltt", ce);
  return HRG_OK;
}

int hrg_shield_set_ltt_batch(HrgShield* s, int n_traj, const int32_t* env_index, const double* traj, int max_len,
                             const int32_t* length, double sample_time, const int32_t* starting_index,
                             const double* alpha_i, int32_t* verdict, void* stream) {
  if (!s || !env_index || !traj || !length || !alpha_i || !verdict)
    return fail(HRG_EINVAL, "hrg_shield_set_ltt_batch: NULL argument");
  CK(cudaSetDevice(s->device));
  if (n_traj < 0 || max_len < 1) return fail(HRG_EINVAL, "hrg_shield_set_ltt_batch: n_traj < 0 or max_len < 1");
  if (!(sample_time > 0)) return fail(HRG_EINVAL, "hrg_shield_set_ltt_batch: sample_time must be > 0");
  if (n_traj == 0) return HRG_OK;
  k_set_ltt_batch<<<n_traj, 128, 0, static_cast<cudaStream_t>(stream)>>>(s->S, max_len, env_index, traj, length, sample_time,
                                                                         starting_index, alpha_i, verdict);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_shield_profile(HrgShield* s, int max_steps) {
  if (!s || max_steps < 0) return fail(HRG_EINVAL, "hrg_shield_profile: bad argument");
  CK(cudaSetDevice(s->device));
  while (s->ev.size() < (size_t)max_steps * (HRG_PROF_SLOTS + 1)) {
    cudaEvent_t e;
    CK(cudaEventCreate(&e));
    s->ev.push_back(e);
  }
  s->ev_id.resize(s->ev.size(), -1);
  s->prof = max_steps > 0;
  s->ev_used = 0;
  return HRG_OK;
}

int hrg_shield_profile_read_kernels(HrgShield* s, double* ms_host, int* steps_host) {
  if (!s || !ms_host || !steps_host) return fail(HRG_EINVAL, "hrg_shield_profile_read_kernels: NULL argument");
  CK(cudaSetDevice(s->device));
  for (int k = 0; k < HRG_PROF_SLOTS; ++k) ms_host[k] = 0.0;
  int steps = 0;
  if (s->ev_used) CK(cudaEventSynchronize(s->ev[s->ev_used - 1]));
  for (size_t i = 0; i < s->ev_used; ++i) {
    const int id = s->ev_id[i];
    if (id < 0) {
      ++steps;
      continue;
    }
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, s->ev[i - 1], s->ev[i]));
    if (id < HRG_PROF_SLOTS) ms_host[id] += ms;
  }
  *steps_host = steps;
  s->ev_used = 0;
  return HRG_OK;
}

int hrg_shield_profile_read(HrgShield* s, double* ms_host, int* steps_host) {
  if (!s || !ms_host || !steps_host) return fail(HRG_EINVAL, "hrg_shield_profile_read: NULL argument");
  double k[HRG_PROF_SLOTS];
  const int rc = hrg_shield_profile_read_kernels(s, k, steps_host);
  if (rc) return rc;
  ms_host[0] = k[HRG_PROF_MEASURE] + k[HRG_PROF_PRE] + k[HRG_PROF_PLAN];
  ms_host[1] = k[HRG_PROF_PLAN_LTT] + k[HRG_PROF_DEEP] + k[HRG_PROF_SAMPLE] + k[HRG_PROF_PFL] + k[HRG_PROF_TAIL];
  ms_host[2] = k[HRG_PROF_VERIFY_P0] + k[HRG_PROF_STAGE_VEL] + k[HRG_PROF_STAGE_ACC] + k[HRG_PROF_VERIFY_P1];
  ms_host[3] = k[HRG_PROF_SELECT];
  return HRG_OK;
}

// ---- FP64 peak microbenchmark (roofline denominator of the FP64-bound kernels) -----------------------------------------
int hrg_microbench_fp64(int device, int mode, int iters, double* gflops_host, double* ginst_host, double* ms_host) {
  if (mode < 0 || mode > 2 || iters < 1 || !gflops_host) return fail(HRG_EINVAL, "hrg_microbench_fp64: bad argument");
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev < 1) return fail(HRG_ENODEVICE, "no CUDA device (this library has no CPU fallback)", ce);
  if (device < 0 || device >= ndev) return fail(HRG_EINVAL, "device index out of range");
  CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  const int block = 256, grid = prop.multiProcessorCount * 8;  // 2048 threads per SM: every scheduler has 16 warps to pick from
  double* out = nullptr;
  CK(cudaMalloc(&out, (size_t)grid * block * sizeof(double)));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {  // first launch = warm-up, then best of three
    cudaEventRecord(e0, 0);
    if (mode == 0) hrg::k_fp64_peak<0><<<grid, block>>>(out, iters, 1.0000001, 1e-9);
    if (mode == 1) hrg::k_fp64_peak<1><<<grid, block>>>(out, iters, 1.0000001, 1e-9);
    if (mode == 2) hrg::k_fp64_peak<2><<<grid, block>>>(out, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1, 0);
    ce = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
    if (ce != cudaSuccess) break;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  if (ce != cudaSuccess) return fail(HRG_ECUDA, "hrg_microbench_fp64", ce);
  // per thread and iteration: 4 x 8 chain steps, each one DFMA (2 flop) / one DMUL + one DADD (2 flop) / one DADD (1 flop)
  const double steps = (double)grid * block * (double)iters * 32.0;
  const double inst = steps * (mode == 1 ? 2.0 : 1.0);
  const double flop = steps * (mode == 2 ? 1.0 : 2.0);
  *gflops_host = flop / (best * 1e-3) / 1e9;
  if (ginst_host) *ginst_host = inst / (best * 1e-3) / 1e9;  // thread-level FP64 instructions per second / 1e9
  if (ms_host) *ms_host = best;
  return HRG_OK;
}

int hrg_microbench_fp64_latency(int device, double* cycles_host /*[6]*/) {
  if (!cycles_host) return fail(HRG_EINVAL, "hrg_microbench_fp64_latency: NULL argument");
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev < 1) return fail(HRG_ENODEVICE, "no CUDA device (this library has no CPU fallback)", ce);
  if (device < 0 || device >= ndev) return fail(HRG_EINVAL, "device index out of range");
  CK(cudaSetDevice(device));
  double* out = nullptr;
  long long* cyc = nullptr;
  CK(cudaMalloc(&out, 32 * sizeof(double)));
  CK(cudaMalloc(&cyc, sizeof(long long)));
  const int iters = 256;
  for (int mode = 0; mode < 6 && ce == cudaSuccess; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {  // the second launch has the code in the instruction cache
      if (mode == 0) hrg::k_fp64_latency<0><<<1, 32>>>(out, cyc, iters, 1.0000001, 1e-9);
      if (mode == 1) hrg::k_fp64_latency<1><<<1, 32>>>(out, cyc, iters, 1.0000001, 1e-9);
      if (mode == 2) hrg::k_fp64_latency<2><<<1, 32>>>(out, cyc, iters, 1.0000001, 1e-9);
      if (mode == 3) hrg::k_fp64_latency<3><<<1, 32>>>(out, cyc, iters, 1.0000001, 1e-9);
      if (mode == 4) hrg::k_fp64_latency<4><<<1, 32>>>(out, cyc, iters, 1.0000001, 1e-9);
      if (mode == 5) hrg::k_fp64_latency<5><<<1, 32>>>(out, cyc, iters, 1.0000001, 1e-9);
    }
    long long h = 0;
    ce = cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    cycles_host[mode] = (double)h / (iters * 16.0);
  }
  cudaFree(out);
  cudaFree(cyc);
  if (ce != cudaSuccess) return fail(HRG_ECUDA, "hrg_microbench_fp64_latency", ce);
  return HRG_OK;
}

// ---- checkpoint (state_dict / load_state_dict of the new framework; no counterpart in the reference) ------
// Blob = header {magic, N, J, K, L, pfl, n_segments, bytes} + the persistent arrays back to back.
struct StateHeader {
  uint64_t magic;
  int32_t N, J, K, L, pfl, n_seg;
  int64_t bytes;
};
constexpr uint64_t kStateMagic = 0x4852475348314c44ull;  // "HRGSH1LD"

int64_t hrg_shield_state_bytes(const HrgShield* s) {
  if (!s) return 0;
  int64_t b = sizeof(StateHeader);
  for (const auto& seg : s->persistent) b += (int64_t)seg.second;
  return b;
}

int hrg_shield_state_save(HrgShield* s, void* dst_host, int64_t bytes) {
  if (!s || !dst_host) return fail(HRG_EINVAL, "hrg_shield_state_save: NULL argument");
  if (bytes < hrg_shield_state_bytes(s)) return fail(HRG_EINVAL, "hrg_shield_state_save: buffer smaller than hrg_shield_state_bytes");
  CK(cudaSetDevice(s->device));
  CK(cudaDeviceSynchronize());  // between ticks: nothing of this handle may be in flight
  StateHeader h{kStateMagic, s->S.N, s->S.J, s->S.K, s->S.L, s->pfl ? 1 : 0, (int32_t)s->persistent.size(),
                hrg_shield_state_bytes(s)};
  char* p = static_cast<char*>(dst_host);
  memcpy(p, &h, sizeof h);
  p += sizeof h;
  for (const auto& seg : s->persistent) {
    CK(cudaMemcpy(p, seg.first, seg.second, cudaMemcpyDeviceToHost));
    p += seg.second;
  }
  return HRG_OK;
}

int hrg_shield_state_load(HrgShield* s, const void* src_host, int64_t bytes) {
  if (!s || !src_host) return fail(HRG_EINVAL, "hrg_shield_state_load: NULL argument");
  StateHeader h;
  if (bytes < (int64_t)sizeof h) return fail(HRG_EINVAL, "hrg_shield_state_load: truncated blob");
  memcpy(&h, src_host, sizeof h);
  if (h.magic != kStateMagic) return fail(HRG_EINVAL, "hrg_shield_state_load: not a shield state blob");
  if (h.pfl && !s->pfl) {
    if (int rc = hrg_shield_enable_pfl(s)) return rc;
  }
  if (h.N != s->S.N || h.J != s->S.J || h.K != s->S.K || h.L != s->S.L || h.pfl != (s->pfl ? 1 : 0) ||
      h.n_seg != (int32_t)s->persistent.size() || h.bytes != hrg_shield_state_bytes(s) || bytes < h.bytes)
    return fail(HRG_EINVAL, "hrg_shield_state_load: blob was saved from a handle of another shape (envs, joints, keypoints, "
                            "ltt_capacity or PFL capability differ)");
  CK(cudaSetDevice(s->device));
  CK(cudaDeviceSynchronize());
  const char* p = static_cast<const char*>(src_host) + sizeof h;
  for (const auto& seg : s->persistent) {
    CK(cudaMemcpy(seg.first, p, seg.second, cudaMemcpyHostToDevice));
    p += seg.second;
  }
  return HRG_OK;
}

// ---- per-stage entry points ------------------------------------------------------------------------------------
int hrg_stage_capsule_dist(int n, const double* c1, const double* c2, double* dist, void* stream) {
  if (n < 0 || !c1 || !c2 || !dist) return fail(HRG_EINVAL, "hrg_stage_capsule_dist: bad argument");
  if (n == 0) return HRG_OK;
  k_stage_capsule_dist<<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(n, c1, c2, dist);
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_stage_plan_safety_shield(double sample_time, int n, const double* in, uint8_t* ok, double* phases, double* fin,
                                 void* stream) {
  if (n < 0 || !in || !ok || !phases || !fin) return fail(HRG_EINVAL, "hrg_stage_plan_safety_shield: bad argument");
  if (n == 0) return HRG_OK;
  k_stage_plan_safety_shield<<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(sample_time, n, in, ok,
                                                                                               phases, fin);
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_stage_ltp_plan(const HrgConfig* cfg, int n, const double* goal, const double* q0, const double* v0,
                       const double* a0, double* traj, int32_t* length, int cap, void* stream) {
  int rc = cfg_check(cfg);
  if (rc) return rc;
  if (n < 0 || !goal || !q0 || !v0 || !a0 || !traj || !length || cap < 1)
    return fail(HRG_EINVAL, "hrg_stage_ltp_plan: bad argument");
  if (n == 0) return HRG_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HrgConfig* d = nullptr;
  CK(cudaMalloc(&d, sizeof(HrgConfig)));
  cudaError_t ce = cudaMemcpyAsync(d, cfg, sizeof(HrgConfig), cudaMemcpyHostToDevice, st);
  if (ce == cudaSuccess) {
    k_stage_ltp_plan<<<grid_for(n, 64), 64, 0, st>>>(d, n, goal, q0, v0, a0, traj, length, cap);
    ce = cudaStreamSynchronize(st);
  }
  cudaFree(d);
  if (ce != cudaSuccess) return fail(HRG_ECUDA, "hrg_stage_ltp_plan", ce);
  return HRG_OK;
}

int hrg_stage_reach_verify(const HrgConfig* cfg, int n, const double* base_pose, const double* q_start,
                           const double* q_goal, const double* s_diff, const double* t_reach, const double* alpha_i,
                           const double* kp_pos, const double* kp_vel, double* robot, double* human, double* min_dist,
                           uint8_t* is_safe, void* stream) {
  int rc = cfg_check(cfg);
  if (rc) return rc;
  if (n < 0 || !base_pose || !q_start || !q_goal || !s_diff || !t_reach || !alpha_i || !kp_pos || !kp_vel)
    return fail(HRG_EINVAL, "hrg_stage_reach_verify: bad argument");
  if (n == 0) return HRG_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // scratch field-major state for this call: [cfg | base_T 12 | reach_in 3J+2 | kp 2*K3 | min_dist 3] x n doubles + flags
  const int J = cfg->nb_joints, K3 = cfg->n_keypoints * 3, H = cfg->n_ext + 2 * cfg->n_bodies;
  const size_t per_env = 12 + reach_fields(J) + 2 * (size_t)K3 + 3;
  const size_t cfg_doubles = (sizeof(HrgConfig) + 7) / 8;
  double* d = nullptr;
  CK(cudaMalloc(&d, (cfg_doubles + per_env * n) * sizeof(double) + (size_t)n * sizeof(uint32_t) + 16));
  DevState S;
  memset(&S, 0, sizeof(S));
  S.N = n; S.J = J; S.K = cfg->n_keypoints; S.H = H; S.L = 0;
  S.cfg = reinterpret_cast<const HrgConfig*>(d);
  double* p = d + cfg_doubles;
  S.base_T = p; p += (size_t)12 * n;
  S.reach_in = p; p += (size_t)reach_fields(J) * n;
  S.kp_pos = p; p += (size_t)K3 * n;
  S.kp_vel = p; p += (size_t)K3 * n;
  S.min_dist = p; p += (size_t)3 * n;
  S.flags = reinterpret_cast<uint32_t*>(p);
  S.robot_caps = robot;
  S.human_caps = human;
  cudaError_t ce = cudaMemcpyAsync(d, cfg, sizeof(HrgConfig), cudaMemcpyHostToDevice, st);
  if (ce == cudaSuccess) {
    const size_t smem = (size_t)reach_smem_doubles(J, H) * sizeof(double);
    cudaFuncSetAttribute(k_reach_verify<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_stage_pack<<<grid_for(n, 128), 128, 0, st>>>(S, base_pose, q_start, q_goal, s_diff, t_reach, alpha_i, kp_pos, kp_vel);
    k_reach_verify<1><<<tile_grid(n), 32 * J, smem, st>>>(S, 0);
    k_stage_unpack<<<grid_for(n, 128), 128, 0, st>>>(S, min_dist, is_safe);
    ce = cudaStreamSynchronize(st);
  }
  cudaFree(d);
  if (ce != cudaSuccess) return fail(HRG_ECUDA, "hrg_stage_reach_verify", ce);
  return HRG_OK;
}

// ---- batched FailsafeController -------------------------------------------------------------------
static int ctrl_check(const HrgShield* s, const HrgControllerConfig* k, const char* who) {
  if (!s || !k) return fail(HRG_EINVAL, who);
  if (k->nb_joints != s->S.J) return fail(HRG_EINVAL, "controller nb_joints differs from the shield's");
  return HRG_OK;
}

int hrg_capsule_markers(int n, const double* caps, double* pos, double* size, double* mat, void* stream) {
  if (n < 0 || !caps || !pos || !size || !mat) return fail(HRG_EINVAL, "hrg_capsule_markers: bad argument");
  if (n == 0) return HRG_OK;
  k_capsule_markers<<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(n, caps, pos, size, mat);
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_shield_ltt_max_window(HrgShield* s, const int32_t* index, int window_k, double* max_acc, double* max_jerk,
                              void* stream) {
  if (!s) return fail(HRG_EINVAL, "hrg_shield_ltt_max_window: NULL handle");
  CK(cudaSetDevice(s->device));
  if (window_k <= 0) window_k = (int)floor(s->cfg.max_s_stop / s->cfg.sample_time);  // sliding_window_k_ (safety_shield.cc:88)
  k_ltt_max_window<<<grid_for(s->S.N, 4), 128, 0, static_cast<cudaStream_t>(stream)>>>(s->S, index, window_k, max_acc, max_jerk);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_controller_set_goal(HrgShield* s, const HrgControllerConfig* cfg, const uint8_t* env_mask, const double* action,
                            const double* joint_pos, const double* set_qpos, double* goal_out, void* stream) {
  if (int rc = ctrl_check(s, cfg, "hrg_controller_set_goal: NULL argument")) return rc;
  if (!set_qpos && (!action || !joint_pos)) return fail(HRG_EINVAL, "hrg_controller_set_goal: action / joint_pos NULL");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  k_ctrl_set_goal<<<grid_for(s->S.N, 128), 128, 0, st>>>(s->S, *cfg, env_mask, action, joint_pos, set_qpos, goal_out);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_controller_torques(HrgShield* s, const HrgControllerConfig* cfg, const double* q_des, const double* dq_des,
                           const double* ddq_des, const double* joint_pos, const double* joint_vel,
                           const double* mass_matrix, const double* torque_comp, double* torques, void* stream) {
  if (int rc = ctrl_check(s, cfg, "hrg_controller_torques: NULL argument")) return rc;
  if (!q_des || !dq_des || !ddq_des || !joint_pos || !joint_vel || !mass_matrix || !torque_comp || !torques)
    return fail(HRG_EINVAL, "hrg_controller_torques: NULL argument");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  k_ctrl_torques<<<grid_for(s->S.N, 128), 128, 0, st>>>(s->S.N, s->S.J, *cfg, q_des, dq_des, ddq_des, joint_pos, joint_vel,
                                                        mass_matrix, torque_comp, torques);
  s->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

}  // extern "C"
