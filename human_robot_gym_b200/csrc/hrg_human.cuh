// C ABI of the animation -> keypoints path (include/hrg_shield.h, "animation -> keypoints"). Included by hrg_shield.cu.
#pragma once
#include "kernels_human.cuh"

struct HrgHuman {
  hrg::HumanDev H;
  hrg::HumanTreeDev T;
  int device = 0;
  long long launches = 0;
  int64_t* d_clock = nullptr;  // staging of hrg_shield_tick_replay_host
  std::vector<void*> allocs;

  template <class X>
  int upload(X** p, const X* host, size_t n, size_t n_alloc = 0) {
    void* q = nullptr;
    const size_t bytes = (n_alloc > n ? n_alloc : n) * sizeof(X) + 16;
    cudaError_t e = cudaMalloc(&q, bytes);
    if (e != cudaSuccess) return fail(HRG_ENOMEM, "cudaMalloc", e);
    allocs.push_back(q);
    e = cudaMemset(q, 0, bytes);
    if (e == cudaSuccess && host && n) e = cudaMemcpy(q, host, n * sizeof(X), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return fail(HRG_ECUDA, "cudaMemcpy", e);
    *p = static_cast<X*>(q);
    return HRG_OK;
  }
};

extern "C" {

int hrg_human_create(int n_bodies, const int32_t* parent_host, const double* anchor_host, int n_clips,
                     const int32_t* clip_len_host, const double* frames_host, const double* clip_info_host,
                     double step_length, int n_envs, int list_len, int device, HrgHuman** out) {
  if (!out) return fail(HRG_EINVAL, "hrg_human_create: out is NULL");
  *out = nullptr;
  if (!parent_host || !anchor_host || !clip_len_host || !frames_host || !clip_info_host)
    return fail(HRG_EINVAL, "hrg_human_create: NULL argument");
  if (n_bodies < 1 || n_bodies > HRG_MAX_KEYPOINTS) return fail(HRG_EINVAL, "hrg_human_create: n_bodies out of range");
  if (n_clips < 1 || n_envs < 1 || list_len < 1) return fail(HRG_EINVAL, "hrg_human_create: n_clips / n_envs / list_len < 1");
  if (!(step_length >= 1.0)) return fail(HRG_EINVAL, "hrg_human_create: step_length must be >= 1 (human_env.py:1466)");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(HRG_ENODEVICE, "hrg_human_create: no CUDA device (there is no CPU fallback)");
  }
  if (device < 0 || device >= ndev) return fail(HRG_EINVAL, "hrg_human_create: device out of range");
  CK(cudaSetDevice(device));
  HrgHuman* h = new (std::nothrow) HrgHuman();
  if (!h) return fail(HRG_ENOMEM, "hrg_human_create: host allocation");
  h->device = device;
  int rc = HRG_OK;
  switch (hrg::human_build_tree(n_bodies, parent_host, anchor_host, &h->T)) {
    case 0: break;
    case -1: rc = fail(HRG_EINVAL, "hrg_human_create: parents must precede their children"); break;
    case -2: rc = fail(HRG_EINVAL, "hrg_human_create: tree needs more parked frames than kHumanSlots"); break;
    default: rc = fail(HRG_EINVAL, "hrg_human_create: tree is not connected to the Pelvis");
  }
  std::vector<int64_t> off(n_clips);
  std::vector<double> tab((size_t)n_clips * 12);
  int64_t rows = 0;
  for (int c = 0; c < n_clips && rc == HRG_OK; ++c) {
    if (clip_len_host[c] < 1) rc = fail(HRG_EINVAL, "hrg_human_create: empty clip");
    off[c] = rows;
    rows += clip_len_host[c];
    hrg::human_clip_table(clip_info_host + (size_t)c * 7, &tab[(size_t)c * 12]);
  }
  hrg::HumanDev& H = h->H;
  H.N = n_envs; H.K = n_bodies; H.C = 7 + 3 * n_bodies; H.M = list_len; H.n_clips = n_clips;
  H.step_length = step_length;
  double* d_baked = nullptr; int64_t* d_off = nullptr; int32_t* d_len = nullptr; double* d_tab = nullptr;
  if (rc == HRG_OK) rc = h->upload<double>(&d_baked, nullptr, 0, (size_t)rows * (3 + 3 * n_bodies));
  if (rc == HRG_OK) rc = h->upload(&d_off, off.data(), (size_t)n_clips);
  if (rc == HRG_OK) rc = h->upload(&d_len, clip_len_host, (size_t)n_clips);
  if (rc == HRG_OK) rc = h->upload(&d_tab, tab.data(), tab.size());
  if (rc == HRG_OK) {
    std::vector<double> place((size_t)n_envs * 12, 0.0);
    for (int e = 0; e < n_envs; ++e) place[(size_t)e * 12 + 3] = place[(size_t)e * 12 + 7] = place[(size_t)e * 12 + 11] = 1.0;
    rc = h->upload(&H.placement, place.data(), place.size());
  }
  if (rc == HRG_OK) {  // bake the bank: the tree walk runs once per frame here, not once per env per tick
    std::vector<int32_t> row_clip((size_t)rows);
    for (int c = 0; c < n_clips; ++c)
      for (int64_t r = off[c]; r < off[c] + clip_len_host[c]; ++r) row_clip[(size_t)r] = c;
    double* d_raw = nullptr;
    int32_t* d_row_clip = nullptr;
    cudaError_t e = cudaMalloc(&d_raw, (size_t)rows * H.C * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&d_row_clip, (size_t)rows * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMemcpy(d_raw, frames_host, (size_t)rows * H.C * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_row_clip, row_clip.data(), (size_t)rows * sizeof(int32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
      hrg::k_human_bake<<<(unsigned)((rows + 127) / 128), 128>>>(h->T, H.C, rows, d_raw, d_row_clip, d_tab, d_baked);
      h->launches++;
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaFree(d_raw);
    cudaFree(d_row_clip);
    if (e != cudaSuccess) rc = fail(HRG_ECUDA, "hrg_human_create: baking the clip bank", e);
  }
  if (rc == HRG_OK) rc = h->upload<int32_t>(&H.clip_list, nullptr, 0, (size_t)n_envs * list_len);
  if (rc == HRG_OK) rc = h->upload<int32_t>(&H.list_idx, nullptr, 0, (size_t)n_envs);
  if (rc == HRG_OK) rc = h->upload<int64_t>(&H.anim_start, nullptr, 0, (size_t)n_envs);
  if (rc == HRG_OK) rc = h->upload<int64_t>(&h->d_clock, nullptr, 0, (size_t)n_envs);
  H.baked = d_baked; H.clip_off = d_off; H.clip_len = d_len; H.clip_tab = d_tab;
  if (rc != HRG_OK) {
    for (void* p : h->allocs) cudaFree(p);
    delete h;
    return rc;
  }
  *out = h;
  return HRG_OK;
}

int hrg_human_destroy(HrgHuman* h) {
  if (!h) return HRG_OK;
  cudaSetDevice(h->device);
  for (void* p : h->allocs) cudaFree(p);
  delete h;
  return HRG_OK;
}

int hrg_human_reset(HrgHuman* h, const uint8_t* env_mask, const double* placement, const int32_t* clip_list,
                    void* stream) {
  if (!h || !placement || !clip_list) return fail(HRG_EINVAL, "hrg_human_reset: NULL argument");
  CK(cudaSetDevice(h->device));
  hrg::k_human_reset<<<grid_for(h->H.N, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(h->H, env_mask, placement,
                                                                                             clip_list);
  h->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_human_keypoints(HrgHuman* h, const int64_t* low_level_time, double* keypoints, int32_t* clip_out,
                        int32_t* frame_out, void* stream) {
  if (!h || !low_level_time || !keypoints) return fail(HRG_EINVAL, "hrg_human_keypoints: NULL argument");
  CK(cudaSetDevice(h->device));
  hrg::k_human_keypoints<<<grid_for(h->H.N, hrg::kHumanTileEnvs), hrg::kHumanBlock, 0,
                           static_cast<cudaStream_t>(stream)>>>(h->H, low_level_time, keypoints, clip_out, frame_out);
  h->launches++;
  CK(cudaGetLastError());
  return HRG_OK;
}

int hrg_shield_tick_replay_host(HrgShield* s, HrgHuman* h, const int64_t* low_level_time_host, const double* t_host,
                                const uint8_t* goal_mask_host, const double* goal_q_host, double* q_host, double* dq_host,
                                double* ddq_host, uint8_t* flags_host, void* stream) {
  if (!s || !h || !low_level_time_host || !t_host) return fail(HRG_EINVAL, "hrg_shield_tick_replay_host: NULL argument");
  if (h->H.N != s->S.N || h->H.K != s->S.K || h->device != s->device)
    return fail(HRG_EINVAL, "hrg_shield_tick_replay_host: the human batch and the shield batch differ (envs, keypoints or device)");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t N = s->S.N, J = s->S.J;
  CK(cudaMemcpyAsync(h->d_clock, low_level_time_host, N * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->d_t, t_host, N * sizeof(double), cudaMemcpyHostToDevice, st));
  const bool goal = goal_mask_host && goal_q_host;
  if (goal) {
    CK(cudaMemcpyAsync(s->d_goal_mask, goal_mask_host, N, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(s->d_goal, goal_q_host, N * J * sizeof(double), cudaMemcpyHostToDevice, st));
  }
  int rc = hrg_human_keypoints(h, h->d_clock, s->d_kp, nullptr, nullptr, stream);
  if (rc) return rc;
  rc = hrg_shield_tick_device(s, s->d_kp, s->d_t, goal ? s->d_goal_mask : nullptr, goal ? s->d_goal : nullptr, s->d_q,
                              s->d_dq, s->d_ddq, nullptr, s->d_safe, s->d_branch, nullptr, stream);
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

// Pipelined form of hrg_shield_tick_replay_host: the tick's kernels run on `stream`, its setpoints go into output slot 0/1 and
// travel to the host on `copy_stream`, so the device->host copy of tick k overlaps the kernels of tick k + 1 (the shield never
// reads the robot's measured state - it follows its own trajectory - so the next tick does not have to wait for the host to
// have seen this one's setpoints). The caller waits for `copy_stream` (or the next call on the same slot) before reading the
// host buffers of a slot; a slot is not overwritten before its copy has finished.
int hrg_shield_tick_replay_piped(HrgShield* s, HrgHuman* h, int slot, const int64_t* low_level_time_host, const double* t_host,
                                 const uint8_t* goal_mask_host, const double* goal_q_host, double* q_host, double* dq_host,
                                 double* ddq_host, uint8_t* flags_host, void* stream, void* copy_stream) {
  if (!s || !h || !low_level_time_host || !t_host || slot < 0 || slot > 1)
    return fail(HRG_EINVAL, "hrg_shield_tick_replay_piped: bad argument");
  if (h->H.N != s->S.N || h->H.K != s->S.K || h->device != s->device)
    return fail(HRG_EINVAL, "hrg_shield_tick_replay_piped: the human batch and the shield batch differ (envs, keypoints or device)");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream), cs = static_cast<cudaStream_t>(copy_stream);
  const size_t N = s->S.N, J = s->S.J;
  CK(cudaMemcpyAsync(h->d_clock, low_level_time_host, N * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->d_t, t_host, N * sizeof(double), cudaMemcpyHostToDevice, st));
  const bool goal = goal_mask_host && goal_q_host;
  if (goal) {
    CK(cudaMemcpyAsync(s->d_goal_mask, goal_mask_host, N, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(s->d_goal, goal_q_host, N * J * sizeof(double), cudaMemcpyHostToDevice, st));
  }
  if (s->out_used[slot]) CK(cudaStreamWaitEvent(st, s->ev_copied[slot], 0));  // the slot's previous copy has left it
  int rc = hrg_human_keypoints(h, h->d_clock, s->d_kp, nullptr, nullptr, stream);
  if (rc) return rc;
  double** o = s->d_out2[slot];
  rc = hrg_shield_tick_device(s, s->d_kp, s->d_t, goal ? s->d_goal_mask : nullptr, goal ? s->d_goal : nullptr, o[0], o[1], o[2],
                              nullptr, s->d_safe2[slot], s->d_branch2[slot], nullptr, stream);
  if (rc) return rc;
  if (flags_host) {
    k_pack_flags<<<grid_for((int)N, 256), 256, 0, st>>>((int)N, s->d_safe2[slot], s->d_branch2[slot], s->d_flags82[slot]);
    s->launches++;
  }
  CK(cudaEventRecord(s->ev_out[slot], st));
  CK(cudaStreamWaitEvent(cs, s->ev_out[slot], 0));
  if (q_host) CK(cudaMemcpyAsync(q_host, o[0], N * J * sizeof(double), cudaMemcpyDeviceToHost, cs));
  if (dq_host) CK(cudaMemcpyAsync(dq_host, o[1], N * J * sizeof(double), cudaMemcpyDeviceToHost, cs));
  if (ddq_host) CK(cudaMemcpyAsync(ddq_host, o[2], N * J * sizeof(double), cudaMemcpyDeviceToHost, cs));
  if (flags_host) CK(cudaMemcpyAsync(flags_host, s->d_flags82[slot], N, cudaMemcpyDeviceToHost, cs));
  CK(cudaEventRecord(s->ev_copied[slot], cs));
  s->out_used[slot] = true;
  return HRG_OK;
}

int64_t hrg_human_launch_count(const HrgHuman* h) { return h ? h->launches : 0; }

}  // extern "C"
