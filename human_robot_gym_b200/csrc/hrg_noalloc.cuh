// Allocation-free variant of hrg_stage_reach_verify (included at the end of hrg_shield.cu, i.e. part of that .cu file's
// translation unit). No counterpart in the reference: added for callers that cannot
// call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// hrg_stage_reach_verify packs its inputs into a field-major scratch state ([cfg | base_T 12 | reach_in 3J+2 | kp 2*3K |
// min_dist 3] x n doubles + n flag words) that it cudaMallocs and frees per call. Here that scratch is a file-scope
// __device__ array of HRG_NOALLOC_SCRATCH_BYTES: every GPU that loads the library holds it until exit, every call on a
// device shares it (calls that overlap on the GPU race), and it keeps the previous call's contents.
#pragma once
#include <mutex>

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[HRG_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol, not a device address: the address a kernel
// may dereference comes from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[kNoallocMaxDevices] = {};

int noalloc_scratch(unsigned char** out) {
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoallocMaxDevices) return fail(HRG_EINVAL, "hrg_stage_reach_verify_noalloc: device ordinal above 63");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);  // `dev` is the current device
    if (e != cudaSuccess || !p) return fail(HRG_ECUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", e);
    g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return HRG_OK;
}

// The constant tables travel as a kernel parameter (3.4 KB) instead of a host->device copy, so the call is plain kernel
// launches on `stream` and nothing else.
__global__ void k_noalloc_put_cfg(HrgConfig cfg, HrgConfig* __restrict__ dst, uint32_t* __restrict__ zero_from,
                                  size_t zero_words) {
  if (blockIdx.x == 0) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&cfg);
    uint32_t* d = reinterpret_cast<uint32_t*>(dst);
    for (int i = threadIdx.x; i < (int)(sizeof(HrgConfig) / 4); i += blockDim.x) d[i] = src[i];
  }
  // what the previous call left behind in the result fields (min_dist, flags) is cleared; the input fields are fully
  // rewritten by k_stage_pack
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < zero_words; i += (size_t)gridDim.x * blockDim.x)
    zero_from[i] = 0u;
}

}  // namespace

extern "C" int hrg_stage_reach_verify_noalloc(const HrgConfig* cfg, int n, const double* base_pose, const double* q_start,
                                              const double* q_goal, const double* s_diff, const double* t_reach,
                                              const double* alpha_i, const double* kp_pos, const double* kp_vel,
                                              double* robot, double* human, double* min_dist, uint8_t* is_safe,
                                              void* stream) {
  int rc = cfg_check(cfg);
  if (rc) return rc;
  if (n < 0 || !base_pose || !q_start || !q_goal || !s_diff || !t_reach || !alpha_i || !kp_pos || !kp_vel)
    return fail(HRG_EINVAL, "hrg_stage_reach_verify_noalloc: bad argument");
  if (n == 0) return HRG_OK;
  const int J = cfg->nb_joints, K3 = cfg->n_keypoints * 3, H = cfg->n_ext + 2 * cfg->n_bodies;
  const size_t per_env = 12 + reach_fields(J) + 2 * (size_t)K3 + 3;  // doubles; + one 32-bit flag word
  const size_t cfg_doubles = (sizeof(HrgConfig) + 7) / 8;
  const size_t need = (cfg_doubles + per_env * (size_t)n) * sizeof(double) + (size_t)n * sizeof(uint32_t);
  if (need > (size_t)HRG_NOALLOC_SCRATCH_BYTES) {
    static thread_local char msg[200];
    snprintf(msg, sizeof msg,
             "hrg_stage_reach_verify_noalloc: n = %d needs %zu B of scratch, the static array holds %zu B (at most %zu envs "
             "with this config); use hrg_stage_reach_verify",
             n, need, (size_t)HRG_NOALLOC_SCRATCH_BYTES,
             ((size_t)HRG_NOALLOC_SCRATCH_BYTES - cfg_doubles * 8) / (per_env * 8 + 4));
    return fail(HRG_EINVAL, msg);
  }
  unsigned char* base = nullptr;
  if ((rc = noalloc_scratch(&base))) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  double* d = reinterpret_cast<double*>(base);
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
  const size_t zero_words = (size_t)3 * n * 2 + n;  // min_dist (as 32-bit words) + flags, contiguous
  k_noalloc_put_cfg<<<grid_for((int)((zero_words + 3) / 4), 256), 256, 0, st>>>(
      *cfg, reinterpret_cast<HrgConfig*>(d), reinterpret_cast<uint32_t*>(S.min_dist), zero_words);
  const size_t smem = (size_t)reach_smem_doubles(J, H) * sizeof(double);
  CK(cudaFuncSetAttribute(k_reach_verify<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));  // a function attribute, not an allocation
  k_stage_pack<<<grid_for(n, 128), 128, 0, st>>>(S, base_pose, q_start, q_goal, s_diff, t_reach, alpha_i, kp_pos, kp_vel);
  k_reach_verify<1><<<tile_grid(n), 32 * J, smem, st>>>(S, 0);
  k_stage_unpack<<<grid_for(n, 128), 128, 0, st>>>(S, min_dist, is_safe);
  CK(cudaGetLastError());
  return HRG_OK;
}
