// Warp-per-env kernels: human measurement, and robot reach + human reach + ISO verification.
// One warp owns one env at a time (grid-stride over envs); the env's keypoints, the two FK chains, the robot and the
// human capsule sets live in that warp's slice of shared memory; the J x H capsule distances are spread over the 32
// lanes and reduced with ballots / shuffles. FP64 throughout, no tensor cores (nothing here is a contraction).
#pragma once
#include "state.cuh"

namespace hrg {

constexpr int kWarpsPerBlock = 8;

// HumanReach::measurement (SS/src/human_reach.cc:52-74): vel = (new - old) * (1 / dt) when a previous measurement
// exists and dt >= 1e-7; pos = new; last_meas = t. kp_in/kp_pos/kp_vel are env-major [N][K*3] (coalesced rows).
__global__ void k_measure(DevState S, const uint8_t* __restrict__ mask, const double* __restrict__ kp_in,
                          const double* __restrict__ t) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int K3 = S.K * 3;
  for (int e = warp; e < S.N; e += nwarps) {
    if (mask && !mask[e]) continue;
    const double last = S.last_meas[e];
    const double now = t[e];
    __syncwarp();
    const double* in = kp_in + (size_t)e * K3;
    double* pos = S.kp_pos + (size_t)e * K3;
    double* vel = S.kp_vel + (size_t)e * K3;
    bool diff = false;
    double inv = 0.0;
    if (last != -1) {
      double dt = now - last;
      if (!(dt < 1e-7)) {
        diff = true;
        inv = 1 / dt;
      }
    }
    for (int i = lane; i < K3; i += 32) {
      double p = in[i];
      if (diff) vel[i] = (p - pos[i]) * inv;
      pos[i] = p;
    }
    if (lane == 0) S.last_meas[e] = now;
  }
}

struct ReachSmem {  // per-warp slice, sizes for the configured K / H
  double* kp;   // [K*3] positions
  double* kv;   // [K*3] velocities
  double* Tm;   // [2][16] FK chains: 0 = start config, 1 = goal config
  double* sc;   // [2][8][2] cos, sin per chain and joint
  double* P;    // [2][8][2][3] transformed capsule end points per chain and link
  double* rob;  // [8][7]
  double* hum;  // [H][7]
};
__host__ __device__ inline int reach_smem_doubles(int K, int H) { return 2 * K * 3 + 32 + 32 + 96 + 56 + H * 7; }

// Explicit-input form used by both the tick kernel and the stage entry point.
// base [12] rows 0..2 of the base matrix, ri = {q_start[J], q_goal[J], alpha[J], s_diff, t_reach}.
// Returns (all lanes): bit m of `coll` set iff human model m collides; md[3] per-model minimum distance; near flag.
__device__ __forceinline__ void reach_verify_warp(const HrgConfig& c, const ReachSmem& W, int lane,
                                                  const double* __restrict__ base, const double* __restrict__ ri,
                                                  const double* __restrict__ kp_g, const double* __restrict__ kv_g,
                                                  unsigned& coll, double* md, bool& near) {
  const int J = c.nb_joints, K3 = c.n_keypoints * 3, H = c.n_ext + 2 * c.n_bodies;
  // ---- stage inputs -------------------------------------------------------------------------------------------
  for (int i = lane; i < K3; i += 32) {
    W.kp[i] = kp_g[i];
    W.kv[i] = kv_g[i];
  }
  const int ch = lane >> 4, r = (lane >> 2) & 3, k = lane & 3;  // FK role: chain, row, column
  W.Tm[ch * 16 + r * 4 + k] = (r < 3) ? base[r * 4 + k] : (k == 3 ? 1.0 : 0.0);
  if (lane < 2 * J) {
    const int cc = lane / J, jj = lane - cc * J;
    double sn, cs;
    sincos(ri[cc * J + jj], &sn, &cs);
    W.sc[(cc * 8 + jj) * 2] = cs;
    W.sc[(cc * 8 + jj) * 2 + 1] = sn;
  }
  const double s_diff = ri[3 * J], t_reach = ri[3 * J + 1];
  __syncwarp();
  // ---- human capsules: one lane per capsule (HumanReach::humanReachabilityAnalysis, human_reach.cc:76-87) --------
  for (int h = lane; h < H; h += 32) {
    Cap hc = human_capsule(c, h, W.kp, W.kv, t_reach);
    cap_store(hc, W.hum + h * 7);
  }
  // ---- forward kinematics of both chains, 16 lanes per chain, one matrix element per lane -----------------------
  // RobotReach::forwardKinematic (robot_reach.h:136-142): T = T * T_fixed[i] * Rz(q_i); transformCapsule (robot_reach.cc:54-59)
  double* Tc = W.Tm + ch * 16;
  for (int i = 0; i < J; ++i) {
    const double* F = c.link_T[i];
    double t0 = Tc[r * 4 + 0] * F[0 * 4 + k] + Tc[r * 4 + 1] * F[1 * 4 + k] + Tc[r * 4 + 2] * F[2 * 4 + k] +
                Tc[r * 4 + 3] * F[3 * 4 + k];
    __syncwarp();
    Tc[r * 4 + k] = t0;
    __syncwarp();
    const double cs = W.sc[(ch * 8 + i) * 2], sn = W.sc[(ch * 8 + i) * 2 + 1];
    const double a = Tc[r * 4 + 0], b = Tc[r * 4 + 1];
    double t1 = (k == 0) ? a * cs + b * sn : ((k == 1) ? b * cs - a * sn : t0);
    __syncwarp();
    Tc[r * 4 + k] = t1;
    __syncwarp();
    if (r < 3 && k < 2) {  // k = 0: end point p1 of link i, k = 1: p2 (identical op sequence => spheres stay spheres)
      const double* lp = c.link_caps[i] + 3 * k;
      W.P[((ch * 8 + i) * 2 + k) * 3 + r] = Tc[r * 4 + 0] * lp[0] + Tc[r * 4 + 1] * lp[1] + Tc[r * 4 + 2] * lp[2] + Tc[r * 4 + 3];
    }
  }
  __syncwarp();
  // ---- swept link capsules: one lane per link (RobotReach::reach, robot_reach.cc:61-98) -------------------------
  if (lane < J) {
    const double* b1 = W.P + ((0 * 8 + lane) * 2 + 0) * 3;  // before p1
    const double* b2 = b1 + 3;                              // before p2
    const double* a1 = W.P + ((1 * 8 + lane) * 2 + 0) * 3;  // after p1
    const double* a2 = a1 + 3;
    const double rl = c.link_caps[lane][6];
    const double grow = ri[2 * J + lane] * s_diff * s_diff / 8.0;
    V3 d1 = v3(b1[0] - a1[0], b1[1] - a1[1], b1[2] - a1[2]);
    V3 d2 = v3(b2[0] - a2[0], b2[1] - a2[1], b2[2] - a2[2]);
    double r_1 = vnorm(d1) / 2.0 + grow + rl;
    double r_2 = vnorm(d2) / 2.0 + grow + rl;
    double* o = W.rob + lane * 7;
    o[0] = (b1[0] + a1[0]) * 0.5; o[1] = (b1[1] + a1[1]) * 0.5; o[2] = (b1[2] + a1[2]) * 0.5;
    o[3] = (b2[0] + a2[0]) * 0.5; o[4] = (b2[1] + a2[1]) * 0.5; o[5] = (b2[2] + a2[2]) * 0.5;
    o[6] = fmax(r_1, r_2) + c.secure_radius;
  }
  __syncwarp();
  // ---- verification: J x H capsule pairs over the lanes (VerifyISO::verify_human_reach, verify_iso.cc:5-33) ------
  unsigned my_coll = 0;
  double m0 = INFINITY, m1 = INFINITY, m2 = INFINITY;
  bool my_near = false;
  const int npairs = J * H;
  for (int p = lane; p < npairs; p += 32) {
    const int h = p / J, rr = p - h * J;
    Cap rc = cap_load(W.rob + rr * 7);
    Cap hc = cap_load(W.hum + h * 7);
    double d = capsule_capsule_dist(rc, hc);
    const int m = human_model_of(c, h);
    if (d < 0.0) my_coll |= 1u << m;
    if (fabs(d) < 1e-6) my_near = true;
    if (m == 0) m0 = fmin(m0, d);
    else if (m == 1) m1 = fmin(m1, d);
    else m2 = fmin(m2, d);
  }
  coll = 0;
  for (int m = 0; m < 3; ++m)
    if (__ballot_sync(0xffffffffu, (my_coll >> m) & 1u)) coll |= 1u << m;
  near = __ballot_sync(0xffffffffu, my_near) != 0;
  for (int o = 16; o > 0; o >>= 1) {
    m0 = fmin(m0, __shfl_xor_sync(0xffffffffu, m0, o));
    m1 = fmin(m1, __shfl_xor_sync(0xffffffffu, m1, o));
    m2 = fmin(m2, __shfl_xor_sync(0xffffffffu, m2, o));
  }
  md[0] = m0; md[1] = m1; md[2] = m2;
}

__device__ __forceinline__ ReachSmem reach_smem_slice(double* base, int warp_in_block, int K, int H) {
  double* p = base + (size_t)warp_in_block * reach_smem_doubles(K, H);
  ReachSmem W;
  W.kp = p; p += K * 3;
  W.kv = p; p += K * 3;
  W.Tm = p; p += 32;
  W.sc = p; p += 32;
  W.P = p; p += 96;
  W.rob = p; p += 56;
  W.hum = p;
  return W;
}

// Tick kernel: reach + verify for every env whose plan step asked for it (F_VERIFY); writes is_safe_ into the flag
// word, the per-model minimum distances, and (optionally) the capsules for the drop-in getters.
__global__ void __launch_bounds__(kWarpsPerBlock * 32) k_reach_verify(DevState S) {
  extern __shared__ double smem[];
  __shared__ HrgConfig cfg_s;
  {
    const int nw = sizeof(HrgConfig) / 8;
    const double* src = reinterpret_cast<const double*>(S.cfg);
    double* dst = reinterpret_cast<double*>(&cfg_s);
    for (int i = threadIdx.x; i < nw; i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const HrgConfig& c = cfg_s;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp = blockIdx.x * kWarpsPerBlock + wib, nwarps = gridDim.x * kWarpsPerBlock;
  const ReachSmem W = reach_smem_slice(smem, wib, S.K, S.H);
  const int J = S.J, K3 = S.K * 3, H = S.H;
  for (int e = warp; e < S.N; e += nwarps) {
    uint32_t f = S.flags[e];
    if (!(f & F_VERIFY)) {
      if (lane < 3) S.min_dist[(size_t)e * 3 + lane] = INFINITY;
      continue;
    }
    unsigned coll;
    double md[3];
    bool near;
    __syncwarp();
    reach_verify_warp(c, W, lane, S.base_T + (size_t)e * 12, S.reach_in + (size_t)e * reach_stride(J),
                      S.kp_pos + (size_t)e * K3, S.kp_vel + (size_t)e * K3, coll, md, near);
    // safe iff ANY human model is collision-free (verify_iso.cc:22-27)
    const bool safe = (coll != 7u);
    if (lane == 0) {
      f = safe ? (f | F_IS_SAFE) : (f & ~F_IS_SAFE);
      if (near) f |= (uint32_t)HRG_ST_NEAR_THRESHOLD << F_STATUS_SHIFT;
      S.flags[e] = f;
    }
    if (lane < 3) S.min_dist[(size_t)e * 3 + lane] = md[lane];
    if (S.robot_caps)
      for (int i = lane; i < J * 7; i += 32) S.robot_caps[(size_t)e * J * 7 + i] = W.rob[i];
    if (S.human_caps)
      for (int i = lane; i < H * 7; i += 32) S.human_caps[(size_t)e * H * 7 + i] = W.hum[i];
    __syncwarp();
  }
}

// Stage entry point on explicit inputs (parity tests, ncu).
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
    k_stage_reach_verify(const HrgConfig* __restrict__ cfg, int n, const double* base_pose, const double* q_start,
                         const double* q_goal, const double* s_diff, const double* t_reach, const double* alpha,
                         const double* kp_pos, const double* kp_vel, double* robot, double* human, double* min_dist,
                         uint8_t* is_safe) {
  extern __shared__ double smem[];
  __shared__ HrgConfig cfg_s;
  __shared__ double in_s[kWarpsPerBlock][12 + 3 * HRG_MAX_JOINTS + 2];
  {
    const int nw = sizeof(HrgConfig) / 8;
    const double* src = reinterpret_cast<const double*>(cfg);
    double* dst = reinterpret_cast<double*>(&cfg_s);
    for (int i = threadIdx.x; i < nw; i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const HrgConfig& c = cfg_s;
  const int J = c.nb_joints, K = c.n_keypoints, H = c.n_ext + 2 * c.n_bodies;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int warp = blockIdx.x * kWarpsPerBlock + wib, nwarps = gridDim.x * kWarpsPerBlock;
  const ReachSmem W = reach_smem_slice(smem, wib, K, H);
  for (int e = warp; e < n; e += nwarps) {
    double* in = in_s[wib];
    __syncwarp();
    if (lane == 0) {
      const double* p = base_pose + (size_t)e * 6;
      base_pose_matrix(p[0], p[1], p[2], p[3], p[4], p[5], in);
    }
    if (lane < J) {
      in[12 + lane] = q_start[(size_t)e * J + lane];
      in[12 + J + lane] = q_goal[(size_t)e * J + lane];
      in[12 + 2 * J + lane] = alpha[(size_t)e * J + lane];
    }
    if (lane == 0) {
      in[12 + 3 * J] = s_diff[e];
      in[12 + 3 * J + 1] = t_reach[e];
    }
    __syncwarp();
    unsigned coll;
    double md[3];
    bool near;
    reach_verify_warp(c, W, lane, in, in + 12, kp_pos + (size_t)e * K * 3, kp_vel + (size_t)e * K * 3, coll, md, near);
    if (lane == 0 && is_safe) is_safe[e] = (coll != 7u) ? 1 : 0;
    if (lane < 3 && min_dist) min_dist[(size_t)e * 3 + lane] = md[lane];
    if (robot)
      for (int i = lane; i < J * 7; i += 32) robot[(size_t)e * J * 7 + i] = W.rob[i];
    if (human)
      for (int i = lane; i < H * 7; i += 32) human[(size_t)e * H * 7 + i] = W.hum[i];
    __syncwarp();
  }
}

}  // namespace hrg
