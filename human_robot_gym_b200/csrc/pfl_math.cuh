// PFL (power-and-force-limiting) shields: per-sample Cartesian capsule velocities of a long-term trajectory.
// Restates the PFL LongTermTraj constructor (SS/src/long_term_traj.cc:5-39) and the RobotReach velocity helpers it uses
// (SS/src/robot_reach.cc:110-199: calculateAllTransformationMatricesAndCapsules, getJacobian, getVelocityOfCapsule,
// approximateVelOfCapsule). exactVelOfCapsule is never reached from this constructor (SURVEY.md §8(a) a18 note).
#pragma once
#include "state.cuh"

namespace hrg {

// One LTT sample (q, qd): forward kinematics from the base pose (rows 0..2 at base[k * bs]), world end points of every
// link capsule, the geometric Jacobian at p1 of link j (columns i <= j: z_{i+1} x (p1_j - p1_i) ; z_{i+1}) applied to
// qd[0..j], v2 = v1 + w x (p2 - p1). Outputs |v1|, |v2| per link (for alpha_i) and returns the approximate maximum
// surface speed of the LAST link: long_term_traj.cc:19-35 overwrites motion_vel per link, so the value stored per
// sample is the last link's, not the maximum over links - mirrored on purpose.
HRG_HD_CALL double pfl_sample_velocity(const HrgConfig& c, const double* base, size_t bs, const double* q,
                                       const double* qd, double* n1 /*[J]*/, double* n2 /*[J]*/) {
  const int J = c.nb_joints;
  double T[12];
  for (int k = 0; k < 12; ++k) T[k] = base[(size_t)k * bs];
  double z[HRG_MAX_JOINTS + 1][3], p1[HRG_MAX_JOINTS][3], p2[HRG_MAX_JOINTS][3];
  z[0][0] = T[2]; z[0][1] = T[6]; z[0][2] = T[10];
  for (int i = 0; i < J; ++i) {  // RobotReach::forwardKinematic (robot_reach.h:136-142) + transformCapsule (:54-59)
    const double* F = c.link_T[i];
    double sn, cs;
    sincos(q[i], &sn, &cs);  // one range reduction for both (20 % of k_pfl_ltt's instructions were the two separate calls)
    for (int r = 0; r < 3; ++r) {
      const double r0 = T[r * 4 + 0], r1 = T[r * 4 + 1], r2 = T[r * 4 + 2], r3 = T[r * 4 + 3];
      const double t0 = r0 * F[0] + r1 * F[4] + r2 * F[8] + r3 * F[12];
      const double t1 = r0 * F[1] + r1 * F[5] + r2 * F[9] + r3 * F[13];
      const double t2 = r0 * F[2] + r1 * F[6] + r2 * F[10] + r3 * F[14];
      const double t3 = r0 * F[3] + r1 * F[7] + r2 * F[11] + r3 * F[15];
      T[r * 4 + 0] = t0 * cs + t1 * sn;
      T[r * 4 + 1] = t1 * cs - t0 * sn;
      T[r * 4 + 2] = t2;
      T[r * 4 + 3] = t3;
      const double* lp = c.link_caps[i];
      p1[i][r] = T[r * 4 + 0] * lp[0] + T[r * 4 + 1] * lp[1] + T[r * 4 + 2] * lp[2] + T[r * 4 + 3];
      p2[i][r] = T[r * 4 + 0] * lp[3] + T[r * 4 + 1] * lp[4] + T[r * 4 + 2] * lp[5] + T[r * 4 + 3];
      z[i + 1][r] = T[r * 4 + 2];
    }
  }
  double motion_vel = 0.0;
  for (int j = 0; j < J; ++j) {
    // result1 = jacobian1 * velocity.segment(0, j + 1): s = col_0 * qd_0, then += col_k * qd_k (k ascending)
    double v1[3], w[3];
    for (int i = 0; i <= j; ++i) {
      const V3 zi = v3(z[i + 1][0], z[i + 1][1], z[i + 1][2]);
      const V3 up = vcross(zi, v3(p1[j][0] - p1[i][0], p1[j][1] - p1[i][1], p1[j][2] - p1[i][2]));
      if (i == 0) {
        v1[0] = up.x * qd[0]; v1[1] = up.y * qd[0]; v1[2] = up.z * qd[0];
        w[0] = zi.x * qd[0]; w[1] = zi.y * qd[0]; w[2] = zi.z * qd[0];
      } else {
        v1[0] += up.x * qd[i]; v1[1] += up.y * qd[i]; v1[2] += up.z * qd[i];
        w[0] += zi.x * qd[i]; w[1] += zi.y * qd[i]; w[2] += zi.z * qd[i];
      }
    }
    const V3 W = v3(w[0], w[1], w[2]);
    const V3 d21 = v3(p2[j][0] - p1[j][0], p2[j][1] - p1[j][1], p2[j][2] - p1[j][2]);
    const V3 V1 = v3(v1[0], v1[1], v1[2]);
    const V3 V2 = vadd(V1, vcross(W, d21));  // v_2 = v_1 + w x (p_2 - p_1)
    n1[j] = vnorm(V1);
    n2[j] = vnorm(V2);
    // approximateVelOfCapsule(j, v2, w) (robot_reach.cc:165-177)
    const V3 link = v3(p1[j][0] - p2[j][0], p1[j][1] - p2[j][1], p1[j][2] - p2[j][2]);
    const double r = c.link_caps[j][6];
    const double wn = vnorm(W);
    const double s2 = vnorm(V2) + wn * r;
    const double s1 = vnorm(vadd(V2, vcross(W, link))) + wn * r;
    motion_vel = fmax(s1, s2);
  }
  return motion_vel;
}

// The same arithmetic, operation for operation, with the joint count a compile-time constant and the forward kinematics
// FUSED with the velocity loop: link j's velocities only need the p1 / z of links <= j, so they are evaluated right after FK
// step j and p2, T, q never become arrays. Everything is unrolled, every index a constant: the routine lives in registers
// (the generic form keeps z / p1 / p2 / n1 / n2 in 1.7 KB of local memory per thread - 650 KB per SM at three resident
// blocks, far beyond L1, which is what bounded k_pfl_ltt: profiles/r01s_pfl_ncu_summary.md).
// q, qd: pointers to the sample's J angles / velocities; n12: n1[j] at n12[j * ns], n2[j] at n12[(JT + j) * ns].
template <int JT>
HRG_HD double pfl_sample_velocity_t(const HrgConfig& c, const double* base, size_t bs, const double* q, const double* qd,
                                    double* n12, size_t ns) {
  double T[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) T[k] = base[(size_t)k * bs];
  double z[JT + 1][3], p1[JT][3], vq[JT];
  z[0][0] = T[2]; z[0][1] = T[6]; z[0][2] = T[10];
#pragma unroll
  for (int i = 0; i < JT; ++i) vq[i] = qd[i];
  double motion_vel = 0.0;
#pragma unroll
  for (int j = 0; j < JT; ++j) {
    double p2[3];
    {  // RobotReach::forwardKinematic (robot_reach.h:136-142) + transformCapsule (:54-59)
      const double* F = c.link_T[j];
      double sn, cs;
      sincos(q[j], &sn, &cs);
      const double* lp = c.link_caps[j];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const double r0 = T[r * 4 + 0], r1 = T[r * 4 + 1], r2 = T[r * 4 + 2], r3 = T[r * 4 + 3];
        const double t0 = r0 * F[0] + r1 * F[4] + r2 * F[8] + r3 * F[12];
        const double t1 = r0 * F[1] + r1 * F[5] + r2 * F[9] + r3 * F[13];
        const double t2 = r0 * F[2] + r1 * F[6] + r2 * F[10] + r3 * F[14];
        const double t3 = r0 * F[3] + r1 * F[7] + r2 * F[11] + r3 * F[15];
        T[r * 4 + 0] = t0 * cs + t1 * sn;
        T[r * 4 + 1] = t1 * cs - t0 * sn;
        T[r * 4 + 2] = t2;
        T[r * 4 + 3] = t3;
        p1[j][r] = T[r * 4 + 0] * lp[0] + T[r * 4 + 1] * lp[1] + T[r * 4 + 2] * lp[2] + T[r * 4 + 3];
        p2[r] = T[r * 4 + 0] * lp[3] + T[r * 4 + 1] * lp[4] + T[r * 4 + 2] * lp[5] + T[r * 4 + 3];
        z[j + 1][r] = T[r * 4 + 2];
      }
    }
    // result1 = jacobian1 * velocity.segment(0, j + 1): s = col_0 * qd_0, then += col_k * qd_k (k ascending)
    double v1[3], w[3];
#pragma unroll
    for (int i = 0; i <= j; ++i) {
      const V3 zi = v3(z[i + 1][0], z[i + 1][1], z[i + 1][2]);
      const V3 up = vcross(zi, v3(p1[j][0] - p1[i][0], p1[j][1] - p1[i][1], p1[j][2] - p1[i][2]));
      if (i == 0) {
        v1[0] = up.x * vq[0]; v1[1] = up.y * vq[0]; v1[2] = up.z * vq[0];
        w[0] = zi.x * vq[0]; w[1] = zi.y * vq[0]; w[2] = zi.z * vq[0];
      } else {
        v1[0] += up.x * vq[i]; v1[1] += up.y * vq[i]; v1[2] += up.z * vq[i];
        w[0] += zi.x * vq[i]; w[1] += zi.y * vq[i]; w[2] += zi.z * vq[i];
      }
    }
    const V3 W = v3(w[0], w[1], w[2]);
    const V3 d21 = v3(p2[0] - p1[j][0], p2[1] - p1[j][1], p2[2] - p1[j][2]);
    const V3 V1 = v3(v1[0], v1[1], v1[2]);
    const V3 V2 = vadd(V1, vcross(W, d21));  // v_2 = v_1 + w x (p_2 - p_1)
    const double nv2 = vnorm(V2);
    n12[(size_t)j * ns] = vnorm(V1);
    n12[(size_t)(JT + j) * ns] = nv2;
    if (j == JT - 1) {  // approximateVelOfCapsule of the LAST link is what the sample keeps (long_term_traj.cc:19-35)
      const V3 link = v3(p1[j][0] - p2[0], p1[j][1] - p2[1], p1[j][2] - p2[2]);
      const double r = c.link_caps[j][6];
      const double wn = vnorm(W);
      const double s2 = nv2 + wn * r;
      const double s1 = vnorm(vadd(V2, vcross(W, link))) + wn * r;
      motion_vel = fmax(s1, s2);
    }
  }
  return motion_vel;
}

// Sequential form of the whole constructor for one env (host emulation, single-thread use): fills time[i] (the Motion
// times of calculateLongTermTrajectory, safety_shield.cc:739-752: new_time accumulates), cart[i] and alpha[j].
HRG_HD void pfl_ltt_env_seq(const HrgConfig& c, const double* base, size_t bs, const double* traj /*[len][4][J]*/,
                            int len, double t0, double* time, double* cart, double* alpha /*[J]*/) {
  const int J = c.nb_joints;
  double prev1[HRG_MAX_JOINTS], prev2[HRG_MAX_JOINTS];
  for (int j = 0; j < J; ++j) alpha[j] = 0.0;
  double t = t0;
  for (int i = 0; i < len; ++i) {
    time[i] = t;
    t += c.sample_time;
    double n1[HRG_MAX_JOINTS], n2[HRG_MAX_JOINTS];
    const double* m = traj + (size_t)i * 4 * J;
    cart[i] = pfl_sample_velocity(c, base, bs, m, m + J, n1, n2);
    for (int j = 0; j < J; ++j) {
      if (i > 0) {
        const double dt = time[i] - time[i - 1];
        const double a1 = fabs(n1[j] - prev1[j]) / dt, a2 = fabs(n2[j] - prev2[j]) / dt;
        alpha[j] = fmax(alpha[j], fmax(a1, a2));
      }
      prev1[j] = n1[j];
      prev2[j] = n2[j];
    }
  }
}

#if defined(__CUDACC__)
// The PFL LongTermTraj constructor (long_term_traj.cc:5-39) of the trajectories k_replan_sample has just
// written: block = replanning env (grid-stride over the replan list), thread = sample within a chunk of kPflBlock. The
// samples are independent except for the neighbour difference of alpha_i, which goes through shared memory (row 0 carries
// the last sample of the previous chunk). The first form ran inside the sampler with lane = sample, one warp per env:
// ~7 sequential rounds of this ~2 500-instruction routine per warp added 0.43 ms to the replan chain
// (profiles/r01j_summary.md); here the ~480 k samples of a tick spread over the whole GPU.
constexpr int kPflBlock = 128;
#ifndef HRG_PFL_MINBLOCKS
#define HRG_PFL_MINBLOCKS 4
#endif
__global__ void __launch_bounds__(kPflBlock, HRG_PFL_MINBLOCKS) k_pfl_ltt(DevState S) {
  __shared__ double n_s[kPflBlock + 1][2 * HRG_MAX_JOINTS];
  __shared__ double amax_s[kPflBlock / 32][HRG_MAX_JOINTS];
  const HrgConfig& c = *S.cfg;
  const int J = S.J, N = S.N, tid = threadIdx.x;
  const int count = *S.replan_count;
  for (int idx = blockIdx.x; idx < count; idx += gridDim.x) {
    const int e = S.replan_list[idx];
    const uint32_t f = S.flags[e];
    const PlanHead h = S.plan_head[idx];
    if (((f & F_TYPE_MASK) >> F_TYPE_SHIFT) != HRG_SHIELD_PFL || !(h.ok & 1)) continue;  // block-uniform
    const int slot = (f & F_LIVE_SLOT) ? 0 : 1, len = h.slowest;
    double* tm = S.ltt_time + ((size_t)e * 2 + slot) * S.L;
    double* cv = S.ltt_cart + ((size_t)e * 2 + slot) * S.L;
    if (tid == 0) {  // Motion times of calculateLongTermTrajectory (safety_shield.cc:739-752): new_time accumulates
      double t = S.path_s[e];
      for (int i = 0; i < len; ++i) {
        tm[i] = t;
        t += c.sample_time;
      }
    }
    __syncthreads();
    double amax[HRG_MAX_JOINTS];
    for (int j = 0; j < HRG_MAX_JOINTS; ++j) amax[j] = 0.0;
    for (int c0 = 0; c0 < len; c0 += kPflBlock) {
      const int i = c0 + tid;
      const bool valid = i < len;
      if (valid) {
        const double* m = ltt_sample(S, e, slot, i);
        if (J == 6) {  // the shipped robots: register-resident forms (block-uniform branch)
          cv[i] = pfl_sample_velocity_t<6>(c, S.base_T + e, (size_t)N, m, m + 6, n_s[tid + 1], 1);
        } else if (J == 7) {
          cv[i] = pfl_sample_velocity_t<7>(c, S.base_T + e, (size_t)N, m, m + 7, n_s[tid + 1], 1);
        } else {
          double q[HRG_MAX_JOINTS], qd[HRG_MAX_JOINTS], n1[HRG_MAX_JOINTS], n2[HRG_MAX_JOINTS];
          for (int j = 0; j < J; ++j) {
            q[j] = m[j];
            qd[j] = m[J + j];
          }
          cv[i] = pfl_sample_velocity(c, S.base_T + e, (size_t)N, q, qd, n1, n2);
          for (int j = 0; j < J; ++j) {
            n_s[tid + 1][j] = n1[j];
            n_s[tid + 1][J + j] = n2[j];
          }
        }
      }
      __syncthreads();
      if (valid && i > 0) {
        const double dt = tm[i] - tm[i - 1];
        for (int j = 0; j < J; ++j)
          amax[j] = fmax(amax[j], fmax(fabs(n_s[tid + 1][j] - n_s[tid][j]) / dt, fabs(n_s[tid + 1][J + j] - n_s[tid][J + j]) / dt));
      }
      __syncthreads();
      if (tid == kPflBlock - 1 && valid)
        for (int j = 0; j < 2 * J; ++j) n_s[0][j] = n_s[kPflBlock][j];
      __syncthreads();
    }
    for (int j = 0; j < J; ++j) {
      double a = amax[j];
      for (int o = 16; o > 0; o >>= 1) a = fmax(a, __shfl_xor_sync(0xffffffffu, a, o));
      if ((tid & 31) == 0) amax_s[tid >> 5][j] = a;
    }
    __syncthreads();
    if (tid < J) {
      double a = amax_s[0][tid];
      for (int w = 1; w < kPflBlock / 32; ++w) a = fmax(a, amax_s[w][tid]);
      S.ltt_alpha[((size_t)slot * J + tid) * N + e] = a;
    }
    __syncthreads();
  }
}
#endif

}  // namespace hrg
