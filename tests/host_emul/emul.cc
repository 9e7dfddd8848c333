// TEST-ONLY host emulation of the tick kernels (never loaded by the product; the product has no CPU path).
//
// The per-env arithmetic of the CUDA kernels lives in host/device inline functions (human_robot_gym_b200/csrc/*.cuh).
// This file compiles those SAME functions with g++ and drives them env by env in the kernels' order, so that the
// control logic can be checked against oracle/_ref in the build container, which has no GPU. The warp-cooperative
// parts of k_reach_verify / k_replan (shared-memory FK, shuffles) are replaced by straightforward sequential glue here
// and are only exercised by the `-m gpu` tests.
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#if defined(HRG_POLY_LOG)
#include <vector>
static std::vector<double> g_poly_log;  // rows of 8: ncoef, coef[0..6]
static bool g_poly_log_on = false;
namespace hrg {
void hrg_poly_log(const double* coef, int ncoef) {
  if (!g_poly_log_on || ncoef > 7) return;
  g_poly_log.push_back((double)ncoef);
  for (int i = 0; i < 7; ++i) g_poly_log.push_back(i < ncoef ? coef[i] : 0.0);
}
}  // namespace hrg
#endif
#include "../../human_robot_gym_b200/csrc/ltp_math.cuh"
#include "../../human_robot_gym_b200/csrc/step_logic.cuh"
#include "../../human_robot_gym_b200/csrc/pfl_math.cuh"
#include "../../human_robot_gym_b200/csrc/kernels_human.cuh"

#if defined(HRG_CAND_LOG)
#include "../../profiles/candlog.inc"  // host experiment of profiles/rest_start_candidates.py
#endif
using namespace hrg;

struct Emul {
  DevState S;
  HrgConfig cfg;
  std::vector<void*> allocs;
  std::vector<double> robot, human;  // last capsules [N][J][7], [N][H][7]
  template <class T>
  void al(T** p, size_t n) {
    *p = static_cast<T*>(calloc(n + 2, sizeof(T)));
    allocs.push_back(*p);
  }
};

static void emul_reset_env(Emul* E, int e, const double* pose, const double* q0, int type) {
  DevState& S = E->S;
  const HrgConfig& c = E->cfg;
  const int J = S.J, N = S.N;
  uint32_t f = S.flags[e] & F_UNDER_V_LIMIT;
  f |= ((uint32_t)type << F_TYPE_SHIFT) & F_TYPE_MASK;
  if (type == HRG_SHIELD_OFF) f |= F_IS_SAFE;
  double B[12];
  base_pose_matrix(pose[0], pose[1], pose[2], pose[3], pose[4], pose[5], B);
  for (int k = 0; k < 12; ++k) S.base_T[(size_t)k * N + e] = B[k];
  S.last_meas[e] = -1;
  PathS z;
  path_clear(z);
  for (int w = 0; w < N_PATHS; ++w) path_store(S, w, e, z);
  S.path_s[e] = 0;
  S.path_sd[e] = 0;
  double* m = ltt_sample_w(S, e, 0, 0);
  for (int j = 0; j < J; ++j) {
    m[j] = q0[j];
    m[J + j] = m[2 * J + j] = m[3 * J + j] = 0.0;
    S.prev_dq[(size_t)j * N + e] = 0.0;
    S.ltt_alpha[((size_t)0 * J + j) * N + e] = (type == HRG_SHIELD_PFL) ? 0.0 : c.alpha_i_max;
  }
  S.ltt_cart[((size_t)e * 2 + 0) * S.L] = 0.0;
  S.ltt_time[((size_t)e * 2 + 0) * S.L] = 0.0;
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
}

// sequential stand-in for the warp kernel: FK chains, swept capsules, human capsules, pair distances
static void emul_reach_verify(Emul* E, int e, uint32_t& f) {
  DevState& S = E->S;
  const HrgConfig& c = E->cfg;
  const int J = S.J, H = S.H;
  double ri[3 * HRG_MAX_JOINTS + 2], base[12];
  for (int k = 0; k < reach_fields(J); ++k) ri[k] = ri_at(S, k, e);
  for (int k = 0; k < 12; ++k) base[k] = S.base_T[(size_t)k * S.N + e];
  double T[2][16];
  for (int ch = 0; ch < 2; ++ch) {
    for (int i = 0; i < 12; ++i) T[ch][i] = base[i];
    T[ch][12] = T[ch][13] = T[ch][14] = 0.0;
    T[ch][15] = 1.0;
  }
  double* rob = E->robot.data() + (size_t)e * J * 7;
  double* hum = E->human.data() + (size_t)e * H * 7;
  const double s_diff = ri[3 * J], t_reach = ri[3 * J + 1];
  for (int i = 0; i < J; ++i) {
    double P[2][2][3];
    for (int ch = 0; ch < 2; ++ch) {
      double A[16], B[16];
      const double* F = c.link_T[i];
      for (int r = 0; r < 4; ++r)
        for (int k = 0; k < 4; ++k)
          A[r * 4 + k] = T[ch][r * 4 + 0] * F[0 * 4 + k] + T[ch][r * 4 + 1] * F[1 * 4 + k] +
                         T[ch][r * 4 + 2] * F[2 * 4 + k] + T[ch][r * 4 + 3] * F[3 * 4 + k];
      const double cs = cos(ri[ch * J + i]), sn = sin(ri[ch * J + i]);
      for (int r = 0; r < 4; ++r) {
        B[r * 4 + 0] = A[r * 4 + 0] * cs + A[r * 4 + 1] * sn;
        B[r * 4 + 1] = A[r * 4 + 1] * cs - A[r * 4 + 0] * sn;
        B[r * 4 + 2] = A[r * 4 + 2];
        B[r * 4 + 3] = A[r * 4 + 3];
      }
      memcpy(T[ch], B, sizeof(B));
      for (int k = 0; k < 2; ++k) {
        const double* lp = c.link_caps[i] + 3 * k;
        for (int r = 0; r < 3; ++r)
          P[ch][k][r] = B[r * 4 + 0] * lp[0] + B[r * 4 + 1] * lp[1] + B[r * 4 + 2] * lp[2] + B[r * 4 + 3];
      }
    }
    const double rl = c.link_caps[i][6];
    const double grow = ri[2 * J + i] * s_diff * s_diff / 8.0;
    V3 d1 = v3(P[0][0][0] - P[1][0][0], P[0][0][1] - P[1][0][1], P[0][0][2] - P[1][0][2]);
    V3 d2 = v3(P[0][1][0] - P[1][1][0], P[0][1][1] - P[1][1][1], P[0][1][2] - P[1][1][2]);
    double r_1 = vnorm(d1) / 2.0 + grow + rl, r_2 = vnorm(d2) / 2.0 + grow + rl;
    double* o = rob + i * 7;
    for (int r = 0; r < 3; ++r) {
      o[r] = (P[0][0][r] + P[1][0][r]) * 0.5;
      o[3 + r] = (P[0][1][r] + P[1][1][r]) * 0.5;
    }
    o[6] = fmax(r_1, r_2) + c.secure_radius;
  }
  for (int h = 0; h < H; ++h)
    cap_store(human_capsule(c, h, S.kp_pos + e, S.kp_vel + e, (size_t)S.N, t_reach), hum + h * 7);
  unsigned coll = 0;
  double md[3] = {INFINITY, INFINITY, INFINITY};
  bool near = false;
  for (int h = 0; h < H; ++h)
    for (int r = 0; r < J; ++r) {
      double d = capsule_capsule_dist(cap_load(rob + r * 7), cap_load(hum + h * 7));
      int m = human_model_of(c, h);
      if (d < 0.0) coll |= 1u << m;
      if (fabs(d) < 1e-6) near = true;
      md[m] = fmin(md[m], d);
    }
  for (int m = 0; m < 3; ++m) S.min_dist[(size_t)m * S.N + e] = md[m];
  f = (coll != 7u) ? (f | F_IS_SAFE) : (f & ~F_IS_SAFE);
  if (near) f |= (uint32_t)HRG_ST_NEAR_THRESHOLD << F_STATUS_SHIFT;
}

extern "C" {

void* emul_create(const HrgConfig* cfg, int n, const double* pose, const double* q0, const int32_t* type) {
  Emul* E = new Emul();
  E->cfg = *cfg;
  DevState& S = E->S;
  const size_t N = n;
  const int J = cfg->nb_joints, K = cfg->n_keypoints, H = cfg->n_ext + 2 * cfg->n_bodies, L = cfg->ltt_capacity;
  memset(&S, 0, sizeof(S));
  S.N = n; S.J = J; S.K = K; S.H = H; S.L = L;
  S.cfg = &E->cfg;
  E->al(&S.paths, (size_t)N_PATHS * PATH_FIELDS * N);
  E->al(&S.path_s, N); E->al(&S.path_sd, N); E->al(&S.flags, N);
  E->al(&S.prev_dq, J * N); E->al(&S.goal_q, J * N); E->al(&S.replan_q, J * N);
  E->al(&S.ltt_len, 2 * N); E->al(&S.ltt_cur, 2 * N); E->al(&S.ltt_start, 2 * N);
  E->al(&S.ltt_t0, 2 * N); E->al(&S.ltt_ts, 2 * N); E->al(&S.ltt_alpha, 2 * J * N);
  E->al(&S.base_T, 12 * N); E->al(&S.kp_pos, K * 3 * N); E->al(&S.kp_vel, K * 3 * N); E->al(&S.last_meas, N);
  E->al(&S.ltt, N * 2 * L * 4 * J); E->al(&S.ltt_cart, N * 2 * L); E->al(&S.ltt_time, N * 2 * L);
  E->al(&S.cur, N * cur_fields(J)); E->al(&S.reach_in, N * reach_fields(J)); E->al(&S.min_dist, 3 * N);
  E->robot.assign(N * J * 7, 0.0);
  E->human.assign(N * H * 7, 0.0);
  for (int e = 0; e < n; ++e) emul_reset_env(E, e, pose + 6 * (size_t)e, q0 + (size_t)J * e, type ? type[e] : HRG_SHIELD_SSM);
  return E;
}
void emul_destroy(void* h) {
  Emul* E = static_cast<Emul*>(h);
  for (void* p : E->allocs) free(p);
  delete E;
}
void emul_reset(void* h, const uint8_t* mask, const double* pose, const double* q0, const int32_t* type) {
  Emul* E = static_cast<Emul*>(h);
  for (int e = 0; e < E->S.N; ++e)
    if (!mask || mask[e]) emul_reset_env(E, e, pose + 6 * (size_t)e, q0 + (size_t)E->S.J * e, type ? type[e] : HRG_SHIELD_SSM);
}
void emul_human_measurement(void* h, const uint8_t* mask, const double* kp, const double* t) {
  Emul* E = static_cast<Emul*>(h);
  DevState& S = E->S;
  const int K3 = S.K * 3;
  for (int e = 0; e < S.N; ++e) {
    if (mask && !mask[e]) continue;
    const double last = S.last_meas[e], now = t[e];
    bool diff = false;
    double inv = 0.0;
    if (last != -1) {
      double dt = now - last;
      if (!(dt < 1e-7)) {
        diff = true;
        inv = 1 / dt;
      }
    }
    for (int i = 0; i < K3; ++i) {
      double p = kp[(size_t)e * K3 + i];
      const size_t o = (size_t)i * S.N + e;
      if (diff) S.kp_vel[o] = (p - S.kp_pos[o]) * inv;
      S.kp_pos[o] = p;
    }
    S.last_meas[e] = now;
  }
}
void emul_new_goal(void* h, const uint8_t* mask, const double* goal_q) {
  Emul* E = static_cast<Emul*>(h);
  for (int e = 0; e < E->S.N; ++e)
    if (!mask || mask[e]) new_goal_env(E->S, E->cfg, e, goal_q + (size_t)e * E->S.J);
}
void emul_step(void* h, const double* t, double* q, double* dq, double* ddq, double* dddq, uint8_t* is_safe,
               uint8_t* branch, uint32_t* status) {
  Emul* E = static_cast<Emul*>(h);
  DevState& S = E->S;
  const HrgConfig& c = E->cfg;
  const int J = S.J, N = S.N;
  for (int e = 0; e < N; ++e) {
    uint32_t f = S.flags[e];
    if (pre_env(S, c, e, f)) {
      // k_replan, sequential
      const int pend = (f & F_LIVE_SLOT) ? 0 : 1;
      double cur[3 * HRG_MAX_JOINTS];
      for (int k = 0; k < 3 * J; ++k) cur[k] = cur_at(S, k, e);
      double goal[HRG_MAX_JOINTS];
      for (int j = 0; j < J; ++j) goal[j] = S.goal_q[(size_t)j * N + e];
      uint32_t st = HRG_ST_REPLANNED;
      int len = ltp_plan_env_seq(c, goal, cur, cur + J, cur + 2 * J, ltt_sample_w(S, e, pend, 0), S.L);
      if (len > 0) {
        const bool pfl = ((f & F_TYPE_MASK) >> F_TYPE_SHIFT) == HRG_SHIELD_PFL;
        double alpha[HRG_MAX_JOINTS];
        if (pfl)
          pfl_ltt_env_seq(c, S.base_T + e, (size_t)N, ltt_sample(S, e, pend, 0), len, S.path_s[e],
                          S.ltt_time + ((size_t)e * 2 + pend) * S.L, S.ltt_cart + ((size_t)e * 2 + pend) * S.L, alpha);
        for (int j = 0; j < J; ++j) {
          S.replan_q[(size_t)j * N + e] = cur[j];
          S.ltt_alpha[((size_t)pend * J + j) * N + e] = pfl ? alpha[j] : c.alpha_i_max;
        }
        S.ltt_len[(size_t)pend * N + e] = len;
        S.ltt_cur[(size_t)pend * N + e] = 0;
        S.ltt_start[(size_t)pend * N + e] = S.path_sd[e];
        S.ltt_t0[(size_t)pend * N + e] = S.path_s[e];
        S.ltt_ts[(size_t)pend * N + e] = c.sample_time;
        f |= F_NEW_LTT;
        f &= ~F_NEW_LTT_PROCESSED;
      } else {
        if (len < 0) st |= HRG_ST_LTT_TRUNCATED;
        f &= ~F_NEW_LTT;
        st |= HRG_ST_REPLAN_FAILED;
      }
      f &= ~F_PENDING_REPLAN;
      f |= st << F_STATUS_SHIFT;
    }
    plan_env(S, c, e, t[e], f, S.last_meas[e]);
    if (f & F_VERIFY) emul_reach_verify(E, e, f);
    select_env(S, c, e, f, q, dq, ddq, dddq);
    f |= F_NEW_LTT_PROCESSED | F_STEPPED;
    S.flags[e] = f;
    if (is_safe) is_safe[e] = (f & F_IS_SAFE) ? 1 : 0;
    if (branch) branch[e] = (uint8_t)((f & F_BRANCH_MASK) >> F_BRANCH_SHIFT);
    if (status) status[e] = f >> F_STATUS_SHIFT;
  }
}
void emul_get_capsules(void* h, double* robot, double* human, double* min_dist) {
  Emul* E = static_cast<Emul*>(h);
  if (robot) memcpy(robot, E->robot.data(), E->robot.size() * 8);
  if (human) memcpy(human, E->human.data(), E->human.size() * 8);
  if (min_dist)
    for (int e = 0; e < E->S.N; ++e)
      for (int m = 0; m < 3; ++m) min_dist[(size_t)e * 3 + m] = E->S.min_dist[(size_t)m * E->S.N + e];
}
void emul_get_state(void* h, int e, double* out /*[52]*/) {
  Emul* E = static_cast<Emul*>(h);
  DevState& S = E->S;
  for (int w = 0; w < N_PATHS; ++w) {
    PathS p;
    path_load(S, w, e, p);
    for (int i = 0; i < 6; ++i) out[w * 10 + i] = p.ph[i];
    out[w * 10 + 6] = p.pos; out[w * 10 + 7] = p.vel; out[w * 10 + 8] = p.acc; out[w * 10 + 9] = p.jerk;
  }
  uint32_t f = S.flags[e];
  out[40] = S.path_s[e]; out[41] = S.path_sd[e]; out[42] = (f & F_NEW_LTT) ? 1 : 0; out[43] = (f & F_NEW_GOAL) ? 1 : 0;
  out[44] = (f & F_NEW_LTT_PROCESSED) ? 1 : 0; out[45] = (f & F_REC_CORRECT) ? 1 : 0;
  int live = (f & F_LIVE_SLOT) ? 1 : 0;
  out[46] = S.ltt_len[(size_t)live * S.N + e]; out[47] = S.ltt_len[(size_t)(1 - live) * S.N + e];
  out[48] = (f & F_REC_CURRENT) ? 1 : 0;
}
void emul_capsule_dist(int n, const double* c1, const double* c2, double* d) {
  for (int i = 0; i < n; ++i) d[i] = capsule_capsule_dist(cap_load(c1 + 7 * i), cap_load(c2 + 7 * i));
}
void emul_plan_safety_shield(double ts, int n, const double* in, uint8_t* ok, double* phases, double* fin) {
  for (int i = 0; i < n; ++i) {
    const double* a = in + 6 * i;
    PathS p;
    path_clear(p);
    bool touched;
    ok[i] = plan_safety_shield(a[0], a[1], a[2], a[3], a[4], a[5], ts, p, touched) ? 1 : 0;
    for (int k = 0; k < 6; ++k) phases[6 * i + k] = p.ph[k];
    path_final(p, fin[3 * i], fin[3 * i + 1], fin[3 * i + 2]);
  }
}
int emul_ltp_plan(const HrgConfig* cfg, const double* goal, const double* q0, const double* v0, const double* a0,
                  double* traj, int cap) {
  return ltp_plan_env_seq(*cfg, goal, q0, v0, a0, traj, cap);
}
#if defined(HRG_RR_TRACE)
void emul_rr_trace(long long* out /*[8][64]*/, int clear) {
  for (int i = 0; i < 8 * 64; ++i) {
    out[i] = (&hrg_rr_trace_hist[0][0])[i];
    if (clear) (&hrg_rr_trace_hist[0][0])[i] = 0;
  }
}
#endif
#if defined(HRG_POLY_LOG)
long long emul_poly_log(int on, double* out, long long cap_rows) {  // copies and clears the log; returns the row count
  g_poly_log_on = on != 0;
  const long long rows = (long long)(g_poly_log.size() / 8);
  if (out) {
    memcpy(out, g_poly_log.data(), sizeof(double) * 8 * (size_t)(rows < cap_rows ? rows : cap_rows));
    g_poly_log.clear();
  }
  return rows;
}
#endif
// planner-wide root-finder mode of the emulation (LtpLimits::root_mode): 0 exact QR, 1 fast with QR fallback
static int g_emul_root_mode = 0;
void emul_set_root_mode(int m) { g_emul_root_mode = m; }
// the derivative-chain root finder (csrc/ltp_roots_fast.cuh) next to the exact one, for many polynomials of one degree
void emul_poly_root_fast_batch(long long n, int ncoef, const double* coef, double* out_fast, double* out_margin,
                               double* out_reg) {
  for (long long i = 0; i < n; ++i) {
    double m = 0.0;
    out_fast[i] = rr_has_degree(ncoef) ? fr_root(coef + (size_t)i * ncoef, ncoef, m) : NAN;
    out_margin[i] = m;
    if (rr_has_degree(ncoef)) {  // the mixed-degree form must return the same bits
      double m2 = 0.0;
      const double r2 = fr_root_mixed(coef + (size_t)i * ncoef, ncoef, m2);
      if (memcmp(&r2, &out_fast[i], 8) != 0 || memcmp(&m2, &m, 8) != 0) out_margin[i] = -1.0;
    }
    if (out_reg) out_reg[i] = rr_has_degree(ncoef) ? rr_root(coef + (size_t)i * ncoef, ncoef) : NAN;
  }
}
double emul_poly_root(const double* coef, int ncoef) { LtpLimits L{}; return smallest_positive_real_root(L, coef, ncoef); }
// the register-resident solver the device uses for one-problem-per-lane kernels (csrc/ltp_roots_reg.cuh), degrees 4..6
double emul_poly_root_reg(const double* coef, int ncoef) { return rr_has_degree(ncoef) ? rr_root(coef, ncoef) : NAN; }
void emul_poly_root_batch(int n, int ncoef, const double* coef /*[n, ncoef]*/, double* out_mat, double* out_reg) {
  LtpLimits L{};
  for (int i = 0; i < n; ++i) {
    out_mat[i] = smallest_positive_real_root(L, coef + (size_t)i * ncoef, ncoef);
    out_reg[i] = rr_has_degree(ncoef) ? rr_root(coef + (size_t)i * ncoef, ncoef) : NAN;
  }
}

// ---- single-function entry points for the golden-vector tests -------------------------------------------------------
static LtpLimits lim1(double ts, const double* lim) {
  LtpLimits L{};
  L.v_max = lim[2]; L.a_max = lim[3]; L.j_max = lim[4]; L.t_sample = ts;
  L.root_mode = g_emul_root_mode;  // emul_set_root_mode: 0 exact QR for every polynomial, 1 fast roots + rest-start rule
  return L;
}
void emul_ltp_opt_braking(double ts, const double* lim, double v0, double a0, double* q, double* t_rel, double* dir) {
  LtpLimits L = lim1(ts, lim);
  for (int i = 0; i < 7; ++i) t_rel[i] = 0.0;
  ltp_opt_braking(L, v0, a0, *q, t_rel, *dir);
}
int emul_ltp_opt_switch_times(double ts, const double* lim, double q_goal, double q0, double v0, double a0, double v_drive,
                              double* T, double* dir, int* mod) {
  LtpLimits L = lim1(ts, lim);
  for (int i = 0; i < 7; ++i) T[i] = 0.0;
  return ltp_opt_switch_times(L, q_goal, q0, v0, a0, v_drive, T, *dir, *mod) ? 1 : 0;
}
int emul_ltp_time_scaling(double ts, const double* lim, double q_goal, double q0, double v0, double a0, double dir,
                          double t_required, double* T, double* v_drive, int* mod) {
  LtpLimits L = lim1(ts, lim);
  for (int i = 0; i < 7; ++i) T[i] = 0.0;
  return ltp_time_scaling(L, q_goal, q0, v0, a0, dir, t_required, T, *v_drive, *mod) ? 1 : 0;
}
// LongTermTraj::interpolate on traj [L][4][J] (current_pos = 0)
void emul_ltt_interpolate(int J, int L, const double* traj, double sample_time, int starting_index, int n,
                          const double* sdd, const double* v_max, const double* a_max, double* out) {
  for (int k = 0; k < n; ++k) {
    int sample;
    double dt;
    ltt_locate(sdd[4 * k], sample_time, starting_index, L, 0, sample, dt);
    const double* m = traj + (size_t)sample * 4 * J;
    double* o = out + (size_t)k * 4 * J;
    for (int j = 0; j < J; ++j)
      ltt_interp_joint(m[j], m[J + j], m[2 * J + j], m[3 * J + j], dt, sdd[4 * k + 1], sdd[4 * k + 2], sdd[4 * k + 3],
                       v_max[j], a_max[j], o[j], o[J + j], o[2 * J + j], o[3 * J + j]);
  }
}
// all human capsules (Pos, Vel, Accel) of one env from contiguous keypoints / velocities [K*3]
int emul_human_capsules(const HrgConfig* cfg, const double* kp, const double* kv, double t_reach, double* out) {
  const int H = cfg->n_ext + 2 * cfg->n_bodies;
  for (int h = 0; h < H; ++h) cap_store(human_capsule(*cfg, h, kp, kv, 1, t_reach), out + 7 * h);
  return H;
}
void emul_capsule_markers(int n, const double* caps, double* pos, double* size, double* mat) {
  for (int i = 0; i < n; ++i) capsule_marker(caps + 7 * (size_t)i, pos + 3 * (size_t)i, size + 3 * (size_t)i, mat + 9 * (size_t)i);
}
// Animation -> keypoints for n (frame row, clip info, placement) triples through the functions k_human_keypoints calls.
int emul_human_keypoints(int K, const int32_t* parent, const double* anchor, int n, const double* rows,
                         const double* info, const double* place, double* out) {
  HumanTreeDev T;
  if (int rc = human_build_tree(K, parent, anchor, &T)) return rc;
  const int C = 7 + 3 * K;
  std::vector<double> baked(3 + 3 * K);
  for (int i = 0; i < n; ++i) {  // bake the row (k_human_bake), then place it (k_human_reset + k_human_keypoints)
    double tab[12], P[12];
    human_clip_table(info + (size_t)i * 7, tab);
    human_bake_row(T, rows + (size_t)i * C, tab, baked.data());
    human_placement(place + (size_t)i * 7, P);
    for (int k = 0; k < 3; ++k) P[k] = P[k] + baked[k];
    for (int c = 0; c < 3 * K; ++c) out[(size_t)i * 3 * K + c] = human_apply(P, baked.data() + 3 + 3 * (c / 3), c % 3);
  }
  return 0;
}

}
