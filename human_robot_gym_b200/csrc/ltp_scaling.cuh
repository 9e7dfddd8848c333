// timeScaling of the LongTermPlanner (LTP/src/long_term_planner.cc:358-645): find the cruise velocity v_drive with which
// a non-slowest joint needs exactly t_required, trying the 8 profile shapes in the reference's order.
#pragma once

namespace hrg {

// One candidate: accept if optSwitchTimes succeeds and -0.01 < t_required - T[6] < 0.1 (:398-405 and repeats).
HRG_HD_CALL bool ltp_try_vdrive(const LtpLimits& L, double q_goal, double q_0, double v_orig, double a_orig, double v_drive,
                           double t_required, double* T, int& mod_jerk_profile, OptRootCache* cache = nullptr) {
  const double tol = 0.1;
  if (!isnan(v_drive) && v_drive > 0) {
    double trash;
    bool ok = ltp_opt_switch_times(L, q_goal, q_0, v_orig, a_orig, v_drive, T, trash, mod_jerk_profile, cache);
    if (ok && t_required - T[6] < tol && t_required - T[6] > -tol / 10) return true;
  }
  return false;
}

// Candidate c (0..7 = the reference's profile shapes in its order, :378-629): either the closed-form v_drive (c = 0, 1;
// ncoef = 0) or the coefficients A[0..ncoef) of the polynomial whose smallest positive real root gives v_drive.
// v_0 / a_0 are already mirrored for dir < 0.
HRG_HD_CALL double ltp_scaling_setup(const LtpLimits& L, int cand, double q_goal, double q_0, double v_0, double a_0,
                                     double dir, double tr /*t_required*/, double* A /*[7]*/, int& ncoef) {
  const double jm = L.j_max, am = L.a_max;
  double v_drive = 0.0;
  ncoef = 0;
  switch (cand) {
    case 0: {  // (1) standard profile, phases 2 and 6 exist (:378-396)
  v_drive = (am * jm * tr / 2 - sq(a_0) / 4 + a_0 * am / 2 - sq(am) / 2 + v_0 * jm / 2 -
             sqrt(36 * sq(am) * sq(jm) * sq(tr) - 36 * sq(a_0) * am * jm * tr + 72.0 * a_0 * sq(am) * jm * tr -
                  72.0 * cube(am) * jm * tr + 144 * am * dir * sq(jm) * q_0 - 144 * am * dir * sq(jm) * q_goal +
                  72.0 * am * sq(jm) * v_0 * tr - 9 * pow4(a_0) + 12.0 * cube(a_0) * am + 36 * sq(a_0) * sq(am) +
                  36 * sq(a_0) * jm * v_0 - 72.0 * a_0 * cube(am) - 72.0 * a_0 * am * jm * v_0 + 36 * pow4(am) -
                  36 * sq(jm) * sq(v_0)) /
                 12) /
            jm;
      break;
    }
    case 1: {  // (2) modified profile, phases 2 and 6 exist (:408-436). Common sub-terms of the reference expression:
      //   w  = (v_0 + a_0 (a_0 - am) / (2 jm)) / am,  h  = am / (2 jm),  g = (a_0 - am) / (2 jm),  s = a_0 + am
    const double w = (v_0 + (a_0 * (a_0 - am)) / (2.0 * jm)) / am;
    const double h = am / (2.0 * jm);
    const double g = (a_0 - am) / (2.0 * jm);
    const double s = a_0 + am;
    const double num =
        dir * (q_0 - q_goal) -
        jm * (cube(s) / (6 * cube(jm)) - cube(am) / (6 * cube(jm)) + (sq(am) * s) / (2.0 * cube(jm)) +
              (sq(s) * (w + h + g)) / (2.0 * sq(jm))) +
        a_0 * (sq(s) / (2.0 * sq(jm)) + sq(am) / (2.0 * sq(jm)) + (s * (w + h + g)) / jm) -
        am * (sq(w - h + g) / 2 + (am * (w - h + g)) / jm) + v_0 * (w + s / jm + h + g);
    const double den = h - v_0 / am + am * ((w - h + g) / am + 1.0 / jm) -
                       (sq(a_0) + 2.0 * a_0 * am + 4 * sq(am) - 2.0 * jm * tr * am + 2.0 * jm * v_0) / (2.0 * am * jm) +
                       sq(s) / (2.0 * am * jm) - (a_0 * s) / (am * jm);
    v_drive = -num / den;
      break;
    }
    case 2:  // (3) standard profile, phase 2 does not exist (:449-473)
  A[0] = 3;
  A[1] = 12.0 * am;
  A[2] = -24 * am * jm * tr - 12.0 * sq(a_0) - 24 * a_0 * am + 12.0 * sq(am) + 24 * jm * v_0;
  A[3] = 0;
  A[4] = 48 * sq(a_0) * am * jm * tr - 96 * dir * sq(jm) * am * q_0 + 96 * dir * sq(jm) * am * q_goal -
         96 * am * sq(jm) * v_0 * tr + 12.0 * pow4(a_0) + 16 * cube(a_0) * am - 24 * sq(a_0) * sq(am) -
         48 * sq(a_0) * jm * v_0 + 48 * sq(am) * jm * v_0 + 48 * sq(jm) * sq(v_0);
      ncoef = 5;
      break;
    case 3:  // (4) standard profile, phase 6 does not exist (:485-514)
  A[0] = 12;
  A[1] = 24 * am;
  A[2] = -24 * am * jm * tr + 24 * sq(a_0) - 48 * a_0 * am + 24 * sq(am) - 24 * jm * v_0 + 12.0 * a_0 - 12.0 * am;
  A[3] = 0;
  A[4] = -24 * dir * sq(jm) * am * q_0 + 24 * dir * sq(jm) * am * q_goal + 9 * pow4(a_0) - 12.0 * cube(a_0) * am -
         24 * sq(a_0) * jm * v_0 + 48 * a_0 * am * jm * v_0 + 4 * pow4(am) - 24 * sq(am) * jm * v_0 +
         12.0 * sq(jm) * sq(v_0) + 6 * cube(a_0) + 6 * sq(a_0) * am - 12.0 * a_0 * sq(am) - 12.0 * a_0 * jm * v_0 +
         12.0 * am * jm * v_0 + 4 * a_0 * am - 4 * sq(am);
      ncoef = 5;
      break;
    case 4:  // (5) standard profile, phases 2 and 6 do not exist (:526-541), degree 5
  A[0] = (144 * jm * tr + 144 * a_0);
  A[1] = (-72.0 * sq(jm) * sq(tr) - 144 * a_0 * jm * tr + 36 * sq(a_0) - 216 * jm * v_0);
  A[2] = (144 * dir * sq(jm) * q_0 - 144 * dir * sq(jm) * q_goal + 48 * cube(a_0) - 144 * a_0 * jm * v_0);
  A[3] = (-144 * dir * cube(jm) * q_0 * tr + 144 * dir * cube(jm) * q_goal * tr - 48 * cube(a_0) * jm * tr -
          144 * a_0 * dir * sq(jm) * q_0 + 144 * a_0 * dir * sq(jm) * q_goal + 144 * a_0 * sq(jm) * v_0 * tr +
          6 * pow4(a_0) - 72.0 * sq(a_0) * jm * v_0 + 216 * sq(jm) * sq(v_0));
  A[4] = 0;
  A[5] = -72.0 * sq(dir) * pow4(jm) * sq(q_0) + 144 * sq(dir) * pow4(jm) * q_0 * q_goal -
         72.0 * sq(dir) * pow4(jm) * sq(q_goal) - 48 * cube(a_0) * dir * sq(jm) * q_0 +
         48 * cube(a_0) * dir * sq(jm) * q_goal + 144 * a_0 * dir * cube(jm) * q_0 * v_0 -
         144 * a_0 * dir * cube(jm) * q_goal * v_0 + pow6(a_0) - 6 * pow4(a_0) * jm * v_0 +
         36 * sq(a_0) * sq(jm) * sq(v_0) - 72.0 * cube(jm) * cube(v_0);
      ncoef = 6;
      break;
    case 5:  // (6) modified profile, phase 2 does not exist (:553-567)
  A[0] = 3;
  A[1] = -6 * sqrt(2.0) * am;
  A[2] = (12.0 * am * jm * tr - 6 * sq(a_0) - 12.0 * a_0 * am - 6 * sq(am) - 12.0 * jm * v_0);
  A[3] = 0;
  A[4] = -12.0 * sq(a_0) * am * jm * tr - 24 * dir * sq(jm) * am * q_0 + 24 * dir * sq(jm) * am * q_goal -
         24 * am * sq(jm) * v_0 * tr + 3 * pow4(a_0) + 4 * cube(a_0) * am + 6 * sq(a_0) * sq(am) +
         12.0 * sq(a_0) * jm * v_0 + 12.0 * sq(am) * jm * v_0 + 12.0 * sq(jm) * sq(v_0);
      ncoef = 5;
      break;
    case 6:  // (7) modified profile, phase 6 does not exist (:579-593)
  A[0] = 12;
  A[1] = -24 * am;
  A[2] = (24 * am * jm * tr - 12.0 * sq(a_0) - 24 * a_0 * am - 12.0 * sq(am) - 24 * jm * v_0);
  A[3] = 0;
  A[4] = 24 * dir * sq(jm) * am * q_0 - 24 * dir * sq(jm) * am * q_goal + 3 * pow4(a_0) + 8 * cube(a_0) * am +
         6 * sq(a_0) * sq(am) + 12.0 * sq(a_0) * jm * v_0 + 24 * a_0 * am * jm * v_0 + 12.0 * sq(am) * jm * v_0 +
         12.0 * sq(jm) * sq(v_0);
      ncoef = 5;
      break;
    default:  // (8) modified profile, phases 2 and 6 do not exist (:606-629), degree 6
  A[0] = 144;
  A[1] = (-144 * jm * tr + 144 * a_0);
  A[2] = (72.0 * sq(jm) * sq(tr) - 144 * a_0 * jm * tr - 36 * sq(a_0) - 216 * jm * v_0);
  A[3] = (-144 * dir * sq(jm) * q_0 + 144 * dir * sq(jm) * q_goal - 48 * cube(a_0) - 144 * a_0 * jm * v_0);
  A[4] = (144 * dir * cube(jm) * q_0 * tr - 144 * dir * cube(jm) * q_goal * tr + 48 * cube(a_0) * jm * tr -
          144 * a_0 * dir * sq(jm) * q_0 + 144 * a_0 * dir * sq(jm) * q_goal + 144 * a_0 * sq(jm) * v_0 * tr +
          6 * pow4(a_0) + 72.0 * sq(a_0) * jm * v_0 + 216 * sq(jm) * sq(v_0));
  A[5] = 0;
  A[6] = 72.0 * sq(dir) * pow4(jm) * sq(q_0) - 144 * sq(dir) * pow4(jm) * q_0 * q_goal +
         72.0 * sq(dir) * pow4(jm) * sq(q_goal) + 48 * cube(a_0) * dir * sq(jm) * q_0 -
         48 * cube(a_0) * dir * sq(jm) * q_goal + 144 * a_0 * dir * cube(jm) * q_0 * v_0 -
         144 * a_0 * dir * cube(jm) * q_goal * v_0 - pow6(a_0) - 6 * pow4(a_0) * jm * v_0 -
         36 * sq(a_0) * sq(jm) * sq(v_0) - 72.0 * cube(jm) * cube(v_0);
      ncoef = 7;
      break;
  }
  return v_drive;
}

// v_drive of candidate c >= 2 from its polynomial root (the line after each roots() call of the reference).
HRG_HD double ltp_scaling_vdrive(const LtpLimits& L, int cand, double root, double v_0, double a_0) {
  const double jm = L.j_max;
  if (cand == 2) return (-2.0 * sq(a_0) + 4 * jm * v_0 + sq(root)) / (4 * jm);
  if (cand == 5) return -(sq(root) - sq(a_0) - 2.0 * jm * v_0) / (2.0 * jm);
  return sq(root) / jm;
}

// A timeScaling candidate that is rejected WHATEVER the reference's eigen-solve answers - the one family of close calls
// that is common: a start from exact rest (v_0 = a_0 = 0) makes the candidate polynomials x^k q(x) with k >= 2 EXACT
// trailing zeros. Their companion matrix has a defective k-fold eigenvalue 0, which the QR iteration returns as k
// rounding-noise values of size ~(eps |C|)^(1/k) - up to 1e-3 - in every direction of the complex plane; when one of them
// comes out real and above the 1e-7 cut, roots.h:46 hands it to the planner as "the smallest positive root" in place of
// the true one. Which of the two happens is decided by rounding, so only the exact restatement of the iteration can tell
// (~25 sweeps of a 6x6 matrix: the slowest thing left on the tick's critical path). But the planner only uses the root
// through the trial `optSwitchTimes(v_drive(root))` + acceptance window (:398-405), so if BOTH alternatives are rejected
// the candidate is rejected without knowing which one the iteration picks:
//   * the true root r_t = smallest root > 1e-7 of the deflated q (degree <= 2: closed form), tried as usual;
//   * a noise root g in (1e-7, g_hi]: from rest v_drive(g) = +-g^2 / (c jm) - either <= 0 for every g (candidate 5, rejected
//     by `v_drive > 0`), or a cruise speed of at most g_hi^2 / jm, for which the profile takes T_6 >= |q_goal - q_0| /
//     v_drive; T_6 falls monotonically with the cruise speed, so if the FASTEST noise root already misses the acceptance
//     window by a full second (T_6 - t_required >= 1), every noise root is rejected.
// g_hi = 0.05 is 50x the largest noise root the bound above allows for |C| <= 1e6 and 500x the largest one seen
// (profiles/r02p_rest_start_candidates.md); r_t must lie clear of that range. Anything else (true root accepted, true root
// small, deflated degree > 2, a move so short that the window could be met) returns false and the QR iteration decides.
#if defined(HRG_RULE_DEBUG) && !defined(__CUDA_ARCH__)
static long long hrg_rule_dbg[8];
#endif
HRG_HD_CALL bool ltp_cand_rejected_from_rest(const LtpLimits& L, int cand, const double* A, int ncoef, double q_goal,
                                             double q_0, double v_0, double a_0, double v_orig, double a_orig, double tr,
                                             OptRootCache* cache) {
  if (!(v_0 == 0.0 && a_0 == 0.0)) return false;
  int k = 0;
  while (k < ncoef - 1 && A[ncoef - 1 - k] == 0.0) ++k;
  if (k < 2 || A[0] == 0.0) return false;
  const int nd = ncoef - k;  // coefficients of the deflated polynomial q
  double rt = INFINITY;
  if (nd == 2) {
    const double x = -A[1] / A[0];
    if (x > 1e-7) rt = x;
  } else if (nd == 3) {
    const double disc = A[1] * A[1] - 4.0 * A[0] * A[2];
    if (disc >= 0.0) {
      const double sd = sqrt(disc);
      const double q = -0.5 * (A[1] + (A[1] >= 0.0 ? sd : -sd));
      const double r1 = q / A[0], r2 = (q != 0.0) ? A[2] / q : r1;
      if (r1 > 1e-7) rt = fmin(rt, r1);
      if (r2 > 1e-7) rt = fmin(rt, r2);
      if (!(disc > 1e-6 * A[1] * A[1])) return false;  // q itself has a close call: not ours to decide
    } else if (!(-disc > 1e-6 * A[1] * A[1])) {
      return false;
    }
  } else if (nd > 3) {
    return false;
  }
  const double g_hi = 0.05;
#if defined(HRG_RULE_DEBUG) && !defined(__CUDA_ARCH__)
#define HRG_RD(i) (++hrg_rule_dbg[i])
#else
#define HRG_RD(i) ((void)0)
#endif
  if (!(rt > 10.0 * g_hi)) { HRG_RD(0); return false; }  // (INFINITY passes)
  double T[7];
  int mod = 0;
  ltp_zero7(T);
  if (ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, ltp_scaling_vdrive(L, cand, rt, v_0, a_0), tr, T, mod, cache)) { HRG_RD(1); return false; }
  const double vd_g = ltp_scaling_vdrive(L, cand, g_hi, v_0, a_0);
  if (vd_g <= 0.0) return true;  // -g^2 / (2 jm): no noise root passes `v_drive > 0`
  if (!(vd_g > 0.0)) return false;
  double trash;
  ltp_zero7(T);
  if (!ltp_opt_switch_times(L, q_goal, q_0, v_orig, a_orig, vd_g, T, trash, mod, cache)) { HRG_RD(2); return false; }
  if (T[6] - tr >= 1.0) return true;
  // a joint that is (almost) at its goal: optSwitchTimes returns the braking times before it ever looks at v_drive (:95-99,
  // |q_goal - (q_0 + q_stop)| < 4e-3), so every positive v_drive - noise root or true root - gets the SAME verdict, which
  // the trial of the true root above has already shown to be a rejection
  {
    double tb[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    double q_stop = 0.0, d;
    ltp_opt_braking(L, v_orig, a_orig, q_stop, tb, d);
    const double vd_t = ltp_scaling_vdrive(L, cand, rt, v_0, a_0);
    if (fabs(q_goal - (q_0 + q_stop)) < 4e-3 && vd_t > 0.0) return true;
  }
  HRG_RD(3);
  return false;
}

#if defined(HRG_CAND_LOG) && !defined(__CUDA_ARCH__)
void hrg_cand_log(const LtpLimits& L, int cand, const double* A, int ncoef, double q_goal, double q_0, double v_0, double a_0,
                  double v_orig, double a_orig, double tr, OptRootCache* cache);
#endif
// Sequential form: the candidates one after the other, first accepted wins (host tests, stage kernel).
HRG_HD_CALL bool ltp_time_scaling(const LtpLimits& L, double q_goal, double q_0, double v_0, double a_0, double dir,
                                  double tr /*t_required*/, double* T /*scaled_t*/, double& v_drive,
                                  int& mod_jerk_profile, OptRootCache* cache = nullptr) {
  if (dir < 0) {
    v_0 = -v_0;
    a_0 = -a_0;
  }
  const double v_orig = dir * v_0, a_orig = dir * a_0;  // what the reference passes on to optSwitchTimes
  for (int cand = 0; cand < 8; ++cand) {
    double A[7];
    int ncoef;
    v_drive = ltp_scaling_setup(L, cand, q_goal, q_0, v_0, a_0, dir, tr, A, ncoef);
#if defined(HRG_CAND_LOG) && !defined(__CUDA_ARCH__)
    if (ncoef) hrg_cand_log(L, cand, A, ncoef, q_goal, q_0, v_0, a_0, v_orig, a_orig, tr, cache);
#endif
    if (ncoef) {
      if (L.root_mode == 1 && rr_has_degree(ncoef)) {
        double margin;
        const double r = fr_root(A, ncoef, margin);
        if (margin >= kRootMargin) {
          v_drive = ltp_scaling_vdrive(L, cand, r, v_0, a_0);
        } else if (ltp_cand_rejected_from_rest(L, cand, A, ncoef, q_goal, q_0, v_0, a_0, v_orig, a_orig, tr, cache)) {
          continue;  // rejected whichever way the QR iteration's rounding falls (a rejected trial leaves nothing behind)
        } else {
          LtpLimits Lx = L;
          Lx.root_mode = 0;
          v_drive = ltp_scaling_vdrive(L, cand, smallest_positive_real_root(Lx, A, ncoef), v_0, a_0);
        }
      } else {
        v_drive = ltp_scaling_vdrive(L, cand, smallest_positive_real_root(L, A, ncoef), v_0, a_0);
      }
    }
    if (ltp_try_vdrive(L, q_goal, q_0, v_orig, a_orig, v_drive, tr, T, mod_jerk_profile, cache)) return true;
  }
  // no valid solution (:640-644)
  mod_jerk_profile = 0;
  ltp_zero7(T);
  v_drive = L.v_max;
  return false;
}


}  // namespace hrg
#include "ltp_traj.cuh"
