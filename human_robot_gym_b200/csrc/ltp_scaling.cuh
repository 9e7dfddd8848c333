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
    if (ncoef) v_drive = ltp_scaling_vdrive(L, cand, smallest_positive_real_root(L, A, ncoef), v_0, a_0);
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
