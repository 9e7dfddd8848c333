// Per-joint jerk-profile arithmetic of the LongTermPlanner (included from ltp_math.cuh).
// Notation: jm = j_max_[joint], am = a_max_[joint]; t[0..6] relative phase durations, T[0..6] cumulative switch times.
#pragma once

namespace hrg {

// optBraking (LTP/src/long_term_planner.cc:650-701): time-optimal stop from (v_0, a_0).
HRG_HD void ltp_opt_braking(const LtpLimits& L, double v_0, double a_0, double& q, double* t /*[7], 0..2 written*/,
                            double& dir) {
  const double jm = L.j_max, am = L.a_max;
  if (v_0 * a_0 > 0) {
    dir = -isign(v_0);
  } else {
    if (fabs(v_0) > 1.0 / 2.0 * sq(a_0) / jm)
      dir = -isign(v_0);
    else
      dir = -isign(a_0);
  }
  if (dir < 0) {
    a_0 = -a_0;
    v_0 = -v_0;
  }
  t[0] = (am - a_0) / jm;
  t[2] = am / jm;
  t[1] = (-v_0 - 1.0 / 2.0 * t[0] * a_0) / am - 1.0 / 2.0 * (t[0] + t[2]);
  if (t[1] < -L.t_sample) {
    t[0] = -a_0 / jm + sqrt(sq(a_0) / (2 * sq(jm)) - v_0 / jm);
    t[2] = t[0] + a_0 / jm;
    t[1] = 0;
  }
  q = v_0 * (t[0] + t[1] + t[2]) + a_0 * (1.0 / 2.0 * sq(t[0]) + t[0] * (t[1] + t[2]) + 1.0 / 2.0 * sq(t[2])) +
      jm * (1.0 / 6.0 * cube(t[0]) + 1.0 / 2.0 * sq(t[0]) * (t[1] + t[2]) - 1.0 / 6.0 * cube(t[2]) +
            1.0 / 2.0 * t[0] * sq(t[2])) +
      am * (1.0 / 2.0 * sq(t[1]) + t[1] * t[2]);
  q = dir * q;
}

HRG_HD void ltp_zero7(double* T) {
  for (int i = 0; i < 7; ++i) T[i] = 0.0;
}
HRG_HD void ltp_cumsum7(const double* t, double* T) {
  double s = t[0];
  T[0] = s;
  for (int i = 1; i < 7; ++i) {
    s = s + t[i];
    T[i] = s;
  }
}

// The two quartics optSwitchTimes solves in its "neither max velocity nor max acceleration reached" branch (:245-337) have
// coefficients in (q_goal, q_0, v_0, a_0, dir, j_max, a_max) only - NOT in v_drive - and dir is itself a function of
// (q_goal, q_0, v_0, a_0). planTrajectory calls optSwitchTimes for one joint up to nine times with the same four values and
// different v_drive (once with v_max, then once per timeScaling candidate), so the roots are computed at most once per
// (env, joint) and handed on: same inputs, same deterministic routine, hence the same bits as solving again. The cache is
// keyed by the exact bit patterns of the four inputs; a call with any other input (e.g. the -0.0 a mirrored zero velocity
// turns into) solves afresh and leaves the cache alone.
struct OptRootCache {
  double q_goal, q_0, v_0, a_0;  // key
  double root[2];                // :261, :321
  int have;                      // bit 0 / 1: root[0] / root[1] valid; bit 2: key set
};
HRG_HD void opt_cache_clear(OptRootCache& c) {
  c.q_goal = c.q_0 = c.v_0 = c.a_0 = 0.0;
  c.root[0] = c.root[1] = 0.0;
  c.have = 0;
}
HRG_HD bool same_bits(double a, double b) {
#if defined(__CUDA_ARCH__)
  return __double_as_longlong(a) == __double_as_longlong(b);
#else
  long long x, y;
  memcpy(&x, &a, 8);
  memcpy(&y, &b, 8);
  return x == y;
#endif
}
HRG_HD double opt_cached_root(const LtpLimits& L, OptRootCache* c, bool key_ok, int which, const double* A) {
  if (c && key_ok && (c->have & (1 << which))) return c->root[which];
  const double r = smallest_positive_real_root(L, A, 5);
  if (c && key_ok) {
    c->root[which] = r;
    c->have |= 1 << which;
  }
  return r;
}

// Coefficients of the two quartics (:246-261 and :301-321); v_0 / a_0 already mirrored for dir < 0.
HRG_HD void opt_quartic_coef(const LtpLimits& L, int which, double q_goal, double q_0, double v_0, double a_0, double dir,
                             double* A /*[5]*/) {
  const double jm = L.j_max, am = L.a_max;
  if (which == 0) {
    A[0] = 12;
    A[1] = 0;
    A[2] = -24 * sq(a_0) + 48 * jm * v_0;
    A[3] = 48 * dir * sq(jm) * q_0 - 48 * dir * sq(jm) * q_goal + 16 * cube(a_0) - 48 * a_0 * jm * v_0;
    A[4] = -3 * pow4(a_0) + 12.0 * sq(a_0) * jm * v_0 - 12.0 * sq(jm) * sq(v_0);
  } else {
    A[0] = 12;
    A[1] = -24 * am;
    A[2] = -12.0 * sq(a_0) + 12.0 * sq(am) + 24 * jm * v_0;
    A[3] = 0;
    A[4] = 24 * dir * sq(jm) * q_0 * am - 24 * dir * sq(jm) * q_goal * am + 3 * pow4(a_0) + 8 * cube(a_0) * am +
           6 * sq(a_0) * sq(am) - 12.0 * sq(a_0) * jm * v_0 - 24 * a_0 * jm * v_0 * am - 12.0 * sq(am) * jm * v_0 +
           12.0 * sq(jm) * sq(v_0);
  }
}

// Solves quartic `which` of optSwitchTimes for the inputs (q_goal, q_0, v_0, a_0) AHEAD of the call, whether or not the
// call will reach the branch that needs it: the head of optSwitchTimes (:95-112) up to dir and the mirrored v_0 / a_0,
// then the same coefficients and the same solver as the branch itself, so the cached root has the bits the branch would
// compute. Lets every lane of a warp run its eigen-solve at ONE program point (converged) instead of wherever its
// branches lead it.
HRG_HD void opt_prefill_root(const LtpLimits& L, double q_goal, double q_0, double v_0, double a_0, int which,
                             OptRootCache& c) {
  if (!(c.have & 4)) {
    c.q_goal = q_goal; c.q_0 = q_0; c.v_0 = v_0; c.a_0 = a_0;
    c.have |= 4;
  }
  double t[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  double q_stop = 0, dir;
  ltp_opt_braking(L, v_0, a_0, q_stop, t, dir);
  const double q_diff = q_goal - (q_0 + q_stop);
  dir = isign(q_diff);
  if (dir < 0) {
    v_0 = -v_0;
    a_0 = -a_0;
  }
  double A[5];
  opt_quartic_coef(L, which, q_goal, q_0, v_0, a_0, dir, A);
  c.root[which] = smallest_positive_real_root(L, A, 5);
  c.have |= 1 << which;
}

// optSwitchTimes (LTP/src/long_term_planner.cc:82-353). Returns the reference's bool; T = cumulative switch times.
HRG_HD_CALL bool ltp_opt_switch_times(const LtpLimits& L, double q_goal, double q_0, double v_0, double a_0, double v_drive,
                                 double* T, double& dir, int& mod_jerk_profile, OptRootCache* cache = nullptr) {
  const double jm = L.j_max, am = L.a_max;
  bool key_ok = false;
  if (cache) {
    if (!(cache->have & 4)) {
      cache->q_goal = q_goal; cache->q_0 = q_0; cache->v_0 = v_0; cache->a_0 = a_0;
      cache->have |= 4;
    }
    key_ok = same_bits(cache->q_goal, q_goal) && same_bits(cache->q_0, q_0) && same_bits(cache->v_0, v_0) &&
             same_bits(cache->a_0, a_0);
  }
  double t[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  mod_jerk_profile = 0;
  const double eps = 4e-3;
  double q_stop = 0;
  ltp_opt_braking(L, v_0, a_0, q_stop, t, dir);
  double q_diff = q_goal - (q_0 + q_stop);
  if (fabs(q_diff) < eps) {
    ltp_cumsum7(t, T);
    return true;
  }
  dir = isign(q_diff);
  if (dir < 0) {
    v_0 = -v_0;
    a_0 = -a_0;
  }
  double q_brake = 0.0, emp;
  if (v_0 + 0.5 * a_0 * fabs(a_0) / jm > v_drive) {
    mod_jerk_profile = 1;
    ltp_opt_braking(L, v_0 - v_drive, a_0, q_brake, t, emp);
  } else {
    t[0] = (am - a_0) / jm;
    t[2] = am / jm;
    t[1] = (v_drive - v_0 - 0.5 * t[0] * a_0) / am - 0.5 * (t[0] + t[2]);
    if (t[1] < -eps) {
      double root = jm * (v_drive - v_0) + 0.5 * sq(a_0);
      if (root > 0) {
        t[2] = sqrt(root) / jm;
        t[0] = t[2] - a_0 / jm;
        t[1] = 0;
      } else {
        ltp_zero7(T);
        return true;
      }
    }
  }
  t[4] = am / jm;
  t[6] = t[4];
  t[5] = v_drive / am - 1.0 / 2.0 * (t[4] + t[6]);
  if (t[5] < -eps) {
    double root = v_drive / jm;
    if (root > 0) {
      t[4] = sqrt(root);
      t[6] = t[4];
      t[5] = 0;
    } else {
      ltp_zero7(T);
      return true;
    }
  }
  double q_part1 = 0;
  if (mod_jerk_profile == 1) {
    q_part1 = q_brake + v_drive * (t[0] + t[1] + t[2]);
  } else {
    q_part1 = v_0 * (t[0] + t[1] + t[2]) +
              a_0 * (1.0 / 2.0 * sq(t[0]) + t[0] * (t[1] + t[2]) + 1.0 / 2.0 * sq(t[2])) +
              jm * (1.0 / 6.0 * cube(t[0]) + 1.0 / 2.0 * sq(t[0]) * (t[1] + t[2]) - 1.0 / 6.0 * cube(t[2]) +
                    1.0 / 2.0 * t[0] * sq(t[2])) +
              am * (1.0 / 2.0 * sq(t[1]) + t[1] * t[2]);
  }
  double q_part2 = jm * (1.0 / 6.0 * cube(t[6]) + 1.0 / 2.0 * sq(t[6]) * (t[5] + t[4]) - 1.0 / 6.0 * cube(t[4]) +
                         1.0 / 2.0 * t[6] * sq(t[4])) +
                   am * (1.0 / 2.0 * sq(t[5]) + t[5] * t[4]);
  t[3] = ((q_goal - q_0) * dir - q_part1 - q_part2) / v_drive;

  if (t[3] < -eps) {
    if (mod_jerk_profile == 1) {
      ltp_zero7(T);
      return false;
    }
    // max velocity is not reached (:202-223)
    double root = (sq(jm) * pow4(t[0])) / 2 - (sq(jm) * pow4(t[2])) / 4 + (sq(jm) * sq(t[2]) * sq(t[4])) / 2 -
                  (sq(jm) * pow4(t[4])) / 4 + (sq(jm) * pow4(t[6])) / 2 + 2.0 * jm * a_0 * cube(t[0]) -
                  (2.0 * jm * am * cube(t[0])) / 3 - 2.0 * jm * am * t[0] * sq(t[2]) + (2.0 * jm * am * cube(t[2])) / 3 +
                  (2.0 * jm * am * cube(t[4])) / 3 - 2.0 * jm * am * sq(t[4]) * t[6] - (2.0 * jm * am * cube(t[6])) / 3 +
                  2.0 * jm * v_0 * sq(t[0]) + 2.0 * sq(a_0) * sq(t[0]) - 2.0 * a_0 * am * sq(t[0]) -
                  2.0 * a_0 * am * sq(t[2]) + 4 * a_0 * v_0 * t[0] + 2.0 * sq(am) * sq(t[2]) + 2.0 * sq(am) * sq(t[4]) -
                  4 * am * v_0 * t[0] + 4 * dir * (q_goal - q_0) * am + 2.0 * sq(v_0);
    if (root > 0) {
      t[5] = -(4 * am * t[4] - 2.0 * sqrt(root) + jm * sq(t[2]) - jm * sq(t[4]) + 2.0 * jm * sq(t[6])) / (4 * am);
      t[1] = (-v_0 - a_0 * t[0] - 1.0 / 2.0 * jm * sq(t[0]) + 1.0 / 2.0 * jm * sq(t[2]) + 1.0 / 2.0 * jm * sq(t[6]) -
              1.0 / 2.0 * jm * sq(t[4])) / am -
             t[2] + t[5] + t[4];
      t[3] = 0;
    } else {
      ltp_zero7(T);
      return true;
    }
    // neither max velocity nor max acceleration reached (:245-337)
    if (t[5] < -eps || t[1] < -eps) {
      {
        double A[5];
        opt_quartic_coef(L, 0, q_goal, q_0, v_0, a_0, dir, A);
        root = opt_cached_root(L, cache, key_ok, 0, A);
      }
      t[0] = (2.0 * sq(root) - 4 * a_0 * root + sq(a_0) - 2.0 * v_0 * jm) / (4 * jm * root);
      t[6] = sqrt(4 * sq(jm) * sq(t[0]) + 8 * a_0 * jm * t[0] + 2.0 * sq(a_0) + 4 * jm * v_0) / (2.0 * jm);
      t[4] = a_0 / jm + t[0] + t[6];
      t[1] = 0;
      t[5] = 0;
      if (a_0 + t[0] * jm > am) {  // phase 2 exists (:273-296)
        t[0] = (am - a_0) / jm;
        t[6] = 1.0 / jm *
               (am / 2 +
                sqrt(9 * sq(am) +
                     6 * sqrt(-12.0 * am * cube(jm) * cube(t[0]) + 9 * sq(a_0) * sq(jm) * sq(t[0]) -
                              18 * a_0 * am * sq(jm) * sq(t[0]) + 9 * sq(am) * sq(jm) * sq(t[0]) +
                              36 * a_0 * sq(jm) * t[0] * v_0 - 72.0 * am * dir * sq(jm) * q_0 +
                              72.0 * am * dir * sq(jm) * q_goal - 36 * am * sq(jm) * t[0] * v_0 + 3 * pow4(am) +
                              36 * sq(jm) * sq(v_0))) /
                    6.0 -
                am);
        t[4] = t[6] + am / jm;
        t[1] = -(-jm * sq(t[4]) - 2.0 * jm * t[4] * t[6] + jm * sq(t[6]) + a_0 * t[0] + am * t[0] + 2.0 * am * t[4] +
                 2.0 * am * t[6] + 2.0 * v_0) /
               (2.0 * am);
        t[5] = 0;
      }
      if (t[6] * jm > am) {  // phase 6 exists (:299-333)
        t[6] = am / jm;
        double A[5];
        opt_quartic_coef(L, 1, q_goal, q_0, v_0, a_0, dir, A);
        root = opt_cached_root(L, cache, key_ok, 1, A);
        t[0] = (root - a_0 - am) / jm;
        t[4] = (a_0 + am) / jm + t[0];
        t[5] = (sq(jm) * sq(t[0]) + 2.0 * sq(jm) * t[0] * t[4] - sq(jm) * sq(t[4]) + 2.0 * a_0 * jm * t[0] +
                2.0 * a_0 * jm * t[4] - sq(am) + 2.0 * jm * v_0) /
               (2.0 * jm * am);
        t[1] = 0;
      }
      t[2] = 0;
      t[3] = 0;
    }
  }
  // safety checks (:340-348)
  for (int i = 0; i < 7; ++i) {
    if (t[i] < -eps) return false;  // T keeps the caller's previous content, like the reference
    if (t[i] < 0.0 && t[i] >= -eps) t[i] = 0.0;
  }
  ltp_cumsum7(t, T);
  return true;
}

}  // namespace hrg
#include "ltp_scaling.cuh"
