// LongTermPlanner on the device: per-joint time-optimal 7-phase jerk profiles, synchronisation to the slowest joint,
// sampled trajectory. LTP = SS/external/LongTermPlanner (reference tree); everything here follows
// LTP/src/long_term_planner.cc and LTP/include/long_term_planner/roots.h. One lane plans one (env, joint).
#pragma once
#include "path_math.cuh"

namespace hrg {

struct LtpLimits {  // limits of ONE joint (LongTermPlanner members, long_term_planner.h)
  double v_max, a_max, j_max, t_sample;
  // 36 doubles of scratch for the <= 6x6 companion matrix of the matrix-in-memory root finder (host build; on the device
  // only for degrees other than the planner's 4..6, which go to the register-resident solver of ltp_roots_reg.cuh).
  // NULL = a local array on the host; the device needs shared memory here.
  double* mat = nullptr;
  // 0: every root through the exact restatement of the reference's QR iteration; 1: the derivative-chain root finder of
  // ltp_roots_fast.cuh, QR only where its margin is small (the product default, HRG_ROOT_MODE)
  int root_mode = 0;
};

HRG_HD int isign(double v) { return (0.0 < v) - (v < 0.0); }  // long_term_planner.h sign()
HRG_HD double sq(double x) { return x * x; }
HRG_HD double cube(double x) { return x * x * x; }   // gcc folds pow(x,3) only with -ffast-math; glibc pow(x,3) is
HRG_HD double pow4(double x) { return sq(sq(x)); }   // correctly rounded to <1ulp, so differences stay ~1e-16 relative
HRG_HD double pow6(double x) { return sq(cube(x)); }

// ---- polynomial roots: eigenvalues of the companion matrix (roots.h:22-34) ------------------------------------
// Same algorithm family as Eigen::EigenSolver (real Schur form by Francis double-shift QR on the — already upper
// Hessenberg — companion matrix); real eigenvalues are those whose 1x1 block was deflated or whose 2x2 block has a
// non-negative discriminant, which is exactly when Eigen reports imag() == 0 (roots.h:46).
struct SmallMat {  // column-major with a fixed leading dimension of 6
  double* a;
  int n;
  HRG_HD double& operator()(int i, int j) { return a[j * 6 + i]; }
};

HRG_HD void mk_householder(const double* x, int m, double* ess, double& tau, double& beta) {
  double tail = 0.0;
  for (int i = 1; i < m; ++i) tail += x[i] * x[i];
  const double c0 = x[0];
  if (tail <= 2.2250738585072014e-308) {
    tau = 0.0;
    beta = c0;
    for (int i = 1; i < m; ++i) ess[i - 1] = 0.0;
  } else {
    beta = sqrt(c0 * c0 + tail);
    if (c0 >= 0.0) beta = -beta;
    for (int i = 1; i < m; ++i) ess[i - 1] = x[i] / (c0 - beta);
    tau = (beta - c0) / beta;
  }
}
HRG_HD void hh_left(SmallMat& M, int r0, int c0, int nr, int nc, const double* ess, double tau) {
  if (nr == 1) {
    for (int j = 0; j < nc; ++j) M(r0, c0 + j) *= (1.0 - tau);
  } else if (tau != 0.0) {
    for (int j = 0; j < nc; ++j) {
      double tmp = 0.0;
      for (int i = 1; i < nr; ++i) tmp += ess[i - 1] * M(r0 + i, c0 + j);
      tmp += M(r0, c0 + j);
      M(r0, c0 + j) -= tau * tmp;
      for (int i = 1; i < nr; ++i) M(r0 + i, c0 + j) -= tau * ess[i - 1] * tmp;
    }
  }
}
HRG_HD void hh_right(SmallMat& M, int r0, int c0, int nr, int nc, const double* ess, double tau) {
  if (nc == 1) {
    for (int i = 0; i < nr; ++i) M(r0 + i, c0) *= (1.0 - tau);
  } else if (tau != 0.0) {
    for (int i = 0; i < nr; ++i) {
      double tmp = 0.0;
      for (int j = 1; j < nc; ++j) tmp += M(r0 + i, c0 + j) * ess[j - 1];
      tmp += M(r0 + i, c0);
      M(r0 + i, c0) -= tau * tmp;
      for (int j = 1; j < nc; ++j) M(r0 + i, c0 + j) -= tau * tmp * ess[j - 1];
    }
  }
}

// In-place real Schur form of an upper Hessenberg matrix (eigenvalues only).
HRG_HD void real_schur_hessenberg(SmallMat& T) {
  const int n = T.n;
  const double eps = 2.220446049250313e-16, dmin = 2.2250738585072014e-308;
  int iu = n - 1, iter = 0, total_iter = 0;
  const int max_iters = 40 * n;
  double exshift = 0.0, norm = 0.0;
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < (n < j + 2 ? n : j + 2); ++i) norm += fabs(T(i, j));
  double caz = norm * eps * eps;
  if (caz < dmin) caz = dmin;
  if (norm == 0.0) return;
  while (iu >= 0) {
    int il = iu;
    while (il > 0) {
      double s = fabs(T(il - 1, il - 1)) + fabs(T(il, il));
      s = s * eps > caz ? s * eps : caz;
      if (fabs(T(il, il - 1)) <= s) break;
      il--;
    }
    if (il == iu) {
      T(iu, iu) += exshift;
      if (iu > 0) T(iu, iu - 1) = 0.0;
      iu--;
      iter = 0;
    } else if (il == iu - 1) {
      double p = 0.5 * (T(iu - 1, iu - 1) - T(iu, iu));
      double q = p * p + T(iu, iu - 1) * T(iu - 1, iu);
      T(iu, iu) += exshift;
      T(iu - 1, iu - 1) += exshift;
      if (q >= 0.0) {
        double z = sqrt(fabs(q));
        double gp = (p >= 0.0) ? p + z : p - z, gq = T(iu, iu - 1), c, s;
        if (gq == 0.0) {
          c = gp < 0.0 ? -1.0 : 1.0;
          s = 0.0;
        } else if (gp == 0.0) {
          c = 0.0;
          s = gq < 0.0 ? 1.0 : -1.0;
        } else if (fabs(gp) > fabs(gq)) {
          double t = gq / gp, u = sqrt(1.0 + t * t);
          if (gp < 0.0) u = -u;
          c = 1.0 / u;
          s = -t * c;
        } else {
          double t = gp / gq, u = sqrt(1.0 + t * t);
          if (gq < 0.0) u = -u;
          s = -1.0 / u;
          c = -t * s;
        }
        for (int j = iu - 1; j < n; ++j) {
          double x = T(iu - 1, j), y = T(iu, j);
          T(iu - 1, j) = c * x - s * y;
          T(iu, j) = s * x + c * y;
        }
        for (int i = 0; i <= iu; ++i) {
          double x = T(i, iu - 1), y = T(i, iu);
          T(i, iu - 1) = c * x - s * y;
          T(i, iu) = s * x + c * y;
        }
        T(iu, iu - 1) = 0.0;
      }
      if (iu > 1) T(iu - 1, iu - 2) = 0.0;
      iu -= 2;
      iter = 0;
    } else {
      double sh0 = T(iu, iu), sh1 = T(iu - 1, iu - 1), sh2 = T(iu, iu - 1) * T(iu - 1, iu);
      if (iter == 10) {
        exshift += sh0;
        for (int i = 0; i <= iu; ++i) T(i, i) -= sh0;
        double s = fabs(T(iu, iu - 1)) + fabs(T(iu - 1, iu - 2));
        sh0 = 0.75 * s;
        sh1 = 0.75 * s;
        sh2 = -0.4375 * s * s;
      }
      if (iter == 30) {
        double s = (sh1 - sh0) / 2.0;
        s = s * s + sh2;
        if (s > 0.0) {
          s = sqrt(s);
          if (sh1 < sh0) s = -s;
          s = s + (sh1 - sh0) / 2.0;
          s = sh0 - sh2 / s;
          exshift += s;
          for (int i = 0; i <= iu; ++i) T(i, i) -= s;
          sh0 = sh1 = sh2 = 0.964;
        }
      }
      iter++;
      total_iter++;
      if (total_iter > max_iters) break;
      int im;
      double v[3] = {0.0, 0.0, 0.0};
      for (im = iu - 2; im >= il; --im) {
        const double Tmm = T(im, im);
        const double r = sh0 - Tmm, s = sh1 - Tmm;
        v[0] = (r * s - sh2) / T(im + 1, im) + T(im, im + 1);
        v[1] = T(im + 1, im + 1) - Tmm - r - s;
        v[2] = T(im + 2, im + 1);
        if (im == il) break;
        const double lhs = T(im, im - 1) * (fabs(v[1]) + fabs(v[2]));
        const double rhs = v[0] * (fabs(T(im - 1, im - 1)) + fabs(Tmm) + fabs(T(im + 1, im + 1)));
        if (fabs(lhs) < eps * rhs) break;
      }
      for (int k = im; k <= iu - 2; ++k) {
        const bool first = (k == im);
        double w[3];
        if (first) {
          w[0] = v[0]; w[1] = v[1]; w[2] = v[2];
        } else {
          w[0] = T(k, k - 1); w[1] = T(k + 1, k - 1); w[2] = T(k + 2, k - 1);
        }
        double tau, beta, ess[2];
        mk_householder(w, 3, ess, tau, beta);
        if (beta != 0.0) {
          if (first && k > il)
            T(k, k - 1) = -T(k, k - 1);
          else if (!first)
            T(k, k - 1) = beta;
          hh_left(T, k, k, 3, n - k, ess, tau);
          hh_right(T, 0, k, (iu < k + 3 ? iu : k + 3) + 1, 3, ess, tau);
        }
      }
      {
        double w[2] = {T(iu - 1, iu - 2), T(iu, iu - 2)};
        double tau, beta, ess[1];
        mk_householder(w, 2, ess, tau, beta);
        if (beta != 0.0) {
          T(iu - 1, iu - 2) = beta;
          hh_left(T, iu - 1, iu - 1, 2, n - iu + 1, ess, tau);
          hh_right(T, 0, iu - 1, iu + 1, 2, ess, tau);
        }
      }
      for (int i = im + 2; i <= iu; ++i) {
        T(i, i - 2) = 0.0;
        if (i > im + 2) T(i, i - 3) = 0.0;
      }
    }
  }
}

}  // namespace hrg
#include "ltp_roots_reg.cuh"
#include "ltp_roots_fast.cuh"
namespace hrg {

#if defined(HRG_POLY_LOG) && !defined(__CUDA_ARCH__)
void hrg_poly_log(const double* coef, int ncoef);
#endif
// roots() + getSmallestPositiveNonComplexRoot (roots.h:22-50): coef[0] is the leading coefficient, ncoef = degree+1.
// Returns +inf if there is no real root > 1e-7.
HRG_HD_CALL double smallest_positive_real_root(const LtpLimits& L, const double* coef, int ncoef) {
#if defined(HRG_POLY_LOG) && !defined(__CUDA_ARCH__)
  hrg_poly_log(coef, ncoef);  // host-only instrumentation (profiles/harvest_polys.py)
#endif
  if (rr_has_degree(ncoef)) {
    if (L.root_mode == 1) {  // derivative-chain finder; the QR iteration below stays the arbiter of close calls
      double margin;
      const double r = fr_root(coef, ncoef, margin);
      if (margin >= kRootMargin) return r;
#if defined(__CUDA_ARCH__) && defined(HRG_TASK_TRACE)
      HRG_TRACE_COUNT(1);
#endif
    }
#if defined(__CUDA_ARCH__) || defined(HRG_HOST_USE_RR)
    // one problem per lane: the register-resident solver (bit-identical, ltp_roots_reg.cuh) for the planner's degrees
    // (HRG_HOST_USE_RR: the host emulation can be built on it too, tests/host_emul/build.sh)
    return rr_root(coef, ncoef);
#endif
  }
  SmallMat C;
#if defined(__CUDA_ARCH__)
  // every device caller hands in per-lane shared-memory scratch; telling the compiler so turns the matrix accesses of the
  // QR sweeps from generic loads (long-scoreboard latency) into LDS / STS
  C.a = L.mat;
  __builtin_assume(__isShared(C.a));
#else
  double local_mat[36];
  C.a = L.mat ? L.mat : local_mat;
#endif
  const int n = ncoef - 1;
  C.n = n;
  for (int i = 0; i < 36; ++i) C.a[i] = 0.0;
  for (int i = 0; i + 1 < n; ++i) C(i + 1, i) = 1.0;
  // poly = -1 * poly / poly[0]; last column = reversed poly, first n entries (constant term on top)
  for (int i = 0; i < n; ++i) C(i, n - 1) = (-1 * coef[ncoef - 1 - i]) / coef[0];
  // EigenSolver: scale by the largest |coefficient|, (Hessenberg reduction is the identity here), Schur, scale back
  double scale = 0.0;
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < n; ++i) scale = fabs(C(i, j)) > scale ? fabs(C(i, j)) : scale;
  double best = INFINITY;
  if (scale < 2.2250738585072014e-308) return best;  // all eigenvalues 0
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < n; ++i)
      if (C(i, j) != 0.0) C(i, j) /= scale;  // (+-0 / scale = +-0: skipping keeps the bits and the device's fast division path)
  real_schur_hessenberg(C);
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < n; ++i) C(i, j) *= scale;
  int i = 0;
  while (i < n) {
    if (i == n - 1 || C(i + 1, i) == 0.0) {
      double re = C(i, i);
      if (re > 1e-7) best = fmin(best, re);
      ++i;
    } else {
      // complex pair unless the discriminant is exactly 0 (then z == 0 and Eigen reports imag() == +-0)
      double p = 0.5 * (C(i, i) - C(i + 1, i + 1));
      double t0 = C(i + 1, i), t1 = C(i, i + 1);
      double maxval = fmax(fabs(p), fmax(fabs(t0), fabs(t1)));
      t0 /= maxval;
      t1 /= maxval;
      double p0 = p / maxval;
      double z = maxval * sqrt(fabs(p0 * p0 + t0 * t1));
      if (z == 0.0) {
        double re = C(i + 1, i + 1) + p;
        if (re > 1e-7) best = fmin(best, re);
      }
      i += 2;
    }
  }
  return best;
}

}  // namespace hrg
#include "ltp_profiles.cuh"
