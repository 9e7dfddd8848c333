// Smallest real root > 1e-7 of a degree 4..6 polynomial WITHOUT the companion-matrix eigen-solve.
//
// roots() + getSmallestPositiveNonComplexRoot (LTP/include/long_term_planner/roots.h:22-50) go through
// Eigen::EigenSolver: a Francis double-shift QR iteration that costs 8..25 sweeps of ~700 (4x4) to ~1500 (6x6) dependent
// FP64 instructions - 85 % of the replan chain, which is the critical path of a tick (profiles/r02i, r02j). The planner
// only ever asks for ONE number: the smallest real root above 1e-7. That number is found here by isolating the real roots
// through the derivative chain:
//   the roots of p' inside (a, B) cut [a, B] (a = 1e-7, B = Cauchy bound) into intervals on which p is monotone, so a sign
//   change of p between two neighbouring cut points brackets exactly one simple root, which a bracketed Newton iteration
//   finds to rounding accuracy; the roots of p' come from p'' the same way, down to the linear derivative.
// Every root this returns is a true root of the polynomial to a few ulps of its condition - the QR iteration's eigenvalues
// carry the same kind of error, so the two agree to ~1e-13 relative (north_star tolerance on setpoints: 1e-5).
//
// Where the two could DISAGREE is the classification, not the value: at a (near-)double root the QR iteration's verdict
// "two real roots / a complex pair" (roots.h:46 `root.imag() == 0`) is decided by its rounding noise, and a root within
// noise of the 1e-7 cut is in or out by noise. The QR iteration is norm-wise backward stable: it returns the eigenvalues
// of C + E with |E| ~ eps |C|, which moves p(x) = det(xI - C) at x by about eps |C| |u(x)| |w(x)| (u = (1, x, .., x^(n-1)),
// w = the Horner partial sums: the right / left eigenvector shapes of a companion matrix). So |p| at a positive critical
// point, or at the cut, is measured in units of that noise; the smallest such ratio is returned as `margin`, and the
// caller runs the exact restatement of the QR iteration (ltp_roots_reg.cuh) whenever margin < kRootMargin, so decisions
// stay those of the reference. The same noise bounds how far the iteration's VALUE of an ill-conditioned root can be
// from the true one (noise / |p'(r)|): a root it could move by more than 1e-11 relative (1e-7 at kRootMargin units of
// noise) is handed over as well.
// Validation: profiles/r02k_root_check.md (device check mode over > 1e7 polynomials of the bench workload) and
// tests/test_golden_vectors.py::test_fast_roots_*.
#pragma once

namespace hrg {

#define HRG_FR_UNROLL _Pragma("unroll")
constexpr double kRootCut = 1e-7;        // roots.h:46
constexpr double kRootMargin = 1e4;      // |p| below this many units of the QR iteration's noise: the exact path decides
constexpr double kRootValueTol = 1e-7;   // relative distance the iteration's noise may move a root that is still trusted

// NMAX = compile-time bound on the degree (array sizes, unrolling), n = the degree at run time (n <= NMAX). With n == NMAX a
// literal the guards fold away; with NMAX = 6 and n = 4..6 one instruction stream serves lanes of different degrees
// (k_replan_deep: six candidates of degrees 4, 4, 5, 4, 4, 6 side by side) - the arithmetic per polynomial is the same
// either way, so both forms return the same bits.
template <int NMAX>
struct FastRootPoly {
  double D[NMAX + 1][NMAX + 1];  // D[k][0..k]: coefficients (leading first) of the (n-k)-th derivative, degree k
};

// value and derivative of level K at x (Horner)
template <int NMAX, int K>
HRG_HD void fr_eval(const FastRootPoly<NMAX>& P, double x, double& g, double& dg) {
  g = P.D[K][0];
  dg = 0.0;
  HRG_FR_UNROLL
  for (int i = 1; i <= K; ++i) {
    dg = dg * x + g;
    g = g * x + P.D[K][i];
  }
}
// |p(x)| (p = level K, the polynomial itself) in units of the noise eps |C| |u(x)| |w(x)| a backward-stable eigen-solve of
// the companion matrix C puts on it (normC >= |C|_F of the monic companion matrix; the partial sums are those of the raw
// coefficients, so p needs no normalisation). Also never better than |p| / (eps * sum |c_i| |x|^(n-i)), the noise of
// evaluating p itself.
template <int NMAX, int K>
HRG_HD double fr_margin(const FastRootPoly<NMAX>& P, double x, double px, double normC) {
  const double ax = fabs(x);
  double w = P.D[K][0], ww = w * w, uu = 1.0, xp = 1.0, s = fabs(w);
  HRG_FR_UNROLL
  for (int i = 1; i < K; ++i) {
    w = w * x + P.D[K][i];
    ww += w * w;
    xp *= x;
    uu += xp * xp;
    s = s * ax + fabs(P.D[K][i]);
  }
  s = s * ax + fabs(P.D[K][K]);
  const double noise_qr = 2.220446049250313e-16 * normC * sqrt(uu * ww), noise_eval = 2.220446049250313e-16 * s;
  const double noise = noise_qr > noise_eval ? noise_qr : noise_eval;
  const double m = fabs(px) / noise;
  return (m == m) ? m : 0.0;  // NaN -> 0
}

// The root of level K inside (lo, hi), where g is monotone and g(lo), g(hi) have opposite signs (flo = g(lo) != 0):
// Newton steps kept inside the bracket, bisection (geometric while the bracket spans more than a factor 4) when a step
// leaves it or does not halve the previous step. Newton converges quadratically, so once a step is below `tol` relative
// the point it leads to is accurate to ~tol^2: the top level (the root that is returned) stops at 1e-11, the derivative
// levels at 1e-8 - a critical point only has to be good enough for the SIGN of the next level there, and that value is
// stationary in the critical point's position.
template <int NMAX, int K>
HRG_HD double fr_bracket(const FastRootPoly<NMAX>& P, double lo, double hi, double flo, double fhi, bool top) {
  const double tol = top ? 1e-11 : 1e-8;
  const bool up = flo < 0.0;  // g rises through the root
  // first guess: secant through the end points when it is meaningful, else the (geometric) middle
  double x;
  {
    const double w = flo / (flo - fhi);
    x = lo + w * (hi - lo);
    if (!(x > lo && x < hi) || hi > 4.0 * lo) x = (hi > 4.0 * lo) ? sqrt(lo * hi) : 0.5 * (lo + hi);
  }
  double dx_old = hi - lo;
  for (int it = 0; it < 100; ++it) {
    double g, dg;
    fr_eval<NMAX, K>(P, x, g, dg);
    if (g == 0.0) return x;
    if ((g < 0.0) == up)
      lo = x;
    else
      hi = x;
    const double step = g / dg;
    double xn = x - step;
    const bool inside = xn >= lo && xn <= hi;
    if (inside && fabs(step) <= tol * fabs(x)) return xn;
    const bool newton_ok = inside && xn != x && fabs(2.0 * step) <= fabs(dx_old);
    if (!newton_ok) {
      xn = (hi > 4.0 * lo) ? sqrt(lo * hi) : 0.5 * (lo + hi);
      if (!(xn > lo && xn < hi)) return x;  // the bracket has collapsed to neighbouring doubles
    }
    dx_old = xn - x;
    x = xn;
  }
  return x;
}

template <int NMAX>
HRG_HD double fr_pick(const double (&R)[NMAX], int i) {
  double v = R[0];
  HRG_FR_UNROLL
  for (int t = 1; t < NMAX; ++t)
    if (i == t) v = R[t];
  return v;
}
template <int NMAX>
HRG_HD void fr_put(double (&R)[NMAX], int i, double v) {
  HRG_FR_UNROLL
  for (int t = 0; t < NMAX; ++t)
    if (i == t) R[t] = v;
}

// One level of the chain: the roots of level K inside (a, B), in ascending order, from the m roots of level K - 1.
// top (K == n): stops at the first root and checks the margins; returns the root (or INFINITY) in `first`.
template <int NMAX, int K>
HRG_HD int fr_level(const FastRootPoly<NMAX>& P, bool top, double a, double B, const double (&Rp)[NMAX], int m,
                    double (&Rn)[NMAX], double& first, double& margin) {
  int cnt = 0;
  double lo = a, flo, dflo;
  fr_eval<NMAX, K>(P, lo, flo, dflo);
  const double normC = sqrt((double)K) * B;
  if (top) {
    first = INFINITY;
    margin = fmin(margin, fr_margin<NMAX, K>(P, lo, flo, normC));  // a root within noise of the 1e-7 cut?
  }
  for (int iv = 0; iv <= m; ++iv) {
    const double hi = (iv == m) ? B : fr_pick<NMAX>(Rp, iv);
    double fhi, dfhi;
    fr_eval<NMAX, K>(P, hi, fhi, dfhi);
    if (top && iv < m)  // hi is a critical point of p: a (near-)double root there is the QR iteration's call, not ours
      margin = fmin(margin, fr_margin<NMAX, K>(P, hi, fhi, normC));
    if ((flo < 0.0) != (fhi < 0.0) && flo != 0.0 && fhi != 0.0) {
      const double r = fr_bracket<NMAX, K>(P, lo, hi, flo, fhi, top);
      if (top) {
        // the VALUE is a close call too when the root is ill conditioned: the QR iteration's noise moves it by
        // noise / |p'(r)|; trusted only while kRootMargin units of noise stay below kRootValueTol relative (the bound is
        // generous: the largest difference seen on 9 M planner polynomials is 8e-10, profiles/r02k_root_check.md)
        double gr, dgr;
        fr_eval<NMAX, K>(P, r, gr, dgr);
        margin = fmin(margin, fr_margin<NMAX, K>(P, r, dgr * r, normC) * (kRootMargin * kRootValueTol));
        first = r;
        return 1;
      }
      fr_put<NMAX>(Rn, cnt, r);
      ++cnt;
    }
    lo = hi;
    flo = fhi;
  }
  return cnt;
}

template <int NMAX, int K>
struct FrChain {
  HRG_HD static int run(const FastRootPoly<NMAX>& P, int n, double a, double B, double (&R)[NMAX], double& first,
                        double& margin) {
    double Rp[NMAX];
    const int m = FrChain<NMAX, K - 1>::run(P, n, a, B, Rp, first, margin);
    if (K > n) {  // (a lane of lower degree in a mixed-degree warp: nothing left to do)
      HRG_FR_UNROLL
      for (int t = 0; t < NMAX; ++t) R[t] = Rp[t];
      return m;
    }
    return fr_level<NMAX, K>(P, K == n, a, B, Rp, m, R, first, margin);
  }
};
// The chain starts at the quadratic derivative, whose roots are closed form (the cancellation-free pair q / a, c / q).
template <int NMAX>
struct FrChain<NMAX, 2> {
  HRG_HD static int run(const FastRootPoly<NMAX>& P, int, double a, double B, double (&R)[NMAX], double&, double&) {
    HRG_FR_UNROLL
    for (int t = 0; t < NMAX; ++t) R[t] = 0.0;
    const double qa = P.D[2][0], qb = P.D[2][1], qc = P.D[2][2];
    const double disc = qb * qb - 4.0 * qa * qc;
    if (!(disc >= 0.0)) return 0;
    const double sd = sqrt(disc);
    const double q = -0.5 * (qb + (qb >= 0.0 ? sd : -sd));
    double r1 = q / qa, r2 = (q != 0.0) ? qc / q : r1;
    if (r2 < r1) {
      const double t = r1;
      r1 = r2;
      r2 = t;
    }
    int cnt = 0;
    if (r1 > a && r1 < B) R[cnt++] = r1;
    if (r2 > a && r2 < B && r2 != r1) {
      if (cnt == 0)
        R[0] = r2;
      else
        R[1] = r2;
      ++cnt;
    }
    return cnt;
  }
};

// coef[0] x^n + ... + coef[n], 3 <= n <= NMAX. Returns the smallest real root > cut (1e-7: roots.h:46; ltp_scaling.cuh also
// asks for the smallest root above the noise range of a root cluster at 0) - INFINITY if none - and, in `margin`,
// how many units of the QR iteration's noise the closest call was away (0: degenerate input) - below kRootMargin the
// exact path must decide.
template <int NMAX>
HRG_HD double fr_solve(const double* coef, int n, double& margin, double cut = kRootCut) {
  margin = INFINITY;
  FastRootPoly<NMAX> P;
  double cmax = 0.0;
  bool finite = true;
  HRG_FR_UNROLL
  for (int k = 0; k <= NMAX; ++k) {
    HRG_FR_UNROLL
    for (int i = 0; i <= NMAX; ++i) P.D[k][i] = 0.0;
  }
  HRG_FR_UNROLL
  for (int i = 0; i <= NMAX; ++i) {
    if (i <= n) {
      const double ci = coef[i];
      HRG_FR_UNROLL
      for (int k = 3; k <= NMAX; ++k)
        if (k == n) P.D[k][i] = ci;
      finite = finite && (fabs(ci) <= 1.7976931348623157e308);  // false for NaN / inf
      if (i > 0) cmax = fabs(ci) > cmax ? fabs(ci) : cmax;
    }
  }
  if (!finite || coef[0] == 0.0) {
    margin = 0.0;
    return INFINITY;
  }
  // Cauchy: every root (and, Gauss-Lucas, every root of every derivative) has modulus below 1 + max |c_i / c_0|
  const double B = (1.0 + cmax / fabs(coef[0])) * 1.0000001;
  if (!(B <= 1.7976931348623157e308)) {
    margin = 0.0;
    return INFINITY;
  }
  HRG_FR_UNROLL
  for (int k = NMAX; k >= 3; --k) {
    if (k <= n) {
      HRG_FR_UNROLL
      for (int i = 0; i < k; ++i) P.D[k - 1][i] = P.D[k][i] * (double)(k - i);
    }
  }
  double R[NMAX], first = INFINITY;
  if (!(B > cut)) return INFINITY;  // no root above the cut
  FrChain<NMAX, NMAX>::run(P, n, cut, B, R, first, margin);
  return first;
}

HRG_HD_CALL double fr_root4(const double* coef, double& margin) { return fr_solve<4>(coef, 4, margin); }
HRG_HD_CALL double fr_root5(const double* coef, double& margin) { return fr_solve<5>(coef, 5, margin); }
HRG_HD_CALL double fr_root6(const double* coef, double& margin) { return fr_solve<6>(coef, 6, margin); }
// one instruction stream for lanes of different degrees (same bits as the three above)
HRG_HD_CALL double fr_root_mixed(const double* coef, int ncoef, double& margin) { return fr_solve<6>(coef, ncoef - 1, margin); }
HRG_HD double fr_root(const double* coef, int ncoef, double& margin) {
  if (ncoef == 5) return fr_root4(coef, margin);
  if (ncoef == 6) return fr_root5(coef, margin);
  return fr_root6(coef, margin);
}

#undef HRG_FR_UNROLL
}  // namespace hrg
