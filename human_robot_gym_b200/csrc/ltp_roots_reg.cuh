// Register-resident polynomial root finder: one problem per LANE, the <= 6x6 companion matrix held in registers.
//
// Same algorithm, same operations on every matrix element in the same order as real_schur_hessenberg (ltp_math.cuh), i.e.
// as Eigen::EigenSolver's real Schur form of the companion matrix (LTP/include/long_term_planner/roots.h:22-50) - the roots
// are bit-identical. What differs is the shape of the code: the matrix size N is a template parameter and every loop over
// matrix rows / columns is unrolled over ALL positions and guarded by the run-time window (il, im, iu), so that every
// matrix index is a compile-time constant and the matrix never leaves the register file. The shared-memory form needs an
// address computation and an LDS/STS round trip per element access and 8 lanes per problem; this one needs neither, and
// the 32 lanes of a warp solve 32 different polynomials in one instruction stream (lanes only diverge on the three-way
// case split of a QR step and on the number of steps).
#pragma once

namespace hrg {

#define HRG_RR_UNROLL _Pragma("unroll")
#if defined(HRG_RR_TRACE) && !defined(__CUDA_ARCH__)
static long long hrg_rr_trace_hist[8][64];  // [N][Francis step number] -> how many solves reached it
#endif

// Householder vector of (x0, x1, x2) / (x0, x1): mk_householder (ltp_math.cuh) for m = 3 / m = 2.
HRG_HD void rr_householder3(double x0, double x1, double x2, double& e0, double& e1, double& tau, double& beta) {
  double tail = 0.0;
  tail += x1 * x1;
  tail += x2 * x2;
  if (tail <= 2.2250738585072014e-308) {
    tau = 0.0;
    beta = x0;
    e0 = 0.0;
    e1 = 0.0;
  } else {
    beta = sqrt(x0 * x0 + tail);
    if (x0 >= 0.0) beta = -beta;
    e0 = x1 / (x0 - beta);
    e1 = x2 / (x0 - beta);
    tau = (beta - x0) / beta;
  }
}
HRG_HD void rr_householder2(double x0, double x1, double& e0, double& tau, double& beta) {
  double tail = 0.0;
  tail += x1 * x1;
  if (tail <= 2.2250738585072014e-308) {
    tau = 0.0;
    beta = x0;
    e0 = 0.0;
  } else {
    beta = sqrt(x0 * x0 + tail);
    if (x0 >= 0.0) beta = -beta;
    e0 = x1 / (x0 - beta);
    tau = (beta - x0) / beta;
  }
}

// In-place real Schur form (eigenvalues only) of the upper Hessenberg matrix T[row][col]; see real_schur_hessenberg.
template <int N>
HRG_HD void rr_real_schur(double (&T)[N][N]) {
  const double eps = 2.220446049250313e-16, dmin = 2.2250738585072014e-308;
  int iu = N - 1, iter = 0, total_iter = 0;
  const int max_iters = 40 * N;
  double exshift = 0.0, norm = 0.0;
  HRG_RR_UNROLL
  for (int j = 0; j < N; ++j) {
    HRG_RR_UNROLL
    for (int i = 0; i < N; ++i)
      if (i < j + 2) norm += fabs(T[i][j]);
  }
  double caz = norm * eps * eps;
  if (caz < dmin) caz = dmin;
  if (norm == 0.0) return;
  while (iu >= 0) {
    // il: walk up the sub-diagonal from iu until a negligible entry
    int il = iu;
    {
      bool go = true;
      HRG_RR_UNROLL
      for (int l = N - 1; l >= 1; --l) {
        if (go && l <= iu) {
          double s = fabs(T[l - 1][l - 1]) + fabs(T[l][l]);
          s = s * eps > caz ? s * eps : caz;
          if (fabs(T[l][l - 1]) <= s)
            go = false;
          else
            il = l - 1;
        }
      }
    }
    if (il == iu) {  // one root found
      HRG_RR_UNROLL
      for (int u = 0; u < N; ++u)
        if (iu == u) {
          T[u][u] += exshift;
          if (u > 0) T[u][u > 0 ? u - 1 : 0] = 0.0;
        }
      iu--;
      iter = 0;
    } else if (il == iu - 1) {  // two roots found
      HRG_RR_UNROLL
      for (int u = 1; u < N; ++u)
        if (iu == u) {
          const double p = 0.5 * (T[u - 1][u - 1] - T[u][u]);
          const double q = p * p + T[u][u - 1] * T[u - 1][u];
          T[u][u] += exshift;
          T[u - 1][u - 1] += exshift;
          if (q >= 0.0) {
            const double z = sqrt(fabs(q));
            const double gp = (p >= 0.0) ? p + z : p - z, gq = T[u][u - 1];
            double c, s;
            if (gq == 0.0) {
              c = gp < 0.0 ? -1.0 : 1.0;
              s = 0.0;
            } else if (gp == 0.0) {
              c = 0.0;
              s = gq < 0.0 ? 1.0 : -1.0;
            } else if (fabs(gp) > fabs(gq)) {
              double t = gq / gp, w = sqrt(1.0 + t * t);
              if (gp < 0.0) w = -w;
              c = 1.0 / w;
              s = -t * c;
            } else {
              double t = gp / gq, w = sqrt(1.0 + t * t);
              if (gq < 0.0) w = -w;
              s = -1.0 / w;
              c = -t * s;
            }
            HRG_RR_UNROLL
            for (int j = u - 1; j < N; ++j) {
              const double x = T[u - 1][j], y = T[u][j];
              T[u - 1][j] = c * x - s * y;
              T[u][j] = s * x + c * y;
            }
            HRG_RR_UNROLL
            for (int i = 0; i <= u; ++i) {
              const double x = T[i][u - 1], y = T[i][u];
              T[i][u - 1] = c * x - s * y;
              T[i][u] = s * x + c * y;
            }
            T[u][u - 1] = 0.0;
          }
          if (u > 1) T[u - 1][u > 1 ? u - 2 : 0] = 0.0;
        }
      iu -= 2;
      iter = 0;
    } else {  // no convergence yet: one Francis double-shift QR step on rows / columns il .. iu   (iu >= il + 2)
      double sh0 = 0.0, sh1 = 0.0, sh2 = 0.0, sub1 = 0.0, sub2 = 0.0;
      HRG_RR_UNROLL
      for (int u = 2; u < N; ++u)
        if (iu == u) {
          sh0 = T[u][u];
          sh1 = T[u - 1][u - 1];
          sh2 = T[u][u - 1] * T[u - 1][u];
          sub1 = T[u][u - 1];
          sub2 = T[u - 1][u - 2];
        }
      if (iter == 10) {  // Wilkinson's original ad hoc shift
        exshift += sh0;
        HRG_RR_UNROLL
        for (int i = 0; i < N; ++i)
          if (i <= iu) T[i][i] -= sh0;
        const double s = fabs(sub1) + fabs(sub2);
        sh0 = 0.75 * s;
        sh1 = 0.75 * s;
        sh2 = -0.4375 * s * s;
      }
      if (iter == 30) {  // MATLAB's new ad hoc shift
        double s = (sh1 - sh0) / 2.0;
        s = s * s + sh2;
        if (s > 0.0) {
          s = sqrt(s);
          if (sh1 < sh0) s = -s;
          s = s + (sh1 - sh0) / 2.0;
          s = sh0 - sh2 / s;
          exshift += s;
          HRG_RR_UNROLL
          for (int i = 0; i < N; ++i)
            if (i <= iu) T[i][i] -= s;
          sh0 = sh1 = sh2 = 0.964;
        }
      }
      iter++;
      total_iter++;
#if defined(HRG_RR_TRACE) && !defined(__CUDA_ARCH__)
      ++hrg_rr_trace_hist[N][total_iter < 63 ? total_iter : 63];  // host-only instrumentation (tests/host_emul)
#endif
      if (total_iter > max_iters) break;
      // im: two consecutive small sub-diagonal elements
      int im = il;
      double v0 = 0.0, v1 = 0.0, v2 = 0.0;
      {
        bool act = true;
        HRG_RR_UNROLL
        for (int m = N - 3; m >= 0; --m) {
          if (act && m <= iu - 2 && m >= il) {
            const double Tmm = T[m][m];
            const double r = sh0 - Tmm, s = sh1 - Tmm;
            v0 = (r * s - sh2) / T[m + 1][m] + T[m][m + 1];
            v1 = T[m + 1][m + 1] - Tmm - r - s;
            v2 = T[m + 2][m + 1];
            if (m == il) {
              im = m;
              act = false;
            } else if (m > 0) {  // (m > il >= 0)
              const double lhs = T[m][m > 0 ? m - 1 : 0] * (fabs(v1) + fabs(v2));
              const double rhs = v0 * (fabs(T[m > 0 ? m - 1 : 0][m > 0 ? m - 1 : 0]) + fabs(Tmm) + fabs(T[m + 1][m + 1]));
              if (fabs(lhs) < eps * rhs) {
                im = m;
                act = false;
              }
            }
          }
        }
      }
      // the 3 x 3 reflectors chase the bulge from row im to row iu - 2
      HRG_RR_UNROLL
      for (int k = 0; k <= N - 3; ++k) {
        if (k >= im && k <= iu - 2) {
          const bool first = (k == im);
          const int km1 = k > 0 ? k - 1 : 0;
          double w0, w1, w2;
          if (first) {
            w0 = v0; w1 = v1; w2 = v2;
          } else {
            w0 = T[k][km1]; w1 = T[k + 1][km1]; w2 = T[k + 2][km1];
          }
          double tau, beta, e0, e1;
          rr_householder3(w0, w1, w2, e0, e1, tau, beta);
          if (beta != 0.0) {
            if (k > 0) {
              if (first && k > il)
                T[k][km1] = -T[k][km1];
              else if (!first)
                T[k][km1] = beta;
            }
            if (tau != 0.0) {
              HRG_RR_UNROLL
              for (int j = k; j < N; ++j) {  // hh_left(T, k, k, 3, n - k)
                double tmp = 0.0;
                tmp += e0 * T[k + 1][j];
                tmp += e1 * T[k + 2][j];
                tmp += T[k][j];
                T[k][j] -= tau * tmp;
                T[k + 1][j] -= tau * e0 * tmp;
                T[k + 2][j] -= tau * e1 * tmp;
              }
              HRG_RR_UNROLL
              for (int i = 0; i < N; ++i) {  // hh_right(T, 0, k, min(iu, k + 3) + 1, 3)
                if (i <= k + 3 && i <= iu) {
                  double tmp = 0.0;
                  tmp += T[i][k + 1] * e0;
                  tmp += T[i][k + 2] * e1;
                  tmp += T[i][k];
                  T[i][k] -= tau * tmp;
                  T[i][k + 1] -= tau * tmp * e0;
                  T[i][k + 2] -= tau * tmp * e1;
                }
              }
            }
          }
        }
      }
      // the last, 2 x 2 reflector on rows iu - 1, iu
      HRG_RR_UNROLL
      for (int u = 2; u < N; ++u)
        if (iu == u) {
          double tau, beta, e0;
          rr_householder2(T[u - 1][u - 2], T[u][u - 2], e0, tau, beta);
          if (beta != 0.0) {
            T[u - 1][u - 2] = beta;
            if (tau != 0.0) {
              HRG_RR_UNROLL
              for (int j = u - 1; j < N; ++j) {  // hh_left(T, iu - 1, iu - 1, 2, n - iu + 1)
                double tmp = 0.0;
                tmp += e0 * T[u][j];
                tmp += T[u - 1][j];
                T[u - 1][j] -= tau * tmp;
                T[u][j] -= tau * e0 * tmp;
              }
              HRG_RR_UNROLL
              for (int i = 0; i <= u; ++i) {  // hh_right(T, 0, iu - 1, iu + 1, 2)
                double tmp = 0.0;
                tmp += T[i][u] * e0;
                tmp += T[i][u - 1];
                T[i][u - 1] -= tau * tmp;
                T[i][u] -= tau * tmp * e0;
              }
            }
          }
        }
      // clean up the pollution below the sub-diagonal
      HRG_RR_UNROLL
      for (int i = 2; i < N; ++i)
        if (i >= im + 2 && i <= iu) {
          T[i][i - 2] = 0.0;
          if (i >= 3 && i > im + 2) T[i][i >= 3 ? i - 3 : 0] = 0.0;
        }
    }
  }
}

// roots() + getSmallestPositiveNonComplexRoot (roots.h:22-50) of the degree-N polynomial coef[0] x^N + ... + coef[N]:
// the register form of smallest_positive_real_root (ltp_math.cuh), bit-identical to it.
template <int N>
HRG_HD double rr_smallest_positive_real_root(const double* coef) {
  double C[N][N];
  HRG_RR_UNROLL
  for (int i = 0; i < N; ++i) {
    HRG_RR_UNROLL
    for (int j = 0; j < N; ++j) C[i][j] = 0.0;
  }
  HRG_RR_UNROLL
  for (int i = 0; i + 1 < N; ++i) C[i + 1][i] = 1.0;
  HRG_RR_UNROLL
  for (int i = 0; i < N; ++i) C[i][N - 1] = (-1 * coef[N - i]) / coef[0];
  double scale = 0.0;
  HRG_RR_UNROLL
  for (int i = 0; i < N; ++i) scale = fabs(C[i][N - 1]) > scale ? fabs(C[i][N - 1]) : scale;
  if (N > 1 && 1.0 > scale) scale = 1.0;  // the sub-diagonal ones
  double best = INFINITY;
  if (scale < 2.2250738585072014e-308) return best;
  // C /= scale: only the sub-diagonal and the last column can be non-zero (+-0 / scale = +-0 keeps its bits)
  HRG_RR_UNROLL
  for (int i = 0; i + 1 < N; ++i) C[i + 1][i] /= scale;
  HRG_RR_UNROLL
  for (int i = 0; i < N; ++i)
    if (C[i][N - 1] != 0.0) C[i][N - 1] /= scale;
  rr_real_schur<N>(C);
  HRG_RR_UNROLL
  for (int i = 0; i < N; ++i) {
    HRG_RR_UNROLL
    for (int j = 0; j < N; ++j) C[i][j] *= scale;
  }
  bool skip = false;
  HRG_RR_UNROLL
  for (int i = 0; i < N; ++i) {
    if (skip) {
      skip = false;
    } else if (i == N - 1 || C[i + 1 < N ? i + 1 : i][i] == 0.0) {
      const double re = C[i][i];
      if (re > 1e-7) best = fmin(best, re);
    } else {
      const int i1 = i + 1 < N ? i + 1 : i;
      const double p = 0.5 * (C[i][i] - C[i1][i1]);
      double t0 = C[i1][i], t1 = C[i][i1];
      const double maxval = fmax(fabs(p), fmax(fabs(t0), fabs(t1)));
      t0 /= maxval;
      t1 /= maxval;
      const double p0 = p / maxval;
      const double z = maxval * sqrt(fabs(p0 * p0 + t0 * t1));
      if (z == 0.0) {
        const double re = C[i1][i1] + p;
        if (re > 1e-7) best = fmin(best, re);
      }
      skip = true;
    }
  }
  return best;
}

// The three degrees the planner solves (long_term_planner.cc:257,317 quartics of optSwitchTimes; :468-625 degree 4, 5, 6 of
// timeScaling) as real functions: each is a few thousand instructions of straight-line code, one copy per kernel.
HRG_HD_CALL double rr_root4(const double* coef) { return rr_smallest_positive_real_root<4>(coef); }
HRG_HD_CALL double rr_root5(const double* coef) { return rr_smallest_positive_real_root<5>(coef); }
HRG_HD_CALL double rr_root6(const double* coef) { return rr_smallest_positive_real_root<6>(coef); }
HRG_HD bool rr_has_degree(int ncoef) { return ncoef >= 5 && ncoef <= 7; }
HRG_HD double rr_root(const double* coef, int ncoef) {
  if (ncoef == 5) return rr_root4(coef);
  if (ncoef == 6) return rr_root5(coef);
  return rr_root6(coef);
}

#undef HRG_RR_UNROLL
}  // namespace hrg
