// Scalar kernels of the divide-and-conquer merge (fp64), usable from host and device so the logic can be
// unit-tested on the CPU (tests/test_dc_core.py builds a host harness with g++) against oracle/eigh_model.py.
//
// One merge joins two solved halves of a symmetric tridiagonal matrix:  D + rho z z^T  with D diagonal.
//   1. deflate_scan      : drop components with negligible z, rotate away (near-)equal poles
//   2. secular_root(_lane): roots of 1 + rho sum z_i^2 / (d_i - lambda) = 0, each as (origin pole, offset)
//   3. Loewner z-hat + eigenvectors are assembled by the callers (eigh.cuh)
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define KFAC_HD __host__ __device__ __forceinline__
#else
#define KFAC_HD inline
#endif

namespace kfac {

constexpr double DC_TOL_SCALE = 7.5e-9;  // deflation tolerance relative to max(|d|, rho): below fp32 resolution

struct DeflateCounts {
  int k;      // entries that stay in the secular problem
  int ndefl;  // deflated entries
  int nrot;   // Givens rotations recorded
};

// Sequential scan over the m entries sorted by pole (ds ascending, zs the matching z components, |z| = 1).
// Outputs (all in sorted-position space):
//   keep[0..k)       positions that stay, ascending pole order
//   defl[0..ndefl)   deflated positions
//   rot_i/rot_j/rot_c/rot_s[0..nrot): columns (i, j) of Q are replaced by (c q_i + s q_j, -s q_i + c q_j)
// ds / zs are updated in place for rotated pairs.
KFAC_HD DeflateCounts deflate_scan(int m, double rho, double* ds, double* zs, int* keep, int* defl, int* rot_i,
                                   int* rot_j, double* rot_c, double* rot_s) {
  double dmax = 0.0;
  for (int i = 0; i < m; ++i) {
    const double a = fabs(ds[i]);
    if (a > dmax) dmax = a;
  }
  const double tol = DC_TOL_SCALE * (dmax > rho ? dmax : rho);
  DeflateCounts cnt = {0, 0, 0};
  int prev = -1;
  for (int i = 0; i < m; ++i) {
    if (rho * fabs(zs[i]) <= tol) {
      defl[cnt.ndefl++] = i;
      continue;
    }
    if (prev >= 0) {
      double s = zs[prev], c = zs[i];
      const double tau = hypot(c, s);
      const double t = ds[i] - ds[prev];
      c /= tau;
      s = -s / tau;
      if (fabs(t * c * s) <= tol) {
        zs[i] = tau;
        zs[prev] = 0.0;
        rot_i[cnt.nrot] = prev;
        rot_j[cnt.nrot] = i;
        rot_c[cnt.nrot] = c;
        rot_s[cnt.nrot] = s;
        ++cnt.nrot;
        const double dp = ds[prev] * c * c + ds[i] * s * s;
        const double di = ds[prev] * s * s + ds[i] * c * c;
        ds[prev] = dp;
        ds[i] = di;
        defl[cnt.ndefl++] = prev;
        prev = i;
        continue;
      }
      keep[cnt.k++] = prev;
    }
    prev = i;
  }
  if (prev >= 0) keep[cnt.k++] = prev;
  return cnt;
}

// Evaluation of g(t) = 1 + rho sum z2_i / (delta_i - sgn t) and its derivative w.r.t. mu = sgn t over a strided
// slice of the poles (lane, nlanes): callers reduce the partial sums (warp shuffle on the GPU, nlanes = 1 on host).
KFAC_HD void secular_partial(const double* dk, const double* z2, int k, double dorigin, double sgn, double t, int lane,
                             int nlanes, double& sum, double& dsum) {
  sum = 0.0;
  dsum = 0.0;
  for (int i = lane; i < k; i += nlanes) {
    const double den = (dk[i] - dorigin) - sgn * t;
    const double q = z2[i] / den;
    sum += q;
    dsum += q / den;
  }
}

// State machine of the safeguarded root finder so that the host (serial) and device (warp-cooperative) drivers
// share every decision.  Usage:
//   SecularState st; secular_init(...);  // needs g(mid) for interior roots: see secular_choose_origin
//   loop: evaluate g, g' at st.t  ->  secular_update(st, g, gprime)  until st.done
struct SecularState {
  int origin;
  double sgn;   // mu = sgn * t, t > 0
  double tlo, thi, t;
  int phase;    // 0: validating the lower bound, 1: iterating
  int iters;
  int done;
};

KFAC_HD double secular_pick(double tlo, double thi) {
  if (tlo > 0.0 && thi / tlo > 16.0) return sqrt(tlo * thi);
  return 0.5 * (tlo + thi);
}

// j < k-1: interior root in (d_j, d_{j+1}); fmid = g evaluated at the midpoint with origin j, t = gap/2.
KFAC_HD void secular_init_interior(SecularState& st, const double* dk, const double* z2, double rho, int j,
                                   double fmid) {
  const double gap = dk[j + 1] - dk[j];
  const double half = 0.5 * gap;
  if (fmid > 0.0) {
    st.origin = j;
    st.sgn = 1.0;
    st.tlo = rho * z2[j] / (1.0 + 2.0 * rho / gap);
  } else {
    st.origin = j + 1;
    st.sgn = -1.0;
    st.tlo = z2[j + 1] * half;
  }
  st.thi = half;
  if (st.tlo > st.thi) st.tlo = st.thi;
  st.t = st.tlo;
  st.phase = 0;
  st.iters = 0;
  st.done = 0;
}
// last root in (d_{k-1}, d_{k-1} + rho |z|^2]
KFAC_HD void secular_init_last(SecularState& st, const double* z2, double rho, int k, double z2sum) {
  st.origin = k - 1;
  st.sgn = 1.0;
  st.thi = rho * z2sum;
  st.tlo = rho * z2[k - 1];
  if (st.tlo > st.thi) st.tlo = st.thi;
  st.t = st.tlo;
  st.phase = 0;
  st.iters = 0;
  st.done = 0;
}

// val = 1 + rho * sum, der = rho * dsum evaluated at st.t
KFAC_HD void secular_update(SecularState& st, double val, double der) {
  if (st.phase == 0) {
    // make sure the root is to the right of tlo (g(mu) increases with mu): sgn * g(tlo) <= 0
    if (st.sgn * val <= 0.0 || st.tlo == 0.0 || st.iters >= 64) {
      st.phase = 1;
      st.iters = 0;
      st.t = secular_pick(st.tlo, st.thi);
    } else {
      st.thi = st.tlo;  // root is below the supposed lower bound: it becomes an upper bound
      st.tlo *= 1.0 / 16.0;
      st.t = st.tlo;
      ++st.iters;
    }
    return;
  }
  if (st.sgn * val > 0.0)
    st.thi = st.t;
  else
    st.tlo = st.t;
  ++st.iters;
  if (st.thi - st.tlo <= 4e-16 * st.thi || st.iters >= 200) {
    st.done = 1;
    return;
  }
  double tn = -1.0;
  if (der > 0.0) {
    // Newton on g in t, and a one-pole rational step (exact when the origin pole dominates); prefer the
    // rational step near the pole where plain Newton crawls.
    const double newton = st.t - st.sgn * val / der;
    const double denom = st.sgn * val + der * st.t;   // from g ~ a - b / t  (sign-adjusted)
    const double rational = denom != 0.0 ? der * st.t * st.t / denom : -1.0;
    if (rational > st.tlo && rational < st.thi)
      tn = rational;
    else if (newton > st.tlo && newton < st.thi)
      tn = newton;
  }
  if (!(tn > st.tlo && tn < st.thi)) tn = secular_pick(st.tlo, st.thi);
  if (tn == st.t) {
    st.done = 1;
    return;
  }
  st.t = tn;
}

// Serial driver (host tests, and k == small cases on the device).
KFAC_HD void secular_root(const double* dk, const double* z2, double rho, int k, int j, int& origin, double& mu) {
  SecularState st;
  double sum, dsum;
  if (j < k - 1) {
    const double half = 0.5 * (dk[j + 1] - dk[j]);
    secular_partial(dk, z2, k, dk[j], 1.0, half, 0, 1, sum, dsum);
    secular_init_interior(st, dk, z2, rho, j, 1.0 + rho * sum);
  } else {
    double z2sum = 0.0;
    for (int i = 0; i < k; ++i) z2sum += z2[i];
    secular_init_last(st, z2, rho, k, z2sum);
  }
  while (!st.done) {
    secular_partial(dk, z2, k, dk[st.origin], st.sgn, st.t, 0, 1, sum, dsum);
    secular_update(st, 1.0 + rho * sum, rho * dsum);
  }
  origin = st.origin;
  mu = st.sgn * st.t;
}

}  // namespace kfac
