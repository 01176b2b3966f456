// Lagrangian-to-Eulerian vertical remapping: MapSingle (Fortran map_single / map_scalar / map1_ppm), the PPM
// profile reconstruction (remap_profile, kord <= 10), MapNTracer + fillz, and the pointwise stencils of
// LagrangianToEulerian.__call__.
//
// Reference: pyFV3/stencils/map_single.py:14-81,155-208; remap_profile.py:20-147 (constraints), :150-250
// (set_initial_vals), :253-337 (apply_constraints), :340-563 (set_interpolation_coefficients);
// mapn_tracer.py:14-87; fillz.py:16-172; remapping.py:43-284 (stencils), :522-731 (__call__);
// moist_cv.py:15-199.
//
// One MapSingle = 7 reference launches through 9 full-size temporaries; here it is ONE column kernel: a thread
// owns a column, the profile (q, gam, a4_1..4, extm) lives in thread-local arrays, pe1/pe2/q are read once and q
// written once.  i is the fastest thread index, so every level access is a coalesced row segment.
#include "core.hpp"

namespace pace {

namespace {

constexpr int KMAX = 96;  // thread-local column length (npz + 1 <= KMAX; 79 and 91 level configurations)

template <class T>
struct ColT {
  T q[KMAX + 1], gam[KMAX + 1], a1[KMAX], a2[KMAX], a3[KMAX], a4[KMAX], dp[KMAX];
  bool extm[KMAX], ext5[KMAX], ext6[KMAX];
};
using Col = ColT<double>;

template <class T>
__device__ __forceinline__ void a4_from(T a1, T a2, T a3, T& a4) { a4 = T(3.0) * (T(2.0) * a1 - (a2 + a3)); }

// remap_profile.py:88-111 posdef_constraint_iv1
template <class T>
__device__ __forceinline__ void posdef_iv1(T& a1, T& a2, T& a3, T& a4) {
  T da1 = a3 - a2, da2 = da1 * da1, a6da = a4 * da1;
  if (((a1 - a2) * (a1 - a3)) >= T(0.0)) { a2 = a1; a3 = a1; a4 = T(0.0); }
  else if (a6da < -T(1.0) * da2) { a4 = T(3.0) * (a2 - a1); a3 = a2 - a4; }
  else if (a6da > da2) { a4 = T(3.0) * (a3 - a1); a2 = a3 - a4; }
}
// remap_profile.py:56-85 posdef_constraint_iv0
template <class T>
__device__ __forceinline__ void posdef_iv0(T& a1, T& a2, T& a3, T& a4) {
  if (a1 <= T(0.0)) { a2 = a1; a3 = a1; a4 = T(0.0); }
  else if (fabs(a3 - a2) < -a4 && (a1 + T(0.25) * ((a3 - a2) * (a3 - a2)) / a4 + a4 * (T(1.0) / T(12.0))) < T(0.0)) {
    if (a1 < a3 && a1 < a2) { a3 = a1; a2 = a1; a4 = T(0.0); }
    else if (a3 > a2) { a4 = T(3.0) * (a2 - a1); a3 = a2 - a4; }
    else { a4 = T(3.0) * (a3 - a1); a2 = a3 - a4; }
  }
}
// remap_profile.py:114-135 remap_constraint
template <class T>
__device__ __forceinline__ void remap_constraint(T& a1, T& a2, T& a3, T& a4, bool extm) {
  T da1 = a3 - a2, da2 = da1 * da1, a6da = a4 * da1;
  if (extm) { a2 = a1; a3 = a1; a4 = T(0.0); }
  else if (a6da < -da2) { a4 = T(3.0) * (a2 - a1); a3 = a2 - a4; }
  else if (a6da > da2) { a4 = T(3.0) * (a3 - a1); a2 = a3 - a4; }
}
template <class T>
__device__ __forceinline__ T lim3_min(T a, T b, T c) {  // a if a<b and a<c else b if b<c else c
  return (a < b && a < c) ? a : (b < c ? b : c);
}
template <class T>
__device__ __forceinline__ T lim3_max(T a, T b, T c) {
  return (a > b && a > c) ? a : (b > c ? b : c);
}

// RemapProfile.__call__ (remap_profile.py:644-697) on one column, in two parts.
//
// The reference runs three stencils through full-size a4_1..a4_4 / gam / extm arrays.  Only the first one
// (set_initial_vals) is a recurrence over the column: a tridiagonal solve for the interface values q.  Everything
// after it - the monotonicity constraints on q (apply_constraints) and the parabola coefficients a4_2..a4_4 of a layer
// (set_interpolation_coefficients) - is a function of q and the cell means a4_1 = a1 within two levels of that layer.
// So the column keeps a1, dp and q only, and the coefficients of a source layer are evaluated when the remap reaches
// it (coef_at), with the same expressions in the same order as the reference's statements: results are bit-identical,
// and the thread-local working set (which does not fit any cache at full occupancy and is DRAM traffic) drops from 7
// arrays passed over ~30 times to 4 arrays passed over ~10 times per MapSingle.

// set_initial_vals (remap_profile.py:150-250): interface values q[0..km] from the cell means c.a1 and thicknesses c.dp
template <class T>
__device__ __forceinline__ void profile_solve(ColT<T>& c, int km, int iv, T qs) {
  T* q = c.q;
  T* gam = c.gam;
  const T *a1 = c.a1, *delp = c.dp;
  if (iv != -2) {
    T gr = delp[1] / delp[0];
    T bet = gr * (gr + T(0.5));
    q[0] = ((gr + gr) * (gr + T(1.0)) * a1[0] + a1[1]) / bet;
    gam[0] = (T(1.0) + gr * (gr + T(1.5))) / bet;
#pragma unroll 4
    for (int k = 1; k < km; ++k) {
      T d4 = delp[k - 1] / delp[k];
      bet = T(2.0) + d4 + d4 - gam[k - 1];
      q[k] = fv_div(T(3.0) * (a1[k - 1] + d4 * a1[k]) - q[k - 1], bet);
      gam[k] = d4 / bet;
    }
    {
      T d4 = delp[km - 2] / delp[km - 1];
      T a_bot = T(1.0) + d4 * (d4 + T(1.5));
      q[km] = (T(2.0) * d4 * (d4 + T(1.0)) * a1[km - 1] + a1[km - 2] - a_bot * q[km - 1]) / (d4 * (d4 + T(0.5)) - a_bot * gam[km - 1]);
    }
#pragma unroll 4
    for (int k = km - 1; k >= 0; --k) q[k] = q[k] - gam[k] * q[k + 1];
  } else {
    q[0] = T(1.5) * a1[0];
    gam[1] = T(0.5);
    {
      T gr = delp[0] / delp[1];
      T bet = T(2.0) + gr + gr - gam[1];
      q[1] = (T(3.0) * (a1[0] + a1[1]) - q[0]) / bet;
    }
    for (int k = 2; k < km; ++k) {
      T old_gr = delp[k - 2] / delp[k - 1];
      T old_bet = T(2.0) + old_gr + old_gr - gam[k - 1];
      gam[k] = old_gr / old_bet;
    }
    for (int k = 2; k < km - 1; ++k) {
      T gr = delp[k - 1] / delp[k];
      T bet = T(2.0) + gr + gr - gam[k];
      q[k] = (T(3.0) * (a1[k - 1] + a1[k]) - q[k - 1]) / bet;
    }
    {
      int k = km - 1;
      T gr = delp[k - 1] / delp[k];
      q[k] = (T(3.0) * (a1[k - 1] + a1[k]) - gr * qs - q[k - 1]) / (T(2.0) + gr + gr - gam[k]);
    }
    q[km] = qs;
    for (int k = km - 2; k >= 0; --k) q[k] = q[k] - gam[k + 1] * q[k + 1];
  }
}

// apply_constraints (remap_profile.py:253-337): gam[k] = a1[k] - a1[k-1] (1 <= k <= km-1), and the interface value k
// limited by its two cell means / clipped according to the signs of the neighbouring differences
template <class T>
__device__ __forceinline__ T gam_at(const ColT<T>& c, int k) { return c.a1[k] - c.a1[k - 1]; }

template <class T>
__device__ __forceinline__ T qc_at(const ColT<T>& c, int km, int iv, int k) {
  T q = c.q[k];
  if (k == 0 || k == km) return q;
  const T tmp = c.a1[k - 1] > c.a1[k] ? c.a1[k - 1] : c.a1[k];   // tmpmax
  const T tmp2 = c.a1[k - 1] < c.a1[k] ? c.a1[k - 1] : c.a1[k];  // tmpmin
  if (k == 1 || k == km - 1) {
    if (q >= tmp) q = tmp;
    if (q <= tmp2) q = tmp2;
    return q;
  }
  const T gm = gam_at(c, k - 1), gp = gam_at(c, k + 1);
  if ((gm * gp) > 0) {
    if (q >= tmp) q = tmp;
    if (q <= tmp2) q = tmp2;
  } else if (gm > 0) {
    if (q <= tmp2) q = tmp2;
  } else {
    if (q >= tmp) q = tmp;
    if (iv == 0 && q < T(0.0)) q = T(0.0);
  }
  return q;
}
// extm of layer k, 1 <= k <= km-2 (the values of layers 0 and km-1 are never read)
template <class T>
__device__ __forceinline__ bool extm_at(const ColT<T>& c, int k) { return gam_at(c, k) * gam_at(c, k + 1) < T(0.0); }

// a4_2, a4_3, a4_4 of source layer k as set_interpolation_coefficients leaves them (remap_profile.py:340-563)
template <class T>
__device__ __forceinline__ void coef_at(const ColT<T>& c, int km, int iv, int kord, T qmin, int k, T& a2, T& a3, T& a4) {
  T a1 = c.a1[k];  // a4_1 is an argument of the constraint functions but comes back unchanged
  a2 = qc_at(c, km, iv, k);
  a3 = qc_at(c, km, iv, k + 1);
  if (k == 0) {
    if (iv == 0 && a2 < T(0.0)) a2 = T(0.0);
    if (iv == -1 && a2 * a1 <= T(0.0)) a2 = T(0.0);
    if (iv == 2) { a2 = a1; a3 = a1; a4 = T(0.0); }
    else { a4_from(a1, a2, a3, a4); posdef_iv1(a1, a2, a3, a4); }
    return;
  }
  if (k == 1 || k == km - 2) {
    a4_from(a1, a2, a3, a4);
    remap_constraint(a1, a2, a3, a4, extm_at(c, k));
    return;
  }
  if (k == km - 1) {
    if (iv == 0 && a3 < T(0.0)) a3 = T(0.0);
    if (iv == -1 && a3 * a1 <= T(0.0)) a3 = T(0.0);
    a4_from(a1, a2, a3, a4);
    posdef_iv1(a1, a2, a3, a4);
    return;
  }
  // 2 <= k <= km-3
  if (kord < 9) {
    T pmp_1 = a1 - gam_at(c, k + 1);
    T lac_1 = pmp_1 + T(1.5) * gam_at(c, k + 2);
    T tmp_min = lim3_min(a1, pmp_1, lac_1);
    T tmp_max0 = a2 > tmp_min ? a2 : tmp_min;
    T tmp_max = lim3_max(a1, pmp_1, lac_1);
    a2 = tmp_max0 < tmp_max ? tmp_max0 : tmp_max;
    T pmp_2 = a1 + T(2.0) * gam_at(c, k + 1);
    T lac_2 = pmp_2 - T(1.5) * gam_at(c, k - 1);
    tmp_min = lim3_min(a1, pmp_2, lac_2);
    tmp_max0 = a3 > tmp_min ? a3 : tmp_min;
    tmp_max = lim3_max(a1, pmp_2, lac_2);
    a3 = tmp_max0 < tmp_max ? tmp_max0 : tmp_max;
    a4_from(a1, a2, a3, a4);
  } else if (kord == 9) {
    T pmp_1 = a1 - T(2.0) * gam_at(c, k + 1);
    T lac_1 = pmp_1 + T(1.5) * gam_at(c, k + 2);
    T pmp_2 = a1 + T(2.0) * gam_at(c, k);
    T lac_2 = pmp_2 - T(1.5) * gam_at(c, k - 1);
    const bool ex = extm_at(c, k);
    if ((ex && extm_at(c, k - 1)) || (ex && extm_at(c, k + 1)) || (ex && (qmin > T(0.0) && a1 < qmin))) {
      a2 = a1; a3 = a1; a4 = T(0.0);
    } else {
      a4 = T(6.0) * a1 - T(3.0) * (a2 + a3);
      if (fabs(a4) > fabs(a2 - a3)) {
        T tmp_min = lim3_min(a1, pmp_1, lac_1);
        T tmp_max0 = a2 > tmp_min ? a2 : tmp_min;
        T tmp_max = lim3_max(a1, pmp_1, lac_1);
        a2 = tmp_max0 < tmp_max ? tmp_max0 : tmp_max;
        tmp_min = lim3_min(a1, pmp_2, lac_2);
        tmp_max0 = a3 > tmp_min ? a3 : tmp_min;
        tmp_max = lim3_max(a1, pmp_2, lac_2);
        a3 = tmp_max0 < tmp_max ? tmp_max0 : tmp_max;
        a4 = T(6.0) * a1 - T(3.0) * (a2 + a3);
      }
    }
  } else {  // kord == 10
    // ext5 / ext6 of layers k-1, k, k+1 from their constrained edge values (apply_constraints, kord > 9 part)
    bool e5[3], e6[3];
    T qe[4];
#pragma unroll
    for (int n = 0; n < 4; ++n) qe[n] = (n == 1) ? a2 : (n == 2) ? a3 : qc_at(c, km, iv, k - 1 + n);
#pragma unroll
    for (int n = 0; n < 3; ++n) {
      T x0 = T(2.0) * c.a1[k - 1 + n] - (qe[n] + qe[n + 1]);
      T x1 = fabs(qe[n] - qe[n + 1]);
      T a40 = T(3.0) * x0;
      e5[n] = fabs(x0) > x1;
      e6[n] = fabs(a40) > x1;
    }
    T pmp_1 = a1 - T(2.0) * gam_at(c, k + 1);
    T lac_1 = pmp_1 + T(1.5) * gam_at(c, k + 2);
    T pmp_2 = a1 + T(2.0) * gam_at(c, k);
    T lac_2 = pmp_2 - T(1.5) * gam_at(c, k - 1);
    T tmp_min2 = lim3_min(a1, pmp_1, lac_1);
    T tmp_max2 = lim3_max(a1, pmp_1, lac_1);
    T tmp2 = a2 > tmp_min2 ? a2 : tmp_min2;
    T tmp_min3 = a1 < pmp_2 ? a1 : pmp_2;
    tmp_min3 = lac_2 < tmp_min3 ? lac_2 : tmp_min3;
    T tmp_max3 = a1 > pmp_2 ? a1 : pmp_2;
    tmp_max3 = lac_2 > tmp_max3 ? lac_2 : tmp_max3;
    T tmp3 = a3 > tmp_min3 ? a3 : tmp_min3;
    bool n5 = e5[0] || e5[2], n6 = e6[0] || e6[2];
    if (e5[1]) {
      if (n5) { a2 = a1; a3 = a1; }
      else if (n6) { a2 = tmp2 < tmp_max2 ? tmp2 : tmp_max2; a3 = tmp3 < tmp_max3 ? tmp3 : tmp_max3; }
    } else if (e6[1]) {
      if (n5) { a2 = tmp2 < tmp_max2 ? tmp2 : tmp_max2; a3 = tmp3 < tmp_max3 ? tmp3 : tmp_max3; }
    }
    a4_from(a1, a2, a3, a4);
  }
  if (iv == 0) posdef_iv0(a1, a2, a3, a4);
}

// MapSingle.__call__ on one column: q (in/out, km levels) from source interfaces pe1 to target interfaces pe2
// keep_dp: c.dp already holds the layer thicknesses of pe1 (previous call with the same pe1).  Returns whether any
// remapped value is negative (the only case in which fillz changes a column).
template <class T>
__device__ bool map_single_column(const Pt& p, ColT<T>& c, int km, T* q, const T* pe1, const T* pe2, int iv,
                                  int kord, T qs, T qmin, bool keep_dp = false, bool zero_paths = true) {
  bool any_negative = false;
  bool all_zero = zero_paths;
  for (int k = 0; k < km; ++k) {
    const T v = A(q, 0, 0, k);
    c.a1[k] = v;
    all_zero = all_zero && v == T(0.0);
    if (!keep_dp) c.dp[k] = A(pe1, 0, 0, k + 1) - A(pe1, 0, 0, k);
  }
  // A column of zeros of a positive-definite tracer (a condensate species outside clouds) remaps to +0 on every
  // level: the tridiagonal solve, the constraints and the integrals only combine zeros.  c.dp is filled above
  // because the following fields of the column reuse it.
  if (iv == 0 && all_zero) {
    for (int k = 0; k < km; ++k) A(q, 0, 0, k) = T(0.0);
    return false;
  }
  profile_solve(c, km, iv, qs);
  // lagrangian_contributions (map_single.py:21-81); L = k + lev is the absolute source layer.  Variable-k reads
  // are clipped to the allocated levels [0, km] like the numpy backend does.
  auto P1 = [&](int L) { return A(pe1, 0, 0, L > km ? km : L); };
  auto cl = [&](int L) { return L > km - 1 ? km - 1 : L; };  // layer arrays: km entries meaningful
  int L = 0;
  int Lc = 0;
  T a2, a3, a4;  // coefficients of source layer Lc; re-evaluated when the walk moves to another layer
  coef_at(c, km, iv, kord, qmin, Lc, a2, a3, a4);
  // the values the walk needs of its current source layer live in registers: p1L = P1(L), p1L1 = P1(L + 1),
  // dpL = c.dp[Lc]; pe2 is read once per level
  T p1L = P1(0), p1L1 = P1(1), dpL = c.dp[0];
  T pe2k = A(pe2, 0, 0, 0);
  for (int k = 0; k < km; ++k) {
    const T pe2k1 = A(pe2, 0, 0, k + 1);
    T pl = fv_div(pe2k - p1L, dpL);
    T qv;
    if (pe2k1 <= p1L1) {
      T pr = (pe2k1 - p1L) / dpL;
      qv = a2 + T(0.5) * (a4 + a3 - a2) * (pr + pl) - a4 * T(1.0) / T(3.0) * (pr * (pr + pl) + pl * pl);
    } else {
      T qsum = (p1L1 - pe2k) *
                    (a2 + T(0.5) * (a4 + a3 - a2) * (T(1.0) + pl) - a4 * T(1.0) / T(3.0) * (T(1.0) + pl * (T(1.0) + pl)));
      L = L + 1;
      T pa = p1L1, pb = P1(L + 1);  // P1(L), P1(L + 1) of the layer the walk is at
      while (L < km && pb < pe2k1) {
        qsum += c.dp[cl(L)] * c.a1[cl(L)];
        L = L + 1;
        pa = pb;
        pb = P1(L + 1);
      }
      p1L = pa;
      p1L1 = pb;
      if (cl(L) != Lc) {
        Lc = cl(L);
        coef_at(c, km, iv, kord, qmin, Lc, a2, a3, a4);
      }
      dpL = c.dp[Lc];
      T dp = pe2k1 - p1L;
      T esl = fv_div(dp, dpL);
      qsum += dp * (a2 + T(0.5) * esl * (a3 - a2 + a4 * (T(1.0) - (T(2.0) / T(3.0)) * esl)));
      qv = fv_div(qsum, pe2k1 - pe2k);
    }
    A(q, 0, 0, k) = qv;
    any_negative = any_negative || qv < T(0.0);
    pe2k = pe2k1;
  }
  return any_negative;
}

// fix_tracer (fillz.py:16-107) on one column
__device__ void fillz_column(const Pt& p, Col& c, int km, double* q, const double* dp) {
  double* lower = c.a2;   // reuse thread-local arrays
  double* upper = c.a3;
  double* dm = c.a4;
  double* dmp = c.gam;
  double* qq = c.a1;
  double* d = c.dp;
  for (int k = 0; k < km; ++k) { qq[k] = A(q, 0, 0, k); d[k] = A(dp, 0, 0, k); lower[k] = 0.0; upper[k] = 0.0; dm[k] = 0.0; dmp[k] = 0.0; }
  int zfix = 0;
  double sum0 = 0.0, sum1 = 0.0;
  if (qq[0] < 0.0) qq[1] = qq[1] + qq[0] * d[0] / d[1];
  if (qq[0] < 0) qq[0] = 0;
  dm[0] = qq[0] * d[0];
  for (int k = 1; k < km - 1; ++k) {
    if (lower[k - 1] != 0.0) qq[k] = qq[k] - (lower[k - 1] / d[k]);
    if (qq[k] < 0.0) {
      zfix += 1;
      if (qq[k - 1] > 0.0) {
        double dq = qq[k - 1] * d[k - 1] < -(qq[k] * d[k]) ? qq[k - 1] * d[k - 1] : -(qq[k] * d[k]);
        qq[k] = qq[k] + dq / d[k];
        upper[k] = dq;
      }
      if (qq[k] < 0.0 && qq[k + 1] > 0.0) {
        double dq = qq[k + 1] * d[k + 1] < -(qq[k] * d[k]) ? qq[k + 1] * d[k + 1] : -(qq[k] * d[k]);
        qq[k] = qq[k] + dq / d[k];
        lower[k] = dq;
      }
    }
  }
  for (int k = 0; k < km - 1; ++k) {
    if (upper[k + 1] != 0.0) qq[k] = qq[k] - upper[k + 1] / d[k];
    dm[k] = qq[k] * d[k];
    dmp[k] = dm[k] > 0.0 ? dm[k] : 0.0;
  }
  {
    int k = km - 1;
    if (lower[k - 1] != 0.0) qq[k] = qq[k] - (lower[k - 1] / d[k]);
    double qup = qq[k - 1] * d[k - 1];
    double qly = -qq[k] * d[k];
    double dup = qup < qly ? qup : qly;
    if (qq[k] < 0.0 && qq[k - 1] > 0.0) {
      zfix += 1;
      qq[k] = qq[k] + (dup / d[k]);
      upper[k] = dup;
    }
    dm[k] = qq[k] * d[k];
    dmp[k] = dm[k] > 0.0 ? dm[k] : 0.0;
  }
  {
    int k = km - 2;
    if (upper[k + 1] != 0.0) {
      qq[k] = qq[k] - (upper[k + 1] / d[k]);
      dm[k] = qq[k] * d[k];
      dmp[k] = dm[k] > 0.0 ? dm[k] : 0.0;
    }
  }
  for (int k = 1; k < km; ++k) { sum0 += dm[k]; sum1 += dmp[k]; }
  double fac = sum0 > 0.0 ? sum0 / sum1 : 0.0;
  if (zfix > 0 && fac > 0.0)
    for (int k = 1; k < km; ++k) qq[k] = fac * dm[k] / d[k] > 0.0 ? fac * dm[k] / d[k] : 0.0;
  for (int k = 0; k < km; ++k) A(q, 0, 0, k) = qq[k];
}

struct TrPtrs { double* p[9]; };

#ifndef PACE_HOST_EMU
#define COL_DECL Col col
#define COLT_DECL ColT<T> col
#else
#define COL_DECL static thread_local Col col
#define COLT_DECL static thread_local ColT<T> col
#endif

}  // namespace

const char* tracer_name(int n);
const int* tracer_zero_fields(Core& c, int nq);   // tracer.cu

// Resident CTAs per SM the scalar column kernel is compiled for (8, 10 or 12; see launch_ij_occ).
static int remap_occupancy() {
  static const int n = [] { const char* v = getenv("PACE_B200_REMAP_OCC"); return v ? atoi(v) : 12; }();
  return n;
}

// MapSingle.__call__ exposed on registered fields (also used by the test-suite on synthetic columns).
// xi / yi: the field's horizontal staggering (dims [X(_INTERFACE), Y(_INTERFACE), Z]).
// T = double is the reference's default precision, T = float its PACE_FLOAT_PRECISION=32 mode (Float = np.float32,
// ndsl/dsl/typing.py:25-45): fields, literals and every intermediate are T, as in the GT4Py numpy code.
template <class T>
int map_single_impl(Core& c, T* q, const T* pe1, const T* pe2, const T* qs, int iv, int kord, bool xi, bool yi, T qmin) {
  const Geo g = c.g;
  const int km = g.nz;
  if (km + 1 > KMAX) { c.err = "map_single: npz + 1 exceeds the thread-local column length (96)"; return 1; }
  if (kord > 10) { c.err = "Remap Profile: kord > 10 not implemented. kord <= 10 available."; return 1; }
  launch_ij(c.stream, g, box_dims(g, xi, yi, false), [=] __device__(const Pt& p) {
    COLT_DECL;
    map_single_column<T>(p, col, km, q, pe1, pe2, iv, kord, qs ? A20(qs) : T(0.0), qmin);
  }, sizeof(T) == 4 ? "map_single_f32" : "map_single");
  return 0;
}
int map_single(Core& c, double* q, const double* pe1, const double* pe2, const double* qs, int iv, int kord, bool xi,
               bool yi, double qmin) {
  return map_single_impl<double>(c, q, pe1, pe2, qs, iv, kord, xi, yi, qmin);
}
int map_single_f32(Core& c, float* q, const float* pe1, const float* pe2, const float* qs, int iv, int kord, bool xi,
                   bool yi, float qmin) {
  return map_single_impl<float>(c, q, pe1, pe2, qs, iv, kord, xi, yi, qmin);
}

// LagrangianToEulerian.__call__ up to (not including) the saturation adjustment (remapping.py:590-668)
int remap_main(Core& c, int nq) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const Config& cf = c.cfg;
  if (cf.kord_tm >= 0) { c.err = "map ppm, untested mode where kord_tm >= 0"; return 1; }
  if (cf.hydrostatic) { c.err = "Hydrostatic is not implemented"; return 1; }
  const int km = g.nz;
  if (km + 1 > KMAX) { c.err = "remapping: npz + 1 exceeds the thread-local column length (96)"; return 1; }
  double *pt = c.F("pt"), *delp = c.F("delp"), *delz = c.F("delz"), *peln = c.F("peln"), *u = c.F("u"), *v = c.F("v"),
         *w = c.F("w"), *cappa = c.F("ad.cappa"), *q_con = c.F("q_con"), *pkz = c.F("pkz"), *pk = c.F("pk"),
         *pe = c.F("pe"), *ps = c.F("ps");
  const double *wsd = c.F("ad.wsd"), *ak = c.F("grid.ak"), *bk = c.F("grid.bk");
  const double *qvapor = c.F("qvapor"), *qliquid = c.F("qliquid"), *qrain = c.F("qrain"), *qsnow = c.F("qsnow"),
               *qice = c.F("qice"), *qgraupel = c.F("qgraupel");
  double *pe1 = c.FS("rm.pe1"), *pe2 = c.FS("rm.pe2"), *pe3 = c.FS("rm.pe3"), *dp2 = c.FS("rm.dp2"), *pn2 = c.FS("rm.pn2"),
         *pe0 = c.FS("rm.pe0"), *gz = c.FS("rm.gz"), *cvm = c.FS("rm.cvm");
  if (!c.err.empty()) return 1;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;
  const int kord_tm = cf.kord_tm < 0 ? -cf.kord_tm : cf.kord_tm;
  const int kord_tr = cf.kord_tr < 0 ? -cf.kord_tr : cf.kord_tr;
  const int kord_wz = cf.kord_wz, kord_mt = cf.kord_mt;
  const double r_vir = ZVIR, akap = KAPPA, ptop = cf.ptop, t_min = 184.0;

  // init_pe (domain nx, ny+1, km+1) + moist_cv_pt_pressure + pn2_pk_delp (compute domain)
  launch_ij(st, g, Box{is, js, 0, g.nx, g.ny + 1, 1}, [=] __device__(const Pt& p) {
    for (int k = 0; k <= km; ++k) A(pe1, 0, 0, k) = A(pe, 0, 0, k);
    A(pe2, 0, 0, 0) = ptop;
    A(pe2, 0, 0, km) = A(pe, 0, 0, km);
    if (p.j > je) return;
    for (int k = 0; k < km; ++k) {
      // moist_pt_func (moist_cv.py:62-86)
      double ql = A(qliquid, 0, 0, k) + A(qrain, 0, 0, k);
      double qs_ = A(qice, 0, 0, k) + A(qsnow, 0, 0, k) + A(qgraupel, 0, 0, k);
      double gzv = ql + qs_;
      double qv = A(qvapor, 0, 0, k);
      double cv = (1.0 - (qv + gzv)) * CV_AIR + qv * CV_VAP + ql * C_LIQ + qs_ * C_ICE;
      A(q_con, 0, 0, k) = gzv;
      double cap = RDGAS / (RDGAS + cv / (1.0 + r_vir * qv));
      A(cappa, 0, 0, k) = cap;
      double ptv = A(pt, 0, 0, k), dpv = A(delp, 0, 0, k), dzv = A(delz, 0, 0, k);
      A(pt, 0, 0, k) = ptv * fv_exp(cap / (1.0 - cap) * fv_log(RDG * dpv / dzv * ptv));
      A(delz, 0, 0, k) = -dzv / dpv;
    }
    double psv = A(pe, 0, 0, km);
    A20(ps) = psv;
    A(pn2, 0, 0, 0) = A(peln, 0, 0, 0);
    for (int k = 1; k < km; ++k) A(pe2, 0, 0, k) = ak[k] + bk[k] * psv;
    A(pn2, 0, 0, km) = A(peln, 0, 0, km);
    for (int k = km - 1; k >= 0; --k) {
      double d = A(pe2, 0, 0, k + 1) - A(pe2, 0, 0, k);
      A(dp2, 0, 0, k) = d;
      A(delp, 0, 0, k) = d;
    }
    // pn2_pk_delp (km levels)
    for (int k = 0; k < km; ++k) {
      double l = fv_log(A(pe2, 0, 0, k));
      A(pn2, 0, 0, k) = l;
      A(pk, 0, 0, k) = fv_exp(akap * l);
    }
  });
  // map_single_pt (iv 1, log-p coordinates), MapNTracer (iv 0) + fillz, map_single_w (iv -2, qs = wsd),
  // map_single_delz (iv 1): all on the compute domain -> one column kernel
  double* tr[9];
  for (int n = 0; n < nq; ++n) tr[n] = c.F(tracer_name(n));
  if (!c.err.empty()) return 1;
  TrPtrs trp;
  for (int n = 0; n < 9; ++n) trp.p[n] = n < nq ? tr[n] : nullptr;
  const bool fill = cf.fill != 0;
  const bool zero_paths = zero_paths_enabled();
  // tracers that are zero on a whole subdomain (known from TracerAdvection's level flags inside step()): their columns
  // would be read, found all zero and rewritten as zeros - they are left alone
  const int* zero_field = tracer_zero_fields(c, nq);
  if (!c.err.empty()) return 1;
  auto scalars_column = [=] __device__(const Pt& p) {
    COL_DECL;
    map_single_column(p, col, km, pt, peln, pn2, 1, kord_tm, 0.0, t_min);
    unsigned negative = 0;  // tracers whose remapped column holds a negative value
    bool have_dp = false;   // col.dp holds the thicknesses of pe1 (the first tracer that is remapped fills it)
    for (int n = 0; n < nq; ++n) {
      if (zero_field != nullptr && zero_field[p.t * 9 + n]) continue;
      // mapn_tracer.py:40-41: every tracer uses kord_tr, entry 5 is forced to 9
      int kord = n == 5 ? 9 : kord_tr;
      if (map_single_column(p, col, km, trp.p[n], pe1, pe2, 0, kord, 0.0, 0.0, have_dp, zero_paths)) negative |= 1u << n;
      have_dp = true;
    }
    // fillz leaves a column without negative values exactly as it is (no borrowing, zfix == 0): only the others
    // go through it.  It uses col.dp as scratch, so the thicknesses are rebuilt afterwards if it ran.
    if (fill)
      for (int n = 0; n < nq; ++n)
        if (negative & (1u << n)) fillz_column(p, col, km, trp.p[n], dp2);
    map_single_column(p, col, km, w, pe1, pe2, -2, kord_wz, A20(wsd), 0.0, have_dp && !(fill && negative));
    map_single_column(p, col, km, delz, pe1, pe2, 1, kord_wz, 0.0, 0.0, true);
    // undo_delz_adjust_and_copy_peln
    for (int k = 0; k < km; ++k) A(delz, 0, 0, k) = -A(delz, 0, 0, k) * A(delp, 0, 0, k);
    for (int k = 0; k <= km; ++k) {
      A(pe0, 0, 0, k) = A(peln, 0, 0, k);
      A(peln, 0, 0, k) = A(pn2, 0, 0, k);
    }
    // moist_pkz (moist_cv.py:170-198)
    for (int k = 0; k < km; ++k) {
      double ql = A(qliquid, 0, 0, k) + A(qrain, 0, 0, k);
      double qs_ = A(qice, 0, 0, k) + A(qsnow, 0, 0, k) + A(qgraupel, 0, 0, k);
      double gzv = ql + qs_;
      double qv = A(qvapor, 0, 0, k);
      double cv = (1.0 - (qv + gzv)) * CV_AIR + qv * CV_VAP + ql * C_LIQ + qs_ * C_ICE;
      A(cvm, 0, 0, k) = cv;
      A(gz, 0, 0, k) = gzv;
      A(q_con, 0, 0, k) = gzv;
      double cap = RDGAS / (RDGAS + cv / (1.0 + r_vir * qv));
      A(cappa, 0, 0, k) = cap;
      A(pkz, 0, 0, k) = fv_exp(cap * fv_log(RDG * A(delp, 0, 0, k) / A(delz, 0, 0, k) * A(pt, 0, 0, k)));
    }
  };
  switch (remap_occupancy()) {
    case 8: launch_ij_occ<8>(st, g, box_compute(g), scalars_column, "remap_scalars_column"); break;
    case 10: launch_ij_occ<10>(st, g, box_compute(g), scalars_column, "remap_scalars_column"); break;
    default: launch_ij_occ<12>(st, g, box_compute(g), scalars_column, "remap_scalars_column"); break;
  }
  // pressures_mapu + map_single_u on [is..ie] x [js..je+1]
  launch_ij_occ<12>(st, g, Box{is, js, 0, g.nx, g.ny + 1, 1}, [=] __device__(const Pt& p) {
    COL_DECL;
    double pe_bottom_m = A(pe, 0, -1, km), pe1_bottom = A(pe, 0, 0, km);
    A(pe0, 0, 0, 0) = A(pe, 0, 0, 0);
    for (int k = 1; k <= km; ++k) A(pe0, 0, 0, k) = 0.5 * (A(pe, 0, -1, k) + A(pe1, 0, 0, k));
    for (int k = 0; k <= km; ++k) {
      double bkh = 0.5 * bk[k];
      A(pe3, 0, 0, k) = ak[k] + bkh * (pe_bottom_m + pe1_bottom);
    }
    map_single_column(p, col, km, u, pe0, pe3, -1, kord_mt, 0.0, 0.0);
  });
  // pressures_mapv + map_single_v on [is..ie+1] x [js..je]
  launch_ij_occ<12>(st, g, Box{is, js, 0, g.nx + 1, g.ny, 1}, [=] __device__(const Pt& p) {
    COL_DECL;
    double pe_bottom_m = A(pe, -1, 0, km), pe_bottom = A(pe, 0, 0, km);
    A(pe3, 0, 0, 0) = ak[0];
    A(pe0, 0, 0, 0) = A(pe, 0, 0, 0);
    for (int k = 1; k <= km; ++k) {
      double bkh = 0.5 * bk[k];
      A(pe0, 0, 0, k) = 0.5 * (A(pe, -1, 0, k) + A(pe, 0, 0, k));
      A(pe3, 0, 0, k) = ak[k] + bkh * (pe_bottom_m + pe_bottom);
    }
    map_single_column(p, col, km, v, pe0, pe3, -1, kord_mt, 0.0, 0.0);
  });
  // update_ua(pe2, pe3) on [is..ie] x [js..je+1], then copy_from_below(pe3 -> pe) on the compute domain
  launch_ij(st, g, Box{is, js, 0, g.nx, g.ny + 1, 1}, [=] __device__(const Pt& p) {
    for (int k = 0; k < km; ++k) {
      double val = A(pe2, 0, 0, k + 1);
      if (k < km - 1 && p.j == je + 1) val = A(pe2, 0, -1, k + 1);
      A(pe3, 0, 0, k) = val;
    }
    if (p.j <= je)
      for (int k = 1; k < km; ++k) A(pe, 0, 0, k) = A(pe3, 0, 0, k - 1);
  });
  return 0;
}

// the tail of LagrangianToEulerian.__call__ after the saturation adjustment (remapping.py:715-731)
int remap_finish(Core& c, int last_step) {
  const Geo g = c.g;
  const int km = g.nz;
  double *pt = c.F("pt"), *gz = c.FS("rm.gz");
  const double *pkz = c.F("pkz"), *qvapor = c.F("qvapor"), *qliquid = c.F("qliquid"), *qrain = c.F("qrain"),
               *qsnow = c.F("qsnow"), *qice = c.F("qice"), *qgraupel = c.F("qgraupel");
  if (!c.err.empty()) return 1;
  const double dtmp = 0.0, r_vir = ZVIR;
  if (last_step) {
    if (c.cfg.consv_te > 0.001) { c.err = "We do not support consv_te > 0.001 because that would trigger an allReduce"; return 1; }
    // moist_pt_last_step (moist_cv.py:101-117); the reference's domain has km+1 levels, the extra one is not meaningful
    launch_ijk(c.stream, g, box_compute(g), [=] __device__(const Pt& p) {
      double gzv = A0(qliquid) + A0(qrain) + A0(qice) + A0(qsnow) + A0(qgraupel);
      A0(gz) = gzv;
      A0(pt) = (A0(pt) + dtmp * A0(pkz)) / ((1.0 + r_vir * A0(qvapor)) * (1.0 - gzv));
    });
  } else {
    launch_ijk(c.stream, g, box_compute(g), [=] __device__(const Pt& p) { A0(pt) = A0(pt) / A0(pkz); });
  }
  (void)km;
  return 0;
}

}  // namespace pace
