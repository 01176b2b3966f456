// Piecewise-parabolic 1-D flux operators (Fortran xppm/yppm), direction-generic.
//
// Reference: pyFV3/stencils/xppm.py:18-272 and its mirror yppm.py, ppm.py:1-59.
// `Q(d)` returns the transported scalar at cell offset d along the sweep direction from the cell whose
// lower interface is being computed; `D(d)` the A-grid spacing (dxa / dya) likewise.  `pos` is the
// absolute index of that interface/cell along the direction, lo/hi the first/last compute cell of the
// TILE edge (is/ie or js/je) and elo/ehi whether this subdomain touches that edge — the reference's
// i_start/i_end region externals (NDSL dsl/stencil.py:724-766).
#pragma once
#include "common.cuh"

namespace pace {
namespace ppmc {
constexpr double c1 = -2.0 / 14.0, c2 = 11.0 / 14.0, c3 = 5.0 / 14.0;
constexpr double p1 = 7.0 / 12.0, p2 = -1.0 / 12.0;
constexpr double s11 = 11.0 / 14.0, s14 = 4.0 / 7.0, s15 = 3.0 / 14.0;
}  // namespace ppmc

__device__ __forceinline__ double fv_sign(double a, double b) {  // basic_operations.py:32-39: |a| if b > 0 else -|a|
#ifdef PACE_HOST_EMU
  double r = fabs(a);
  return b > 0 ? r : -r;
#else
  // same value, built from the sign bit: one compare + two integer ops instead of two fp64 moves and two selects
  const unsigned hi = ((unsigned)__double2hiint(a) & 0x7fffffffu) | (b > 0 ? 0u : 0x80000000u);
  return __hiloint2double((int)hi, __double2loint(a));
#endif
}

// compute_al (xppm.py:147-185), al at interface offset d
template <class QF, class DF>
__device__ __forceinline__ double ppm_al(QF Q, DF D, int pos, int lo, int hi, bool elo, bool ehi, int d) {
  using namespace ppmc;
  const int x = pos + d;
  if ((elo && x == lo + 1) || (ehi && x == hi + 2)) return c3 * Q(d - 1) + c2 * Q(d) + c1 * Q(d + 1);
  if ((elo && x == lo) || (ehi && x == hi + 1))
    return 0.5 * (((2.0 * D(d - 1) + D(d - 2)) * Q(d - 1) - D(d - 1) * Q(d - 2)) / (D(d - 2) + D(d - 1)) +
                  ((2.0 * D(d) + D(d + 1)) * Q(d) - D(d) * Q(d + 1)) / (D(d) + D(d + 1)));
  if ((elo && x == lo - 1) || (ehi && x == hi)) return c1 * Q(d - 2) + c2 * Q(d - 1) + c3 * Q(d);
  return p1 * (Q(d - 1) + Q(d)) + p2 * (Q(d - 2) + Q(d + 1));
}

// compute_x_flux / compute_y_flux for mord < 8 (xppm.py:64-82, 257-272)
template <class QF, class DF>
__device__ __forceinline__ double ppm_flux_lt8(QF Q, DF D, double c, int pos, int lo, int hi, bool elo, bool ehi,
                                               int mord) {
  double alm = ppm_al(Q, D, pos, lo, hi, elo, ehi, -1);
  double al0 = ppm_al(Q, D, pos, lo, hi, elo, ehi, 0);
  double alp = ppm_al(Q, D, pos, lo, hi, elo, ehi, 1);
  double q0 = Q(0), qm = Q(-1);
  double bl = al0 - q0, br = alp - q0, b0 = bl + br;
  double blm = alm - qm, brm = al0 - qm, b0m = blm + brm;
  bool s0, sm;
  if (mord == 5) {
    s0 = bl * br < 0;
    sm = blm * brm < 0;
  } else {
    s0 = (3.0 * fabs(b0)) < fabs(bl - br);
    sm = (3.0 * fabs(b0m)) < fabs(blm - brm);
  }
  double mask = (sm || s0) ? 1.0 : 0.0;
  double fx1 = c > 0.0 ? (1.0 - c) * (brm - c * b0m) : (1.0 + c) * (bl + c * b0);
  return c > 0.0 ? qm + fx1 * mask : q0 + fx1 * mask;
}

// ---- mord == 8 (xppm.py:84-102, 188-254; ppm.py:21-33) ----------------------------------------------
template <class QF>
__device__ __forceinline__ double ppm_dm8(QF Q, int d) {  // dm_iord8plus at cell offset d
  double q = Q(d), qm = Q(d - 1), qp = Q(d + 1);
  double xt = 0.25 * (qp - qm);
  double dqr = fv_max(fv_max(q, qm), qp) - q;
  double dql = q - fv_min(fv_min(q, qm), qp);
  return fv_sign(fv_min(fv_min(fabs(xt), dqr), dql), xt);
}

template <class QF, class DF>
__device__ __forceinline__ void ppm_blbr8(QF Q, DF D, int pos, int lo, int hi, bool elo, bool ehi, int d, double& bl,
                                          double& br) {
  using namespace ppmc;
  const int x = pos + d;  // cell index
  const double q = Q(d);
  auto dm = [&](int dd) { return ppm_dm8(Q, dd); };
  auto al = [&](int dd) { return 0.5 * (Q(dd - 1) + Q(dd)) + 1.0 / 3.0 * (dm(dd - 1) - dm(dd)); };
  auto edge0 = [&](int dd) {  // xt_dxa_edge_0 with xt_minmax=True, centred on cell offset dd
    double xt = 0.5 * (((2.0 * D(dd) + D(dd - 1)) * Q(dd) - D(dd) * Q(dd - 1)) / (D(dd - 1) + D(dd)) +
                       ((2.0 * D(dd + 1) + D(dd + 2)) * Q(dd + 1) - D(dd + 1) * Q(dd + 2)) / (D(dd + 1) + D(dd + 2)));
    double minq = fv_min(fv_min(fv_min(Q(dd - 1), Q(dd)), Q(dd + 1)), Q(dd + 2));
    double maxq = fv_max(fv_max(fv_max(Q(dd - 1), Q(dd)), Q(dd + 1)), Q(dd + 2));
    return fv_min(fv_max(xt, minq), maxq);
  };
  auto edge1 = [&](int dd) {  // xt_dxa_edge_1
    double xt = 0.5 * (((2.0 * D(dd - 1) + D(dd - 2)) * Q(dd - 1) - D(dd - 1) * Q(dd - 2)) / (D(dd - 2) + D(dd - 1)) +
                       ((2.0 * D(dd) + D(dd + 1)) * Q(dd) - D(dd) * Q(dd + 1)) / (D(dd) + D(dd + 1)));
    double minq = fv_min(fv_min(fv_min(Q(dd - 2), Q(dd - 1)), Q(dd)), Q(dd + 1));
    double maxq = fv_max(fv_max(fv_max(Q(dd - 2), Q(dd - 1)), Q(dd)), Q(dd + 1));
    return fv_min(fv_max(xt, minq), maxq);
  };
  auto dm_left = [&](int dd) {  // the "dm_left" of bl_br_edges: limited slope of cell dd-1
    double xt = 0.25 * (Q(dd) - Q(dd - 2));
    double dqr = fv_max(fv_max(Q(dd - 1), Q(dd - 2)), Q(dd)) - Q(dd - 1);
    double dql = Q(dd - 1) - fv_min(fv_min(Q(dd - 1), Q(dd - 2)), Q(dd));
    return fv_sign(fv_min(fv_min(fabs(xt), dqr), dql), xt);
  };
  auto dm_right = [&](int dd) {  // limited slope of cell dd+1
    double xt = 0.25 * (Q(dd + 2) - Q(dd));
    double dqr = fv_max(fv_max(Q(dd + 1), Q(dd)), Q(dd + 2)) - Q(dd + 1);
    double dql = Q(dd + 1) - fv_min(fv_min(Q(dd + 1), Q(dd)), Q(dd + 2));
    return fv_sign(fv_min(fv_min(fabs(xt), dqr), dql), xt);
  };
  bool edge = true;
  double xt_bl, xt_br;
  if (elo && x == lo - 1) {
    xt_bl = s14 * dm_left(d) + s11 * (Q(d - 1) - q) + q;
    xt_br = edge0(d);
  } else if (elo && x == lo) {
    xt_bl = edge1(d);
    xt_br = s15 * q + s11 * Q(d + 1) - s14 * dm_right(d);
  } else if (elo && x == lo + 1) {
    xt_bl = s15 * Q(d - 1) + s11 * q - s14 * dm(d);
    xt_br = al(d + 1);
  } else if (ehi && x == hi - 1) {
    xt_bl = al(d);
    xt_br = s15 * Q(d + 1) + s11 * q + s14 * dm(d);
  } else if (ehi && x == hi) {
    xt_bl = s15 * q + s11 * Q(d - 1) + s14 * dm_left(d);
    xt_br = edge0(d);
  } else if (ehi && x == hi + 1) {
    xt_bl = edge1(d);
    xt_br = s11 * (Q(d + 1) - q) - s14 * dm_right(d) + q;
  } else {
    edge = false;
    double xt = 2.0 * dm(d);
    bl = -1.0 * fv_sign(fv_min(fabs(xt), fabs(al(d) - q)), xt);
    br = fv_sign(fv_min(fabs(xt), fabs(al(d + 1) - q)), xt);
  }
  if (edge) {
    bl = xt_bl - q;
    br = xt_br - q;
    // pert_ppm_standard_constraint_fcn (ppm.py:21-33)
    if (bl * br < 0.0) {
      double da1 = bl - br;
      double da2 = da1 * da1;
      double a6da = 3.0 * (bl + br) * da1;
      if (a6da < -da2) br = -2.0 * bl;
      else if (a6da > da2) bl = -2.0 * br;
    } else {
      bl = 0.0;
      br = 0.0;
    }
  }
}

// ---- out-of-line copies of the tile-edge formulas -------------------------------------------------------
// The kernels that keep a whole transport call in one loop body (fvtp2d_march.cu) reach the edge formulas from
// several places; inlined there they made up most of a 70 KB kernel whose loop no longer fitted the instruction
// cache.  These are the same functions on values (q[n] = Q(n - 2), dx[n] = D(n - 2) around the cell / interface at
// index x), compiled once and called only by the lanes next to a tile edge.
struct BlBr {
  double bl, br;
};
static __device__ __noinline__ BlBr ppm_blbr8_vals(double qa, double qb, double qc, double qd, double qe, double da,
                                                   double db, double dc, double dd, double de, int x, int lo, int hi,
                                                   bool elo, bool ehi) {
  auto Q = [&](int d) { return d == -2 ? qa : d == -1 ? qb : d == 0 ? qc : d == 1 ? qd : qe; };
  auto D = [&](int d) { return d == -2 ? da : d == -1 ? db : d == 0 ? dc : d == 1 ? dd : de; };
  BlBr r;
  ppm_blbr8(Q, D, x, lo, hi, elo, ehi, 0, r.bl, r.br);
  return r;
}
// compute_al at interface x from Q(-2..1), D(-2..1)
static __device__ __noinline__ double ppm_al_vals(double qa, double qb, double qc, double qd, double da, double db,
                                                  double dc, double dd, int x, int lo, int hi, bool elo, bool ehi) {
  auto Q = [&](int d) { return d == -2 ? qa : d == -1 ? qb : d == 0 ? qc : qd; };
  auto D = [&](int d) { return d == -2 ? da : d == -1 ? db : d == 0 ? dc : dd; };
  return ppm_al(Q, D, x, lo, hi, elo, ehi, 0);
}

template <class QF, class DF>
__device__ __forceinline__ double ppm_flux_8(QF Q, DF D, double c, int pos, int lo, int hi, bool elo, bool ehi) {
  double bl, br;
  if (c > 0.0) {
    ppm_blbr8(Q, D, pos, lo, hi, elo, ehi, -1, bl, br);
    double b0 = bl + br;
    return Q(-1) + (1.0 - c) * (br - c * b0) * 1.0;
  }
  ppm_blbr8(Q, D, pos, lo, hi, elo, ehi, 0, bl, br);
  double b0 = bl + br;
  return Q(0) + (1.0 + c) * (bl + c * b0) * 1.0;
}

// ---- interior fast paths ------------------------------------------------------------------------------
// For a flux point whose stencil touches no tile edge the functions above reduce to the expressions below (same
// operations in the same order, hence the same bits) without the edge tests and without the grid-spacing reads.
template <class QF>
__device__ __forceinline__ double ppm_flux_lt8_interior(QF Q, double c, int mord) {
  using namespace ppmc;
  const double q3 = Q(-3), q2 = Q(-2), qm = Q(-1), q0 = Q(0), qp = Q(1), qpp = Q(2);
  const double alm = p1 * (q2 + qm) + p2 * (q3 + q0);
  const double al0 = p1 * (qm + q0) + p2 * (q2 + qp);
  const double alp = p1 * (q0 + qp) + p2 * (qm + qpp);
  double bl = al0 - q0, br = alp - q0, b0 = bl + br;
  double blm = alm - qm, brm = al0 - qm, b0m = blm + brm;
  bool s0, sm;
  if (mord == 5) {
    s0 = bl * br < 0;
    sm = blm * brm < 0;
  } else {
    s0 = (3.0 * fabs(b0)) < fabs(bl - br);
    sm = (3.0 * fabs(b0m)) < fabs(blm - brm);
  }
  double mask = (sm || s0) ? 1.0 : 0.0;
  double fx1 = c > 0.0 ? (1.0 - c) * (brm - c * b0m) : (1.0 + c) * (bl + c * b0);
  return c > 0.0 ? qm + fx1 * mask : q0 + fx1 * mask;
}

__device__ __forceinline__ double ppm_dm8_v(double qm, double q, double qp) {  // ppm_dm8 on values
  double xt = 0.25 * (qp - qm);
  double dqr = fv_max(fv_max(q, qm), qp) - q;
  double dql = q - fv_min(fv_min(q, qm), qp);
  return fv_sign(fv_min(fv_min(fabs(xt), dqr), dql), xt);
}

template <class QF>
__device__ __forceinline__ double ppm_flux_8_interior(QF Q, double c) {
  const int d = c > 0.0 ? -1 : 0;  // upwind cell
  const double qa = Q(d - 2), qb = Q(d - 1), q = Q(d), qd = Q(d + 1), qe = Q(d + 2);
  const double dmm = ppm_dm8_v(qa, qb, q), dm0 = ppm_dm8_v(qb, q, qd), dmp = ppm_dm8_v(q, qd, qe);
  const double al_d = 0.5 * (qb + q) + 1.0 / 3.0 * (dmm - dm0);
  const double al_d1 = 0.5 * (q + qd) + 1.0 / 3.0 * (dm0 - dmp);
  const double xt = 2.0 * dm0;
  const double bl = -1.0 * fv_sign(fv_min(fabs(xt), fabs(al_d - q)), xt);
  const double br = fv_sign(fv_min(fabs(xt), fabs(al_d1 - q)), xt);
  const double b0 = bl + br;
  if (c > 0.0) return q + (1.0 - c) * (br - c * b0) * 1.0;
  return q + (1.0 + c) * (bl + c * b0) * 1.0;
}

// ppm_flux with the interior fast path; `pos >= lo` is assumed (flux points of the compute domain)
template <class QF, class DF>
__device__ __forceinline__ double ppm_flux_fast(QF Q, DF D, double c, int pos, int lo, int hi, bool elo, bool ehi,
                                                int ord) {
  const int mord = ord < 0 ? -ord : ord;
  const bool near_edge = (elo && pos <= lo + 2) || (ehi && pos >= hi - 1);
  if (!near_edge) return mord < 8 ? ppm_flux_lt8_interior(Q, c, mord) : ppm_flux_8_interior(Q, c);
  if (mord < 8) return ppm_flux_lt8(Q, D, c, pos, lo, hi, elo, ehi, mord);
  return ppm_flux_8(Q, D, c, pos, lo, hi, elo, ehi);
}

template <class QF, class DF>
__device__ __forceinline__ double ppm_flux(QF Q, DF D, double c, int pos, int lo, int hi, bool elo, bool ehi,
                                           int ord) {
  int mord = ord < 0 ? -ord : ord;
  if (mord < 8) return ppm_flux_lt8(Q, D, c, pos, lo, hi, elo, ehi, mord);
  return ppm_flux_8(Q, D, c, pos, lo, hi, elo, ehi);
}

}  // namespace pace
