// End-of-step pieces of DynamicalCore._compute: AdjustNegativeTracerMixingRatio (Fortran neg_adj3) and
// CubedToLatLon (c2l_ord4).
//
// Reference: pyFV3/stencils/neg_adj3.py:20-313 (stencils), :316-431 (class);
//            NDSL/ndsl/stencils/c2l_ord.py:110-157 (ord4_transform), :249-282 (__call__).
#include "core.hpp"

namespace pace {

namespace {
constexpr double HLV = 2.5e6, HLF = 3.3358e5;
constexpr double DC_ICE = C_LIQ - C_ICE, LI0 = HLF - DC_ICE * TICE;
constexpr int KMAXN = 96;
}  // namespace

int neg_adj3(Core& c) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  if (c.cfg.hydrostatic) { c.err = "Adjust Negative Tracer Mixing Ratio (neg_adj3): Hydrostatic is not implemented"; return 1; }
  double *qvapor = c.F("qvapor"), *qliquid = c.F("qliquid"), *qrain = c.F("qrain"), *qsnow = c.F("qsnow"),
         *qice = c.F("qice"), *qgraupel = c.F("qgraupel"), *qcld = c.F("qcld"), *pt = c.F("pt");
  const double* delp = c.F("delp");
  if (!c.err.empty()) return 1;
  const int km = g.nz;
  if (km > KMAXN) { c.err = "neg_adj3: npz exceeds the thread-local column length (96)"; return 1; }
  const double d0_vap = CV_VAP - C_LIQ;
  const double lv00 = HLV - d0_vap * TICE;
  // fix_neg_water (neg_adj3.py:99-131), pointwise
  launch_ijk(st, g, box_compute(g), [=] __device__(const Pt& p) {
    double qv = A0(qvapor), ql = A0(qliquid), qr = A0(qrain), qs = A0(qsnow), qi = A0(qice), qg = A0(qgraupel), t = A0(pt);
    double q_liq = 0.0 > ql + qr ? 0.0 : ql + qr;
    double q_sol = 0.0 > qi + qs ? 0.0 : qi + qs;
    double cpm = (1.0 - (qv + q_liq + q_sol)) * CV_AIR + qv * CV_VAP + q_liq * C_LIQ + q_sol * C_ICE;
    double lcpk = (lv00 + d0_vap * t) / cpm;
    double icpk = (LI0 + DC_ICE * t) / cpm;
    double dq;
    // fix_negative_ice
    {
      double qsum = qi + qs;
      if (qsum > 0.0) {
        if (qi < 0.0) { qi = 0.0; qs = qsum; }
        else if (qs < 0.0) { qs = 0.0; qi = qsum; }
      } else {
        qi = 0.0; qs = 0.0; qg = qg + qsum;
      }
      if (qg < 0.0) {
        dq = qs < -qg ? qs : -qg;
        qs = qs - dq; qg = qg + dq;
        if (qg < 0.0) {
          dq = qi < -qg ? qi : -qg;
          qi = qi - dq; qg = qg + dq;
        }
      }
      if (qg < 0.0 && qr > 0.0) {
        dq = qr < -qg ? qr : -qg;
        qg = qg + dq; ql = ql - dq; t = t + dq * icpk;
      }
      if (qg < 0.0 && ql > 0.0) {
        dq = ql < -qg ? ql : -qg;
        qg = qg + dq; ql = ql - dq; t = t + dq * icpk;
      }
      if (qg < 0.0 && qv > 0.0) {
        dq = 0.999 * qv < -qg ? 0.999 * qv : -qg;
        qg = qg + dq; qv = qv - dq; t = t + dq * (icpk + lcpk);
      }
    }
    // fix_negative_liq
    {
      double qsum = ql + qr;
      double pos_qg = 0.0 > qg ? 0.0 : qg;
      double qrain_tmp = 0.0, dq1 = 0.0;
      if (qsum > 0.0) {
        if (qr < 0.0) { qr = 0.0; ql = qsum; }
        else if (ql < 0.0) { ql = 0.0; qr = qsum; }
      } else {
        ql = 0.0;
        qrain_tmp = qsum;
        dq = pos_qg < -qrain_tmp ? pos_qg : -qrain_tmp;
        qrain_tmp = qrain_tmp + dq;
        qg = qg - dq;
        t = t - dq * icpk;
        if (qr < 0.0) {
          dq = (qi + qs) < -qrain_tmp ? qi + qs : -qrain_tmp;
          qrain_tmp = qrain_tmp + dq;
          dq1 = dq < qs ? dq : qs;
          qs = qs - dq1;
          qi = qi + dq1 - dq;
          t = t - dq * icpk;
        }
        qr = qrain_tmp;
        if (qr < 0.0 && qv > 0.0) {
          dq = 0.999 * qv < -qr ? 0.999 * qv : -qr;
          qv = qv - dq; qr = qr + dq; t = t + dq * lcpk;
        }
      }
    }
    A0(qvapor) = qv; A0(qliquid) = ql; A0(qrain) = qr; A0(qsnow) = qs; A0(qice) = qi; A0(qgraupel) = qg; A0(pt) = t;
  });
  // fillq(qgraupel), fillq(qrain), fix_water_vapor_down(qvapor), fix_neg_cloud(qcld): column sweeps
  launch_ij(st, g, box_compute(g), [=] __device__(const Pt& p) {
    double dp[KMAXN], q[KMAXN], upper[KMAXN], lower[KMAXN];
    for (int k = 0; k < km; ++k) dp[k] = A(delp, 0, 0, k);
    // fillq (neg_adj3.py:134-153)
    for (int pass = 0; pass < 2; ++pass) {
      double* f = pass == 0 ? qgraupel : qrain;
      double sum1 = 0.0, sum2 = 0.0;
      for (int k = 0; k < km; ++k) {
        q[k] = A(f, 0, 0, k);
        if (q[k] > 0) sum1 = sum1 + q[k] * dp[k];
      }
      for (int k = km - 1; k >= 0; --k)
        if (q[k] < 0.0 && sum1 >= 0) {
          double dq = sum1 < -q[k] * dp[k] ? sum1 : -q[k] * dp[k];
          sum1 = sum1 - dq; sum2 = sum2 + dq;
          q[k] = q[k] + dq / dp[k];
        }
      for (int k = km - 1; k >= 0; --k)
        if (q[k] > 0.0 && sum1 >= 1e-12 && sum2 > 0) {
          double dq = sum2 < q[k] * dp[k] ? sum2 : q[k] * dp[k];
          sum2 = sum2 - dq;
          q[k] = q[k] - dq / dp[k];
        }
      for (int k = 0; k < km; ++k) A(f, 0, 0, k) = q[k];
    }
    // fix_water_vapor_down (neg_adj3.py:156-208)
    for (int k = 0; k < km; ++k) { q[k] = A(qvapor, 0, 0, k); upper[k] = 0.0; lower[k] = 0.0; }
    if (q[0] < 0) q[1] = q[1] + q[0] * dp[0] / dp[1];
    if (q[0] < 0.0) q[0] = 0.0;
    for (int k = 1; k < km - 1; ++k) {
      double dq = q[k - 1] * dp[k - 1];
      if (lower[k - 1] != 0) q[k] += lower[k - 1] / dp[k];
      if (q[k] < 0 && q[k - 1] > 0) {
        dq = dq < -q[k] * dp[k] ? dq : -q[k] * dp[k];
        upper[k] = dq;
        q[k] += dq / dp[k];
      }
      if (q[k] < 0) { lower[k] = q[k] * dp[k]; q[k] = 0; }
    }
    for (int k = 0; k < km - 2; ++k)
      if (upper[k + 1] != 0) q[k] = q[k] - upper[k + 1] / dp[k];
    {
      int k = km - 1;
      if (lower[k - 1] > 0) q[k] = q[k] + lower[k] / dp[k];
      upper[k] = q[k];
    }
    const double dp_bottom = dp[km - 1];
    for (int k = km - 2; k >= 0; --k) {
      double dq = q[k] * dp[k];
      if (upper[k + 1] < 0 && q[k] > 0) {
        if (dq >= -upper[k + 1] * dp_bottom) dq = -upper[k + 1] * dp_bottom;
        q[k] = q[k] - dq / dp[k];
        upper[k] = upper[k + 1] + dq / dp_bottom;
      } else {
        upper[k] = upper[k + 1];
      }
    }
    q[km - 1] = upper[0];
    for (int k = 0; k < km; ++k) A(qvapor, 0, 0, k) = q[k];
    // fix_neg_cloud (neg_adj3.py:211-241)
    for (int k = 0; k < km; ++k) q[k] = A(qcld, 0, 0, k);
    for (int k = 1; k < km - 1; ++k)
      if (q[k - 1] < 0.0) q[k] = q[k] + q[k - 1] * dp[k - 1] / dp[k];
    for (int k = 1; k < km - 1; ++k)
      if (q[k] < 0.0) q[k] = 0.0;
    {
      int k = km - 2;
      if (q[k + 1] < 0.0 && q[k] > 0) {
        double dq = -q[k] * dp[k] < q[k + 1] * dp[k + 1] ? -q[k] * dp[k] : q[k + 1] * dp[k + 1];
        q[k] = q[k] - dq / dp[k];
      }
      k = km - 1;
      if (q[k] < 0 && q[k - 1] > 0.0) {
        double dq = -q[k] * dp[k] < q[k - 1] * dp[k - 1] ? -q[k] * dp[k] : q[k - 1] * dp[k - 1];
        q[k] = q[k] + dq / dp[k];
        q[k] = 0.0 > q[k] ? 0.0 : q[k];
      }
    }
    for (int k = 0; k < km; ++k) A(qcld, 0, 0, k) = q[k];
  });
  return 0;
}

// ord4_transform (c2l_ord.py:110-157), grid_type < 4, compute domain; the u/v halo update is the caller's
int c2l_ord4(Core& c) {
  const Geo g = c.g;
  const double *u = c.F("u"), *v = c.F("v"), *dx = c.F("grid.dx"), *dy = c.F("grid.dy"), *a11 = c.F("grid.a11"),
               *a12 = c.F("grid.a12"), *a21 = c.F("grid.a21"), *a22 = c.F("grid.a22");
  double *ua = c.F("ua"), *va = c.F("va");
  if (!c.err.empty()) return 1;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;
  constexpr double C1 = 1.125, C2 = -0.125;
  launch_ijk(c.stream, g, box_compute(g), [=] __device__(const Pt& p) {
    double utmp = C2 * (A(u, 0, -1, 0) + A(u, 0, 2, 0)) + C1 * (A0(u) + A(u, 0, 1, 0));
    double vtmp = C2 * (A(v, -1, 0, 0) + A(v, 2, 0, 0)) + C1 * (A0(v) + A(v, 1, 0, 0));
    if ((p.S && p.j == js) || (p.N && p.j == je) || (p.W && p.i == is) || (p.E && p.i == ie)) {
      vtmp = 2.0 * ((A0(v) * A20(dy)) + (A(v, 1, 0, 0) * A2(dy, 1, 0))) / (A20(dy) + A2(dy, 1, 0));
      utmp = 2.0 * (A0(u) * A20(dx) + A(u, 0, 1, 0) * A2(dx, 0, 1)) / (A20(dx) + A2(dx, 0, 1));
    }
    A0(ua) = A20(a11) * utmp + A20(a12) * vtmp;
    A0(va) = A20(a21) * utmp + A20(a22) * vtmp;
  });
  return 0;
}

}  // namespace pace
