// Fast saturation adjustment inside the vertical remapping (Fortran fv_sat_adj), pointwise.
//
// Reference: pyFV3/stencils/saturation_adjustment.py:37-163 (saturation tables, evaluated analytically per point
// exactly as the reference's gtscript functions do), :165-534 (process functions), :561-942 (satadjust),
// :945-1108 (SatAdjust3d).
#include "core.hpp"

namespace pace {


namespace {
constexpr double HLV = 2.5e6, HLF = 3.3358e5, SAT_ADJUST_THRESHOLD = 1.0e-8;  // GFS constants
constexpr double DC_ICE = C_LIQ - C_ICE, DC_VAP = CP_VAP - C_LIQ, D2ICE = DC_VAP + DC_ICE;
constexpr double LV0 = HLV - DC_VAP * TICE, LI00 = HLF - DC_ICE * TICE, LI2 = LV0 + LI00;
constexpr double E00 = 611.21, T_WFR = TICE - 40.0, TICE0 = TICE - 0.01, T_SAT_MIN = TICE - 160.0;
constexpr double DELT = 0.1;
constexpr double QS_LENGTH = 2621;

__device__ __forceinline__ double dimf(double a, double b) { return a - b > 0 ? a - b : 0; }
__device__ __forceinline__ double tem_lower(double i) { return T_SAT_MIN + DELT * i; }
__device__ __forceinline__ double tem_upper(double i) { return 253.16 + DELT * i; }
__device__ __forceinline__ double q_table_oneline(double dhc, double lhc, double tem) {
  return E00 * fv_exp((dhc * fv_log(tem / TICE) + (tem - TICE) / (tem * TICE) * lhc) / RVGAS);
}
__device__ __forceinline__ double table_vapor(double tem) { return q_table_oneline(DC_VAP, LV0, tem); }
__device__ __forceinline__ double table_ice(double tem) { return q_table_oneline(D2ICE, LI2, tem); }

__device__ double qs_table2_fn(double i) {
  double tem0 = tem_lower(i);
  double table2 = i < 1600 ? table_ice(tem0) : table_vapor(tem0);
  if (i == 1599) {
    double table = table_ice(tem0);
    tem0 = tem_upper(i - 1400);
    table = (0.05 * (TICE - tem0)) * table + (0.05 * (tem0 - 253.16)) * table_vapor(tem0);
    double table2_m1 = table_ice(tem_lower(1598));
    double table2_p1 = table_vapor(tem_lower(1600));
    table2 = 0.25 * (table2_m1 + 2.0 * table + table2_p1);
  }
  if (i == 1600) {
    tem0 = tem_upper(i - 1400);
    double table = table_vapor(tem0);
    double table2_m1 = table_ice(tem_lower(1599));
    double table2_p1 = table_vapor(tem_lower(1601));
    table2 = 0.25 * (table2_m1 + 2.0 * table + table2_p1);
  }
  return table2;
}
__device__ __forceinline__ double qs_tablew_fn(double i) { return table_vapor(tem_lower(i)); }
__device__ __forceinline__ double des_end(double t, double i, double des2) {
  if (i == QS_LENGTH - 1) {
    double t_m1 = qs_table2_fn(i - 1);
    des2 = fv_max(0.0, t - t_m1);
  }
  return des2;
}
__device__ double des2_table(double i) {
  double t = qs_table2_fn(i);
  double des2 = fv_max(0.0, qs_table2_fn(i + 1) - t);
  return des_end(t, i, des2);
}
__device__ double desw_table(double i) {
  double t = qs_tablew_fn(i);
  double desw = fv_max(0.0, qs_tablew_fn(i + 1) - t);
  return des_end(t, i, desw);
}
__device__ __forceinline__ double ap1_for_wqs2(double ta) {
  double ap1 = 10.0 * dimf(ta, T_SAT_MIN) + 1.0;
  return fv_min(ap1, QS_LENGTH) - 1;
}
// The reference evaluates its saturation tables as functions (no stored table): table(i) costs an exp and a log, and
// wqs2 asks for table(it), des(it), des(it2), des(it2+1) with des(i) = max(0, table(i+1) - table(i)) — 7 table
// evaluations of which only 3 are distinct, because it is it2 or it2 + 1.  Each distinct value is computed once
// here; a pure function evaluated once instead of twice gives the same bits.
template <bool W>
__device__ __forceinline__ double sat_table(double i) { return W ? qs_tablew_fn(i) : qs_table2_fn(i); }
// des2_table / desw_table given table(i) and table(i+1) (the last table entry differences backwards: des_end)
template <bool W>
__device__ __forceinline__ double sat_des(double i, double ti, double tip1) {
  double d = fv_max(0.0, tip1 - ti);
  if (i == QS_LENGTH - 1) d = fv_max(0.0, ti - sat_table<W>(i - 1));
  return d;
}
// wqs2_fn_w / wqs2_fn_2 (saturation_adjustment.py:501-520)
template <bool W>
__device__ void wqs2(double ta, double den, double& wqsat, double& dqdt) {
  double ap1 = ap1_for_wqs2(ta);
  double it = floor(ap1), it2 = floor(ap1 - 0.5);
  const double t0 = sat_table<W>(it2), t1 = sat_table<W>(it2 + 1), t2 = sat_table<W>(it2 + 2);
  const double des2 = sat_des<W>(it2, t0, t1);
  const double des_p1 = sat_des<W>(it2 + 1, t1, t2);
  const bool same = it == it2;  // otherwise it == it2 + 1
  const double tab = same ? t0 : t1;
  const double des = same ? des2 : des_p1;
  double es = tab + (ap1 - it) * des;
  double denom = RVGAS * ta * den;
  wqsat = es / denom;
  dqdt = 10.0 * (des2 + (ap1 - it2) * (des_p1 - des2));
  dqdt = dqdt / denom;
}
template <bool W>
__device__ double wqs1(double it, double ap1, double ta, double den) {
  const double tab = sat_table<W>(it);
  const double des = sat_des<W>(it, tab, sat_table<W>(it + 1));
  double es = tab + (ap1 - it) * des;
  return es / (RVGAS * ta * den);
}
__device__ __forceinline__ double compute_cvm(double mc_air, double qv, double c_vap, double q_liq, double q_sol) {
  return mc_air + qv * c_vap + q_liq * C_LIQ + q_sol * C_ICE;
}
__device__ __forceinline__ double minmax_tmp_h20(double qa, double qb) { return fv_min(-qa, fv_max(qb, 0.0)); }
}  // namespace

int sat_adjust(Core& c, const SatCfg& sc, double mdt, int last_step_i, int kmp) {
  const Geo g = c.g;
  if (c.cfg.hydrostatic) { c.err = "Hydrostatic is not implemented"; return 1; }
  double *qv_ = c.F("qvapor"), *ql_ = c.F("qliquid"), *qi_ = c.F("qice"), *qr_ = c.F("qrain"), *qs_ = c.F("qsnow"),
         *qg_ = c.F("qgraupel"), *qa_ = c.F("qcld"), *cappa_ = c.F("ad.cappa"), *pt_ = c.F("pt"), *q_con_ = c.F("q_con"),
         *pkz_ = c.F("pkz");
  const double *dp_ = c.F("delp"), *delz_ = c.F("delz"), *area = c.F("grid.area_64"), *hs = c.F("phis");
  if (!c.err.empty()) return 1;
  // SatAdjust3d.__call__ (saturation_adjustment.py:1007-1030): host scalars, math.exp == std::exp
  const double sdt = 0.5 * mdt;
  const double fac_i2s = 1.0 - std::exp(-mdt / sc.tau_i2s);
  const double fac_v2l = 1.0 - std::exp(-sdt / sc.tau_v2l);
  const double fac_r2g = 1.0 - std::exp(-mdt / sc.tau_r2g);
  const double fac_l2r = 1.0 - std::exp(-mdt / sc.tau_l2r);
  const double fac_l2v = std::min(sc.sat_adj0, 1.0 - std::exp(-sdt / sc.tau_l2v));
  const double fac_imlt = 1.0 - std::exp(-sdt / sc.tau_imlt);
  const double fac_smlt = 1.0 - std::exp(-mdt / sc.tau_smlt);
  const double c_air = CV_AIR, c_vap = CV_VAP;
  const double d0_vap = c_vap - C_LIQ;
  const double lv00 = HLV - d0_vap * TICE;
  const double LAT2 = std::pow(HLV + HLF, 2.0);  // constants.LAT2 = (HLV + HLF) ** 2 (Python float pow)
  const double zvir = ZVIR;
  const bool last_step = last_step_i != 0, do_qa = sc.do_qa != 0;
  const SatCfg s = sc;
  Box b = box_compute(g);
  b.k0 = kmp;
  b.nk = g.nz - kmp;
  launch_ijk_occ<4>(c.stream, g, b, [=] __device__(const Pt& p) {
    double qv = A0(qv_), ql = A0(ql_), qi = A0(qi_), qr = A0(qr_), qs = A0(qs_), qg = A0(qg_);
    const double pt = A0(pt_), dp = A0(dp_), delz = A0(delz_);
    double q_liq = ql + qr;
    double q_sol = qi + qs + qg;
    double qpz = q_liq + q_sol;
    double pt1 = pt / ((1.0 + zvir * qv) * (1.0 - qpz));
    qpz = qpz + qv;
    const double den = -dp / (GRAV * delz);
    const double mc_air = (1.0 - qpz) * c_air;
    double cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
    double lhi = LI00 + DC_ICE * pt1, icp2 = lhi / cvm, lhl, lcp2;
    auto upd_i = [&]() { lhi = LI00 + DC_ICE * pt1; icp2 = lhi / cvm; };
    auto upd = [&]() { lhl = lv00 + d0_vap * pt1; lcp2 = lhl / cvm; upd_i(); };
    if (qi < 0.0) { qs = qs + qi; qi = 0.0; }
    // melt_cloud_ice
    if (qi > 1.0e-8 && pt1 > TICE) {
      double factmp = fac_imlt * (pt1 - TICE) / icp2;
      double sink = qi < factmp ? qi : factmp;
      qi = qi - sink; ql = ql + sink; q_liq = q_liq + sink; q_sol = q_sol - sink;
      cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
      sink = -sink;
      pt1 = pt1 + sink * lhi / cvm;
    }
    upd_i();
    // fix_negative_snow / fix_negative_cloud_water
    if (qs < 0.0) { qg = qg + qs; qs = 0.0; }
    else if (qg < 0.0) { double t = minmax_tmp_h20(qg, qs); qg = qg + t; qs = qs - t; }
    if (ql < 0.0) { double t = minmax_tmp_h20(ql, qr); ql = ql + t; qr = qr - t; }
    else if (qr < 0.0) { double t = minmax_tmp_h20(qr, ql); ql = ql - t; qr = qr + t; }
    // complete_freezing
    {
      double dtmp = TICE - 48.0 - pt1;
      if (ql > 0.0 && dtmp > 0.0) {
        double sink = fv_min(ql, dtmp / icp2);
        ql = ql - sink; qi = qi + sink; q_liq = q_liq - sink; q_sol = q_sol + sink;
        cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
        pt1 = pt1 + sink * lhi / cvm;
      }
    }
    double wqsat, dq2dt;
    wqs2<true>(pt1, den, wqsat, dq2dt);
    upd();
    double tcp3 = lcp2 + icp2 * fv_min(1.0, dimf(TICE, pt1) / 48.0);
    double dq0 = (qv - wqsat) / (1.0 + tcp3 * dq2dt);
    double src;
    auto evap = [&]() {
      double factor = -fv_min(1.0, fac_l2v * 10.0 * (1.0 - qv / wqsat));
      return -fv_min(ql, factor * dq0);
    };
    if (dq0 > 0) src = fv_min(s.sat_adj0 * dq0, fv_max(s.ql_gen - ql, fac_v2l * dq0));
    else src = evap();
    auto wqsat_correct = [&]() {
      qv = qv - src; ql = ql + src; q_liq = q_liq + src;
      cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
      pt1 = pt1 + src * lhl / cvm;
    };
    wqsat_correct();
    upd();
    tcp3 = lcp2 + icp2 * fv_min(1.0, dimf(TICE, pt1) / 48.0);
    if (last_step) {
      wqs2<true>(pt1, den, wqsat, dq2dt);
      dq0 = (qv - wqsat) / (1.0 + tcp3 * dq2dt);
      if (dq0 > 0) src = dq0;
      else src = evap();
      wqsat_correct();
      upd();
    }
    // homogenous_freezing
    {
      double dtmp = T_WFR - pt1;
      if (ql > 0.0 && dtmp > 0.0) {
        double sink = fv_min(ql, dtmp / icp2);
        sink = fv_min(sink, ql * dtmp * 0.125);
        ql = ql - sink; qi = qi + sink; q_liq = q_liq - sink; q_sol = q_sol + sink;
        cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
        pt1 = pt1 + sink * lhi / cvm;
      }
    }
    upd_i();
    const double exptc = fv_exp(0.66 * (TICE0 - pt1));
    // heterogeneous_freezing (dt_bigg = mdt)
    {
      double tc = TICE0 - pt1;
      if (ql > 0.0 && tc > 0.0) {
        double sink = 3.3333e-10 * mdt * (exptc - 1.0) * den * (ql * ql);
        sink = fv_min(ql, sink);
        sink = fv_min(sink, tc / icp2);
        ql = ql - sink; qi = qi + sink; q_liq = q_liq - sink; q_sol = q_sol + sink;
        cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
        pt1 = pt1 + sink * lhi / cvm;
      }
    }
    upd_i();
    // make_graupel
    {
      double dtmp = (TICE - 0.1) - pt1;
      if (qr > 1e-7 && dtmp > 0.0) {
        double rainfac = (dtmp * 0.025) * (dtmp * 0.025);
        double tmp = 1.0 < rainfac ? qr : rainfac * qr;
        double sink = fv_min(tmp, fac_r2g * dtmp / icp2);
        qr = qr - sink; qg = qg + sink; q_liq = q_liq - sink; q_sol = q_sol + sink;
        cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
        pt1 = pt1 + sink * lhi / cvm;
      }
    }
    upd_i();
    // melt_snow
    {
      double dtmp = pt1 - (TICE + 0.1);
      double dimqs = dimf(s.qs_mlt, ql);
      if (qs > 1e-7 && dtmp > 0.0) {
        double snowfac = (dtmp * 0.1) * (dtmp * 0.1);
        double tmp = 1.0 < snowfac ? qs : snowfac * qs;
        double sink = fv_min(tmp, fac_smlt * dtmp / icp2);
        tmp = fv_min(sink, dimqs);
        qs = qs - sink; ql = ql + tmp; qr = qr + sink - tmp; q_liq = q_liq + sink; q_sol = q_sol - sink;
        cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
        pt1 = pt1 - sink * lhi / cvm;
      }
    }
    // autoconversion_cloud_to_rain
    if (ql > s.ql0_max) {
      double sink = fac_l2r * (ql - s.ql0_max);
      qr = qr + sink; ql = ql - sink;
    }
    double iqs2, dqsdt;
    wqs2<false>(pt1, den, iqs2, dqsdt);
    const double expsubl = fv_exp(0.875 * fv_log(qi * den));
    upd();
    const double tcp2 = lcp2 + icp2;
    const double adj_fac = last_step ? 1.0 : s.sat_adj0;
    // sublimation
    {
      double srcs = 0.0;
      if (pt1 < s.t_sub) {
        srcs = dimf(qv, 1e-6);
      } else if (pt1 < TICE0) {
        double dq = qv - iqs2;
        double sink = adj_fac * dq / (1.0 + tcp2 * dqsdt);
        double pidep;
        if (qi > 1.0e-8)
          pidep = sdt * dq * 349138.78 * expsubl / (iqs2 * den * LAT2 / (0.0243 * RVGAS * (pt1 * pt1)) + 4.42478e4);
        else
          pidep = 0.0;
        if (dq > 0.0) {
          double tmp = TICE - pt1;
          double qi_crt = s.qi_lim < 0.1 * tmp ? s.qi_gen * s.qi_lim / den : s.qi_gen * 0.1 * tmp / den;
          double maxtmp = qi_crt - qi > pidep ? qi_crt - qi : pidep;
          srcs = sink < maxtmp ? sink : maxtmp;
          srcs = srcs < tmp / tcp2 ? srcs : tmp / tcp2;
        } else {
          double dimtmp = dimf(pt1, s.t_sub);
          pidep = 1.0 < (dimtmp * 0.2) ? pidep : pidep * dimtmp * 0.2;
          srcs = pidep > sink ? pidep : sink;
          srcs = srcs > -qi ? srcs : -qi;
        }
      }
      qv = qv - srcs; qi = qi + srcs; q_sol = q_sol + srcs;
      cvm = compute_cvm(mc_air, qv, c_vap, q_liq, q_sol);
      double lh = lhl + lhi;
      pt1 = pt1 + srcs * lh / cvm;
    }
    const double q_con = q_liq + q_sol;
    double tmpv = 1.0 + zvir * qv;
    const double pt_new = pt1 * tmpv * (1.0 - q_con);
    tmpv *= RDGAS;
    const double cappa = tmpv / (tmpv + cvm);
    if (qg < 0) {
      double mintmp = fv_min(-qg, fv_max(0.0, qi));
      qg = qg + mintmp; qi = qi - mintmp;
    }
    const double qim = s.qi0_max / den;
    if (qi > qim) {
      double sink = fac_i2s * (qi - qim);
      qi = qi - sink; qs = qs + sink;
    }
    cvm = mc_air + (qv + q_liq + q_sol) * c_vap;
    upd();
    if (do_qa && last_step) {
      q_sol = s.rad_snow ? (s.rad_graupel ? qi + qs + qg : qi + qs) : qi;
      q_liq = s.rad_rain ? ql + qr : ql;
      const double q_cond = q_sol + q_liq;
      const double tin = s.tintqs ? pt1 : pt1 - (lcp2 * q_cond + icp2 * q_sol);
      double ap1 = ap1_for_wqs2(tin);
      double it = floor(ap1);
      double wq = wqs1<true>(it, ap1, tin, den);
      double iq = wqs1<false>(it, ap1, tin, den);
      double qstar;
      if (tin < T_WFR) qstar = iq;
      else if (tin >= TICE) qstar = wq;
      else {
        double rqi = q_cond > 1e-6 ? q_sol / q_cond : (TICE - tin) / (TICE - T_WFR);
        qstar = rqi * iq + (1.0 - rqi) * wq;
      }
      double mindw = fv_min(1.0, fabs(A20(hs)) / (10.0 * GRAV));
      double dw = s.dw_ocean + (s.dw_land - s.dw_ocean) * mindw;
      double dbl_sqrt_area = dw * fv_sqrt(fv_sqrt(A20(area)) / 100.0e3);
      double hvar = fv_min(0.2, fv_max(0.01, dbl_sqrt_area));
      double rh = qpz / qstar;
      double qa = 0.0;
      if (rh > 0.75 && qpz > SAT_ADJUST_THRESHOLD) {
        double dq = hvar * qpz;
        double q_plus = qpz + dq, q_minus = qpz - dq;
        if (s.icloud_f == 2) {
          if (qpz > qstar) qa = 1.0;
          else if (qstar < q_plus && q_cond > SAT_ADJUST_THRESHOLD) {
            double r = (q_plus - qstar) / dq;
            qa = fv_min(1.0, r * r);
          } else qa = 0.0;
        } else {
          if (qstar < q_minus) qa = 1.0;
          else {
            if (qstar < q_plus) qa = s.icloud_f == 0 ? (q_plus - qstar) / (dq + dq) : (q_plus - qstar) / (2.0 * dq * (1.0 - q_cond));
            else qa = 0.0;
            if (q_cond > SAT_ADJUST_THRESHOLD) qa = fv_max(s.cld_min, qa);
            qa = fv_min(1.0, qa);
          }
        }
      }
      A0(qa_) = qa;
    }
    A0(qv_) = qv; A0(ql_) = ql; A0(qi_) = qi; A0(qr_) = qr; A0(qs_) = qs; A0(qg_) = qg;
    A0(q_con_) = q_con;
    A0(pt_) = pt_new;
    A0(cappa_) = cappa;
    A0(pkz_) = fv_exp(cappa * fv_log(RDG * dp / delz * pt_new));
  });
  return 0;
}

}  // namespace pace
