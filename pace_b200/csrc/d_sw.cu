// D-grid shallow-water Lagrangian dynamics (Fortran d_sw): the full-step update of delp, pt, w, q_con, u, v.
//
// Reference: pyFV3/stencils/d_sw.py:34-741 (stencils), :997-1303 (__call__ order);
//            pyFV3/stencils/xtp_u.py:8-95, ytp_v.py:8-94 (PPM advection of u, v inside the KE stencil).
// 178 reference launches; the transport operators are in transport.cu, the flux prep in fxadv.cu, the
// divergence damping in divergence_damping.cu, the rest is grouped here (comments name the reference
// stencil each kernel covers).
#include <cstdlib>

#include "ppm.cuh"
#include "transport.hpp"

namespace pace {

int divergence_damping(Core& c, double* vort_b, double* delpc, double* ke, const double* vort_a, double dt);

namespace {
// bl/br of xtp_u.get_bl_br / ytp_v.get_bl_br for iord < 8, cell offset d along the sweep direction.
// zero_lo/zero_hi: cell index ranges {lo-1,lo} / {hi,hi+1} get bl = br = 0 on the rows (columns) that
// touch the tile edge in the OTHER direction (xtp_u.py:46-54).
template <class QF, class DF>
__device__ __forceinline__ void adv_blbr(QF Q, DF D, int pos, int lo, int hi, bool elo, bool ehi, bool zero_row,
                                         int d, double& bl, double& br) {
  double al0 = ppm_al(Q, D, pos, lo, hi, elo, ehi, d);
  double al1 = ppm_al(Q, D, pos, lo, hi, elo, ehi, d + 1);
  bl = al0 - Q(d);
  br = al1 - Q(d);
  const int x = pos + d;
  if (zero_row && ((elo && (x == lo - 1 || x == lo)) || (ehi && (x == hi || x == hi + 1)))) {
    bl = 0.0;
    br = 0.0;
  }
}
// advect_u_along_x / advect_v_along_y (xtp_u.py:57-95) for iord < 8
template <class QF, class DF>
__device__ __forceinline__ double adv_along(QF Q, DF D, double ub, double cfl, int pos, int lo, int hi, bool elo,
                                            bool ehi, bool zero_row, int mord) {
  double bl, br, blm, brm;
  if ((elo && pos <= lo + 2) || (ehi && pos >= hi - 1)) {  // an edge formula or the zeroed rows may apply
    adv_blbr(Q, D, pos, lo, hi, elo, ehi, zero_row, 0, bl, br);
    adv_blbr(Q, D, pos, lo, hi, elo, ehi, zero_row, -1, blm, brm);
  } else {
    // interior: the same edge values, each computed once and without the edge case analysis
    using namespace ppmc;
    const double q3 = Q(-3), q2 = Q(-2), qm = Q(-1), q0 = Q(0), qp = Q(1), qpp = Q(2);
    const double alm = p1 * (q2 + qm) + p2 * (q3 + q0);
    const double al0 = p1 * (qm + q0) + p2 * (q2 + qp);
    const double alp = p1 * (q0 + qp) + p2 * (qm + qpp);
    bl = al0 - q0;
    br = alp - q0;
    blm = alm - qm;
    brm = al0 - qm;
  }
  double b0 = bl + br, b0m = blm + brm;
  double fx0 = cfl > 0.0 ? (1.0 - cfl) * (brm - cfl * b0m) : (1.0 + cfl) * (bl + cfl * b0);
  bool s0, sm;
  if (mord == 5) {
    s0 = bl * br < 0;
    sm = blm * brm < 0;
  } else {
    s0 = (3.0 * fabs(b0)) < fabs(bl - br);
    sm = (3.0 * fabs(b0m)) < fabs(blm - brm);
  }
  double mask = (sm || s0) ? 1.0 : 0.0;
  return ub > 0.0 ? Q(-1) + fx0 * mask : Q(0) + fx0 * mask;
}
}  // namespace

int d_sw(Core& c, double dt) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const Config& cf = c.cfg;
  if (cf.inline_q) { c.err = "D-Grid Shallow Water Lagrangian Dynamics (D_SW): inline_q not implemented."; return 1; }
  if (cf.d_ext > 0) { c.err = "D-Grid Shallow Water Lagrangian Dynamics (D_SW): untested d_ext > 0. need to call a2b_ord2, not implemented."; return 1; }
  if (cf.hydrostatic) { c.err = "D-Grid Shallow Water Lagrangian Dynamics (D_SW): Hydrostatic is not implemented"; return 1; }
  if (cf.do_f3d) { c.err = "do_f3d is not implemented"; return 1; }
  if (cf.hord_mt >= 8 || cf.hord_mt < 0) { c.err = "d_sw: hord_mt >= 8 (xtp_u/ytp_v ord-8 branch) is not implemented"; return 1; }

  double *delpc = c.F("ad.vt");  // the reference passes AcousticDynamics._vt as d_sw's delpc (dyn_core.py:867-868)
  double *delp = c.F("delp"), *pt = c.F("pt"), *u = c.F("u"), *v = c.F("v"), *w = c.F("w"), *uc = c.F("uc"),
         *vc = c.F("vc"), *q_con = c.F("q_con"), *mfx = c.F("mfxd"), *mfy = c.F("mfyd"), *cx = c.F("cxd"),
         *cy = c.F("cyd"), *crx = c.F("ad.crx"), *cry = c.F("ad.cry"), *xfx = c.F("ad.xfx"), *yfx = c.F("ad.yfx"),
         *heat_source = c.F("ad.heat_source"), *diss_est = c.F("diss_estd");
  double *ucc = c.FS("dsw.uc_contra"), *vcc = c.FS("dsw.vc_contra");
  const double *rarea = c.F("grid.rarea"), *dx = c.F("grid.dx"), *dy = c.F("grid.dy"), *rdx = c.F("grid.rdx"),
               *rdy = c.F("grid.rdy"), *cosa = c.F("grid.cosa"), *rsina = c.F("grid.rsina"), *f0 = c.F("grid.fC_agrid"),
               *rsin2 = c.F("grid.rsin2"), *cosa_s = c.F("grid.cosa_s");
  const double *col_nord_v = c.F("col.nord_v"), *col_nord_w = c.F("col.nord_w"), *col_nord_t = c.F("col.nord_t"),
               *col_damp_vt = c.F("col.damp_vt"), *col_damp_w = c.F("col.damp_w"), *col_d_con = c.F("col.d_con"),
               *col_ke_bg = c.F("col.ke_bg"), *damp_vt_fv = c.F("col.delnflux_damp_vt_fv"),
               *damp_t_fv = c.F("col.delnflux_damp_t_fv"), *damp_vt_c = c.F("col.delnflux_damp_vt"),
               *damp_w_c = c.F("col.delnflux_damp_w");
  double *tfx = c.FS("dsw.tmp_fx"), *tfy = c.FS("dsw.tmp_fy"), *gx = c.FS("dsw.tmp_gx"), *gy = c.FS("dsw.tmp_gy"),
         *fx2 = c.FS("dsw.tmp_fx2"), *fy2 = c.FS("dsw.tmp_fy2"), *dw = c.FS("dsw.tmp_dw"), *wk = c.FS("dsw.tmp_wk"),
         *heat_s = c.FS("dsw.tmp_heat_s"), *ke = c.FS("dsw.dt_kinetic_energy_on_cell_corners"),
         *vort_a = c.FS("dsw.vorticity_agrid"), *vort_b = c.FS("dsw.vorticity_bgrid_damped"),
         *abs_vort = c.FS("dsw.abs_vorticity_agrid"), *tut = c.FS("dsw.tmp_ut"), *tvt = c.FS("dsw.tmp_vt"),
         *damped_vort = c.FS("dsw.damped_rel_vorticity_agrid");
  if (!c.err.empty()) return 1;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, nz = g.nz;
  const int nmax_v = c.col_max("col.nord_v"), nmax_w = c.col_max("col.nord_w"), nmax_t = c.col_max("col.nord_t");
  if (!c.err.empty()) return 1;

  // 1. fv_prep; its last kernel also does the Courant-number half of flux_capacitor (3.): cx += crx, cy += cry.
  // Outside the columns / rows fv_prep writes, crx / cry are zero for good (nothing else writes them), so adding them
  // there, as the reference's full-box stencil does, changes nothing.
  if (fv_prep(c, dt, true)) return 1;
  // 2. fvtp2d_dp(delp) with delnflux(nord_v, damp_vt)
  Fvtp2dArgs a;
  a.crx = crx; a.cry = cry; a.xfx = xfx; a.yfx = yfx; a.nk = nz;
  a.q = delp; a.fx = tfx; a.fy = tfy; a.hord = cf.hord_dp; a.nord_col = col_nord_v; a.damp = damp_vt_fv; a.nmax = nmax_v;
  if (fvtp2d(c, a)) return 1;
  // 3. flux_capacitor, mass-flux half: mfx += tfx, mfy += tfy are done by the kernel of steps 11./12., which reads the
  // same fluxes anyway (tfx / tfy are zero outside the boxes fvtp2d writes, so the reference's additions on the rest of
  // its compute + 3 halo box change nothing: same argument as for crx / cry)
  (void)cx; (void)cy;
  // 4. delnflux_nosg_w(w)
  DelnHeat heat;
  heat.dw = dw; heat.heat_s = heat_s; heat.col_damp_w = col_damp_w; heat.col_ke_bg = col_ke_bg; heat.absdt = fabs(dt);
  {
    DelnArgs d;
    d.q = w; d.fx = fx2; d.fy = fy2; d.nord_col = col_nord_w; d.damp = damp_w_c; d.d2 = wk; d.nmax = nmax_w; d.nk = nz;
    d.heat = &heat;
    if (delnflux(c, d)) return 1;
  }
  // 5. heat_diss (evaluated inside the DelnFlux kernels when they ran: heat.done)
  if (!heat.done) launch_ijk(st, g, box_dims(g, 0, 0, 0), [=] __device__(const Pt& p) {
    double hs = 0.0;
    if (col_damp_w[p.k] > 1e-5) {
      double dd8 = col_ke_bg[p.k] * fabs(dt);
      double dwv = (A0(fx2) - A(fx2, 1, 0, 0) + A0(fy2) - A(fy2, 0, 1, 0)) * A20(rarea);
      A0(dw) = dwv;
      hs = dd8 - dwv * (A0(w) + 0.5 * dwv);
    }
    A0(heat_s) = hs;  // the reference's tmp_diss_e holds the same value: it is taken from here below
  });
  // 6./7. fvtp2d_vt_nodelnflux(w, mass fluxes) ; apply_fluxes(w)
  a = Fvtp2dArgs();
  a.crx = crx; a.cry = cry; a.xfx = xfx; a.yfx = yfx; a.nk = nz;
  // The fluxes of w go to their own pair of fields: apply_fluxes(w) is then evaluated inside the kernel of steps
  // 11./12. (nothing in between reads w), which saves a pass over w, delp and the fluxes.
  double *gxw = c.FS("dsw.tmp_gx_w"), *gyw = c.FS("dsw.tmp_gy_w");
  if (!c.err.empty()) return 1;
  a.q = w; a.fx = gxw; a.fy = gyw; a.hord = cf.hord_vt; a.x_mass_flux = tfx; a.y_mass_flux = tfy;
  if (fvtp2d(c, a)) return 1;
  a.fx = gx; a.fy = gy;
  // 8./9. fvtp2d_dp_t(q_con, mass=delp) ; apply_fluxes(q_con)
  // levels without condensate (all of them in dry air): q_con stays zero there, its transport is skipped
  const int* qcon_levels = level_nonzero_flags(c, q_con, nz, 8, c.in_acoustic && c.qcon_flags_fresh);
  c.qcon_flags_fresh = c.in_acoustic;
  if (!c.err.empty()) return 1;
  a.q = q_con; a.hord = cf.hord_dp; a.mass = delp; a.nord_col = col_nord_t; a.damp = damp_t_fv; a.nmax = nmax_t;
  a.level_nonzero = qcon_levels;
  if (fvtp2d(c, a)) return 1;
  a.level_nonzero = nullptr;
  launch_ijk(st, g, box_dims(g, 0, 0, 0), [=] __device__(const Pt& p) {
    if (qcon_levels != nullptr && qcon_levels[p.t * nz + p.k] == 0) return;  // 0 * delp + 0
    A0(q_con) = A0(q_con) * A0(delp) + (A0(gx) - A(gx, 1, 0, 0) + A0(gy) - A(gy, 0, 1, 0)) * A20(rarea);
  });
  // 10. fvtp2d_tm(pt, mass=delp)
  a.q = pt; a.hord = cf.hord_tm; a.nord_col = col_nord_v; a.damp = damp_vt_fv; a.nmax = nmax_v;
  if (fvtp2d(c, a)) return 1;
  // 11./12. apply_pt_delp_fluxes ; adjust_w_and_qcon on the compute domain; 3. mfx += tfx on [is, ie+1] x [js, je],
  // mfy += tfy on [is, ie] x [js, je+1] (the boxes fvtp2d writes tfx / tfy on)
  launch_ijk(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    const bool in_i = p.i <= ie, in_j = p.j <= je;
    if (in_j) A0(mfx) = A0(mfx) + A0(tfx);
    if (in_i) A0(mfy) = A0(mfy) + A0(tfy);
    if (!(in_i && in_j)) return;
    const double delp0 = A0(delp);  // delp before this step: the weight of pt and w
    double ptv = A0(pt) * delp0 + (A0(gx) - A(gx, 1, 0, 0) + A0(gy) - A(gy, 0, 1, 0)) * A20(rarea);
    double dp = delp0 + (A0(tfx) - A(tfx, 1, 0, 0) + A0(tfy) - A(tfy, 0, 1, 0)) * A20(rarea);
    A0(delp) = dp;
    A0(pt) = ptv / dp;
    const double wdp = A0(w) * delp0 + (A0(gxw) - A(gxw, 1, 0, 0) + A0(gyw) - A(gyw, 0, 1, 0)) * A20(rarea);  // 7. apply_fluxes(w)
    double wv = wdp / dp;
    wv = col_damp_w[p.k] > 1e-5 ? wv + A0(dw) : wv;
    A0(w) = wv;
    // zero wherever there is no condensate: levels flagged all-zero keep their zeros (0 / dp), no read, no write
    if (qcon_levels == nullptr || qcon_levels[p.t * nz + p.k] != 0) A0(q_con) = fv_div(A0(q_con), dp);
  });
  // 13. compute_kinetic_energy on the corner compute domain
  const int mord_mt = cf.hord_mt < 0 ? -cf.hord_mt : cf.hord_mt;
  const double *dxa = c.F("grid.dxa"), *dya = c.F("grid.dya");
  (void)dxa; (void)dya;
  launch_ijk_occ<6>(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    const bool iW = p.W && i == is, iE = p.E && i == ie + 1, jS = p.S && j == js, jN = p.N && j == je + 1;
    // interpolate_uc_vc_to_cell_corners (d_sw.py:694-741)
    double ub_cov = 0.5 * (A(uc, 0, -1, 0) + A0(uc));
    double vb_cov = 0.5 * (A(vc, -1, 0, 0) + A0(vc));
    double ub = (ub_cov - vb_cov * A20(cosa)) * A20(rsina);
    double vb = (vb_cov - ub_cov * A20(cosa)) * A20(rsina);
    if (jS || jN) ub = 0.25 * (-A(ucc, 0, -2, 0) + 3.0 * (A(ucc, 0, -1, 0) + A0(ucc)) - A(ucc, 0, 1, 0));
    if (iW || iE) ub = 0.5 * (A(ucc, 0, -1, 0) + A0(ucc));
    if (iW || iE) vb = 0.25 * (-A(vcc, -2, 0, 0) + 3.0 * (A(vcc, -1, 0, 0) + A0(vcc)) - A(vcc, 1, 0, 0));
    if (jS || jN) vb = 0.5 * (A(vcc, -1, 0, 0) + A0(vcc));
    // advect_v_along_y(v, vb, rdy, dy, dya, dt): zero corners on columns i == is / ie+1
    double adv_v, adv_u;
    {
      auto Q = [&](int d) { return A(v, 0, d, 0); };
      auto D = [&](int d) { return A2(dy, 0, d); };
      double cfl = vb > 0 ? vb * dt * A2(rdy, 0, -1) : vb * dt * A20(rdy);
      adv_v = adv_along(Q, D, vb, cfl, j, js, je, p.S, p.N, iW || iE, mord_mt);
    }
    {
      auto Q = [&](int d) { return A(u, d, 0, 0); };
      auto D = [&](int d) { return A2(dx, d, 0); };
      double cfl = ub > 0 ? ub * dt * A2(rdx, -1, 0) : ub * dt * A20(rdx);
      adv_u = adv_along(Q, D, ub, cfl, i, is, ie, p.W, p.E, jS || jN, mord_mt);
    }
    double kev = 0.5 * dt * (ub * adv_u + vb * adv_v);
    // all_corners_ke (d_sw.py:266-305)
    if ((iW || iE) && (jS || jN)) {
      int io1, jo1, io2;
      double vsign;
      if (iW && jS) { io1 = 0; jo1 = 0; io2 = -1; vsign = 1; }
      else if (iE && jS) { io1 = -1; jo1 = 0; io2 = 0; vsign = -1; }
      else if (iE && jN) { io1 = -1; jo1 = -1; io2 = 0; vsign = 1; }
      else { io1 = 0; jo1 = -1; io2 = -1; vsign = -1; }
      double dt6 = dt / 6.0;
      kev = dt6 * ((A0(ucc) + A(ucc, 0, -1, 0)) * ((io1 + 1) * A0(u) - (io1 * A(u, -1, 0, 0))) +
                   (A0(vcc) + A(vcc, -1, 0, 0)) * ((jo1 + 1) * A0(v) - (jo1 * A(v, 0, -1, 0))) +
                   (((jo1 + 1) * A0(ucc) - (jo1 * A(ucc, 0, -1, 0))) +
                    vsign * ((io1 + 1) * A0(vcc) - (io1 * A(vcc, -1, 0, 0)))) *
                       ((io2 + 1) * A0(u) - (io2 * A(u, -1, 0, 0))));
    }
    A0(ke) = kev;
  }, "d_sw_ke");
  // 14. compute_vorticity (compute + 3 halo)
  launch_ijk(st, g, box_dims(g, 0, 0, 0, 3, 3), [=] __device__(const Pt& p) {
    double rdy_tmp = A20(rarea) * A20(dx);
    double rdx_tmp = A20(rarea) * A20(dy);
    const double vo = (A0(u) - A(u, 0, 1, 0) * A2(dx, 0, 1) / A20(dx)) * rdy_tmp +
                      (A(v, 1, 0, 0) * A2(dy, 1, 0) / A20(dy) - A0(v)) * rdx_tmp;
    A0(vort_a) = vo;
    A0(abs_vort) = vo + A20(f0);  // 16. rel_vorticity_to_abs (same box; divergence_damping only reads vort_a)
  });
  // 15. divergence_damping
  if (divergence_damping(c, vort_b, delpc, ke, vort_a, dt)) return 1;
  // 17. fvtp2d_vt_nodelnflux(abs vorticity) -> tmp_fx, tmp_fy
  a = Fvtp2dArgs();
  a.crx = crx; a.cry = cry; a.xfx = xfx; a.yfx = yfx; a.nk = nz;
  a.q = abs_vort; a.fx = tfx; a.fy = tfy; a.hord = cf.hord_vt;
  if (fvtp2d(c, a)) return 1;
  // 18. u_and_v_from_ke.  The winds of this stage go to two work fields (us, vs) unless PACE_B200_DSW_UV=split: the
  // heat-source kernel below reads them at neighbouring points, so it can only write the final winds (23.) in the
  // same pass if its inputs live elsewhere.
  static const bool uv_split_env = [] { const char* e = getenv("PACE_B200_DSW_UV"); return e != nullptr && e[0] == 's'; }();
  const bool uv_split = uv_split_env;
  double *us = u, *vs = v;
  if (!uv_split) {
    us = c.S("dsw.u_from_ke"); vs = c.S("dsw.v_from_ke");
    if (!c.err.empty()) return 1;
  }
  launch_ijk(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    if (p.i <= ie) A0(us) = A0(u) * A20(dx) + A0(ke) - A(ke, 1, 0, 0) + A0(tfy);
    if (p.j <= je) A0(vs) = A0(v) * A20(dy) + A0(ke) - A(ke, 0, 1, 0) - A0(tfx);
  });
  // 19. delnflux_nosg_v(vorticity_agrid) -> tmp_ut, tmp_vt
  {
    DelnArgs d;
    d.q = vort_a; d.fx = tut; d.fy = tvt; d.nord_col = col_nord_v; d.damp = damp_vt_c; d.d2 = damped_vort; d.nmax = nmax_v; d.nk = nz;
    if (delnflux(c, d)) return 1;
  }
  // 20.-22. vort_differencing + heat_source_from_vorticity_damping + accumulate_heat_source_and_dissipation_estimate
  // (compute domain).  vort_x_delta / vort_y_delta are differences of the damped B-grid vorticity and only enter here:
  // they are formed in registers (same expression) instead of two scratch fields.
  const bool do_accum = cf.d_con > 1.0e-5;
  launch_ijk(st, g, uv_split ? box_dims(g, 0, 0, 0) : box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    const bool in_i = p.i <= ie, in_j = p.j <= je;
    if (in_i && in_j) {
      const double hs0 = A0(heat_s);
      double hs = hs0;
      const double frac = col_d_con[p.k];
      if (frac > 1e-5) {
        auto vxdf = [&](int dj) { return A(vort_b, 0, dj, 0) - A(vort_b, 1, dj, 0); };   // vort_x_delta(i, j+dj)
        auto vydf = [&](int di) { return A(vort_b, di, 0, 0) - A(vort_b, di, 1, 0); };   // vort_y_delta(i+di, j)
        auto ubt = [&](int dj) { return (vxdf(dj) + A(tvt, 0, dj, 0)) * A2(rdx, 0, dj); };
        auto fyf = [&](int dj) { return A(us, 0, dj, 0) * A2(rdx, 0, dj); };
        auto vbt = [&](int di) { return (vydf(di) - A(tut, di, 0, 0)) * A2(rdy, di, 0); };
        auto fxf = [&](int di) { return A(vs, di, 0, 0) * A2(rdy, di, 0); };
        double ub0 = ubt(0), ub1 = ubt(1), fy0 = fyf(0), fy1 = fyf(1);
        double vb0 = vbt(0), vb1 = vbt(1), fx0 = fxf(0), fx1 = fxf(1);
        double gy0 = fy0 * ub0, gy1 = fy1 * ub1, gx0 = fx0 * vb0, gx1 = fx1 * vb1;
        double u2 = fy0 + fy1, du2 = ub0 + ub1, v2 = fx0 + fx1, dv2 = vb0 + vb1;
        double dampterm = A20(rsin2) * 0.25 *
                          ((ub0 * ub0 + ub1 * ub1 + vb0 * vb0 + vb1 * vb1) + 2.0 * (gy0 + gy1 + gx0 + gx1) -
                           A20(cosa_s) * (u2 * dv2 + v2 * du2 + du2 * dv2));
        hs = A0(delp) * (hs - frac * dampterm);
        A0(heat_s) = hs;
      }
      if (do_accum) {
        A0(heat_source) = A0(heat_source) + hs;
        A0(diss_est) = A0(diss_est) + hs0;
      }
    }
    if (!uv_split) {
      // 23. update_u_and_v, from the winds of stage 18 in the work fields
      const bool damp = col_damp_vt[p.k] > 1e-5;
      if (in_i) A0(u) = damp ? A0(us) + A0(tvt) : A0(us);
      if (in_j) A0(v) = damp ? A0(vs) - A0(tut) : A0(vs);
    }
  });
  // 23. update_u_and_v
  if (uv_split) launch_ijk(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    if (col_damp_vt[p.k] > 1e-5) {
      if (p.i <= ie) A0(u) = A0(u) + A0(tvt);
      if (p.j <= je) A0(v) = A0(v) - A0(tut);
    }
  });
  return 0;
}

}  // namespace pace
