// C-grid shallow-water half step (Fortran c_sw) incl. the D->A->C wind interpolation.
//
// Reference: pyFV3/stencils/c_sw.py:21-500 (stencils), :644-807 (__call__ order);
//            pyFV3/stencils/d2a2c_vect.py:20-377 (stencils/functions), :588-699 (__call__ order).
// The reference runs 23 stencil launches here; the same statements are grouped into 9 kernels:
//   A  utmp/vtmp (set_tmps + lagrange_*_p1 + avg_box) and ua/va (contravariant_components)
//   c1 corner fills of utmp, ua (x) and vtmp, va (y)
//   B  uc, ut (ut_main + east_west_edges + geoadjust_ut) and vc, vt (north_south_edges + vt_main +
//      geoadjust_vt)
//   C  divergence_corner
//   c2 fill_corners_2cells_x(delp, pt, w);  D  compute_nonhydrostatic_fluxes_x;  c3 ..._y
//   E  transportdelp_update_vorticity_and_kineticenergy
//   F  circulation_cgrid + absolute_vorticity;  G  update_y_velocity + update_x_velocity
// Per element every expression tree is the reference's (see Appendix B of SURVEY.md).
#include "core.hpp"
#include "corners.cuh"

namespace pace {

namespace {
constexpr double C1 = -2.0 / 14.0, C2 = 11.0 / 14.0, C3 = 5.0 / 14.0;  // d2a2c_vect.py:14-16
constexpr double A1 = 9.0 / 16.0, A2C = -1.0 / 16.0;                   // a2b_ord4.py:32-33
constexpr double BIG_NUMBER = 1e30;                                    // d2a2c_vect.py:414
}  // namespace

int c_sw_unfused(Core& c, double dt2) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  if (c.cfg.grid_type != 0) {
    c.err = "unimplemented grid_type (only 0)";  // d2a2c_vect.py:394 allows 0 and 4
    return 1;
  }
  double *delp = c.F("delp"), *pt = c.F("pt"), *u = c.F("u"), *v = c.F("v"), *w = c.F("w");
  double *uc = c.F("uc"), *vc = c.F("vc"), *ua = c.F("ua"), *va = c.F("va");
  double *ut = c.F("ad.ut"), *vt = c.F("ad.vt"), *divgd = c.F("ad.divgd"), *omga = c.F("omga");
  double *delpc = c.F("csw.delpc"), *ptc = c.F("csw.ptc");
  const double *cosa_s = c.F("grid.cosa_s"), *rsin2 = c.F("grid.rsin2"), *cosa_u = c.F("grid.cosa_u"),
               *cosa_v = c.F("grid.cosa_v"), *rsin_u = c.F("grid.rsin_u"), *rsin_v = c.F("grid.rsin_v"),
               *sina_u = c.F("grid.sina_u"), *sina_v = c.F("grid.sina_v"), *dxa = c.F("grid.dxa"),
               *dya = c.F("grid.dya"), *dx = c.F("grid.dx"), *dy = c.F("grid.dy"), *dxc = c.F("grid.dxc"),
               *dyc = c.F("grid.dyc"), *rdxc = c.F("grid.rdxc"), *rdyc = c.F("grid.rdyc"),
               *sin_sg1 = c.F("grid.sin_sg1"), *sin_sg2 = c.F("grid.sin_sg2"), *sin_sg3 = c.F("grid.sin_sg3"),
               *sin_sg4 = c.F("grid.sin_sg4"), *cos_sg1 = c.F("grid.cos_sg1"), *cos_sg2 = c.F("grid.cos_sg2"),
               *cos_sg3 = c.F("grid.cos_sg3"), *cos_sg4 = c.F("grid.cos_sg4"), *rarea = c.F("grid.rarea"),
               *rarea_c = c.F("grid.rarea_c"), *fC = c.F("grid.fC");
  double *utmp = c.S("csw.utmp"), *vtmp = c.S("csw.vtmp"), *ke = c.S("csw.ke"), *vort = c.S("csw.vort"),
         *fx = c.S("csw.fx"), *fx1 = c.S("csw.fx1"), *fx2 = c.S("csw.fx2");
  if (!c.err.empty()) return 1;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;
  const int nord = c.cfg.nord;

  // zero_delpc_ptc (c_sw.py:21-30): compute domain + 3 halo
  launch_ijk(st, g, box_dims(g, 0, 0, 0, 3, 3), [=] __device__(const Pt& p) {
    A0(delpc) = 0.0;
    A0(ptc) = 0.0;
  });

  // ---- A: utmp, vtmp, ua, va (d2a2c_vect.py:28-95, 421-540) ---------------------------------------
  int npt = 4;
  if (npt > g.nx - 1 || npt > g.ny - 1) npt = 0;
  const int avg_off = (npt == 0) ? -1 : 3;
  const int npt_ = npt;
  launch_ijk(st, g, box_dims(g, 0, 0, 0, 3, 3), [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    // set_tmps
    double ut_ = BIG_NUMBER, vt_ = BIG_NUMBER;
    // lagrange_interpolation_y_p1 on [is1..ie1] x [js1..je1]
    int js1 = p.S ? npt_ + 2 : js - 1, je1 = p.N ? (je + 1) - npt_ : je + 1;
    int is1 = p.W ? npt_ + 2 : 0, ie1 = p.E ? (ie + 1) - npt_ : ie + 3;
    if (i >= is1 && i <= ie1 && j >= js1 && j <= je1)
      ut_ = A2C * (A(u, 0, -1, 0) + A(u, 0, 2, 0)) + A1 * (A0(u) + A(u, 0, 1, 0));
    // lagrange_interpolation_x_p1 on [is2..ie2] x [js2..je2]
    int is2 = p.W ? npt_ + 2 : is - 1, ie2 = p.E ? (ie + 1) - npt_ : ie + 1;
    int js2 = p.S ? npt_ + 2 : 0, je2 = p.N ? (je + 1) - npt_ : je + 3;
    if (i >= is2 && i <= ie2 && j >= js2 && j <= je2)
      vt_ = A2C * (A(v, -1, 0, 0) + A(v, 2, 0, 0)) + A1 * (A0(v) + A(v, 1, 0, 0));
    // avg_box
    bool edge = (p.S && j < js + avg_off) || (p.N && j > je - avg_off) || (p.W && i < is + avg_off) ||
                (p.E && i > ie - avg_off);
    if (edge) {
      ut_ = 0.5 * (A0(u) + A(u, 0, 1, 0));
      vt_ = 0.5 * (A0(v) + A(v, 1, 0, 0));
    }
    A0(utmp) = ut_;
    A0(vtmp) = vt_;
    // contravariant_components on compute(-2,-2)+(4,4)
    if (i >= is - 2 && i <= ie + 2 && j >= js - 2 && j <= je + 2) {
      A0(ua) = (ut_ - vt_ * A20(cosa_s)) * A20(rsin2);
      A0(va) = (vt_ - ut_ * A20(cosa_s)) * A20(rsin2);
    }
  });

  // ---- c1: fill_corners_x (utmp<-vtmp 3 cells, ua<-va 2 cells), fill_corners_y (vtmp<-utmp, va<-ua)
  {
    CornerOps ops;
    ops.n = 4; ops.nk = g.nz;
    ops.op[0] = CornerOp{utmp, vtmp, 3, 0, -1.0, 1.0, 1.0, -1.0};
    ops.op[1] = CornerOp{ua, va, 2, 0, -1.0, 1.0, 1.0, -1.0};
    ops.op[2] = CornerOp{vtmp, utmp, 3, 1, -1.0, 1.0, 1.0, -1.0};
    ops.op[3] = CornerOp{va, ua, 2, 1, -1.0, 1.0, 1.0, -1.0};
    run_corner_ops(st, g, ops);
  }

  // ---- B: uc/ut on [is-1..ie+2]x[js-1..je+1], vc/vt on [is-1..ie+1]x[js-1..je+2] -----------------
  launch_ijk(st, g, Box{is - 1, js - 1, 0, g.nx + 3, g.ny + 3, g.nz}, [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    if (j <= je + 1) {
      double ucv, utv;
      auto contra = [&](double ucx) { return (ucx - A0(v) * A20(cosa_u)) * A20(rsin_u); };
      auto edge4 = [&]() {  // edge_interpolate4_x (d2a2c_vect.py:360-366)
        double t1 = A2(dxa, -2, 0) + A2(dxa, -1, 0);
        double t2 = A2(dxa, 0, 0) + A2(dxa, 1, 0);
        double n1 = (t1 + A2(dxa, -1, 0)) * A(ua, -1, 0, 0) - A2(dxa, -1, 0) * A(ua, -2, 0, 0);
        double n2 = (t1 + A2(dxa, 0, 0)) * A(ua, 0, 0, 0) - A2(dxa, 0, 0) * A(ua, 1, 0, 0);
        return 0.5 * (n1 / t1 + n2 / t2);
      };
      if ((p.W && i == is - 1) || (p.E && i == ie)) {
        ucv = C1 * A(utmp, -2, 0, 0) + C2 * A(utmp, -1, 0, 0) + C3 * A0(utmp);
        utv = contra(ucv);
      } else if ((p.W && i == is) || (p.E && i == ie + 1)) {
        utv = edge4();
        ucv = utv > 0 ? utv * A2(sin_sg3, -1, 0) : utv * A20(sin_sg1);
      } else if ((p.W && i == is + 1) || (p.E && i == ie + 2)) {
        ucv = C1 * A(utmp, 1, 0, 0) + C2 * A0(utmp) + C3 * A(utmp, -1, 0, 0);
        utv = contra(ucv);
      } else {  // ut_main
        ucv = A2C * (A(utmp, -2, 0, 0) + A(utmp, 1, 0, 0)) + A1 * (A(utmp, -1, 0, 0) + A0(utmp));
        utv = contra(ucv);
      }
      A0(uc) = ucv;
      // geoadjust_ut (c_sw.py:178-200)
      A0(ut) = utv > 0 ? dt2 * utv * A20(dy) * A2(sin_sg3, -1, 0) : dt2 * utv * A20(dy) * A20(sin_sg1);
    }
    if (i <= ie + 1) {
      double vcv, vtv;
      auto contra = [&](double vcx) { return (vcx - A0(u) * A20(cosa_v)) * A20(rsin_v); };
      auto edge4 = [&]() {  // edge_interpolate4_y
        double t1 = A2(dya, 0, -2) + A2(dya, 0, -1);
        double t2 = A2(dya, 0, 0) + A2(dya, 0, 1);
        double n1 = (t1 + A2(dya, 0, -1)) * A(va, 0, -1, 0) - A2(dya, 0, -1) * A(va, 0, -2, 0);
        double n2 = (t1 + A2(dya, 0, 0)) * A(va, 0, 0, 0) - A2(dya, 0, 0) * A(va, 0, 1, 0);
        return 0.5 * (n1 / t1 + n2 / t2);
      };
      if ((p.S && j == js - 1) || (p.N && j == je)) {
        vcv = C1 * A(vtmp, 0, -2, 0) + C2 * A(vtmp, 0, -1, 0) + C3 * A0(vtmp);
        vtv = contra(vcv);
      } else if ((p.S && j == js) || (p.N && j == je + 1)) {
        vtv = edge4();
        vcv = vtv > 0 ? vtv * A2(sin_sg4, 0, -1) : vtv * A20(sin_sg2);
      } else if ((p.S && j == js + 1) || (p.N && j == je + 2)) {
        vcv = C1 * A(vtmp, 0, 1, 0) + C2 * A0(vtmp) + C3 * A(vtmp, 0, -1, 0);
        vtv = contra(vcv);
      } else {  // vt_main / first region of north_south_edges
        vcv = A2C * (A(vtmp, 0, -2, 0) + A(vtmp, 0, 1, 0)) + A1 * (A(vtmp, 0, -1, 0) + A0(vtmp));
        vtv = contra(vcv);
      }
      A0(vc) = vcv;
      // geoadjust_vt (c_sw.py:203-222)
      A0(vt) = vtv > 0 ? dt2 * vtv * A20(dx) * A2(sin_sg4, 0, -1) : dt2 * vtv * A20(dx) * A20(sin_sg2);
    }
  });

  // ---- C: divergence_corner (c_sw.py:33-175) on the corner (interface x interface) compute domain
  if (nord > 0) {
    launch_ijk(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
      const int i = p.i, j = p.j;
      // uf/vf at offsets (di,dj) from this point, general and edge forms
      auto uf_gen = [&](int di, int dj) {
        return (A(u, di, dj, 0) - 0.25 * (A(va, di, dj - 1, 0) + A(va, di, dj, 0)) *
                                      (A2(cos_sg4, di, dj - 1) + A2(cos_sg2, di, dj))) *
               A2(dyc, di, dj) * 0.5 * (A2(sin_sg4, di, dj - 1) + A2(sin_sg2, di, dj));
      };
      auto vf_gen = [&](int di, int dj) {
        return (A(v, di, dj, 0) - 0.25 * (A(ua, di - 1, dj, 0) + A(ua, di, dj, 0)) *
                                      (A2(cos_sg3, di - 1, dj) + A2(cos_sg1, di, dj))) *
               A2(dxc, di, dj) * 0.5 * (A2(sin_sg3, di - 1, dj) + A2(sin_sg1, di, dj));
      };
      auto uf_edge = [&](int di, int dj) {
        return A(u, di, dj, 0) * A2(dyc, di, dj) * 0.5 * (A2(sin_sg4, di, dj - 1) + A2(sin_sg2, di, dj));
      };
      auto vf_edge = [&](int di, int dj) {
        return A(v, di, dj, 0) * A2(dxc, di, dj) * 0.5 * (A2(sin_sg3, di - 1, dj) + A2(sin_sg1, di, dj));
      };
      const bool iw = p.W && i == is, iE = p.E && i == ie + 1, jS = p.S && j == js, jN = p.N && j == je + 1;
      double d;
      // later region blocks overwrite earlier ones: corners > j-edges > i-edges > general
      if ((iw || iE) && jS) {
        d = (-vf_edge(0, 0) + uf_edge(-1, 0) - uf_edge(0, 0)) * A20(rarea_c);
      } else if ((iw || iE) && jN) {
        d = (vf_edge(0, -1) + uf_edge(-1, 0) - uf_edge(0, 0)) * A20(rarea_c);
      } else if (jS || jN) {
        d = (vf_gen(0, -1) - vf_gen(0, 0) + uf_edge(-1, 0) - uf_edge(0, 0)) * A20(rarea_c);
      } else if (iw || iE) {
        d = (vf_edge(0, -1) - vf_edge(0, 0) + uf_gen(-1, 0) - uf_gen(0, 0)) * A20(rarea_c);
      } else {
        d = (vf_gen(0, -1) - vf_gen(0, 0) + uf_gen(-1, 0) - uf_gen(0, 0)) * A20(rarea_c);
      }
      A0(divgd) = d;
    });
  }

  // ---- c2, D: fill_corners_2cells_x(delp,pt,w); compute_nonhydrostatic_fluxes_x (c_sw.py:252-284)
  {
    CornerOps ops;
    ops.n = 3; ops.nk = g.nz;
    ops.op[0] = CornerOp{delp, delp, 2, 0, 1.0, 1.0, 1.0, 1.0};
    ops.op[1] = CornerOp{pt, pt, 2, 0, 1.0, 1.0, 1.0, 1.0};
    ops.op[2] = CornerOp{w, w, 2, 0, 1.0, 1.0, 1.0, 1.0};
    run_corner_ops(st, g, ops);
  }
  launch_ijk(st, g, box_dims(g, 1, 0, 0, 1, 1), [=] __device__(const Pt& p) {
    double utc = A0(ut);
    double f1 = utc > 0.0 ? A(delp, -1, 0, 0) : A0(delp);
    double f = utc > 0.0 ? A(pt, -1, 0, 0) : A0(pt);
    double f2 = utc > 0.0 ? A(w, -1, 0, 0) : A0(w);
    f1 = utc * f1;
    A0(fx1) = f1;
    A0(fx) = f1 * f;
    A0(fx2) = f1 * f2;
  });
  {
    CornerOps ops;
    ops.n = 3; ops.nk = g.nz;
    ops.op[0] = CornerOp{delp, delp, 2, 1, 1.0, 1.0, 1.0, 1.0};
    ops.op[1] = CornerOp{pt, pt, 2, 1, 1.0, 1.0, 1.0, 1.0};
    ops.op[2] = CornerOp{w, w, 2, 1, 1.0, 1.0, 1.0, 1.0};
    run_corner_ops(st, g, ops);
  }

  // ---- E: transportdelp_update_vorticity_and_kineticenergy (c_sw.py:287-391), compute + 1 halo
  launch_ijk_occ<6>(st, g, box_dims(g, 0, 0, 0, 1, 1), [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    double fy1_[2], fy_[2], fy2_[2];
    for (int dj = 0; dj < 2; ++dj) {
      double vtc = A(vt, 0, dj, 0);
      double f1 = vtc > 0.0 ? A(delp, 0, dj - 1, 0) : A(delp, 0, dj, 0);
      double f = vtc > 0.0 ? A(pt, 0, dj - 1, 0) : A(pt, 0, dj, 0);
      double f2 = vtc > 0.0 ? A(w, 0, dj - 1, 0) : A(w, 0, dj, 0);
      f1 = vtc * f1;
      fy1_[dj] = f1;
      fy_[dj] = f1 * f;
      fy2_[dj] = f1 * f2;
    }
    double dpc = A0(delp) + (A0(fx1) - A(fx1, 1, 0, 0) + fy1_[0] - fy1_[1]) * A20(rarea);
    A0(delpc) = dpc;
    A0(ptc) = (A0(pt) * A0(delp) + (A0(fx) - A(fx, 1, 0, 0) + fy_[0] - fy_[1]) * A20(rarea)) / dpc;
    A0(omga) = (A0(w) * A0(delp) + (A0(fx2) - A(fx2, 1, 0, 0) + fy2_[0] - fy2_[1]) * A20(rarea)) / dpc;
    // vorticity / kinetic energy with upstream bias
    double uav = A0(ua), vav = A0(va);
    double kev = uav > 0.0 ? A0(uc) : A(uc, 1, 0, 0);
    double vo = vav > 0.0 ? A0(vc) : A(vc, 0, 1, 0);
    if ((p.S && j == js - 1) || (p.N && j == je))
      vo = (vav <= 0.0) ? vo * A20(sin_sg4) + A(u, 0, 1, 0) * A20(cos_sg4) : vo;
    if ((p.S && j == js) || (p.N && j == je + 1))
      vo = (vav > 0.0) ? vo * A20(sin_sg2) + A0(u) * A20(cos_sg2) : vo;
    if ((p.E && i == ie) || (p.W && i == is - 1))
      kev = (uav <= 0.0) ? kev * A20(sin_sg3) + A(v, 1, 0, 0) * A20(cos_sg3) : kev;
    if ((p.E && i == ie + 1) || (p.W && i == is))
      kev = (uav > 0.0) ? kev * A20(sin_sg1) + A0(v) * A20(cos_sg1) : kev;
    A0(ke) = 0.5 * dt2 * (uav * kev + vav * vo);
  });

  // ---- F: circulation_cgrid + absolute_vorticity (c_sw.py:394-436) on corners
  launch_ijk(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    double fxv = A20(dxc) * A0(uc);
    double fyv = A20(dyc) * A0(vc);
    double fx1v = A2(dxc, 0, -1) * A(uc, 0, -1, 0);
    double fy1v = A2(dyc, -1, 0) * A(vc, -1, 0, 0);
    double vc_ = fx1v - fxv - fy1v + fyv;
    if (p.W && i == is && ((p.S && j == js) || (p.N && j == je + 1))) vc_ = fx1v - fxv + fyv;
    if (p.E && i == ie + 1 && ((p.S && j == js) || (p.N && j == je + 1))) vc_ = fx1v - fxv - fy1v;
    A0(vort) = A20(fC) + A20(rarea_c) * vc_;
  });

  // ---- G: update_y_velocity on [X, Y_INTERFACE], update_x_velocity on [X_INTERFACE, Y]
  launch_ijk(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    if (i <= ie) {  // update_y_velocity(vorticity, ke, velocity=u, velocity_c=vc, cosa_v, sina_v, rdyc)
      double tmp = dt2 * (A0(u) - A0(vc) * A20(cosa_v)) / A20(sina_v);
      if ((p.S && j == js) || (p.N && j == je + 1)) tmp = dt2 * A0(u);
      double flux = tmp > 0.0 ? A0(vort) : A(vort, 1, 0, 0);
      A0(vc) = A0(vc) - tmp * flux + A20(rdyc) * (A(ke, 0, -1, 0) - A0(ke));
    }
    if (j <= je) {  // update_x_velocity(vorticity, ke, velocity=v, velocity_c=uc, cosa_u, sina_u, rdxc)
      double tmp = dt2 * (A0(v) - A0(uc) * A20(cosa_u)) / A20(sina_u);
      if ((p.W && i == is) || (p.E && i == ie + 1)) tmp = dt2 * A0(v);
      double flux = tmp > 0.0 ? A0(vort) : A(vort, 0, 1, 0);
      A0(uc) = A0(uc) + tmp * flux + A20(rdxc) * (A(ke, -1, 0, 0) - A0(ke));
    }
  });
  return 0;
}

}  // namespace pace
