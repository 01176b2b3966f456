// FiniteVolumeFluxPrep (Fortran fxadv): contravariant C-grid winds with tile-edge / corner fixes, then
// Courant numbers and area fluxes.
//
// Reference: pyFV3/stencils/fxadv.py:17-506 (8 stencils), :590-691 (__call__ order).
// Launch grouping (8 reference launches -> 4):
//   K1   main_uc_vc_contra + uc_contra_y_edge + vc_contra_x_edge   (per-point, only read uc/vc)
//   K35  vc_contra_y_edge + uc_contra_x_edge on the four 2-wide edge strips (disjoint read/write sets)
//   K67  uc_contra_corners + vc_contra_corners (16 single cells; disjoint read/write sets)
//   K8   fxadv_fluxes_stencil
// The "restore utmp / vtmp" regions of the reference mean: those cells keep the value they had before
// the stencil — here they are simply not written.
#include "core.hpp"

namespace pace {

// accumulate_courant: also add the new Courant numbers to cxd / cyd (the first half of flux_capacitor, d_sw.py:34-60),
// which d_sw otherwise does in a separate pass over crx / cry
int fv_prep(Core& c, double dt, bool accumulate_courant) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const double *uc = c.F("uc"), *vc = c.F("vc");
  double *crx = c.F("ad.crx"), *cry = c.F("ad.cry"), *xfx = c.F("ad.xfx"), *yfx = c.F("ad.yfx");
  double *ucc = c.FS("dsw.uc_contra"), *vcc = c.FS("dsw.vc_contra");
  const double *cosa_u = c.F("grid.cosa_u"), *cosa_v = c.F("grid.cosa_v"), *rsin_u = c.F("grid.rsin_u"),
               *rsin_v = c.F("grid.rsin_v"), *sin_sg1 = c.F("grid.sin_sg1"), *sin_sg2 = c.F("grid.sin_sg2"),
               *sin_sg3 = c.F("grid.sin_sg3"), *sin_sg4 = c.F("grid.sin_sg4"), *rdxa = c.F("grid.rdxa"),
               *rdya = c.F("grid.rdya"), *dx = c.F("grid.dx"), *dy = c.F("grid.dy");
  if (!c.err.empty()) return 1;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, nz = g.nz;

  // K1 on the full domain
  launch_ijk(st, g, box_full(g), [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    const bool any_edge = p.W || p.E || p.S || p.N;  // edge stencils only run when not tile-interior
    if (i >= is - 1 && i <= ie + 2) {
      bool keep = (p.S && (j == js - 1 || j == js)) || (p.N && (j == je || j == je + 1));
      if (any_edge && ((p.W && i == is) || (p.E && i == ie + 1))) {
        // uc_contra_y_edge (all j, overrides both the main value and the kept rows)
        double u = A0(uc);
        A0(ucc) = u > 0 ? u / A2(sin_sg3, -1, 0) : u / A20(sin_sg1);
      } else if (!keep) {
        double v = 0.25 * (A(vc, -1, 0, 0) + A0(vc) + A(vc, -1, 1, 0) + A(vc, 0, 1, 0));
        A0(ucc) = (A0(uc) - v * A20(cosa_u)) * A20(rsin_u);
      }
    } else if (any_edge && ((p.W && i == is) || (p.E && i == ie + 1))) {
      double u = A0(uc);
      A0(ucc) = u > 0 ? u / A2(sin_sg3, -1, 0) : u / A20(sin_sg1);
    }
    if (any_edge && ((p.S && j == js) || (p.N && j == je + 1))) {
      // vc_contra_x_edge (all i)
      double v = A0(vc);
      A0(vcc) = v > 0 ? v / A2(sin_sg4, 0, -1) : v / A20(sin_sg2);
    } else if (j >= js - 1 && j <= je + 2) {
      double u = 0.25 * (A(uc, 0, -1, 0) + A(uc, 1, -1, 0) + A0(uc) + A(uc, 1, 0, 0));
      A0(vcc) = (A0(vc) - u * A20(cosa_v)) * A20(rsin_v);
    }
  });

  // K35: edge strips.  strip 0/1: columns {is-1,is} (W) / {ie,ie+1} (E), rows [js..je+1];
  //                    strip 2/3: rows {js-1,js} (S) / {je,je+1} (N), columns [is..ie+1]
  const int nlong = (g.nx > g.ny ? g.nx : g.ny) + 2;
  launch_ijk(st, g, Box{0, 0, 0, nlong, 8, nz}, [=] __device__(const Pt& p0) {
    const int s = p0.j >> 1, w = p0.j & 1, m = p0.i;
    if (s < 2) {  // vc_contra_y_edge
      if (!(s == 0 ? p0.W : p0.E) || m >= g.ny + 1) return;
      const int i = (s == 0 ? is - 1 : ie) + w, j = js + m;
      if ((p0.S && (j == js || j == js + 1)) || (p0.N && (j == je || j == je + 1))) return;  // vtmp restore
      Pt p = make_pt(g, p0.t, i, j, p0.k);
      double u_contra = 0.25 * (A(ucc, 0, -1, 0) + A(ucc, 1, -1, 0) + A0(ucc) + A(ucc, 1, 0, 0));
      A0(vcc) = (A0(vc) - u_contra * A20(cosa_v)) * 1.0;
    } else {  // uc_contra_x_edge
      if (!(s == 2 ? p0.S : p0.N) || m >= g.nx + 1) return;
      const int j = (s == 2 ? js - 1 : je) + w, i = is + m;
      if ((p0.W && (i == is || i == is + 1)) || (p0.E && (i == ie || i == ie + 1))) return;  // utmp restore
      Pt p = make_pt(g, p0.t, i, j, p0.k);
      double v_contra = 0.25 * (A(vcc, -1, 0, 0) + A0(vcc) + A(vcc, -1, 1, 0) + A(vcc, 0, 1, 0));
      A0(ucc) = (A0(uc) - v_contra * A20(cosa_u)) * 1.0;
    }
  });

  // K67: corner cells
  launch_ijk(st, g, Box{0, 0, 0, 8, 1, nz}, [=] __device__(const Pt& p0) {
    const int id = p0.i;
    const bool W = p0.W, E = p0.E, S = p0.S, N = p0.N;
    if (id < 8) {  // uc_contra_corners (fxadv.py:165-266): blocks b=0..3, each at a south and a north row
      const int b = id >> 1, north = id & 1;
      int i, j;
      bool on;
      switch (b) {
        case 0: i = is + 1; j = north ? je : js - 1; on = W; break;
        case 1: i = is + 1; j = north ? je + 1 : js; on = W; break;
        case 2: i = ie; j = north ? je : js - 1; on = E; break;
        default: i = ie; j = north ? je + 1 : js; on = E; break;
      }
      on = on && (north ? N : S);
      if (!on) return;
      Pt p = make_pt(g, p0.t, i, j, p0.k);
      double r;
      if (b == 0) {
        double damp = 1.0 / (1.0 - 0.0625 * A20(cosa_u) * A2(cosa_v, -1, 0));
        r = (A0(uc) - 0.25 * A20(cosa_u) *
                          (A(vcc, -1, 1, 0) + A(vcc, 0, 1, 0) + A0(vcc) + A(vc, -1, 0, 0) -
                           0.25 * A2(cosa_v, -1, 0) * (A(ucc, -1, 0, 0) + A(ucc, -1, -1, 0) + A(ucc, 0, -1, 0)))) *
            damp;
      } else if (b == 1) {
        double damp = 1.0 / (1.0 - 0.0625 * A20(cosa_u) * A2(cosa_v, -1, 1));
        r = (A0(uc) - 0.25 * A20(cosa_u) *
                          (A(vcc, -1, 0, 0) + A0(vcc) + A(vcc, 0, 1, 0) + A(vc, -1, 1, 0) -
                           0.25 * A2(cosa_v, -1, 1) * (A(ucc, -1, 0, 0) + A(ucc, -1, 1, 0) + A(ucc, 0, 1, 0)))) *
            damp;
      } else if (b == 2) {
        double damp = 1.0 / (1.0 - 0.0625 * A20(cosa_u) * A20(cosa_v));
        r = (A0(uc) - 0.25 * A20(cosa_u) *
                          (A(vcc, 0, 1, 0) + A(vcc, -1, 1, 0) + A(vcc, -1, 0, 0) + A0(vc) -
                           0.25 * A20(cosa_v) * (A(ucc, 1, 0, 0) + A(ucc, 1, -1, 0) + A(ucc, 0, -1, 0)))) *
            damp;
      } else {
        double damp = 1.0 / (1.0 - 0.0625 * A20(cosa_u) * A2(cosa_v, 0, 1));
        r = (A0(uc) - 0.25 * A20(cosa_u) *
                          (A0(vcc) + A(vcc, -1, 0, 0) + A(vcc, -1, 1, 0) + A(vc, 0, 1, 0) -
                           0.25 * A2(cosa_v, 0, 1) * (A(ucc, 1, 0, 0) + A(ucc, 1, 1, 0) + A(ucc, 0, 1, 0)))) *
            damp;
      }
      A0(ucc) = r;
    }
  });
  launch_ijk(st, g, Box{0, 0, 0, 8, 1, nz}, [=] __device__(const Pt& p0) {
    // vc_contra_corners (fxadv.py:269-370), ut = uc_contra after uc_contra_corners
    const int id = p0.i, b = id >> 1, second = id & 1;
    const bool W = p0.W, E = p0.E, S = p0.S, N = p0.N;
    int i, j;
    bool on;
    switch (b) {
      case 0: j = js + 1; i = second ? ie : is - 1; on = S && (second ? E : W); break;
      case 1: j = js + 1; i = second ? ie + 1 : is; on = S && (second ? E : W); break;
      case 2: j = je; i = second ? is : ie + 1; on = N && (second ? W : E); break;
      default: j = je; i = second ? is - 1 : ie; on = N && (second ? W : E); break;
    }
    if (!on) return;
    Pt p = make_pt(g, p0.t, i, j, p0.k);
    double r;
    if (b == 0) {
      double damp = 1.0 / (1.0 - 0.0625 * A2(cosa_u, 0, -1) * A20(cosa_v));
      r = (A0(vc) - 0.25 * A20(cosa_v) *
                        (A(ucc, 1, -1, 0) + A(ucc, 1, 0, 0) + A0(ucc) + A(uc, 0, -1, 0) -
                         0.25 * A2(cosa_u, 0, -1) * (A(vcc, 0, -1, 0) + A(vcc, -1, -1, 0) + A(vcc, -1, 0, 0)))) *
          damp;
    } else if (b == 1) {
      double damp = 1.0 / (1.0 - 0.0625 * A2(cosa_u, 1, -1) * A20(cosa_v));
      r = (A0(vc) - 0.25 * A20(cosa_v) *
                        (A(ucc, 0, -1, 0) + A0(ucc) + A(ucc, 1, 0, 0) + A(uc, 1, -1, 0) -
                         0.25 * A2(cosa_u, 1, -1) * (A(vcc, 0, -1, 0) + A(vcc, 1, -1, 0) + A(vcc, 1, 0, 0)))) *
          damp;
    } else if (b == 2) {
      double damp = 1.0 / (1.0 - 0.0625 * A2(cosa_u, 1, 0) * A20(cosa_v));
      r = (A0(vc) - 0.25 * A20(cosa_v) *
                        (A0(ucc) + A(ucc, 0, -1, 0) + A(ucc, 1, -1, 0) + A(uc, 1, 0, 0) -
                         0.25 * A2(cosa_u, 1, 0) * (A(vcc, 0, 1, 0) + A(vcc, 1, 1, 0) + A(vcc, 1, 0, 0)))) *
          damp;
    } else {
      double damp = 1.0 / (1.0 - 0.0625 * A20(cosa_u) * A20(cosa_v));
      r = (A0(vc) - 0.25 * A20(cosa_v) *
                        (A(ucc, 1, 0, 0) + A(ucc, 1, -1, 0) + A(ucc, 0, -1, 0) + A0(uc) -
                         0.25 * A20(cosa_u) * (A(vcc, 0, 1, 0) + A(vcc, -1, 1, 0) + A(vcc, -1, 0, 0)))) *
          damp;
    }
    A0(vcc) = r;
  });

  // K8: fxadv_fluxes_stencil (fxadv.py:373-408) on the full domain
  double *cx = nullptr, *cy = nullptr;
  if (accumulate_courant) {
    cx = c.F("cxd"); cy = c.F("cyd");
    if (!c.err.empty()) return 1;
  }
  const int icap = g.nx + 5, jcap = g.ny + 5;  // flux_capacitor runs on the compute domain + 3 halo cells
  launch_ijk(st, g, box_full(g), [=] __device__(const Pt& p) {
    const bool cap = cx != nullptr && p.i <= icap && p.j <= jcap;
    if (p.i >= is && p.i <= ie + 1) {
      double u = A0(ucc);
      double cr;
      if (u > 0) {
        cr = dt * u * A2(rdxa, -1, 0);
        A0(xfx) = A20(dy) * dt * u * A2(sin_sg3, -1, 0);
      } else {
        cr = dt * u * A20(rdxa);
        A0(xfx) = A20(dy) * dt * u * A20(sin_sg1);
      }
      A0(crx) = cr;
      if (cap) A0(cx) = A0(cx) + cr;
    }
    if (p.j >= js && p.j <= je + 1) {
      double v = A0(vcc);
      double cr;
      if (v > 0) {
        cr = dt * v * A2(rdya, 0, -1);
        A0(yfx) = A20(dx) * dt * v * A2(sin_sg4, 0, -1);
      } else {
        cr = dt * v * A20(rdya);
        A0(yfx) = A20(dx) * dt * v * A20(sin_sg2);
      }
      A0(cry) = cr;
      if (cap) A0(cy) = A0(cy) + cr;
    }
  });
  return 0;
}

}  // namespace pace
