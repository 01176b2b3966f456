// A-grid -> B-grid (cell centres -> cell corners) 4th-order interpolation.
//
// Reference: pyFV3/stencils/a2b_ord4.py:36-526 (stencils/functions), :529-808 (AGrid2BGridFourthOrder).
// The reference launches 11 stencils; here 2 kernels:
//   K1  qx, qy (ppm_volume_mean_x/y) + the tile-corner extrapolations + the tile-edge values of qout
//   K2  a2b_interpolation on the points that are not on a tile edge
// The corner extrapolation weights x1/(x2-x1) of extrap_corner depend only on the grid (great-circle
// distances, a2b_ord4.py:36-59); they are evaluated once on the host with the same libm calls the
// reference makes and passed in as "grid.a2b_corner_coef"[t][corner(SW,SE,NE,NW)][3].
#include "core.hpp"

namespace pace {

namespace {
constexpr double c1 = 2.0 / 3.0, c2 = -1.0 / 6.0;
constexpr double b1 = 7.0 / 12.0, b2 = -1.0 / 12.0;
constexpr double a1 = 9.0 / 16.0, a2 = -1.0 / 16.0;
}  // namespace

int a2b_ord4_fused_launch(Core& c, const double* qin, double* qout, int k0, int nk);  // a2b_ord4_fused.cu

int a2b_ord4(Core& c, const double* qin_c, double* qout, int k0, int nk, bool replace) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
#ifndef PACE_HOST_EMU
  // production path: one shared-memory tiled kernel; PACE_B200_A2B=unfused keeps the two-kernel version (A/B runs)
  static const bool fused = [] { const char* v = getenv("PACE_B200_A2B"); return !(v && std::string(v) == "unfused"); }();
  if (fused) {
    if (a2b_ord4_fused_launch(c, qin_c, qout, k0, nk)) return 1;
    if (replace) {
      double* qin_w = const_cast<double*>(qin_c);
      launch_ijk(st, g, Box{g.is, g.js, k0, g.nx + 1, g.ny + 1, nk}, [=] __device__(const Pt& p) { A0(qin_w) = A0(qout); });
    }
    return 0;
  }
#endif
  const double *dxa = c.F("grid.dxa"), *dya = c.F("grid.dya"), *edge_w = c.F("grid.edge_w"),
               *edge_e = c.F("grid.edge_e"), *edge_s = c.F("grid.edge_s"), *edge_n = c.F("grid.edge_n"),
               *coef = c.F("grid.a2b_corner_coef");
  double *qx = c.S("a2b.qx"), *qy = c.S("a2b.qy");
  if (!c.err.empty()) return 1;
  const double* qin = qin_c;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;

  // K1 over [is-2..ie+2]x[js-2..je+2]
  launch_ijk(st, g, Box{is - 2, js - 2, k0, g.nx + 5, g.ny + 5, nk}, [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    // ppm_volume_mean_x on [is..ie+1]x[js-2..je+2]
    if (i >= is && i <= ie + 1) {
      double v;
      if (p.W && i == is) {  // qx_edge_west
        double g_in = A2(dxa, 1, 0) / A20(dxa);
        double g_ou = A2(dxa, -2, 0) / A2(dxa, -1, 0);
        v = 0.5 * (((2.0 + g_in) * A0(qin) - A(qin, 1, 0, 0)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * A(qin, -1, 0, 0) - A(qin, -2, 0, 0)) / (1.0 + g_ou));
      } else if (p.W && i == is + 1) {  // qx_edge_west2
        double g_in = A20(dxa) / A2(dxa, -1, 0);
        double g_ou = A2(dxa, -3, 0) / A2(dxa, -2, 0);
        double qxleft = 0.5 * (((2.0 + g_in) * A(qin, -1, 0, 0) - A0(qin)) / (1.0 + g_in) +
                               ((2.0 + g_ou) * A(qin, -2, 0, 0) - A(qin, -3, 0, 0)) / (1.0 + g_ou));
        double qxright = b2 * (A(qin, -1, 0, 0) + A(qin, 2, 0, 0)) + b1 * (A0(qin) + A(qin, 1, 0, 0));
        v = (3.0 * (g_in * A(qin, -1, 0, 0) + A0(qin)) - (g_in * qxleft + qxright)) / (2.0 + 2.0 * g_in);
      } else if (p.E && i == ie + 1) {  // qx_edge_east
        double g_in = A2(dxa, -2, 0) / A2(dxa, -1, 0);
        double g_ou = A2(dxa, 1, 0) / A20(dxa);
        v = 0.5 * (((2.0 + g_in) * A(qin, -1, 0, 0) - A(qin, -2, 0, 0)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * A0(qin) - A(qin, 1, 0, 0)) / (1.0 + g_ou));
      } else if (p.E && i == ie) {  // qx_edge_east2
        double g_in = A2(dxa, -1, 0) / A20(dxa);
        double g_ou = A2(dxa, 2, 0) / A2(dxa, 1, 0);
        double qxright = 0.5 * (((2.0 + g_in) * A0(qin) - A(qin, -1, 0, 0)) / (1.0 + g_in) +
                                ((2.0 + g_ou) * A(qin, 1, 0, 0) - A(qin, 2, 0, 0)) / (1.0 + g_ou));
        double qxleft = b2 * (A(qin, -3, 0, 0) + A0(qin)) + b1 * (A(qin, -2, 0, 0) + A(qin, -1, 0, 0));
        v = (3.0 * (A(qin, -1, 0, 0) + g_in * A0(qin)) - (g_in * qxright + qxleft)) / (2.0 + 2.0 * g_in);
      } else {
        v = b2 * (A(qin, -2, 0, 0) + A(qin, 1, 0, 0)) + b1 * (A(qin, -1, 0, 0) + A0(qin));
      }
      A0(qx) = v;
    }
    // ppm_volume_mean_y on [is-2..ie+2]x[js..je+1]
    if (j >= js && j <= je + 1) {
      double v;
      if (p.S && j == js) {  // qy_edge_south
        double g_in = A2(dya, 0, 1) / A20(dya);
        double g_ou = A2(dya, 0, -2) / A2(dya, 0, -1);
        v = 0.5 * (((2.0 + g_in) * A0(qin) - A(qin, 0, 1, 0)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * A(qin, 0, -1, 0) - A(qin, 0, -2, 0)) / (1.0 + g_ou));
      } else if (p.S && j == js + 1) {  // qy_edge_south2
        double g_in = A20(dya) / A2(dya, 0, -1);
        double g_ou = A2(dya, 0, -3) / A2(dya, 0, -2);
        double qy_lower = 0.5 * (((2.0 + g_in) * A(qin, 0, -1, 0) - A0(qin)) / (1.0 + g_in) +
                                 ((2.0 + g_ou) * A(qin, 0, -2, 0) - A(qin, 0, -3, 0)) / (1.0 + g_ou));
        double qy_upper = b2 * (A(qin, 0, -1, 0) + A(qin, 0, 2, 0)) + b1 * (A0(qin) + A(qin, 0, 1, 0));
        v = (3.0 * (g_in * A(qin, 0, -1, 0) + A0(qin)) - (g_in * qy_lower + qy_upper)) / (2.0 + 2.0 * g_in);
      } else if (p.N && j == je + 1) {  // qy_edge_north
        double g_in = A2(dya, 0, -2) / A2(dya, 0, -1);
        double g_ou = A2(dya, 0, 1) / A20(dya);
        v = 0.5 * (((2.0 + g_in) * A(qin, 0, -1, 0) - A(qin, 0, -2, 0)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * A0(qin) - A(qin, 0, 1, 0)) / (1.0 + g_ou));
      } else if (p.N && j == je) {  // qy_edge_north2
        double g_in = A2(dya, 0, -1) / A20(dya);
        double g_ou = A2(dya, 0, 2) / A2(dya, 0, 1);
        double qy_lower = b2 * (A(qin, 0, -3, 0) + A(qin, 0, 0, 0)) + b1 * (A(qin, 0, -2, 0) + A(qin, 0, -1, 0));
        double qy_upper = 0.5 * (((2.0 + g_in) * A0(qin) - A(qin, 0, -1, 0)) / (1.0 + g_in) +
                                 ((2.0 + g_ou) * A(qin, 0, 1, 0) - A(qin, 0, 2, 0)) / (1.0 + g_ou));
        v = (3.0 * (A(qin, 0, -1, 0) + g_in * A0(qin)) - (g_in * qy_upper + qy_lower)) / (2.0 + 2.0 * g_in);
      } else {
        v = b2 * (A(qin, 0, -2, 0) + A(qin, 0, 1, 0)) + b1 * (A(qin, 0, -1, 0) + A0(qin));
      }
      A0(qy) = v;
    }
    // tile corners and tile edges of qout
    const bool onW = p.W && i == is, onE = p.E && i == ie + 1, onS = p.S && j == js, onN = p.N && j == je + 1;
    if (!(i >= is && i <= ie + 1 && j >= js && j <= je + 1)) return;
    if ((onW || onE) && (onS || onN)) {
      // extrap_corner: qa + coef*(qa - qb); corner order SW, SE, NE, NW (a2b_ord4.py:62-247; the reference
      // names the SE stencil "_nw_corner" and the NW one "_se_corner")
      const double* cf = coef + (p.t * 4 + (onS ? (onW ? 0 : 1) : (onE ? 2 : 3))) * 3;
      auto ec = [&](int n, int ai, int aj, int bi, int bj) {
        double qa = A(qin, ai, aj, 0), qb = A(qin, bi, bj, 0);
        return qa + cf[n] * (qa - qb);
      };
      double e1, e2, e3;
      if (onS && onW) {
        e1 = ec(0, 0, 0, 1, 1); e2 = ec(1, -1, 0, -2, 1); e3 = ec(2, 0, -1, 1, -2);
      } else if (onS && onE) {
        e1 = ec(0, -1, 0, -2, 1); e2 = ec(1, -1, -1, -2, -2); e3 = ec(2, 0, 0, 1, 1);
      } else if (onN && onE) {
        e1 = ec(0, -1, -1, -2, -2); e2 = ec(1, 0, -1, 1, -2); e3 = ec(2, -1, 0, -2, 1);
      } else {
        e1 = ec(0, 0, -1, 1, -2); e2 = ec(1, -1, -1, -2, -2); e3 = ec(2, 0, 0, 1, 1);
      }
      A0(qout) = (e1 + e2 + e3) * (1.0 / 3.0);
    } else if (onW || onE) {  // qout_x_edge
      const double* ew = onW ? edge_w : edge_e;
      auto q2 = [&](int dj) {
        return (A(qin, -1, dj, 0) * A2(dxa, 0, dj) + A(qin, 0, dj, 0) * A2(dxa, -1, dj)) /
               (A2(dxa, -1, dj) + A2(dxa, 0, dj));
      };
      A0(qout) = A20(ew) * q2(-1) + (1.0 - A20(ew)) * q2(0);
    } else if (onS || onN) {  // qout_y_edge
      const double* es = onS ? edge_s : edge_n;
      auto q1 = [&](int di) {
        return (A(qin, di, -1, 0) * A2(dya, di, 0) + A(qin, di, 0, 0) * A2(dya, di, -1)) /
               (A2(dya, di, -1) + A2(dya, di, 0));
      };
      A0(qout) = A20(es) * q1(-1) + (1.0 - A20(es)) * q1(0);
    }
  });
  // K2: a2b_interpolation on corner points not on a tile edge
  launch_ijk(st, g, Box{is, js, k0, g.nx + 1, g.ny + 1, nk}, [=] __device__(const Pt& p) {
    const int i = p.i, j = p.j;
    if ((p.W && i == is) || (p.E && i == ie + 1) || (p.S && j == js) || (p.N && j == je + 1)) return;
    double qxx = a2 * (A(qx, 0, -2, 0) + A(qx, 0, 1, 0)) + a1 * (A(qx, 0, -1, 0) + A0(qx));
    double qyy = a2 * (A(qy, -2, 0, 0) + A(qy, 1, 0, 0)) + a1 * (A(qy, -1, 0, 0) + A0(qy));
    if (p.S && j == js + 1) {
      double qxx_upper = a2 * (A(qx, 0, -1, 0) + A(qx, 0, 2, 0)) + a1 * (A0(qx) + A(qx, 0, 1, 0));
      qxx = c1 * (A(qx, 0, -1, 0) + A0(qx)) + c2 * (A(qout, 0, -1, 0) + qxx_upper);
    }
    if (p.N && j == je) {
      double qxx_lower = a2 * (A(qx, 0, -3, 0) + A0(qx)) + a1 * (A(qx, 0, -2, 0) + A(qx, 0, -1, 0));
      qxx = c1 * (A(qx, 0, -1, 0) + A0(qx)) + c2 * (A(qout, 0, 1, 0) + qxx_lower);
    }
    if (p.W && i == is + 1) {
      double qyy_right = a2 * (A(qy, -1, 0, 0) + A(qy, 2, 0, 0)) + a1 * (A0(qy) + A(qy, 1, 0, 0));
      qyy = c1 * (A(qy, -1, 0, 0) + A0(qy)) + c2 * (A(qout, -1, 0, 0) + qyy_right);
    }
    if (p.E && i == ie) {
      double qyy_left = a2 * (A(qy, -3, 0, 0) + A0(qy)) + a1 * (A(qy, -2, 0, 0) + A(qy, -1, 0, 0));
      qyy = c1 * (A(qy, -1, 0, 0) + A0(qy)) + c2 * (A(qout, 1, 0, 0) + qyy_left);
    }
    A0(qout) = 0.5 * (qxx + qyy);
  });
  if (replace) {
    double* qin_w = const_cast<double*>(qin);
    launch_ijk(st, g, Box{is, js, k0, g.nx + 1, g.ny + 1, nk}, [=] __device__(const Pt& p) { A0(qin_w) = A0(qout); });
  }
  return 0;
}

}  // namespace pace
