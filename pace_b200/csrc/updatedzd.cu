// UpdateHeightOnDGrid (Fortran update_dz_d): advect the interface heights zh with the D-grid fluxes.
//
// Reference: pyFV3/stencils/updatedzd.py:28-109 (apply_height_fluxes), :130-209 (cubic spline),
//            :316-389 (__call__ order).  29 reference launches -> 1 + fvtp2d(4) + delnflux(5) + 2.
#include "transport.hpp"

namespace pace {

int update_dz_d(Core& c, double dt) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const int km = g.nz;
  const double *zs = c.F("ad.zs"), *crx = c.F("ad.crx"), *cry = c.F("ad.cry"), *xfx = c.F("ad.xfx"), *yfx = c.F("ad.yfx"),
               *area = c.F("grid.area"), *gk = c.F("col.dzd_gk"), *beta = c.F("col.dzd_beta"), *gamma = c.F("col.dzd_gamma"),
               *nord_v = c.F("col.nord_v"), *damp_vt = c.F("col.damp_vt");
  double *zh = c.F("ad.zh"), *ws = c.F("ad.wsd");
  double *crx_i = c.FS("dzd.crx_interface"), *cry_i = c.FS("dzd.cry_interface"), *xfx_i = c.FS("dzd.x_area_flux_interface"),
         *yfx_i = c.FS("dzd.y_area_flux_interface"), *wk = c.FS("dzd.wk"), *hx = c.FS("dzd.height_x_diffusive_flux"),
         *hy = c.FS("dzd.height_y_diffusive_flux"), *fx = c.FS("dzd.fx"), *fy = c.FS("dzd.fy");
  if (!c.err.empty()) return 1;
  const int nmax_v = c.col_max("col.nord_v");
  {
    const std::vector<double>* dv = c.H("col.damp_vt");
    if (!dv) return 1;
    for (int k = 0; k < km; ++k)
      if ((*dv)[k] <= 1e-5) { c.err = "damp <= 1e-5 in column_namelist is untested"; return 1; }
  }

  // cubic_spline_interpolation_from_layer_center_to_interfaces x4, full horizontal domain
  launch_ij(st, g, box_full(g), [=] __device__(const Pt& p) {
    const double* src[4] = {crx, xfx, cry, yfx};
    double* dst[4] = {crx_i, xfx_i, cry_i, yfx_i};
    for (int f = 0; f < 4; ++f) {
      const double* q = src[f];
      double* qi = dst[f];
      double xt1 = 2.0 * gk[0] * (gk[0] + 1.0);
      double prev = (xt1 * A(q, 0, 0, 0) + A(q, 0, 0, 1)) / beta[0];
      A(qi, 0, 0, 0) = prev;
      for (int k = 1; k < km; ++k) {
        prev = (3.0 * (A(q, 0, 0, k - 1) + gk[k] * A(q, 0, 0, k)) - prev) / beta[k];
        A(qi, 0, 0, k) = prev;
      }
      {
        double a_bot = 1.0 + gk[km - 1] * (gk[km - 1] + 1.5);
        double xt1b = 2.0 * gk[km - 1] * (gk[km - 1] + 1.0);
        double xt2 = gk[km - 1] * (gk[km - 1] + 0.5) - a_bot * gamma[km - 1];
        prev = (xt1b * A(q, 0, 0, km - 1) + A(q, 0, 0, km - 2) - a_bot * prev) / xt2;
        A(qi, 0, 0, km) = prev;
      }
      for (int k = km - 1; k >= 0; --k) {
        prev = A(qi, 0, 0, k) - gamma[k] * prev;
        A(qi, 0, 0, k) = prev;
      }
    }
  }, "dzd_spline");
  Fvtp2dArgs a;
  a.q = zh; a.crx = crx_i; a.cry = cry_i; a.xfx = xfx_i; a.yfx = yfx_i; a.fx = fx; a.fy = fy; a.hord = c.cfg.hord_tm; a.nk = km + 1;
  if (fvtp2d(c, a)) return 1;
  DelnArgs d;
  d.q = zh; d.fx = hx; d.fy = hy; d.nord_col = nord_v; d.damp = damp_vt; d.d2 = wk; d.nmax = nmax_v; d.nk = km + 1;
  if (delnflux(c, d)) return 1;
  // apply_height_fluxes: PARALLEL part, then the column part (ws, monotone zh)
  launch_ijk(st, g, box_compute(g, 0, 0, 0, 0, 1), [=] __device__(const Pt& p) {
    double ar = A20(area);
    double area_after_flux = ((ar + A0(xfx_i) - A(xfx_i, 1, 0, 0)) + (ar + A0(yfx_i) - A(yfx_i, 0, 1, 0))) - ar;
    A0(zh) = (A0(zh) * ar + A0(fx) - A(fx, 1, 0, 0) + A0(fy) - A(fy, 0, 1, 0)) / area_after_flux +
             (A0(hx) - A(hx, 1, 0, 0) + A0(hy) - A(hy, 0, 1, 0)) / ar;
  });
  launch_ij(st, g, box_compute(g), [=] __device__(const Pt& p) {
    double below = A(zh, 0, 0, km);
    A20(ws) = (A20(zs) - below) / dt;
    for (int k = km - 1; k >= 0; --k) {
      double other = below + DZ_MIN;
      double h = A(zh, 0, 0, k);
      h = h > other ? h : other;
      A(zh, 0, 0, k) = h;
      below = h;
    }
  });
  return 0;
}

}  // namespace pace
