// The three kernels of UpdateGeopotentialHeightOnCGrid.__call__ (pyFV3/stencils/updatedzc.py:18-118, 185-226), shared by
// pace_b200_update_dz_c (work array from the handle's scratch pool, dz_pgrad.cu) and pace_b200_update_dz_c_static (work
// array inside the library image, update_dz_c_static.cu) so that the two entry points cannot drift apart.
#pragma once
#include <cstdlib>

#include "core.hpp"
#include "corners.cuh"

namespace pace {

// The reference makes two corner-filled copies of gz (gz_x, gz_y); here one plain copy (gz0, caller-provided work array
// of field size) keeps the pre-update values and the corner fills are applied virtually at read time (corners.cuh).
// mode 0: gz0 is a work array; gz is copied into it first (the stand-alone entry points).
// mode 1: gz0 already holds the pre-update heights on the whole array and equals gz (first acoustic substep: zh, just
//         copied from gz): no copy.
// mode 2: gz0 holds the pre-update heights and gz holds anything (later substeps: the reference copies zh to gz and
//         then updates gz): gz is written everywhere — the update on the compute domain + 1, gz0 outside it — which
//         saves the separate copy zh -> gz (one read and one write pass) and the copy gz -> work array (another two).
static int update_dz_c_with(Core& c, double dt, double* gz0, int mode = 0) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const double *zs = c.F("ad.zs"), *ut = c.F("ad.ut"), *vt = c.F("ad.vt"), *area = c.F("grid.area"),
               *dp_ref = c.F("grid.dp_ref");
  double *gz = c.F("ad.gz"), *ws = c.F("ad.ws3");
  if (!c.err.empty() || !gz0) return 1;
  const int km = g.nz;
  if (mode == 0) launch_ijk(st, g, box_full(g, 1), [=] __device__(const Pt& p) { A0(gz0) = A0(gz); });
  const bool everywhere = mode == 2;
  // One column kernel, from the surface up (PACE_B200_DZC=levels: the level-parallel kernel + the column pass below, for
  // A/B runs): the interface winds of a level are read once instead of by the two levels that average them, and the
  // monotonicity pass (which walks the column upward anyway) is applied before the heights are written, not in a second
  // read-modify-write pass over gz.
  static const bool by_levels = [] { const char* e = getenv("PACE_B200_DZC"); return e != nullptr && e[0] == 'l'; }();
  if (!by_levels) {
    launch_ij(st, g, everywhere ? box_full(g, 1) : box_compute(g, -1, -1, 2, 2, 1), [=] __device__(const Pt& p) {
      if (everywhere && (p.i < g.is - 1 || p.i > g.ie + 1 || p.j < g.js - 1 || p.j > g.je + 1)) {
        for (int k = 0; k <= km; ++k) A(gz, 0, 0, k) = A(gz0, 0, 0, k);
        return;
      }
      const EdgeInfo e = edge_info(g, p);
      // cells of the compute domain proper never read a corner-filled value (the fills concern the ring around it)
      const bool plain = p.i >= g.is && p.i <= g.ie && p.j >= g.js && p.j <= g.je;
      // offsets of the five heights the upwind fluxes choose from, seen through the corner fills
      int ox_m = -1, ox_0 = 0, ox_p = 1, oy_m = -p.sj, oy_0 = 0, oy_p = p.sj;
      if (!plain) {
        auto off = [&](bool xdir, int di, int dj) {
          int i = p.i + di, j = p.j + dj;
          if (xdir) xfill_src(e, 2, i, j); else yfill_src(e, 2, i, j);
          return (i - p.i) + (j - p.j) * p.sj;
        };
        ox_m = off(true, -1, 0); ox_0 = off(true, 0, 0); ox_p = off(true, 1, 0);
        oy_m = off(false, 0, -1); oy_0 = off(false, 0, 0); oy_p = off(false, 0, 1);
      }
      const double ar = A20(area);
      const double rdt = 1.0 / dt;
      // winds at the levels k-1 (lo) and k (hi) of the interface in work, at the four faces of the cell
      double u0_lo, u0_hi, u1_lo, u1_hi, v0_lo, v0_hi, v1_lo, v1_hi;
      u0_hi = A(ut, 0, 0, km - 1); u1_hi = A(ut, 1, 0, km - 1); v0_hi = A(vt, 0, 0, km - 1); v1_hi = A(vt, 0, 1, km - 1);
      u0_lo = A(ut, 0, 0, km - 2); u1_lo = A(ut, 1, 0, km - 2); v0_lo = A(vt, 0, 0, km - 2); v1_lo = A(vt, 0, 1, km - 2);
      double below = 0.0;
      for (int k = km; k >= 0; --k) {
#ifndef PACE_HOST_EMU
        if ((p.i & 3) == 0 && k - 7 >= 0) {   // ask L2 for the rows six levels ahead, one request per 32-byte sector
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(ut, 0, 0, k - 7)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(vt, 0, 0, k - 7)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(gz0, 0, 0, k - 6)));
        }
#endif
        double xfx0, xfx1, yfx0, yfx1;
        if (k == km) {          // extrapolated below the lowest layer: hi = level km-1, lo = level km-2
          const double wgt = dp_ref[k - 1] / (dp_ref[k - 2] + dp_ref[k - 1]);
          xfx0 = u0_hi + (u0_hi - u0_lo) * wgt; xfx1 = u1_hi + (u1_hi - u1_lo) * wgt;
          yfx0 = v0_hi + (v0_hi - v0_lo) * wgt; yfx1 = v1_hi + (v1_hi - v1_lo) * wgt;
        } else if (k > 0) {     // hi = level k, lo = level k-1
          if (k < km - 1) {
            u0_hi = u0_lo; u1_hi = u1_lo; v0_hi = v0_lo; v1_hi = v1_lo;
            u0_lo = A(ut, 0, 0, k - 1); u1_lo = A(ut, 1, 0, k - 1); v0_lo = A(vt, 0, 0, k - 1); v1_lo = A(vt, 0, 1, k - 1);
          }
          const double wgt = 1.0 / (dp_ref[k - 1] + dp_ref[k]);
          const double d_hi = dp_ref[k], d_lo = dp_ref[k - 1];
          xfx0 = (d_hi * u0_lo + d_lo * u0_hi) * wgt; xfx1 = (d_hi * u1_lo + d_lo * u1_hi) * wgt;
          yfx0 = (d_hi * v0_lo + d_lo * v0_hi) * wgt; yfx1 = (d_hi * v1_lo + d_lo * v1_hi) * wgt;
        } else {                // extrapolated above the top layer: lo = level 0, hi = level 1
          const double wgt = dp_ref[0] / (dp_ref[0] + dp_ref[1]);
          xfx0 = u0_lo + (u0_lo - u0_hi) * wgt; xfx1 = u1_lo + (u1_lo - u1_hi) * wgt;
          yfx0 = v0_lo + (v0_lo - v0_hi) * wgt; yfx1 = v1_lo + (v1_lo - v1_hi) * wgt;
        }
        const double* z = gz0 + p.o + (long long)k * g.sk;
        const double fx0 = xfx0 * (xfx0 > 0.0 ? z[ox_m] : z[ox_0]);
        const double fx1 = xfx1 * (xfx1 > 0.0 ? z[ox_0] : z[ox_p]);
        const double fy0 = yfx0 * (yfx0 > 0.0 ? z[oy_m] : z[oy_0]);
        const double fy1 = yfx1 * (yfx1 > 0.0 ? z[oy_0] : z[oy_p]);
        double v = (z[0] * ar + fx0 - fx1 + fy0 - fy1) / (ar + xfx0 - xfx1 + yfx0 - yfx1);
        if (k == km) {
          A20(ws) = (A20(zs) - v) * rdt;
        } else {
          const double gz_kp1 = below + DZ_MIN;
          v = v > gz_kp1 ? v : gz_kp1;
        }
        A(gz, 0, 0, k) = v;
        below = v;
      }
    });
    return 0;
  }
  launch_ijk(st, g, everywhere ? box_full(g, 1) : box_compute(g, -1, -1, 2, 2, 1), [=] __device__(const Pt& p) {
    if (everywhere && (p.i < g.is - 1 || p.i > g.ie + 1 || p.j < g.js - 1 || p.j > g.je + 1)) {
      A0(gz) = A0(gz0);
      return;
    }
    const int k = p.k;
    const EdgeInfo e = edge_info(g, p);
    // xfx/yfx at horizontal offset (di,dj): pressure-weighted interface average of ut/vt.  The weight depends on the
    // level only: one division per point instead of one per average.
    const double wgt = k == 0 ? dp_ref[0] / (dp_ref[0] + dp_ref[1])
                     : k < km ? 1.0 / (dp_ref[k - 1] + dp_ref[k])
                              : dp_ref[k - 1] / (dp_ref[k - 2] + dp_ref[k - 1]);
    auto favg = [&](const double* vel, int di, int dj) {
      if (k == 0) {
        return A(vel, di, dj, 0) + (A(vel, di, dj, 0) - A(vel, di, dj, 1)) * wgt;
      } else if (k < km) {
        return (dp_ref[k] * A(vel, di, dj, -1) + dp_ref[k - 1] * A(vel, di, dj, 0)) * wgt;
      } else {
        return A(vel, di, dj, -1) + (A(vel, di, dj, -1) - A(vel, di, dj, -2)) * wgt;
      }
    };
    // cells of the compute domain proper never read a corner-filled value (the fills concern the ring around it)
    const bool plain = p.i >= g.is && p.i <= g.ie && p.j >= g.js && p.j <= g.je;
    auto gzx = [&](int di, int dj) {
      int i = p.i + di, j = p.j + dj;
      if (!plain) xfill_src(e, 2, i, j);
      return gz0[p.o + (i - p.i) + (j - p.j) * p.sj];
    };
    auto gzy = [&](int di, int dj) {
      int i = p.i + di, j = p.j + dj;
      if (!plain) yfill_src(e, 2, i, j);
      return gz0[p.o + (i - p.i) + (j - p.j) * p.sj];
    };
    double xfx0 = favg(ut, 0, 0), xfx1 = favg(ut, 1, 0), yfx0 = favg(vt, 0, 0), yfx1 = favg(vt, 0, 1);
    double fx0 = xfx0 * (xfx0 > 0.0 ? gzx(-1, 0) : gzx(0, 0));
    double fx1 = xfx1 * (xfx1 > 0.0 ? gzx(0, 0) : gzx(1, 0));
    double fy0 = yfx0 * (yfx0 > 0.0 ? gzy(0, -1) : gzy(0, 0));
    double fy1 = yfx1 * (yfx1 > 0.0 ? gzy(0, 0) : gzy(0, 1));
    A0(gz) = (A0(gz0) * A20(area) + fx0 - fx1 + fy0 - fy1) / (A20(area) + xfx0 - xfx1 + yfx0 - yfx1);
  });
  launch_ij(st, g, box_compute(g, -1, -1, 2, 2, 1), [=] __device__(const Pt& p) {
    double rdt = 1.0 / dt;
    double below = A(gz, 0, 0, km);
    A20(ws) = (A20(zs) - below) * rdt;
    for (int k = km - 1; k >= 0; --k) {
      double gz_kp1 = below + DZ_MIN;
      double v = A(gz, 0, 0, k);
      v = v > gz_kp1 ? v : gz_kp1;
      A(gz, 0, 0, k) = v;
      below = v;
    }
  });
  return 0;
}

}  // namespace pace
