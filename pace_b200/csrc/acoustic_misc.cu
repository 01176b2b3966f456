// The small stencils of the acoustic substep around the big modules: pe / pk3 halo rings, geopotential,
// nonhydrostatic pressure gradient, Rayleigh damping, del2cubed hyperdiffusion, diffusive heating, and the
// AcousticDynamics helper stencils.
//
// Reference: pyFV3/stencils/dyn_core.py:53-119 (zero_data, gz_from_surface_height_and_thicknesses,
// interface_pressure_from_toa_pressure_and_thickness, compute_geopotential); pe_halo.py:6-34; pk3_halo.py:11-73;
// nh_p_grad.py:12-113,201-265; ray_fast.py:23-142; del2cubed.py:21-202; temperature_adjust.py:8-42.
#include "core.hpp"
#include "corners.cuh"

#include <algorithm>
#include <vector>

namespace pace {

// zero_data (dyn_core.py:53-82), full domain
int zero_data(Core& c, int first_timestep) {
  const Geo g = c.g;
  double *mfxd = c.F("mfxd"), *mfyd = c.F("mfyd"), *cxd = c.F("cxd"), *cyd = c.F("cyd"),
         *heat_source = c.F("ad.heat_source"), *diss_estd = c.F("diss_estd");
  if (!c.err.empty()) return 1;
  const int nx = g.nx, ny = g.ny;
  launch_ijk(c.stream, g, box_full(g), [=] __device__(const Pt& p) {
    A0(mfxd) = 0.0; A0(mfyd) = 0.0; A0(cxd) = 0.0; A0(cyd) = 0.0;
    // region[3:-3, 3:-3] of the full domain
    if (first_timestep && p.i >= 3 && p.i < nx + 3 && p.j >= 3 && p.j < ny + 3) {
      A0(heat_source) = 0.0;
      A0(diss_estd) = 0.0;
    }
  });
  return 0;
}

// gz_from_surface_height_and_thicknesses (dyn_core.py:85-98), compute domain
int gz_from_zs_delz(Core& c) {
  const Geo g = c.g;
  const double *zs = c.F("ad.zs"), *delz = c.F("delz");
  double* gz = c.F("ad.gz");
  if (!c.err.empty()) return 1;
  const int km = g.nz;
  launch_ij(c.stream, g, box_compute(g), [=] __device__(const Pt& p) {
    double v = A20(zs);
    A(gz, 0, 0, km) = v;
    for (int k = km - 1; k >= 0; --k) {
      v = v - A(delz, 0, 0, k);
      A(gz, 0, 0, k) = v;
    }
  });
  return 0;
}

// interface_pressure_from_toa_pressure_and_thickness (dyn_core.py:101-114), compute(-1,-1)+(2,2)
int pem_from_delp(Core& c) {
  const Geo g = c.g;
  const double* delp = c.F("delp");
  double* pem = c.F("ad.pem");
  if (!c.err.empty()) return 1;
  const int km = g.nz;
  const double ptop = c.cfg.ptop;
  launch_ij(c.stream, g, box_compute(g, -1, -1, 2, 2, 0), [=] __device__(const Pt& p) {
    double v = ptop;
    A(pem, 0, 0, 0) = v;
    // the stencil's domain has npz levels: interval(1, None) = k in [1, npz)
    for (int k = 1; k < km; ++k) {
      v = v + A(delp, 0, 0, k);
      A(pem, 0, 0, k) = v;
    }
  });
  return 0;
}

// basic.copy_defn on the full domain, npz+1 levels (dyn_core.py:603-607): dst = src
int copy_full(Core& c, const char* src_name, const char* dst_name) {
  const Geo g = c.g;
  const double* src = c.F(src_name);
  double* dst = c.F(dst_name);
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_full(g, 1), [=] __device__(const Pt& p) { A0(dst) = A0(src); });
  return 0;
}

// edge_pe (pe_halo.py:6-34): pe on the 1-wide ring around the compute domain
int edge_pe(Core& c) {
  const Geo g = c.g;
  const double* delp = c.F("delp");
  double* pe = c.F("pe");
  if (!c.err.empty()) return 1;
  const int km = g.nz, is = g.is, ie = g.ie, js = g.js, je = g.je;
  const double ptop = c.cfg.ptop;
  launch_ij(c.stream, g, box_compute(g, -1, -1, 2, 2, 0), [=] __device__(const Pt& p) {
    bool ring = ((p.i == is - 1 || p.i == ie + 1) && p.j >= js && p.j <= je) || p.j == js - 1 || p.j == je + 1;
    if (!ring) return;
    double v = ptop;
    A(pe, 0, 0, 0) = v;
    for (int k = 1; k <= km; ++k) {
      v = v + A(delp, 0, 0, k - 1);
      A(pe, 0, 0, k) = v;
    }
  });
  return 0;
}

// PK3Halo (pk3_halo.py:11-73): pk3 = pe**akap on the 2-wide ring
int pk3_halo(Core& c) {
  const Geo g = c.g;
  const double* delp = c.F("delp");
  double* pk3 = c.F("ad.pk3");
  if (!c.err.empty()) return 1;
  const int km = g.nz, is = g.is, ie = g.ie, js = g.js, je = g.je;
  const double ptop = c.cfg.ptop, akap = KAPPA;
  launch_ij(c.stream, g, box_compute(g, -2, -2, 4, 4, 0), [=] __device__(const Pt& p) {
    bool ring = ((p.i < is || p.i > ie) && p.j >= js && p.j <= je) || p.j < js || p.j > je;
    if (!ring) return;
    double pe = ptop;
    for (int k = 1; k <= km; ++k) {
      pe = pe + A(delp, 0, 0, k - 1);
      A(pk3, 0, 0, k) = fv_pow(pe, akap);
    }
  });
  return 0;
}

// compute_geopotential (dyn_core.py:117-119) on [X, Y, Z_INTERFACE] + 2 halo
int compute_geopotential(Core& c) {
  const Geo g = c.g;
  const double* zh = c.F("ad.zh");
  double* gz = c.F("ad.gz");
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_dims(g, 0, 0, 1, 2, 2), [=] __device__(const Pt& p) { A0(gz) = A0(zh) * GRAV; });
  return 0;
}

// NonHydrostaticPressureGradient.__call__ (nh_p_grad.py:201-265)
// keep_agrid: the B-grid interpolations of pp, pk3 and gz go to work arrays instead of replacing the fields (saves the
// three copy-back passes); AcousticDynamics (step.cu) does that on every substep but its last, after which the
// reference's temporaries hold the B-grid values, as here.
int nh_p_grad(Core& c, double dt, bool keep_agrid) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  double *u = c.F("u"), *v = c.F("v"), *pp_a = c.F("ad.pkc"), *gz_a = c.F("ad.gz"), *pk3_a = c.F("ad.pk3");
  const double *delp = c.F("delp"), *rdx = c.F("grid.rdx"), *rdy = c.F("grid.rdy");
  double* wk1 = c.FS("nh.tmp_wk1");
  if (!c.err.empty()) return 1;
  const int km = g.nz, ie = g.ie, je = g.je;
  const double top_value = std::pow(c.cfg.ptop, KAPPA);  // ptk = ptop ** akap
  double *pp = pp_a, *pk3 = pk3_a, *gz = gz_a;
  if (keep_agrid) {
    pp = c.S("nh.pp_b"); pk3 = c.S("nh.pk3_b"); gz = c.S("nh.gz_b");
    if (!c.err.empty()) return 1;
    if (a2b_ord4(c, pp_a, pp, 1, km, false)) return 1;
    if (a2b_ord4(c, pk3_a, pk3, 1, km, false)) return 1;
    if (a2b_ord4(c, gz_a, gz, 0, km + 1, false)) return 1;
  } else {
    if (a2b_ord4(c, pp, wk1, 1, km, true)) return 1;
    if (a2b_ord4(c, pk3, wk1, 1, km, true)) return 1;
    if (a2b_ord4(c, gz, wk1, 0, km + 1, true)) return 1;
  }
  if (a2b_ord4(c, delp, wk1, 0, km, false)) return 1;
  // set_k0_and_calc_wk + calc_u + calc_v (nh_p_grad.py:12-113) in one kernel on compute (+1,+1), npz levels: the
  // reference first sets pp = 0 and pk3 = ptk on level 0 and forms wk = pk3(k+1) - pk3(k) in a work field; here the
  // level-0 values are substituted at read time and wk is formed in registers for the three points that need it (same
  // expression), which saves the work field's write and reads and a launch.  Level 0 of pp / pk3 is still stored.
#ifndef PACE_HOST_EMU
  // Device build: one thread per column, marching down the levels.  Every level of pk3 / pp / gz is then read once per
  // point (the values of interface k+1 become those of interface k of the next level) instead of twice by two CTAs.
  // PACE_B200_NHPG=levels selects the level-parallel kernel below (A/B runs; the host emulation always uses it).
  static const bool by_levels = [] { const char* e = getenv("PACE_B200_NHPG"); return e != nullptr && e[0] == 'l'; }();
  constexpr int NHPG_PF = 6;
  if (!by_levels) {
    launch_ij_occ<5>(st, g, box_compute(g, 0, 0, 1, 1, 0), [=] __device__(const Pt& p) {
      const bool has_u = p.i <= ie, has_v = p.j <= je;
      // interface k: pk3 and pp of level 0 are the prescribed top values, stored here as the reference does
      double pk00 = top_value, pk10 = top_value, pk01 = top_value;
      double pp00 = 0.0, pp10 = 0.0, pp01 = 0.0;
      double gz00 = A(gz, 0, 0, 0), gz10 = A(gz, 1, 0, 0), gz01 = A(gz, 0, 1, 0);
      A(pp, 0, 0, 0) = 0.0;
      A(pk3, 0, 0, 0) = top_value;
      const double rdx0 = A20(rdx), rdy0 = A20(rdy);
#pragma unroll 1
      for (int k = 0; k < km; ++k) {
        if ((p.i & 3) == 0 && k + NHPG_PF < km) {   // ask L2 for the rows a few levels ahead, one request per sector
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(pk3, 0, 0, k + 1 + NHPG_PF)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(pp, 0, 0, k + 1 + NHPG_PF)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(gz, 0, 0, k + 1 + NHPG_PF)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(wk1, 0, 0, k + NHPG_PF)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(u, 0, 0, k + NHPG_PF)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(v, 0, 0, k + NHPG_PF)));
        }
        const double npk00 = A(pk3, 0, 0, k + 1), npk10 = A(pk3, 1, 0, k + 1), npk01 = A(pk3, 0, 1, k + 1);
        const double npp00 = A(pp, 0, 0, k + 1), npp10 = A(pp, 1, 0, k + 1), npp01 = A(pp, 0, 1, k + 1);
        const double ngz00 = A(gz, 0, 0, k + 1), ngz10 = A(gz, 1, 0, k + 1), ngz01 = A(gz, 0, 1, k + 1);
        const double w00 = A(wk1, 0, 0, k), w10 = A(wk1, 1, 0, k), w01 = A(wk1, 0, 1, k);
        const double wk0 = npk00 - pk00;
        if (has_u) {
          double du = dt / (wk0 + (npk10 - pk10)) * ((ngz00 - gz10) * (npk10 - pk00) + (gz00 - ngz10) * (npk00 - pk10));
          A(u, 0, 0, k) = (A(u, 0, 0, k) + du +
                           dt / (w00 + w10) * ((ngz00 - gz10) * (npp10 - pp00) + (gz00 - ngz10) * (npp00 - pp10))) *
                          rdx0;
        }
        if (has_v) {
          double dv = dt / (wk0 + (npk01 - pk01)) * ((ngz00 - gz01) * (npk01 - pk00) + (gz00 - ngz01) * (npk00 - pk01));
          A(v, 0, 0, k) = (A(v, 0, 0, k) + dv +
                           dt / (w00 + w01) * ((ngz00 - gz01) * (npp01 - pp00) + (gz00 - ngz01) * (npp00 - pp01))) *
                          rdy0;
        }
        pk00 = npk00; pk10 = npk10; pk01 = npk01;
        pp00 = npp00; pp10 = npp10; pp01 = npp01;
        gz00 = ngz00; gz10 = ngz10; gz01 = ngz01;
      }
    }, "nh_p_grad_winds");
    return 0;
  }
#endif
  launch_ijk(st, g, box_compute(g, 0, 0, 1, 1, 0), [=] __device__(const Pt& p) {
    const bool top = p.k == 0;
    auto PK = [&](int di, int dj) { return top ? top_value : A(pk3, di, dj, 0); };   // pk3 at this level
    auto PP = [&](int di, int dj) { return top ? 0.0 : A(pp, di, dj, 0); };          // pp at this level
    auto WK = [&](int di, int dj) { return A(pk3, di, dj, 1) - PK(di, dj); };
    if (top) {
      A0(pp) = 0.0;
      A0(pk3) = top_value;
    }
    const double wk0 = WK(0, 0);
    if (p.i <= ie) {
      double du = dt / (wk0 + WK(1, 0)) *
                  ((A(gz, 0, 0, 1) - A(gz, 1, 0, 0)) * (A(pk3, 1, 0, 1) - PK(0, 0)) +
                   (A0(gz) - A(gz, 1, 0, 1)) * (A(pk3, 0, 0, 1) - PK(1, 0)));
      A0(u) = (A0(u) + du +
               dt / (A0(wk1) + A(wk1, 1, 0, 0)) *
                   ((A(gz, 0, 0, 1) - A(gz, 1, 0, 0)) * (A(pp, 1, 0, 1) - PP(0, 0)) +
                    (A0(gz) - A(gz, 1, 0, 1)) * (A(pp, 0, 0, 1) - PP(1, 0)))) *
              A20(rdx);
    }
    if (p.j <= je) {
      double dv = dt / (wk0 + WK(0, 1)) *
                  ((A(gz, 0, 0, 1) - A(gz, 0, 1, 0)) * (A(pk3, 0, 1, 1) - PK(0, 0)) +
                   (A0(gz) - A(gz, 0, 1, 1)) * (A(pk3, 0, 0, 1) - PK(0, 1)));
      A0(v) = (A0(v) + dv +
               dt / (A0(wk1) + A(wk1, 0, 1, 0)) *
                   ((A(gz, 0, 0, 1) - A(gz, 0, 1, 0)) * (A(pp, 0, 1, 1) - PP(0, 0)) +
                    (A0(gz) - A(gz, 0, 1, 1)) * (A(pp, 0, 0, 1) - PP(0, 1)))) *
              A20(rdy);
    }
  });
  return 0;
}

// RayleighDamping.__call__ (ray_fast.py:49-206).  rf[k] depends only on pfull/dt, so it is evaluated on
// the host (same libm calls as the reference) and read from "col.rf".
int ray_fast(Core& c, double dt) {
  const Geo g = c.g;
  double *u = c.F("u"), *v = c.F("v"), *w = c.F("w");
  const double* dp = c.F("grid.dp_ref");
  const std::vector<double>* pf = c.H("grid.p");
  if (!c.err.empty() || !pf) return 1;
  const int km = g.nz, ie = g.ie, je = g.je;
  const double rf_cutoff = c.cfg.rf_cutoff, ptop = c.cfg.ptop;
  const double rf_cutoff_nudge = rf_cutoff + std::min(100.0, 10.0 * ptop);
  constexpr double PI = 3.1415926535897931, SDAY = 86400.0;
  const double tau0 = c.cfg.tau * SDAY;
  // number of leading levels with pfull < rf_cutoff / < rf_cutoff_nudge (pfull increases downward)
  std::vector<double> rf(km + 1, 0.0);
  int n_rf = 0, n_nudge = 0;
  for (int k = 0; k < km; ++k) {
    double pfull = (*pf)[k];
    if (pfull < rf_cutoff) {
      double s = std::sin(0.5 * PI * std::log(rf_cutoff / pfull) / std::log(rf_cutoff / ptop));
      double rfv = dt / tau0 * (s * s);
      rf[k] = 1.0 / (1.0 + rfv);
      if (k != n_rf) { c.err = "ray_fast: pfull < rf_cutoff is not a leading block of levels"; return 1; }
      n_rf = k + 1;
    }
    if (pfull < rf_cutoff_nudge) {
      if (k != n_nudge) { c.err = "ray_fast: pfull < rf_cutoff_nudge is not a leading block of levels"; return 1; }
      n_nudge = k + 1;
    }
  }
  if (set_column(c, "col.rf", rf.data(), km + 1)) return 1;
  const double* rfd = c.F("col.rf");
  if (n_rf == 0) return 0;
  if (n_nudge < n_rf || n_nudge >= km) { c.err = "ray_fast: unsupported rf_cutoff / rf_cutoff_nudge layering"; return 1; }
  const bool hydrostatic = c.cfg.hydrostatic;
  launch_ij(c.stream, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    // p_ref: FORWARD accumulation of dp over the nudge layer, then (BACKWARD) every layer level takes the
    // value of the level below it -> the first level outside the layer, i.e. the layer total.
    // The reference's FORWARD pass also sets p_ref = 0 at k = 0 when level 0 is outside the rf layer; with
    // n_rf >= 1 that branch is not taken.
    double p_ref = 0.0;
    for (int k = 0; k < n_nudge; ++k) p_ref = k == 0 ? dp[0] : p_ref + dp[k];
    for (int pass = 0; pass < 2; ++pass) {
      double* wind = pass == 0 ? u : v;
      if (pass == 0 ? (p.i > ie) : (p.j > je)) continue;
      double dmdir = 0.0;
      for (int k = 0; k < n_rf; ++k) {
        double wv = A(wind, 0, 0, k);
        double dml = (1.0 - rfd[k]) * dp[k] * wv;
        dmdir = k == 0 ? dml : dmdir + dml;
        A(wind, 0, 0, k) = wv * rfd[k];
      }
      // BACKWARD: levels inside the rf layer take dmdir of the level below -> the value after the last
      // rf level, which FORWARD carried on unchanged; levels in [n_rf, n_nudge) keep that same value.
      for (int k = 0; k < n_nudge; ++k) A(wind, 0, 0, k) = A(wind, 0, 0, k) + dmdir / p_ref;
    }
    if (!hydrostatic && p.i <= ie && p.j <= je)
      for (int k = 0; k < n_rf; ++k) A(w, 0, 0, k) = A(w, 0, 0, k) * rfd[k];
  });
  return 0;
}

// HyperdiffusionDamping.__call__ (del2cubed.py:163-202) on field `q` (heat_source or omga)
int del2cubed(Core& c, double* qdel, double cd, int nmax) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const double *del6_u = c.F("damp.del6_u"), *del6_v = c.F("damp.del6_v"), *rarea = c.F("grid.rarea");
  double* q = c.S("del2.q");
  if (!c.err.empty()) return 1;
  const int ntimes = nmax < 3 ? nmax : 3;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, nz = g.nz;
  for (int n = 0; n < ntimes; ++n) {
    const int nt = ntimes - (n + 1);
    // corner_fill (del2cubed.py:33-75) on compute + 3 halo
    launch_ijk(st, g, box_dims(g, 0, 0, 0, 3, 3), [=] __device__(const Pt& p) {
      const double third = 1.0 / 3.0;
      const int i = p.i, j = p.j;
      double r = A0(qdel);
      if (p.W && p.S) {
        if (i == is && j == js) r = (A0(qdel) + A(qdel, -1, 0, 0) + A(qdel, 0, -1, 0)) * third;
        if (i == is - 1 && j == js) r = (A(qdel, 1, 0, 0) + A0(qdel) + A(qdel, 1, -1, 0)) * third;
        if (i == is && j == js - 1) r = (A(qdel, 0, 1, 0) + A(qdel, -1, 1, 0) + A0(qdel)) * third;
      }
      if (p.E && p.S) {
        if (i == ie && j == js) r = (A0(qdel) + A(qdel, 1, 0, 0) + A(qdel, 0, -1, 0)) * third;
        if (i == ie + 1 && j == js) r = (A(qdel, -1, 0, 0) + A0(qdel) + A(qdel, -1, -1, 0)) * third;
        if (i == ie && j == js - 1) r = (A(qdel, 0, 1, 0) + A(qdel, 1, 1, 0) + A0(qdel)) * third;
      }
      if (p.E && p.N) {
        if (i == ie && j == je) r = (A0(qdel) + A(qdel, 1, 0, 0) + A(qdel, 0, 1, 0)) * third;
        if (i == ie + 1 && j == je) r = (A(qdel, -1, 0, 0) + A0(qdel) + A(qdel, -1, 1, 0)) * third;
        if (i == ie && j == je + 1) r = (A(qdel, 0, -1, 0) + A(qdel, 1, -1, 0) + A0(qdel)) * third;
      }
      if (p.W && p.N) {
        if (i == is && j == je) r = (A0(qdel) + A(qdel, -1, 0, 0) + A(qdel, 0, 1, 0)) * third;
        if (i == is - 1 && j == je) r = (A(qdel, 1, 0, 0) + A0(qdel) + A(qdel, 1, 1, 0)) * third;
        if (i == is && j == je + 1) r = (A(qdel, 0, -1, 0) + A(qdel, -1, -1, 0) + A0(qdel)) * third;
      }
      A0(q) = r;
    });
    // fluxes from the (virtually) corner-copied q, then qdel = q (y-copied corners) + update on [is-nt..ie+nt]^2
    launch_ijk(st, g, box_dims(g, 0, 0, 0, 3, 3), [=] __device__(const Pt& p) {
      const EdgeInfo e = edge_info(g, p);
      const bool cc = nt > 0;
      double val = cc ? rd_cy(q, p, e, 0, 0) : A0(q);
      if (p.i >= is - nt && p.i <= ie + nt && p.j >= js - nt && p.j <= je + nt) {
        auto qx = [&](int di) { return cc ? rd_cx(q, p, e, di, 0) : A(q, di, 0, 0); };
        auto qy = [&](int dj) { return cc ? rd_cy(q, p, e, 0, dj) : A(q, 0, dj, 0); };
        double fx0 = A20(del6_v) * (qx(-1) - qx(0));
        double fx1 = A2(del6_v, 1, 0) * (qx(0) - qx(1));
        double fy0 = A20(del6_u) * (qy(-1) - qy(0));
        double fy1 = A2(del6_u, 0, 1) * (qy(0) - qy(1));
        val = val + cd * A20(rarea) * (fx0 - fx1 + fy0 - fy1);
      }
      A0(qdel) = val;
    });
  }
  (void)nz;
  return 0;
}

// apply_diffusive_heating (temperature_adjust.py:8-42) on the compute domain, levels [0, nk)
int apply_diffusive_heating(Core& c, double delt_time_factor, int nk) {
  const Geo g = c.g;
  const double *delp = c.F("delp"), *delz = c.F("delz"), *cappa = c.F("ad.cappa"), *heat_source = c.F("ad.heat_source");
  double* pt = c.F("pt");
  if (!c.err.empty()) return 1;
  Box b = box_compute(g);
  b.nk = nk;
  launch_ijk(c.stream, g, b, [=] __device__(const Pt& p) {
    double pkz = fv_exp(A0(cappa) / (1.0 - A0(cappa)) * fv_log(RDG * A0(delp) / A0(delz) * A0(pt)));
    double dtmp = A0(heat_source) / (CV_AIR * A0(delp));
    double lim = p.k == 0 ? delt_time_factor * 0.1 : (p.k == 1 ? delt_time_factor * 0.5 : delt_time_factor);
    double m = fv_min(lim, fabs(dtmp));
    double deltmin = dtmp > 0 ? fabs(m) : -fabs(m);
    A0(pt) = A0(pt) + deltmin / pkz;
  });
  return 0;
}

}  // namespace pace
