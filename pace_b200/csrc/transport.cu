// 2-D finite-volume transport (Putman & Lin fvtp2d) and del-n damping fluxes (delnflux).
//
// Reference: pyFV3/stencils/fvtp2d.py:34-120 (q_i / q_j / final_fluxes), :251-368 (__call__ order);
//            pyFV3/stencils/delnflux.py:22-941 (stencils), :944-1278 (DelnFlux, DelnFluxNoSG);
//            NDSL/ndsl/stencils/corners.py:307-424 (copy_corners_x/y).
//
// fvtp2d: the reference runs 9 stencils (2 corner copies, 4 PPM sweeps, q_i, q_j, final_fluxes) through 6
// full-size temporaries.  Here: one corner-copy launch (the x-copied corners are part of the observable
// state of q), then 3 kernels; the y-copied view of q is read virtually (corners.cuh).
// delnflux: 1 + 2*nmax kernels instead of ~20; the corner copies between sub-steps are virtual.
#include "transport.hpp"

#include "corners.cuh"
#include "ppm.cuh"

#include <cstdlib>
#include <string>

namespace pace {

int fvtp2d_fused_launch(Core& c, const Fvtp2dArgs& a);
int fvtp2d_march_launch(Core& c, const Fvtp2dArgs& a);
int delnflux_fused_launch(Core& c, const DelnArgs& a);

// ---- materialised copy_corners ---------------------------------------------------------------------
__device__ __forceinline__ void copy_corner_cell(const Geo& g, double* q, bool ydir, int t, int k, int c, int a, int b) {
  const unsigned ed = edge_of(g, t);
  bool W = ed & EDGE_W, E = ed & EDGE_E, S = ed & EDGE_S, N = ed & EDGE_N;
  EdgeInfo e{g.is, g.ie, g.js, g.je, W, E, S, N};
  int i, j;
  bool on;
  switch (c) {
    case 0: on = W && S; i = g.is - a; j = g.js - b; break;
    case 1: on = E && S; i = g.ie + a; j = g.js - b; break;
    case 2: on = W && N; i = g.is - a; j = g.je + b; break;
    default: on = E && N; i = g.ie + a; j = g.je + b; break;
  }
  if (!on) return;
  int si = i, sj = j;
  if (ydir) copyy_src(e, si, sj); else copyx_src(e, si, sj);
  long long base = (long long)t * g.st3 + (long long)k * g.sk;
  q[base + j * g.sj + i] = q[base + sj * g.sj + si];
}
#ifndef PACE_HOST_EMU
static __global__ void k_copy_corners(Geo g, double* q, bool ydir, int nk) {
  int k = blockIdx.x % nk, t = blockIdx.x / nk;
  int id = threadIdx.x;
  if (id >= 36) return;
  copy_corner_cell(g, q, ydir, t, k, id / 9, (id % 9) / 3 + 1, id % 3 + 1);
}
#endif
void copy_corners(Core& c, double* q, bool ydir, int nk) {
#ifndef PACE_HOST_EMU
  ++launch_counter();
  k_copy_corners<<<nk * c.g.nt, 64, 0, c.stream>>>(c.g, q, ydir, nk);
#else
  for (int t = 0; t < c.g.nt; ++t)
    for (int k = 0; k < nk; ++k)
      for (int id = 0; id < 36; ++id) copy_corner_cell(c.g, q, ydir, t, k, id / 9, (id % 9) / 3 + 1, id % 3 + 1);
#endif
}

// ---- fvtp2d ------------------------------------------------------------------------------------------
int fvtp2d(Core& c, const Fvtp2dArgs& a) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const double *area = c.F("grid.area"), *dxa = c.F("grid.dxa"), *dya = c.F("grid.dya");
  if (!c.err.empty()) return 1;
  double* q = a.q;
  const double *crx = a.crx, *cry = a.cry, *xfx = a.xfx, *yfx = a.yfx;
  double *fx = a.fx, *fy = a.fy;
  const double* xunit = a.x_mass_flux ? a.x_mass_flux : a.xfx;
  const double* yunit = a.y_mass_flux ? a.y_mass_flux : a.yfx;
  const int ord_outer = a.hord, ord_inner = (a.hord == 10) ? 8 : a.hord;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, nk = a.nk;

  // copy_corners_x(q): final observable corner state of q (the y-copied view is read virtually below)
  copy_corners(c, q, false, nk);

#ifndef PACE_HOST_EMU
  // production path: one row-marching kernel (fvtp2d_march.cu).  PACE_B200_FVTP2D=tile selects the CTA-tiled fused
  // kernel (fvtp2d_fused.cu), =unfused the three-kernel version below (same arithmetic) for A/B measurements
  static const int mode = [] {
    const char* v = getenv("PACE_B200_FVTP2D");
    std::string m = v ? v : "";
    return m == "unfused" ? 2 : (m == "tile" || m == "fused") ? 1 : 0;
  }();
  const bool unfused = mode == 2;
  if (!unfused) {
    int rc = mode == 0 ? fvtp2d_march_launch(c, a) : -1;
    if (rc < 0) rc = fvtp2d_fused_launch(c, a);
    if (rc) return 1;
    if (a.nord_col) {
      DelnArgs d;
      d.q = q; d.fx = fx; d.fy = fy; d.mass = a.mass; d.nord_col = a.nord_col; d.damp = a.damp; d.nk = nk;
      d.accumulate = true; d.d2 = nullptr; d.nmax = a.nmax;
      d.level_nonzero = a.level_nonzero;
      return delnflux(c, d);
    }
    return 0;
  }
#endif
  double *fy_in = c.S("tp.q_y_advected_mean"), *fx_in = c.S("tp.q_x_advected_mean"), *q_i = c.S("tp.q_advected_y"),
         *q_j = c.S("tp.q_advected_x");
  if (!c.err.empty()) return 1;

  // K1: inner sweeps.  q_y_advected_mean on [is-3..ie+3]x[js..je+1] (y-PPM of y-copied q, courant cry),
  //                    q_x_advected_mean on [is..ie+1]x[js-3..je+3] (x-PPM of x-copied q, courant crx)
  launch_ijk(st, g, Box{0, 0, 0, g.nx + 6, g.ny + 6, nk}, [=] __device__(const Pt& p) {
    const EdgeInfo e = edge_info(g, p);
    if (p.j >= js && p.j <= je + 1) {
      auto Q = [&](int d) { return rd_cy(q, p, e, 0, d); };
      auto D = [&](int d) { return A2(dya, 0, d); };
      A0(fy_in) = ppm_flux(Q, D, A0(cry), p.j, js, je, p.S, p.N, ord_inner);
    }
    if (p.i >= is && p.i <= ie + 1) {
      auto Q = [&](int d) { return A(q, d, 0, 0); };
      auto D = [&](int d) { return A2(dxa, d, 0); };
      A0(fx_in) = ppm_flux(Q, D, A0(crx), p.i, is, ie, p.W, p.E, ord_inner);
    }
  }, "fvtp2d_inner_ppm");
  // K2: q_i on [is-3..ie+3]x[js..je], q_j on [is..ie]x[js-3..je+3]   (fvtp2d.py:34-77)
  launch_ijk(st, g, Box{0, 0, 0, g.nx + 6, g.ny + 6, nk}, [=] __device__(const Pt& p) {
    if (p.j >= js && p.j <= je) {
      double fyy0 = A0(yfx) * A0(fy_in), fyy1 = A(yfx, 0, 1, 0) * A(fy_in, 0, 1, 0);
      A0(q_i) = (A0(q) * A20(area) + fyy0 - fyy1) / (A20(area) + A0(yfx) - A(yfx, 0, 1, 0));
    }
    if (p.i >= is && p.i <= ie) {
      double fx10 = A0(xfx) * A0(fx_in), fx11 = A(xfx, 1, 0, 0) * A(fx_in, 1, 0, 0);
      double area_with_x_flux = A20(area) + A0(xfx) - A(xfx, 1, 0, 0);
      A0(q_j) = (A0(q) * A20(area) + fx10 - fx11) / area_with_x_flux;
    }
  }, "fvtp2d_qi_qj");
  // K3: outer sweeps + final_fluxes (fvtp2d.py:80-119)
  launch_ijk(st, g, Box{is, js, 0, g.nx + 1, g.ny + 1, nk}, [=] __device__(const Pt& p) {
    if (p.j <= je) {
      auto Q = [&](int d) { return A(q_i, d, 0, 0); };
      auto D = [&](int d) { return A2(dxa, d, 0); };
      double outer = ppm_flux(Q, D, A0(crx), p.i, is, ie, p.W, p.E, ord_outer);
      A0(fx) = 0.5 * (outer + A0(fx_in)) * A0(xunit);
    }
    if (p.i <= ie) {
      auto Q = [&](int d) { return A(q_j, 0, d, 0); };
      auto D = [&](int d) { return A2(dya, 0, d); };
      double outer = ppm_flux(Q, D, A0(cry), p.j, js, je, p.S, p.N, ord_outer);
      A0(fy) = 0.5 * (outer + A0(fy_in)) * A0(yunit);
    }
  }, "fvtp2d_outer_ppm");
  if (a.nord_col) {
    DelnArgs d;
    d.q = q; d.fx = fx; d.fy = fy; d.mass = a.mass; d.nord_col = a.nord_col; d.damp = a.damp; d.nk = nk;
    d.accumulate = true; d.d2 = nullptr; d.nmax = a.nmax;
    return delnflux(c, d);
  }
  return 0;
}

// ---- delnflux ----------------------------------------------------------------------------------------
// DelnFluxNoSG.__call__ (delnflux.py:1243-1278), optionally followed by DelnFlux's add_diffusive_component /
// diffusive_damp (delnflux.py:1023-1044) when a.accumulate.
int delnflux(Core& c, const DelnArgs& a) {
#ifndef PACE_HOST_EMU
  // production path: one shared-memory tiled kernel (delnflux_fused.cu); PACE_B200_DELNFLUX=unfused for A/B runs
  static const bool fused = [] { const char* v = getenv("PACE_B200_DELNFLUX"); return !(v && std::string(v) == "unfused"); }();
  if (fused) {
    const int rc = delnflux_fused_launch(c, a);
    if (rc >= 0) return rc;
  }
#endif
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const double *del6_u = c.F("damp.del6_u"), *del6_v = c.F("damp.del6_v"), *rarea = c.F("grid.rarea");
  double* d2 = a.d2 ? a.d2 : c.S("deln.d2");
  double *fx2 = a.accumulate ? c.S("deln.fx2") : a.fx, *fy2 = a.accumulate ? c.S("deln.fy2") : a.fy;
  if (!c.err.empty()) return 1;
  const double *q = a.q, *mass = a.mass, *nord_col = a.nord_col, *damp = a.damp;
  double *fx = a.fx, *fy = a.fy;
  const int nmax = a.nmax, nk = a.nk;
  const bool has_mass = mass != nullptr;
  const bool accumulate = a.accumulate;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;

  // K0: d2 = damp*q (or q) on [is-1-nmax..ie+1+nmax]^2 (nord_k == 0: [is-1..ie+1]^2)
  launch_ijk(st, g, Box{is - 1 - nmax, js - 1 - nmax, 0, g.nx + 2 + 2 * nmax, g.ny + 2 + 2 * nmax, nk},
             [=] __device__(const Pt& p) {
               const int nord = (int)nord_col[p.k < 3 ? p.k : 3];
               if (nord == 0 && !(p.i >= is - 1 && p.i <= ie + 1 && p.j >= js - 1 && p.j <= je + 1)) return;
               A0(d2) = has_mass ? A0(q) : damp[p.k] * A0(q);
             });
  // flux kernels: fx2 from x-copied d2, fy2 from y-copied d2 (corner copies only where nord_k > 0)
  auto flux_pass = [&](int nt, bool first, bool last) {
    // fx box: origin (is-nt, js-nt), (nx+1+2nt, ny+2nt); fy box: (nx+2nt, ny+1+2nt)
    launch_ijk(st, g, Box{is - nt, js - nt, 0, g.nx + 1 + 2 * nt, g.ny + 1 + 2 * nt, nk}, [=] __device__(const Pt& p) {
      const int nord = (int)nord_col[p.k < 3 ? p.k : 3];
      const EdgeInfo e = edge_info(g, p);
      const bool in_fx = p.j <= je + nt, in_fy = p.i <= ie + nt;
      if (first) {
        // fx_calc_stencil_nord / fy_calc_stencil_nord: all levels; nord_k == 0 restricted to compute domain
        bool okx = in_fx && (nord != 0 || (p.i >= is && p.i <= ie + 1 && p.j >= js && p.j <= je));
        bool oky = in_fy && (nord != 0 || (p.i >= is && p.i <= ie && p.j >= js && p.j <= je + 1));
        if (okx) {
          double dm = nord > 0 ? rd_cx(d2, p, e, -1, 0) : A(d2, -1, 0, 0);
          double d0 = nord > 0 ? rd_cx(d2, p, e, 0, 0) : A0(d2);
          A0(fx2) = A20(del6_v) * (dm - d0);
        }
        if (oky) {
          double dm = nord > 0 ? rd_cy(d2, p, e, 0, -1) : A(d2, 0, -1, 0);
          double d0 = nord > 0 ? rd_cy(d2, p, e, 0, 0) : A0(d2);
          A0(fy2) = A20(del6_u) * (dm - d0);
        }
      } else if (nord > 0) {
        if (in_fx) A0(fx2) = -A20(del6_v) * (rd_cx(d2, p, e, -1, 0) - rd_cx(d2, p, e, 0, 0));
        if (in_fy) A0(fy2) = -A20(del6_u) * (rd_cy(d2, p, e, 0, -1) - rd_cy(d2, p, e, 0, 0));
      }
      if (last && accumulate) {
        // DelnFlux: add_diffusive_component / diffusive_damp (the reference also touches the unused row
        // je+1 of fx / column ie+1 of fy; only the meaningful interfaces are updated here)
        if (has_mass) {
          if (in_fx) A0(fx) = A0(fx) + 0.5 * damp[p.k] * (A(mass, -1, 0, 0) + A0(mass)) * A0(fx2);
          if (in_fy) A0(fy) = A0(fy) + 0.5 * damp[p.k] * (A(mass, 0, -1, 0) + A0(mass)) * A0(fy2);
        } else {
          if (in_fx) A0(fx) = A0(fx) + A0(fx2);
          if (in_fy) A0(fy) = A0(fy) + A0(fy2);
        }
      }
    });
  };
  flux_pass(nmax, true, nmax == 0);
  for (int n = 0; n < nmax; ++n) {
    const int nt = nmax - 1 - n;
    // d2_highorder_stencil on [is-nt-1..ie+nt+1]^2, levels with nord_k > 0
    launch_ijk(st, g, Box{is - nt - 1, js - nt - 1, 0, g.nx + 2 + 2 * nt, g.ny + 2 + 2 * nt, nk},
               [=] __device__(const Pt& p) {
                 const int nord = (int)nord_col[p.k < 3 ? p.k : 3];
                 if (nord > 0) A0(d2) = (A0(fx2) - A(fx2, 1, 0, 0) + A0(fy2) - A(fy2, 0, 1, 0)) * A20(rarea);
               });
    flux_pass(nt, false, n == nmax - 1);
  }
  return 0;
}

}  // namespace pace
