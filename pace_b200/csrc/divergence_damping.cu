// Divergence damping on the D grid (sponge-layer 2nd order + higher-order nord loop + Smagorinsky term).
//
// Reference: pyFV3/stencils/divergence_damping.py:25-297 (stencils), :549-713 (__call__ order);
//            NDSL/ndsl/stencils/corners.py:545-717 (FillCornersBGrid), :987-1129 (fill_corners_dgrid_defn).
// Levels k < kstart (first level with nord > 0) take the "zero order" path, levels >= kstart the nord loop
// (divergence_damping.py:372-380).  ~35 reference launches -> 3 + 6*nord + 3 here.
#include "core.hpp"
#include "corners.cuh"

namespace pace {

int a2b_ord4(Core& c, const double* qin, double* qout, int k0, int nk, bool replace);

namespace {

// B-grid (corner-point) halo-corner fills, in absolute indices with ie1 = ie+1, je1 = je+1 the last
// corner point of the compute domain; a, b = 1..3 distances from the tile edge:
//   x:  (is-a, js-b) <- (is-b, js+a)    (ie1+a, js-b) <- (ie1+b, js+a)
//       (is-a, je1+b) <- (is-b, je1-a)  (ie1+a, je1+b) <- (ie1+b, je1-a)
//   y:  (is-a, js-b) <- (is+b, js-a)    (ie1+a, js-b) <- (ie1-b, js-a)
//       (is-a, je1+b) <- (is+b, je1+a)  (ie1+a, je1+b) <- (ie1-b, je1+a)
__device__ __forceinline__ void bgrid_fill_cell(const Geo& g, double* q, bool ydir, int t, int k, int c, int a, int b) {
  const unsigned ed = edge_of(g, t);
  bool W = ed & EDGE_W, E = ed & EDGE_E, S = ed & EDGE_S, N = ed & EDGE_N;
  const int is = g.is, ie1 = g.ie + 1, js = g.js, je1 = g.je + 1;
  int i, j, si, sj;
  bool on;
  switch (c) {
    case 0: on = W && S; i = is - a; j = js - b;
      if (!ydir) { si = is - b; sj = js + a; } else { si = is + b; sj = js - a; }
      break;
    case 1: on = E && S; i = ie1 + a; j = js - b;
      if (!ydir) { si = ie1 + b; sj = js + a; } else { si = ie1 - b; sj = js - a; }
      break;
    case 2: on = W && N; i = is - a; j = je1 + b;
      if (!ydir) { si = is - b; sj = je1 - a; } else { si = is + b; sj = je1 + a; }
      break;
    default: on = E && N; i = ie1 + a; j = je1 + b;
      if (!ydir) { si = ie1 + b; sj = je1 - a; } else { si = ie1 - b; sj = je1 + a; }
      break;
  }
  if (!on) return;
  long long base = (long long)t * g.st3 + (long long)k * g.sk;
  q[base + j * g.sj + i] = q[base + sj * g.sj + si];
}

// fill_corners_dgrid_defn table: {corner (0 SW,1 NE,2 NW,3 SE), target di, dj (from the corner base point),
// which output (0 x_out <- y_in, 1 y_out <- x_in), source offset from the target, signed}
struct DgEntry { signed char corner, ti, tj, out, si, sj, sgn; };
__device__ const DgEntry DG_TABLE[] = {
    // SW, base (is, js), mysign
    {0, -1, -1, 0, 0, 1, 1}, {0, -1, -1, 1, 1, 0, 1}, {0, -1, -2, 0, -1, 2, 1}, {0, -1, -2, 1, 2, 1, 1},
    {0, -1, -3, 0, -2, 3, 1}, {0, -1, -3, 1, 3, 2, 1}, {0, -2, -1, 0, 1, 2, 1}, {0, -2, -1, 1, 2, -1, 1},
    {0, -2, -2, 0, 0, 3, 1}, {0, -2, -2, 1, 3, 0, 1}, {0, -2, -3, 0, -1, 4, 1}, {0, -2, -3, 1, 4, 1, 1},
    {0, -3, -1, 0, 2, 3, 1}, {0, -3, -1, 1, 3, -2, 1}, {0, -3, -2, 0, 1, 4, 1}, {0, -3, -2, 1, 4, -1, 1},
    {0, -3, -3, 0, 0, 5, 1}, {0, -3, -3, 1, 5, 0, 1},
    // NE, base (ie, je), mysign
    {1, 1, 2, 0, 1, -2, 1}, {1, 2, 1, 1, -2, 1, 1}, {1, 1, 3, 0, 2, -3, 1}, {1, 2, 2, 1, -3, 0, 1},
    {1, 2, 2, 0, 0, -3, 1}, {1, 1, 4, 0, 3, -4, 1}, {1, 2, 3, 1, -4, -1, 1}, {1, 2, 3, 0, 1, -4, 1},
    {1, 3, 1, 1, -3, 2, 1}, {1, 3, 2, 1, -4, 1, 1}, {1, 3, 2, 0, -1, -4, 1}, {1, 2, 4, 0, 2, -5, 1},
    {1, 3, 3, 1, -5, 0, 1}, {1, 3, 3, 0, 0, -5, 1}, {1, 4, 1, 1, -4, 3, 1}, {1, 4, 2, 1, -5, 2, 1},
    {1, 3, 4, 0, 1, -6, 1}, {1, 4, 3, 1, -6, 1, 1},
    // NW, base (is, je), unsigned
    {2, -1, 2, 0, 0, -2, 0}, {2, -1, 2, 1, 2, 0, 0}, {2, -1, 1, 1, 1, 1, 0}, {2, -1, 3, 0, -1, -3, 0},
    {2, -1, 3, 1, 3, -1, 0}, {2, -1, 4, 0, -2, -4, 0}, {2, -2, 2, 0, 1, -3, 0}, {2, -2, 2, 1, 3, 1, 0},
    {2, -2, 1, 1, 2, 2, 0}, {2, -2, 3, 0, 0, -4, 0}, {2, -2, 3, 1, 4, 0, 0}, {2, -2, 4, 0, -1, -5, 0},
    {2, -3, 2, 0, 2, -4, 0}, {2, -3, 2, 1, 4, 2, 0}, {2, -3, 1, 1, 3, 3, 0}, {2, -3, 3, 0, 1, -5, 0},
    {2, -3, 3, 1, 5, 1, 0}, {2, -3, 4, 0, 0, -6, 0},
    // SE, base (ie, js), unsigned
    {3, 1, -1, 0, 1, 1, 0}, {3, 2, -1, 1, -2, 0, 0}, {3, 2, -1, 0, 0, 2, 0}, {3, 1, -2, 0, 2, 2, 0},
    {3, 2, -2, 1, -3, 1, 0}, {3, 2, -2, 0, 1, 3, 0}, {3, 1, -3, 0, 3, 3, 0}, {3, 2, -3, 1, -4, 2, 0},
    {3, 2, -3, 0, 2, 4, 0}, {3, 3, -1, 1, -3, -1, 0}, {3, 3, -1, 0, -1, 3, 0}, {3, 3, -2, 1, -4, 0, 0},
    {3, 3, -2, 0, 0, 4, 0}, {3, 3, -3, 1, -5, 1, 0}, {3, 3, -3, 0, 1, 5, 0}, {3, 4, -1, 1, -4, -2, 0},
    {3, 4, -2, 1, -5, -1, 0}, {3, 4, -3, 1, -6, 0, 0},
};
constexpr int DG_N = sizeof(DG_TABLE) / sizeof(DgEntry);

__device__ __forceinline__ void dgrid_fill_entry(const Geo& g, double* x, double* y, double mysign, int t, int k, int n) {
  const DgEntry e = DG_TABLE[n];
  const unsigned ed = edge_of(g, t);
  bool W = ed & EDGE_W, E = ed & EDGE_E, S = ed & EDGE_S, N = ed & EDGE_N;
  int bi, bj;
  bool on;
  switch (e.corner) {
    case 0: on = W && S; bi = g.is; bj = g.js; break;
    case 1: on = E && N; bi = g.ie; bj = g.je; break;
    case 2: on = W && N; bi = g.is; bj = g.je; break;
    default: on = E && S; bi = g.ie; bj = g.js; break;
  }
  if (!on) return;
  const int i = bi + e.ti, j = bj + e.tj;
  long long base = (long long)t * g.st3 + (long long)k * g.sk;
  long long dst = base + j * g.sj + i, src = base + (j + e.sj) * g.sj + (i + e.si);
  const double* in = e.out == 0 ? y : x;
  double* out = e.out == 0 ? x : y;
  out[dst] = e.sgn ? mysign * in[src] : in[src];
}

#ifndef PACE_HOST_EMU
__global__ void k_bgrid_fill(Geo g, double* q, bool ydir, int k0, int nk) {
  int k = k0 + blockIdx.x % nk, t = blockIdx.x / nk, id = threadIdx.x;
  if (id < 36) bgrid_fill_cell(g, q, ydir, t, k, id / 9, (id % 9) / 3 + 1, id % 3 + 1);
}
__global__ void k_dgrid_fill(Geo g, double* x, double* y, double mysign, int k0, int nk) {
  int k = k0 + blockIdx.x % nk, t = blockIdx.x / nk, id = threadIdx.x;
  if (id < DG_N) dgrid_fill_entry(g, x, y, mysign, t, k, id);
}
#endif

void bgrid_fill(Core& c, double* q, bool ydir, int k0, int nk) {
#ifndef PACE_HOST_EMU
  ++launch_counter();
  k_bgrid_fill<<<nk * c.g.nt, 64, 0, c.stream>>>(c.g, q, ydir, k0, nk);
#else
  for (int t = 0; t < c.g.nt; ++t)
    for (int k = k0; k < k0 + nk; ++k)
      for (int id = 0; id < 36; ++id) bgrid_fill_cell(c.g, q, ydir, t, k, id / 9, (id % 9) / 3 + 1, id % 3 + 1);
#endif
}
void dgrid_fill(Core& c, double* x, double* y, double mysign, int k0, int nk) {
#ifndef PACE_HOST_EMU
  ++launch_counter();
  k_dgrid_fill<<<nk * c.g.nt, 96, 0, c.stream>>>(c.g, x, y, mysign, k0, nk);
#else
  for (int t = 0; t < c.g.nt; ++t)
    for (int k = k0; k < k0 + nk; ++k)
      for (int id = 0; id < DG_N; ++id) dgrid_fill_entry(c.g, x, y, mysign, t, k, id);
#endif
}

}  // namespace

// DivergenceDamping.__call__(u, v, va, damped_rel_vort_bgrid, ua, divg_d, vc, uc, delpc, ke, rel_vort_agrid, dt)
int divergence_damping(Core& c, double* vort_b, double* delpc, double* ke, const double* vort_a, double dt) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const double *u = c.F("u"), *v = c.F("v"), *ua = c.F("ua"), *va = c.F("va");
  double *uc = c.F("uc"), *vc = c.F("vc"), *divg_d = c.F("ad.divgd");
  const double *cosa_u = c.F("grid.cosa_u"), *cosa_v = c.F("grid.cosa_v"), *sina_u = c.F("grid.sina_u"),
               *sina_v = c.F("grid.sina_v"), *dxc = c.F("grid.dxc"), *dyc = c.F("grid.dyc"),
               *sin_sg1 = c.F("grid.sin_sg1"), *sin_sg2 = c.F("grid.sin_sg2"), *sin_sg3 = c.F("grid.sin_sg3"),
               *sin_sg4 = c.F("grid.sin_sg4"), *rarea_c = c.F("grid.rarea_c"), *divg_u = c.F("damp.divg_u"),
               *divg_v = c.F("damp.divg_v"), *d2_bg = c.F("col.d2_divg"), *nord_col = c.F("col.nord");
  if (!c.err.empty()) return 1;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, nz = g.nz;
  const double da_min_c = c.cfg.da_min_c, dddmp = c.cfg.dddmp;
  // first level with nord > 0 (host copy of the column is tiny; mirrored from the config instead)
  // DivergenceDamping.__init__ (divergence_damping.py:360-371): first level with nord > 0
  const std::vector<double>* hn = c.H("col.nord");
  if (!hn) return 1;
  int kstart = 0, nonzero_nord = c.cfg.nord;
  for (int k = 0; k < nz && k < (int)hn->size(); ++k)
    if ((*hn)[k] > 0) {
      kstart = k;
      nonzero_nord = (int)(*hn)[k];
      break;
    }
  (void)nord_col;
  const int nkh = nz - kstart;

  if (kstart > 0) {
    // compute_u_contra_dyc, compute_v_contra_dxc, delpc_computation, damping  (k < kstart)
    launch_ijk(st, g, Box{is, js, 0, g.nx + 1, g.ny + 1, kstart}, [=] __device__(const Pt& p) {
      const int i = p.i, j = p.j;
      auto udyc = [&](int di) {  // u_contra_dyc at (i+di, j)
        double vc_from_va = 0.5 * (A(va, di, -1, 0) + A(va, di, 0, 0));
        double uct = (A(u, di, 0, 0) - vc_from_va * A2(cosa_v, di, 0)) * A2(sina_v, di, 0);
        if ((p.S && j == js) || (p.N && j == je + 1))
          uct = A(vc, di, 0, 0) > 0 ? A(u, di, 0, 0) * A2(sin_sg4, di, -1) : A(u, di, 0, 0) * A2(sin_sg2, di, 0);
        return uct * A2(dyc, di, 0);
      };
      auto vdxc = [&](int dj) {  // v_contra_dxc at (i, j+dj)
        double uc_from_ua = 0.5 * (A(ua, -1, dj, 0) + A(ua, 0, dj, 0));
        double vct = (A(v, 0, dj, 0) - uc_from_ua * A2(cosa_u, 0, dj)) * A2(sina_u, 0, dj);
        if ((p.W && i == is) || (p.E && i == ie + 1))
          vct = A(uc, 0, dj, 0) > 0 ? A(v, 0, dj, 0) * A2(sin_sg3, -1, dj) : A(v, 0, dj, 0) * A2(sin_sg1, 0, dj);
        return vct * A2(dxc, 0, dj);
      };
      double vm = vdxc(-1), v0 = vdxc(0), um = udyc(-1), u0 = udyc(0);
      double d = vm - v0 + um - u0;
      if (((p.W && i == is) || (p.E && i == ie + 1)) && (p.S && j == js)) d = d - vm;
      if (((p.W && i == is) || (p.E && i == ie + 1)) && (p.N && j == je + 1)) d = d + v0;
      d = A20(rarea_c) * d;
      A0(delpc) = d;
      double delpcdt = d * dt;
      double damp = da_min_c * fv_max(d2_bg[p.k], fv_min(0.2, dddmp * fabs(delpcdt)));
      double vo = damp * d;
      A0(vort_b) = vo;
      A0(ke) = A0(ke) + vo;
    });
  }
  // copy_computeplus(divg_d -> delpc), k >= kstart
  launch_ijk(st, g, Box{is, js, kstart, g.nx + 1, g.ny + 1, nkh}, [=] __device__(const Pt& p) { A0(delpc) = A0(divg_d); });
  for (int n = 0; n < nonzero_nord; ++n) {
    const int nt = nonzero_nord - (n + 1);
    const bool fillc = (n + 1 != nonzero_nord);
    if (fillc) bgrid_fill(c, divg_d, false, kstart, nkh);
    // vc_from_divg on origin (is-nt-1, js-nt), domain (nx+2nt+2, ny+2nt+1)
    launch_ijk(st, g, Box{is - nt - 1, js - nt, kstart, g.nx + 2 * nt + 2, g.ny + 2 * nt + 1, nkh},
               [=] __device__(const Pt& p) { A0(vc) = (A(divg_d, 1, 0, 0) - A0(divg_d)) * A20(divg_u); });
    if (fillc) bgrid_fill(c, divg_d, true, kstart, nkh);
    // uc_from_divg on origin (is-nt, js-nt-1), domain (nx+2nt+1, ny+2nt+2)
    launch_ijk(st, g, Box{is - nt, js - nt - 1, kstart, g.nx + 2 * nt + 1, g.ny + 2 * nt + 2, nkh},
               [=] __device__(const Pt& p) { A0(uc) = (A(divg_d, 0, 1, 0) - A0(divg_d)) * A20(divg_v); });
    if (fillc) dgrid_fill(c, vc, uc, -1.0, kstart, nkh);
    // redo_divg_d on origin (is-nt, js-nt), domain (nx+2nt+1, ny+2nt+1)
    launch_ijk(st, g, Box{is - nt, js - nt, kstart, g.nx + 2 * nt + 1, g.ny + 2 * nt + 1, nkh}, [=] __device__(const Pt& p) {
      const int i = p.i, j = p.j;
      double d = A(uc, 0, -1, 0) - A0(uc) + A(vc, -1, 0, 0) - A0(vc);
      if (((p.W && i == is) || (p.E && i == ie + 1)) && (p.S && j == js)) d = d - A(uc, 0, -1, 0);
      if (((p.W && i == is) || (p.E && i == ie + 1)) && (p.N && j == je + 1)) d = d + A0(uc);
      A0(divg_d) = d * A20(rarea_c);  // do_adjustment = not stretched_grid
    });
  }
  const double absdt = fabs(dt);
  const double dd8 = std::pow(da_min_c * c.cfg.d4_bg, (double)(nonzero_nord + 1));
  if (dddmp >= 1e-5 && a2b_ord4(c, vort_a, vort_b, kstart, nkh, false)) return 1;
  // smagorinsky_diffusion_approx (x**2.0 and x**0.5 evaluate as x*x and sqrt(x) in the numpy backend; vort = 0
  // where dddmp < 1e-5) and damping_nord_highorder_stencil on the same points: one kernel, vort stays in a register
  const bool smag = dddmp >= 1e-5;
  if (!smag) launch_ijk(st, g, Box{0, 0, kstart, g.nx + 7, g.ny + 7, nkh}, [=] __device__(const Pt& p) { A0(vort_b) = 0.0; });
  launch_ijk(st, g, Box{is, js, kstart, g.nx + 1, g.ny + 1, nkh}, [=] __device__(const Pt& p) {
    const double vort = smag ? absdt * fv_sqrt(A0(delpc) * A0(delpc) + A0(vort_b) * A0(vort_b)) : A0(vort_b);
    double damp = da_min_c * fv_max(d2_bg[p.k], fv_min(0.2, dddmp * fabs(vort)));
    double vo = damp * A0(delpc) + dd8 * A0(divg_d);
    A0(vort_b) = vo;
    A0(ke) = A0(ke) + vo;
  });
  return 0;
}

}  // namespace pace
