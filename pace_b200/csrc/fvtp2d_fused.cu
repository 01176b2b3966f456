// Fused, shared-memory tiled fvtp2d for sm_100a: one kernel per FiniteVolumeTransport call.
//
// The straightforward version (transport.cu: 3 kernels through 4 full-size temporaries, every PPM edge value
// recomputed from global memory) moves ~20 field passes per call and was 43 % of the dycore step
// (profiles/r01_baseline_launches_c128.summary.txt).  Here a CTA owns a TX x TY tile of one level:
//   P0  q tile with a 3(+1)-cell apron -> shared memory, twice: seen through copy_corners_x and through
//       copy_corners_y (identical outside the 3x3 cube-corner blocks)
//   P1  inner sweeps from shared memory: y-PPM flux on (TX+6) x (TY+1), x-PPM flux on (TX+1) x (TY+6)
//   P2  q_i on (TX+6) x TY and q_j on TX x (TY+6)
//   P3  outer sweeps on q_i / q_j + final_fluxes -> global fx, fy
// DRAM traffic: q, crx, cry, xfx, yfx (+ mass fluxes) read once (+ apron re-reads served by L2), fx, fy written
// once.  Every expression is the one of transport.cu / the reference (fvtp2d.py:34-120, xppm.py, yppm.py), so the
// results are bit-identical; the order of evaluation per element is unchanged.
#include "corners.cuh"
#include "ppm.cuh"
#include "transport.hpp"

namespace pace {

#ifndef PACE_HOST_EMU

namespace {
constexpr int TX = 32, TY = 8;           // outputs per CTA
constexpr int QW = TX + 7, QH = TY + 7;  // q tile incl. apron: i0-3 .. i0+TX+3, j0-3 .. j0+TY+3

struct FusedArgs {
  const double *q, *crx, *cry, *xfx, *yfx, *xunit, *yunit, *area, *dxa, *dya;
  double *fx, *fy;
  int ord_inner, ord_outer, nk;
};

// OIN / OOUT: |ord| of the inner / outer sweeps as compile-time constants (0 = read it from the arguments)
template <int OIN, int OOUT>
__global__ void __launch_bounds__(TX* TY) k_fvtp2d_fused(Geo g, FusedArgs a) {
  const int ord_in = OIN ? OIN : a.ord_inner, ord_out = OOUT ? OOUT : a.ord_outer;
  __shared__ double qx[QH][QW];          // q seen through copy_corners_x
  __shared__ double qy[QH][QW];          // q seen through copy_corners_y
  __shared__ double fyi[TY + 1][TX + 6]; // inner y flux at (i0-3+ii, j0+jj)
  __shared__ double fxi[TY + 6][TX + 1]; // inner x flux at (i0+ii, j0-3+jj)
  __shared__ double qi[TY][TX + 6];      // q_i at (i0-3+ii, j0+jj)
  __shared__ double qj[TY + 6][TX];      // q_j at (i0+ii, j0-3+jj)

  const int k = blockIdx.z % a.nk, t = blockIdx.z / a.nk;
  const int i0 = g.is + blockIdx.x * TX, j0 = g.js + blockIdx.y * TY;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;
  const unsigned ed = edge_of(g, t);
  const bool W = ed & EDGE_W, E = ed & EDGE_E, S = ed & EDGE_S, N = ed & EDGE_N;
  const EdgeInfo e{is, ie, js, je, W, E, S, N};
  const long long base3 = (long long)t * g.st3 + (long long)k * g.sk;
  const int base2 = t * g.sk;
  const int tid = threadIdx.y * TX + threadIdx.x;
  const int nthr = TX * TY;
  const int imax = ie + 3, jmax = je + 3;  // last halo cell
  auto G3 = [&](const double* f, int i, int j) { return f[base3 + j * g.sj + i]; };
  auto G2 = [&](const double* f, int i, int j) { return f[base2 + j * g.sj + i]; };

  // P0: q tile
  for (int n = tid; n < QW * QH; n += nthr) {
    const int li = n % QW, lj = n / QW;
    int i = i0 - 3 + li, j = j0 - 3 + lj;
    double vx = 0.0, vy = 0.0;
    if (i <= imax && j <= jmax) {  // imax = ie+3: the corner maps are defined for the 3-cell halo only
      int xi = i, xj = j, yi = i, yj = j;
      copyx_src(e, xi, xj);
      copyy_src(e, yi, yj);
      vx = G3(a.q, xi, xj);
      vy = (yi == xi && yj == xj) ? vx : G3(a.q, yi, yj);
    }
    qx[lj][li] = vx;
    qy[lj][li] = vy;
  }
  __syncthreads();

  // P1: inner sweeps
  for (int n = tid; n < (TX + 6) * (TY + 1); n += nthr) {
    const int li = n % (TX + 6), lj = n / (TX + 6);
    const int i = i0 - 3 + li, j = j0 + lj;
    double v = 0.0;
    if (i <= ie + 3 && j <= je + 1) {
      auto Q = [&](int d) { return qy[lj + 3 + d][li]; };
      auto D = [&](int d) { return G2(a.dya, i, j + d); };
      v = ppm_flux_fast(Q, D, G3(a.cry, i, j), j, js, je, S, N, ord_in);
    }
    fyi[lj][li] = v;
  }
  for (int n = tid; n < (TX + 1) * (TY + 6); n += nthr) {
    const int li = n % (TX + 1), lj = n / (TX + 1);
    const int i = i0 + li, j = j0 - 3 + lj;
    double v = 0.0;
    if (i <= ie + 1 && j <= je + 3) {
      auto Q = [&](int d) { return qx[lj][li + 3 + d]; };
      auto D = [&](int d) { return G2(a.dxa, i + d, j); };
      v = ppm_flux_fast(Q, D, G3(a.crx, i, j), i, is, ie, W, E, ord_in);
    }
    fxi[lj][li] = v;
  }
  __syncthreads();

  // P2: q_i, q_j (fvtp2d.py:34-77); q at non-corner cells: qx == qy == q
  for (int n = tid; n < (TX + 6) * TY; n += nthr) {
    const int li = n % (TX + 6), lj = n / (TX + 6);
    const int i = i0 - 3 + li, j = j0 + lj;
    double v = 0.0;
    if (i <= ie + 3 && j <= je) {
      const double yf0 = G3(a.yfx, i, j), yf1 = G3(a.yfx, i, j + 1), ar = G2(a.area, i, j);
      const double fyy0 = yf0 * fyi[lj][li], fyy1 = yf1 * fyi[lj + 1][li];
      v = (qy[lj + 3][li] * ar + fyy0 - fyy1) / (ar + yf0 - yf1);
    }
    qi[lj][li] = v;
  }
  for (int n = tid; n < TX * (TY + 6); n += nthr) {
    const int li = n % TX, lj = n / TX;
    const int i = i0 + li, j = j0 - 3 + lj;
    double v = 0.0;
    if (i <= ie && j <= je + 3) {
      const double xf0 = G3(a.xfx, i, j), xf1 = G3(a.xfx, i + 1, j), ar = G2(a.area, i, j);
      const double fx10 = xf0 * fxi[lj][li], fx11 = xf1 * fxi[lj][li + 1];
      const double area_with_x_flux = ar + xf0 - xf1;
      v = (qx[lj][li + 3] * ar + fx10 - fx11) / area_with_x_flux;
    }
    qj[lj][li] = v;
  }
  __syncthreads();

  // P3: outer sweeps + final_fluxes (fvtp2d.py:80-119)
  {
    const int li = threadIdx.x, lj = threadIdx.y;
    const int i = i0 + li, j = j0 + lj;
    if (i <= ie + 1 && j <= je) {
      auto Q = [&](int d) { return qi[lj][li + 3 + d]; };
      auto D = [&](int d) { return G2(a.dxa, i + d, j); };
      const double outer = ppm_flux_fast(Q, D, G3(a.crx, i, j), i, is, ie, W, E, ord_out);
      a.fx[base3 + j * g.sj + i] = 0.5 * (outer + fxi[lj + 3][li]) * G3(a.xunit, i, j);
    }
    if (i <= ie && j <= je + 1) {
      auto Q = [&](int d) { return qj[lj + 3 + d][li]; };
      auto D = [&](int d) { return G2(a.dya, i, j + d); };
      const double outer = ppm_flux_fast(Q, D, G3(a.cry, i, j), j, js, je, S, N, ord_out);
      a.fy[base3 + j * g.sj + i] = 0.5 * (outer + fyi[lj][li + 3]) * G3(a.yunit, i, j);
    }
  }
}
}  // namespace

int fvtp2d_fused_launch(Core& c, const Fvtp2dArgs& a) {
  const Geo g = c.g;
  FusedArgs f;
  f.q = a.q; f.crx = a.crx; f.cry = a.cry; f.xfx = a.xfx; f.yfx = a.yfx;
  f.xunit = a.x_mass_flux ? a.x_mass_flux : a.xfx;
  f.yunit = a.y_mass_flux ? a.y_mass_flux : a.yfx;
  f.area = c.F("grid.area"); f.dxa = c.F("grid.dxa"); f.dya = c.F("grid.dya");
  f.fx = a.fx; f.fy = a.fy;
  f.ord_outer = a.hord;
  f.ord_inner = (a.hord == 10) ? 8 : a.hord;
  f.nk = a.nk;
  if (!c.err.empty()) return 1;
  dim3 blk(TX, TY, 1);
  // the fx / fy output domains are (nx+1) x ny and nx x (ny+1): tile the (nx+1) x (ny+1) point set
  dim3 grd((g.nx + 1 + TX - 1) / TX, (g.ny + 1 + TY - 1) / TY, a.nk * g.nt);
  LaunchScope scope(c.stream, "fvtp2d_fused");
  const int oi = f.ord_inner < 0 ? -f.ord_inner : f.ord_inner, oo = f.ord_outer < 0 ? -f.ord_outer : f.ord_outer;
  if (oi == 5 && oo == 5) k_fvtp2d_fused<5, 5><<<grd, blk, 0, c.stream>>>(g, f);
  else if (oi == 6 && oo == 6) k_fvtp2d_fused<6, 6><<<grd, blk, 0, c.stream>>>(g, f);
  else if (oi == 8 && oo == 8) k_fvtp2d_fused<8, 8><<<grd, blk, 0, c.stream>>>(g, f);
  else if (oi == 8 && oo == 10) k_fvtp2d_fused<8, 10><<<grd, blk, 0, c.stream>>>(g, f);
  else k_fvtp2d_fused<0, 0><<<grd, blk, 0, c.stream>>>(g, f);
  return 0;
}

#else  // host emulation keeps the straightforward three-kernel version (transport.cu)
int fvtp2d_fused_launch(Core&, const Fvtp2dArgs&) { return -1; }
#endif

}  // namespace pace
