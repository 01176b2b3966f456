// Divergence damping on the D grid (sponge-layer 2nd order + higher-order nord loop + Smagorinsky term).
//
// Reference: pyFV3/stencils/divergence_damping.py:25-297 (stencils), :549-713 (__call__ order);
//            NDSL/ndsl/stencils/corners.py:545-717 (FillCornersBGrid), :987-1129 (fill_corners_dgrid_defn).
// Levels k < kstart (first level with nord > 0) take the "zero order" path, levels >= kstart the nord loop
// (divergence_damping.py:372-380).  ~35 reference launches -> 3 + 6*nord + 3 here.
#include <cstdlib>
#include <string>

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
template <class At>
__device__ __forceinline__ void bgrid_fill_cell_at(const Geo& g, At at, bool ydir, int t, int c, int a, int b) {
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
  at(i, j) = at(si, sj);
}
__device__ __forceinline__ void bgrid_fill_cell(const Geo& g, double* q, bool ydir, int t, int k, int c, int a, int b) {
  const long long base = (long long)t * g.st3 + (long long)k * g.sk;
  const int row = g.sj;
  bgrid_fill_cell_at(g, [=](int i, int j) -> double& { return q[base + j * row + i]; }, ydir, t, c, a, b);
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

// AtX / AtY: (i, j) -> reference to the x / y component; corner < 0: every corner of the subdomain, else only that one
template <class AtX, class AtY>
__device__ __forceinline__ void dgrid_fill_entry_at(const Geo& g, AtX ax, AtY ay, double mysign, int t, int n, int corner) {
  const DgEntry e = DG_TABLE[n];
  if (corner >= 0 && e.corner != corner) return;
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
  const double in = e.out == 0 ? ay(i + e.si, j + e.sj) : ax(i + e.si, j + e.sj);
  const double val = e.sgn ? mysign * in : in;
  if (e.out == 0) ax(i, j) = val;
  else ay(i, j) = val;
}
__device__ __forceinline__ void dgrid_fill_entry(const Geo& g, double* x, double* y, double mysign, int t, int k, int n) {
  const long long base = (long long)t * g.st3 + (long long)k * g.sk;
  const int row = g.sj;
  dgrid_fill_entry_at(g, [=](int i, int j) -> double& { return x[base + j * row + i]; },
                      [=](int i, int j) -> double& { return y[base + j * row + i]; }, mysign, t, n, -1);
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


#ifndef PACE_HOST_EMU
// ---- the nord loop as one row-marching kernel ---------------------------------------------------------------------
// Reference loop (divergence_damping.py:586-655), per iteration n = 0 .. N-1 with d_0 = divg_d:
//   vc_n(i,j) = (d_n(i+1,j) - d_n(i,j)) * divg_u(i,j)      uc_n(i,j) = (d_n(i,j+1) - d_n(i,j)) * divg_v(i,j)
//   d_{n+1}(i,j) = (uc_n(i,j-1) - uc_n(i,j) + vc_n(i-1,j) - vc_n(i,j)) * rarea_c(i,j)
// 3 launches and 8 field passes per iteration as separate stencils (24 + 2 for the copy to delpc).  Here a WARP owns a
// strip of 26 columns (lane <-> i, 3-column apron each side, neighbours by shuffle) and a chunk of rows and marches:
// when row R of divg_d arrives, iteration n advances to row R-n (vc) / R-n-1 (uc, d_{n+1}); everything is carried in
// registers.  It reads divg_d once and writes delpc (= divg_d, copy_computeplus), the last iteration's vc and uc and
// d_N — the latter to a work array: divg_d itself is still being read by the neighbouring strips and chunks; the
// Smagorinsky kernel that follows reads d_N from there and stores it to divg_d.  Same expressions, same bits.
// The corner fills of iterations n < N-1 (FillCornersBGrid, fill_corners_dgrid) only reach cells within a few points
// of a cube corner: those cells are left to k_dd_corner below.  What the reference leaves in the HALO of divg_d, uc
// and vc (results of the earlier, wider iterations) is not reproduced; nothing reads it.
constexpr int DSW = 26;
constexpr int DD_CORNER_R = 5;      // cells within this many points of a cube-corner point belong to k_dd_corner
constexpr unsigned DFULL = 0xffffffffu;

struct DdMarchArgs {
  const double *divg, *divg_u, *divg_v, *rarea_c;
  double *delpc, *dout, *uc, *vc;
  int k0, nk, N;
  int nstrips, nchunks, rows;
  long long items;
};

// is (i, j) owned by the corner kernel of subdomain t?
__device__ __forceinline__ bool dd_corner_cell(const Geo& g, unsigned ed, int i, int j) {
  const int ie1 = g.ie + 1, je1 = g.je + 1;
  const bool nw = (ed & EDGE_W) && i - g.is <= DD_CORNER_R, ne = (ed & EDGE_E) && ie1 - i <= DD_CORNER_R;
  const bool ns = (ed & EDGE_S) && j - g.js <= DD_CORNER_R, nn = (ed & EDGE_N) && je1 - j <= DD_CORNER_R;
  return (nw || ne) && (ns || nn);
}

__global__ void __launch_bounds__(128, 4) k_dd_march(Geo g, DdMarchArgs a) {
  const long long item = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (item >= a.items) return;
  const int lane = threadIdx.x & 31;
  const int strip = (int)(item % a.nstrips);
  long long r1 = item / a.nstrips;
  const int chunk = (int)(r1 % a.nchunks);
  r1 /= a.nchunks;
  const int k = a.k0 + (int)(r1 % a.nk), t = (int)(r1 / a.nk);
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj, N = a.N;
  const unsigned ed = edge_of(g, t);
  const int i = is - 1 + strip * DSW - 3 + lane;   // strips cover i in [is-1, ie+1] (vc needs is-1)
  const int ic = min(max(i, 0), sj - 1);
  const int j_lo = js - 1 + chunk * a.rows, j_hi = min(j_lo + a.rows - 1, je + 1);  // rows delivered: [js-1, je+1]
  const int base3 = (int)((long long)t * g.st3 + (long long)k * g.sk);
  const int c3 = base3 + ic, c2 = t * g.sk + ic;
  const bool own = lane >= 3 && lane <= 28;
  // delivery of iteration N-1: vc on i in [is-1, ie+1], j in [js, je+1]; uc on i in [is, ie+1], j in [js-1, je+1];
  // d_N and delpc on [is, ie+1] x [js, je+1]
  const bool st_v = own && i >= is - 1 && i <= ie + 1, st_u = own && i >= is && i <= ie + 1;
  // rows: d_0 row R arrives; vc_n at row R-n; uc_n and d_{n+1} at row R-n-1.  Row j of d_N needs d_0 rows j-N .. j+N.
  const int R0 = max(0, j_lo - N - 1), R1 = min(j_hi + N + 1, g.ny + 6);
  double dprev[3] = {0.0, 0.0, 0.0};   // d_n at the row below its newest one (n = 0, 1, 2)
  double vprev[3] = {0.0, 0.0, 0.0};   // vc_n at its previous row
  double uprev[3] = {0.0, 0.0, 0.0};   // uc_n at its previous row
  double mu[4] = {0.0, 0.0, 0.0, 0.0}, mv[4] = {0.0, 0.0, 0.0, 0.0}, mr[4] = {0.0, 0.0, 0.0, 0.0};  // metric rows R .. R-3
  auto row = [&](int R) { return min(R, R1) * sj; };
  double d_a = a.divg[c3 + row(R0)], d_b = a.divg[c3 + row(R0 + 1)];
  double u_a = a.divg_u[c2 + row(R0)], u_b = a.divg_u[c2 + row(R0 + 1)];
  double v_a = a.divg_v[c2 + row(R0)], v_b = a.divg_v[c2 + row(R0 + 1)];
  double r_a = a.rarea_c[c2 + row(R0)], r_b = a.rarea_c[c2 + row(R0 + 1)];
  for (int R = R0; R <= R1; ++R) {
    const double d0 = d_a;
    mu[3] = mu[2]; mu[2] = mu[1]; mu[1] = mu[0]; mu[0] = u_a;
    mv[3] = mv[2]; mv[2] = mv[1]; mv[1] = mv[0]; mv[0] = v_a;
    mr[3] = mr[2]; mr[2] = mr[1]; mr[1] = mr[0]; mr[0] = r_a;
    d_a = d_b; u_a = u_b; v_a = v_b; r_a = r_b;
    {
      const int rin = row(R + 2);
      d_b = a.divg[c3 + rin]; u_b = a.divg_u[c2 + rin]; v_b = a.divg_v[c2 + rin]; r_b = a.rarea_c[c2 + rin];
      if ((lane & 3) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.divg + c3 + row(R + 8)));  // see delnflux_fused.cu
    }
    // iteration n works on d_n at row R - n; its results: vc_n(R-n), uc_n(R-n-1), d_{n+1}(R-n-1)
    double dn = d0;
    double vc_out = 0.0, uc_out = 0.0, d_out = 0.0;
#pragma unroll
    for (int n = 0; n < 3; ++n) {
      if (n < N) {
        const double vc_n = (__shfl_down_sync(DFULL, dn, 1) - dn) * mu[n];            // row R-n
        const double uc_n = (dn - dprev[n]) * mv[n + 1];                               // row R-n-1
        const double vp = vprev[n];                                                    // vc_n at row R-n-1
        const double dnext = (uprev[n] - uc_n + __shfl_up_sync(DFULL, vp, 1) - vp) * mr[n + 1];  // d_{n+1}(R-n-1)
        if (n == N - 1) { vc_out = vc_n; uc_out = uc_n; d_out = dnext; }
        dprev[n] = dn; vprev[n] = vc_n; uprev[n] = uc_n;
        dn = dnext;
      }
    }
    // deliveries of this step: vc at row jv = R-N+1, uc / d_N / delpc at row ju = R-N
    const int jv = R - N + 1, ju = R - N;
    if (jv >= max(j_lo, js) && jv <= j_hi && st_v && !dd_corner_cell(g, ed, i, jv)) a.vc[base3 + jv * sj + i] = vc_out;
    if (ju >= j_lo && ju <= j_hi && st_u && !dd_corner_cell(g, ed, i, ju)) {
      const int o = base3 + ju * sj + i;
      a.uc[o] = uc_out;
      if (ju >= js) a.dout[o] = d_out;
    }
    // copy_computeplus: delpc = divg_d on [is, ie+1] x [js, je+1] (row R itself)
    if (R >= max(j_lo, js) && R <= j_hi && st_u) a.delpc[base3 + R * sj + i] = d0;
  }
}

// The cells next to a cube corner: the same loop on a 24 x 24 window in shared memory, with the corner fills.
constexpr int DCW = 24, DCH = 12;  // window width, distance of its low edge from the corner point

struct DdCornerArgs {
  const double *divg, *divg_u, *divg_v, *rarea_c;
  double *dout, *uc, *vc;
  int k0, nk, N;
};

__global__ void __launch_bounds__(256) k_dd_corner(Geo g, DdCornerArgs a) {
  __shared__ double D[DCW][DCW], U[DCW][DCW], V[DCW][DCW];
  const int c = blockIdx.x & 3;                       // 0 SW, 1 SE, 2 NW, 3 NE (bgrid_fill numbering)
  const int kk = (blockIdx.x >> 2) % a.nk, t = (blockIdx.x >> 2) / a.nk, k = a.k0 + kk;
  const unsigned ed = edge_of(g, t);
  const bool W = ed & EDGE_W, E = ed & EDGE_E, S = ed & EDGE_S, Nn = ed & EDGE_N;
  const bool on = c == 0 ? (W && S) : c == 1 ? (E && S) : c == 2 ? (W && Nn) : (E && Nn);
  if (!on) return;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj, ie1 = ie + 1, je1 = je + 1;
  const int ci = (c & 1) ? ie1 : is, cj = (c & 2) ? je1 : js;   // the corner point
  const int wi0 = ci - DCH, wj0 = cj - DCH;                     // window origin
  const int ni = g.nx + 7, nj = g.ny + 7;
  const long long base3 = (long long)t * g.st3 + (long long)k * g.sk;
  const int base2 = t * g.sk;
  const int tid = threadIdx.x;
  auto inwin = [&](int i, int j) { return i >= wi0 && i < wi0 + DCW && j >= wj0 && j < wj0 + DCW; };
  for (int n = tid; n < DCW * DCW; n += 256) {
    const int lj = n / DCW, li = n % DCW, i = wi0 + li, j = wj0 + lj;
    D[lj][li] = (i >= 0 && i < ni && j >= 0 && j < nj) ? a.divg[base3 + j * sj + i] : 0.0;
    U[lj][li] = 0.0;
    V[lj][li] = 0.0;
  }
  __syncthreads();
  auto atD = [&](int i, int j) -> double& { return D[j - wj0][i - wi0]; };
  auto atU = [&](int i, int j) -> double& { return U[j - wj0][i - wi0]; };
  auto atV = [&](int i, int j) -> double& { return V[j - wj0][i - wi0]; };
  const int dgc = c == 0 ? 0 : c == 1 ? 3 : c == 2 ? 2 : 1;   // the same corner in fill_corners_dgrid's numbering
  const int N = a.N;
  for (int n = 0; n < N; ++n) {
    const int nt = N - (n + 1);
    const bool fillc = (n + 1 != N);
    if (fillc) {
      if (tid < 9) bgrid_fill_cell_at(g, atD, false, t, c, tid / 3 + 1, tid % 3 + 1);
      __syncthreads();
    }
    // vc_from_divg on [is-nt-1, ie+nt+1] x [js-nt, je+nt+1]
    for (int m = tid; m < DCW * DCW; m += 256) {
      const int lj = m / DCW, li = m % DCW, i = wi0 + li, j = wj0 + lj;
      if (li + 1 < DCW && i >= is - nt - 1 && i <= ie + nt + 1 && j >= js - nt && j <= je + nt + 1)
        V[lj][li] = (D[lj][li + 1] - D[lj][li]) * a.divg_u[base2 + j * sj + i];
    }
    __syncthreads();
    if (fillc) {
      if (tid < 9) bgrid_fill_cell_at(g, atD, true, t, c, tid / 3 + 1, tid % 3 + 1);
      __syncthreads();
    }
    // uc_from_divg on [is-nt, ie+nt+1] x [js-nt-1, je+nt+1]
    for (int m = tid; m < DCW * DCW; m += 256) {
      const int lj = m / DCW, li = m % DCW, i = wi0 + li, j = wj0 + lj;
      if (lj + 1 < DCW && i >= is - nt && i <= ie + nt + 1 && j >= js - nt - 1 && j <= je + nt + 1)
        U[lj][li] = (D[lj + 1][li] - D[lj][li]) * a.divg_v[base2 + j * sj + i];
    }
    __syncthreads();
    if (fillc) {
      if (tid < DG_N) dgrid_fill_entry_at(g, atV, atU, -1.0, t, tid, dgc);
      __syncthreads();
    }
    // redo_divg_d on [is-nt, ie+nt+1] x [js-nt, je+nt+1]
    for (int m = tid; m < DCW * DCW; m += 256) {
      const int lj = m / DCW, li = m % DCW, i = wi0 + li, j = wj0 + lj;
      if (li >= 1 && lj >= 1 && i >= is - nt && i <= ie + nt + 1 && j >= js - nt && j <= je + nt + 1) {
        double d = U[lj - 1][li] - U[lj][li] + V[lj][li - 1] - V[lj][li];
        if (((W && i == is) || (E && i == ie1)) && (S && j == js)) d = d - U[lj - 1][li];
        if (((W && i == is) || (E && i == ie1)) && (Nn && j == je1)) d = d + U[lj][li];
        D[lj][li] = d * a.rarea_c[base2 + j * sj + i];
      }
    }
    __syncthreads();
  }
  (void)inwin;
  // deliveries: the cells the marching kernel skipped
  for (int m = tid; m < DCW * DCW; m += 256) {
    const int lj = m / DCW, li = m % DCW, i = wi0 + li, j = wj0 + lj;
    if (i < 0 || j < 0 || i >= ni || j >= nj) continue;
    const bool mine = abs(i - ci) <= DD_CORNER_R && abs(j - cj) <= DD_CORNER_R;
    if (!mine) continue;
    const long long o = base3 + j * sj + i;
    if (i >= is - 1 && i <= ie1 && j >= js && j <= je1) a.vc[o] = V[lj][li];
    if (i >= is && i <= ie1 && j >= js - 1 && j <= je1) a.uc[o] = U[lj][li];
    if (i >= is && i <= ie1 && j >= js && j <= je1) a.dout[o] = D[lj][li];
  }
}
#endif

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
  // k >= kstart: copy_computeplus(divg_d -> delpc) and the nord loop
  const double* d_final = divg_d;   // where the Smagorinsky kernel finds the loop's result
  bool marched = false;
#ifndef PACE_HOST_EMU
  static const bool unfused = [] { const char* v = getenv("PACE_B200_DD"); return v && std::string(v) == "unfused"; }();
  if (!unfused && nonzero_nord >= 1 && nonzero_nord <= 3 && nkh > 0) {
    double* dout = c.S("dd.divg_out");
    if (!c.err.empty()) return 1;
    DdMarchArgs m;
    m.divg = divg_d; m.divg_u = divg_u; m.divg_v = divg_v; m.rarea_c = rarea_c;
    m.delpc = delpc; m.dout = dout; m.uc = uc; m.vc = vc;
    m.k0 = kstart; m.nk = nkh; m.N = nonzero_nord;
    m.nstrips = (g.nx + 3 + DSW - 1) / DSW;
    static const int rows_env = [] { const char* v = getenv("PACE_B200_DD_ROWS"); return v ? atoi(v) : 0; }();
    const int nrows = g.ny + 3, rows = rows_env > 0 ? rows_env : 48;
    m.nchunks = (nrows + rows - 1) / rows;
    m.rows = (nrows + m.nchunks - 1) / m.nchunks;
    m.items = (long long)m.nstrips * m.nchunks * nkh * g.nt;
    {
      LaunchScope scope(st, "dd_march");
      k_dd_march<<<(unsigned)((m.items + 3) / 4), 128, 0, st>>>(g, m);
    }
    if ((g.edge_bits[0] | g.edge_bits[1]) != 0) {
      DdCornerArgs ca;
      ca.divg = divg_d; ca.divg_u = divg_u; ca.divg_v = divg_v; ca.rarea_c = rarea_c;
      ca.dout = dout; ca.uc = uc; ca.vc = vc; ca.k0 = kstart; ca.nk = nkh; ca.N = nonzero_nord;
      LaunchScope scope(st, "dd_corner");
      k_dd_corner<<<4 * nkh * g.nt, 256, 0, st>>>(g, ca);
    }
    d_final = dout;
    marched = true;
  }
#endif
  if (!marched) {
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
    const double dn = A0(d_final);
    double vo = damp * A0(delpc) + dd8 * dn;
    if (marched) A0(divg_d) = dn;  // the marching kernel left the loop's result in a work array
    A0(vort_b) = vo;
    A0(ke) = A0(ke) + vo;
  });
  return 0;
}

}  // namespace pace
