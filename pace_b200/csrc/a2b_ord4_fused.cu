// AGrid2BGridFourthOrder as one shared-memory tiled kernel.
//
// a2b_ord4.cu runs the interpolation as two full-field kernels through the work fields qx, qy (6 field passes, 330
// executed instructions per point: every point evaluates the whole tile-edge case analysis and recomputes 64-bit
// addresses for 16 neighbour reads).  Here a CTA loads a (TX+6) x (TY+6) tile of qin into shared memory, forms qx and
// qy in shared memory and writes qout: 2 field passes.  Tiles that touch no tile edge (all but the rim of the
// subdomains that lie on a tile edge) take a branch-free path; the others evaluate the same case analysis as
// a2b_ord4.cu on shared-memory operands.  Expressions and their order are those of a2b_ord4.cu / the reference
// (pyFV3/stencils/a2b_ord4.py:36-526), so results are bit-identical.
#include "tile.cuh"

namespace pace {

#ifndef PACE_HOST_EMU

namespace {
constexpr double c1 = 2.0 / 3.0, c2 = -1.0 / 6.0;
constexpr double b1 = 7.0 / 12.0, b2 = -1.0 / 12.0;
constexpr double a1 = 9.0 / 16.0, a2 = -1.0 / 16.0;
constexpr int TX = 32, TY = 32, NTY = 8;     // tile of corner points; TX x NTY threads
constexpr int QW = TX + 6, QH = TY + 6;      // qin tile: i0-3 .. i0+TX+2, j0-3 .. j0+TY+2

struct A2bArgs {
  const double *qin, *dxa, *dya, *edge_w, *edge_e, *edge_s, *edge_n, *coef;
  double* qout;
  int k0, nk;
};

// The qin tile arrives by TMA (one elected thread, an mbarrier counts the bytes): its box starts at column i0 - 3 = 32 bx,
// an even index as the 16-byte alignment rule of the TMA unit demands (tile.cuh), and cells outside the array come as zeros.
__global__ void __launch_bounds__(TX * NTY, 6) k_a2b_fused(Geo g, A2bArgs a, const __grid_constant__ CUtensorMap m_qin) {
  __shared__ __align__(128) double sq[QH][QW];           // qin(i0-3+li, j0-3+lj)
  __shared__ __align__(8) unsigned long long bar;
  __shared__ double sqx[QH][TX + 1];      // qx(i0+li, j0-3+lj)
  __shared__ double sqy[TY + 1][QW];      // qy(i0-3+li, j0+lj)
  const int k = a.k0 + blockIdx.z % a.nk, t = blockIdx.z / a.nk;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj;
  const int i0 = is + blockIdx.x * TX, j0 = js + blockIdx.y * TY;
  const unsigned ed = edge_of(g, t);
  const bool W = ed & EDGE_W, E = ed & EDGE_E, S = ed & EDGE_S, N = ed & EDGE_N;
  const int base3 = t * (int)g.st3 + k * g.sk, base2 = t * g.sk;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int imax = g.nx + 6, jmax = g.ny + 6;
  // points this tile owns: [i0, i0+TX) x [j0, j0+TY) plus the closing column / row ie+1 / je+1 in the last tiles
  const int ni = min(TX, ie + 1 - i0) + (i0 + TX > ie ? 1 : 0), nj = min(TY, je + 1 - j0) + (j0 + TY > je ? 1 : 0);

  // flat element loops: the tile widths (TX+6, TX+1) are not multiples of the warp size
  const int tid = ty * TX + tx;
  constexpr int NTHR = TX * NTY;
  if (tid == 0) tma::mbar_init(&bar, 1);
  __syncthreads();
  if (tid == 0) {
    tma::mbar_expect_tx(&bar, (unsigned)sizeof(sq));
    tma::load_tile(&sq[0][0], &m_qin, &bar, i0 - 3, j0 - 3, k, t);
  }
  tma::mbar_wait(&bar, 0);
  auto Q = [&](int i, int j) { return sq[j - j0 + 3][i - i0 + 3]; };
  auto DXA = [&](int i, int j) { return a.dxa[base2 + j * sj + i]; };
  auto DYA = [&](int i, int j) { return a.dya[base2 + j * sj + i]; };

  // does the case analysis of the tile edges concern this tile at all?
  const bool edge_x = (W && i0 - 3 <= is + 1) || (E && i0 + TX + 2 >= ie);
  const bool edge_y = (S && j0 - 3 <= js + 1) || (N && j0 + TY + 2 >= je);

  if (!edge_x && !edge_y) {
    // Interior tile (no tile edge within reach): a thread owns one column and a group of rows, so that the shared
    // memory addresses are a running row index plus constants and nothing but the interpolation is executed.
    {
      const int c = tid % (TX + 1), rg = tid / (TX + 1);  // qx: 33 columns x 7 groups of 6 rows
      if (rg < 7 && c < ni)
        for (int lj = rg * 6; lj < min(rg * 6 + 6, QH); ++lj)
          sqx[lj][c] = b2 * (sq[lj][c + 1] + sq[lj][c + 4]) + b1 * (sq[lj][c + 2] + sq[lj][c + 3]);
    }
    {
      const int c = tid % QW, rg = tid / QW;  // qy: 38 columns x 6 groups of 6 rows
      if (rg < 6)
        for (int lj = rg * 6; lj < min(rg * 6 + 6, nj); ++lj)
          sqy[lj][c] = b2 * (sq[lj + 1][c] + sq[lj + 4][c]) + b1 * (sq[lj + 2][c] + sq[lj + 3][c]);
    }
    __syncthreads();
    {
      const int c = tid % (TX + 1), rg = tid / (TX + 1);  // qout: 33 columns x 7 groups of 5 rows
      if (rg < 7 && c < ni) {
        double* out = a.qout + base3 + j0 * sj + i0 + c;
        for (int lj = rg * 5; lj < min(rg * 5 + 5, nj); ++lj) {
          const double qxx = a2 * (sqx[lj + 1][c] + sqx[lj + 4][c]) + a1 * (sqx[lj + 2][c] + sqx[lj + 3][c]);
          const double qyy = a2 * (sqy[lj][c + 1] + sqy[lj][c + 4]) + a1 * (sqy[lj][c + 2] + sqy[lj][c + 3]);
          out[lj * sj] = 0.5 * (qxx + qyy);
        }
      }
    }
    return;
  }

  // Tiles within reach of a tile edge: the interpolation of the interior path everywhere it applies, and the case
  // analysis only on the (at most four) columns / rows next to the edges.
  auto xspecial = [&](int i) { return edge_x && ((W && (i == is || i == is + 1)) || (E && (i == ie || i == ie + 1))); };
  auto yspecial = [&](int j) { return edge_y && ((S && (j == js || j == js + 1)) || (N && (j == je || j == je + 1))); };
  const int xcand[4] = {is, is + 1, ie, ie + 1}, ycand[4] = {js, js + 1, je, je + 1};
  {
    const int c = tid % (TX + 1), rg = tid / (TX + 1);  // qx, plain columns
    if (rg < 7 && c < ni && !xspecial(i0 + c))
      for (int lj = rg * 6; lj < min(rg * 6 + 6, QH); ++lj)
        if (j0 - 3 + lj <= jmax) sqx[lj][c] = b2 * (sq[lj][c + 1] + sq[lj][c + 4]) + b1 * (sq[lj][c + 2] + sq[lj][c + 3]);
  }
  // qx(i, j) on the special columns, j in [j0-3, j0+TY+2] (rows beyond the arrays are never used)
  for (int n = tid; n < QH * 4; n += NTHR) {
    {
      const int lj = n >> 2, i = xcand[n & 3], li = i - i0;
      const int j = j0 - 3 + lj;
      if (li < 0 || li >= ni || !xspecial(i) || j > jmax) continue;
      double v;
      if (edge_x && W && i == is) {  // qx_edge_west
        double g_in = DXA(i + 1, j) / DXA(i, j);
        double g_ou = DXA(i - 2, j) / DXA(i - 1, j);
        v = 0.5 * (((2.0 + g_in) * Q(i, j) - Q(i + 1, j)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * Q(i - 1, j) - Q(i - 2, j)) / (1.0 + g_ou));
      } else if (edge_x && W && i == is + 1) {  // qx_edge_west2
        double g_in = DXA(i, j) / DXA(i - 1, j);
        double g_ou = DXA(i - 3, j) / DXA(i - 2, j);
        double qxleft = 0.5 * (((2.0 + g_in) * Q(i - 1, j) - Q(i, j)) / (1.0 + g_in) +
                               ((2.0 + g_ou) * Q(i - 2, j) - Q(i - 3, j)) / (1.0 + g_ou));
        double qxright = b2 * (Q(i - 1, j) + Q(i + 2, j)) + b1 * (Q(i, j) + Q(i + 1, j));
        v = (3.0 * (g_in * Q(i - 1, j) + Q(i, j)) - (g_in * qxleft + qxright)) / (2.0 + 2.0 * g_in);
      } else if (edge_x && E && i == ie + 1) {  // qx_edge_east
        double g_in = DXA(i - 2, j) / DXA(i - 1, j);
        double g_ou = DXA(i + 1, j) / DXA(i, j);
        v = 0.5 * (((2.0 + g_in) * Q(i - 1, j) - Q(i - 2, j)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * Q(i, j) - Q(i + 1, j)) / (1.0 + g_ou));
      } else if (edge_x && E && i == ie) {  // qx_edge_east2
        double g_in = DXA(i - 1, j) / DXA(i, j);
        double g_ou = DXA(i + 2, j) / DXA(i + 1, j);
        double qxright = 0.5 * (((2.0 + g_in) * Q(i, j) - Q(i - 1, j)) / (1.0 + g_in) +
                                ((2.0 + g_ou) * Q(i + 1, j) - Q(i + 2, j)) / (1.0 + g_ou));
        double qxleft = b2 * (Q(i - 3, j) + Q(i, j)) + b1 * (Q(i - 2, j) + Q(i - 1, j));
        v = (3.0 * (Q(i - 1, j) + g_in * Q(i, j)) - (g_in * qxright + qxleft)) / (2.0 + 2.0 * g_in);
      } else {
        v = b2 * (Q(i - 2, j) + Q(i + 1, j)) + b1 * (Q(i - 1, j) + Q(i, j));
      }
      sqx[lj][li] = v;
    }
  }
  {
    const int c = tid % QW, rg = tid / QW;  // qy, plain rows
    if (rg < 6 && i0 - 3 + c <= imax)
      for (int lj = rg * 6; lj < min(rg * 6 + 6, nj); ++lj)
        if (!yspecial(j0 + lj)) sqy[lj][c] = b2 * (sq[lj + 1][c] + sq[lj + 4][c]) + b1 * (sq[lj + 2][c] + sq[lj + 3][c]);
  }
  // qy(i, j) on the special rows, i in [i0-3, i0+TX+2]
  for (int n = tid; n < 4 * QW; n += NTHR) {
    {
      const int li = n % QW, j = ycand[n / QW], lj = j - j0;
      const int i = i0 - 3 + li;
      if (lj < 0 || lj >= nj || !yspecial(j) || i > imax) continue;
      double v;
      if (edge_y && S && j == js) {  // qy_edge_south
        double g_in = DYA(i, j + 1) / DYA(i, j);
        double g_ou = DYA(i, j - 2) / DYA(i, j - 1);
        v = 0.5 * (((2.0 + g_in) * Q(i, j) - Q(i, j + 1)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * Q(i, j - 1) - Q(i, j - 2)) / (1.0 + g_ou));
      } else if (edge_y && S && j == js + 1) {  // qy_edge_south2
        double g_in = DYA(i, j) / DYA(i, j - 1);
        double g_ou = DYA(i, j - 3) / DYA(i, j - 2);
        double qy_lower = 0.5 * (((2.0 + g_in) * Q(i, j - 1) - Q(i, j)) / (1.0 + g_in) +
                                 ((2.0 + g_ou) * Q(i, j - 2) - Q(i, j - 3)) / (1.0 + g_ou));
        double qy_upper = b2 * (Q(i, j - 1) + Q(i, j + 2)) + b1 * (Q(i, j) + Q(i, j + 1));
        v = (3.0 * (g_in * Q(i, j - 1) + Q(i, j)) - (g_in * qy_lower + qy_upper)) / (2.0 + 2.0 * g_in);
      } else if (edge_y && N && j == je + 1) {  // qy_edge_north
        double g_in = DYA(i, j - 2) / DYA(i, j - 1);
        double g_ou = DYA(i, j + 1) / DYA(i, j);
        v = 0.5 * (((2.0 + g_in) * Q(i, j - 1) - Q(i, j - 2)) / (1.0 + g_in) +
                   ((2.0 + g_ou) * Q(i, j) - Q(i, j + 1)) / (1.0 + g_ou));
      } else if (edge_y && N && j == je) {  // qy_edge_north2
        double g_in = DYA(i, j - 1) / DYA(i, j);
        double g_ou = DYA(i, j + 2) / DYA(i, j + 1);
        double qy_lower = b2 * (Q(i, j - 3) + Q(i, j)) + b1 * (Q(i, j - 2) + Q(i, j - 1));
        double qy_upper = 0.5 * (((2.0 + g_in) * Q(i, j) - Q(i, j - 1)) / (1.0 + g_in) +
                                 ((2.0 + g_ou) * Q(i, j + 1) - Q(i, j + 2)) / (1.0 + g_ou));
        v = (3.0 * (Q(i, j - 1) + g_in * Q(i, j)) - (g_in * qy_upper + qy_lower)) / (2.0 + 2.0 * g_in);
      } else {
        v = b2 * (Q(i, j - 2) + Q(i, j + 1)) + b1 * (Q(i, j - 1) + Q(i, j));
      }
      sqy[lj][li] = v;
    }
  }
  __syncthreads();
  auto QX = [&](int i, int j) { return sqx[j - j0 + 3][i - i0]; };
  auto QY = [&](int i, int j) { return sqy[j - j0][i - i0 + 3]; };

  // qout on a tile edge / corner (from qin only); also serves the rows / columns next to an edge below
  auto qout_edge = [&](int i, int j) {
    const bool onW = W && i == is, onE = E && i == ie + 1, onS = S && j == js, onN = N && j == je + 1;
    if ((onW || onE) && (onS || onN)) {
      // extrap_corner: qa + coef*(qa - qb); corner order SW, SE, NE, NW
      const double* cf = a.coef + (t * 4 + (onS ? (onW ? 0 : 1) : (onE ? 2 : 3))) * 3;
      auto ec = [&](int n, int ai, int aj, int bi, int bj) {
        double qa = Q(i + ai, j + aj), qb = Q(i + bi, j + bj);
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
      return (e1 + e2 + e3) * (1.0 / 3.0);
    } else if (onW || onE) {  // qout_x_edge
      const double ew = (onW ? a.edge_w : a.edge_e)[base2 + j * sj + i];
      auto q2 = [&](int dj) {
        return (Q(i - 1, j + dj) * DXA(i, j + dj) + Q(i, j + dj) * DXA(i - 1, j + dj)) /
               (DXA(i - 1, j + dj) + DXA(i, j + dj));
      };
      return ew * q2(-1) + (1.0 - ew) * q2(0);
    } else {  // qout_y_edge
      const double es = (onS ? a.edge_s : a.edge_n)[base2 + j * sj + i];
      auto q1 = [&](int di) {
        return (Q(i + di, j - 1) * DYA(i + di, j) + Q(i + di, j) * DYA(i + di, j - 1)) /
               (DYA(i + di, j - 1) + DYA(i + di, j));
      };
      return es * q1(-1) + (1.0 - es) * q1(0);
    }
  };

  {
    const int c = tid % (TX + 1), rg = tid / (TX + 1);  // qout, points on plain columns and rows
    if (rg < 7 && c < ni && !xspecial(i0 + c)) {
      double* outp = a.qout + base3 + j0 * sj + i0 + c;
      for (int lj = rg * 5; lj < min(rg * 5 + 5, nj); ++lj) {
        if (yspecial(j0 + lj)) continue;
        const double qxx = a2 * (sqx[lj + 1][c] + sqx[lj + 4][c]) + a1 * (sqx[lj + 2][c] + sqx[lj + 3][c]);
        const double qyy = a2 * (sqy[lj][c + 1] + sqy[lj][c + 4]) + a1 * (sqy[lj][c + 2] + sqy[lj][c + 3]);
        outp[lj * sj] = 0.5 * (qxx + qyy);
      }
    }
  }
  // the points on special columns (all rows), then those on special rows (plain columns only)
  for (int n = tid; n < 4 * (TY + 1) + 4 * (TX + 1); n += NTHR) {
    {
      int li, lj;
      if (n < 4 * (TY + 1)) {
        li = xcand[n & 3] - i0;
        lj = n >> 2;
        if (li < 0 || li >= ni || lj >= nj || !xspecial(i0 + li)) continue;
      } else {
        const int m = n - 4 * (TY + 1);
        lj = ycand[m & 3] - j0;
        li = m >> 2;
        if (lj < 0 || lj >= nj || li >= ni || !yspecial(j0 + lj) || xspecial(i0 + li)) continue;
      }
      const int i = i0 + li, j = j0 + lj;
      double out;
      if ((edge_x || edge_y) && ((W && i == is) || (E && i == ie + 1) || (S && j == js) || (N && j == je + 1))) {
        out = qout_edge(i, j);
      } else {
        double qxx = a2 * (QX(i, j - 2) + QX(i, j + 1)) + a1 * (QX(i, j - 1) + QX(i, j));
        double qyy = a2 * (QY(i - 2, j) + QY(i + 1, j)) + a1 * (QY(i - 1, j) + QY(i, j));
        if (edge_y && S && j == js + 1) {
          double qxx_upper = a2 * (QX(i, j - 1) + QX(i, j + 2)) + a1 * (QX(i, j) + QX(i, j + 1));
          qxx = c1 * (QX(i, j - 1) + QX(i, j)) + c2 * (qout_edge(i, j - 1) + qxx_upper);
        }
        if (edge_y && N && j == je) {
          double qxx_lower = a2 * (QX(i, j - 3) + QX(i, j)) + a1 * (QX(i, j - 2) + QX(i, j - 1));
          qxx = c1 * (QX(i, j - 1) + QX(i, j)) + c2 * (qout_edge(i, j + 1) + qxx_lower);
        }
        if (edge_x && W && i == is + 1) {
          double qyy_right = a2 * (QY(i - 1, j) + QY(i + 2, j)) + a1 * (QY(i, j) + QY(i + 1, j));
          qyy = c1 * (QY(i - 1, j) + QY(i, j)) + c2 * (qout_edge(i - 1, j) + qyy_right);
        }
        if (edge_x && E && i == ie) {
          double qyy_left = a2 * (QY(i - 3, j) + QY(i, j)) + a1 * (QY(i - 2, j) + QY(i - 1, j));
          qyy = c1 * (QY(i - 1, j) + QY(i, j)) + c2 * (qout_edge(i + 1, j) + qyy_left);
        }
        out = 0.5 * (qxx + qyy);
      }
      a.qout[base3 + j * sj + i] = out;
    }
  }
}
}  // namespace

int a2b_ord4_fused_launch(Core& c, const double* qin, double* qout, int k0, int nk) {
  const Geo g = c.g;
  A2bArgs a;
  a.qin = qin; a.qout = qout; a.k0 = k0; a.nk = nk;
  a.dxa = c.F("grid.dxa"); a.dya = c.F("grid.dya"); a.edge_w = c.F("grid.edge_w"); a.edge_e = c.F("grid.edge_e");
  a.edge_s = c.F("grid.edge_s"); a.edge_n = c.F("grid.edge_n"); a.coef = c.F("grid.a2b_corner_coef");
  if (!c.err.empty()) return 1;
  static_assert(QW % 2 == 0 && (TX % 2) == 0 && HALO == 3, "TMA box start i0 - 3 = 32 bx must be even");
  const CUtensorMap* mq = tile_tensor_map(c, qin, QW, QH);
  if (!mq) return 1;
  dim3 blk(TX, NTY, 1);
  dim3 grd((g.nx + TX - 1) / TX, (g.ny + TY - 1) / TY, nk * g.nt);
  LaunchScope scope(c.stream, "a2b_ord4_fused");
  k_a2b_fused<<<grd, blk, 0, c.stream>>>(g, a, *mq);
  return 0;
}

#else
int a2b_ord4_fused_launch(Core&, const double*, double*, int, int) { return -1; }
#endif

}  // namespace pace
