// DelnFluxNoSG (+ DelnFlux accumulation) as ONE shared-memory tiled kernel.
//
// The straightforward version (transport.cu: delnflux) is 2 + 2*nord launches through three full-size work arrays
// (d2, fx2, fy2): ~28 field passes per call, 216 launches and 13 % of the C384 dycore step
// (profiles/r01b_launches_c384.summary.txt).  Here a CTA owns a TX x TY tile of one level with an apron of
// nord + 1 cells and iterates in shared memory:
//   d2_0 = damp * q (or q)                                                            (delnflux.py:1243-1250)
//   fx2, fy2 from the corner-copied d2 ; for n < nord: d2 = div(fx2, fy2) * rarea ; fx2, fy2 = -del6 * grad(d2)
//   finally fx += fx2 (mass-weighted when given), or fx2 / fy2 / d2 are stored        (delnflux.py:1023-1044)
// Overlapped tiling: every value a tile needs is recomputed from its own apron, with the same expression the
// unfused stencils use, so results are bit-identical.  The cube-corner copies (copy_corners_x/y_nord) only concern
// the four corner tiles, whose sources lie inside the same apron.  DRAM traffic: q (+ mass) read, fx / fy
// read-modify-written (or fx2 / fy2 / d2 written): 5-6 passes.
#include "corners.cuh"
#include "transport.hpp"

namespace pace {

#ifndef PACE_HOST_EMU

namespace {
constexpr int TX = 32, TY = 32, AP = 3;      // supports nord <= AP - 1
constexpr int W = TX + 2 * AP, H = TY + 2 * AP;
constexpr int NTY = 8;                       // thread rows per CTA (TX x NTY threads)

struct DelnFusedArgs {
  const double *q, *mass, *nord_col, *damp, *del6_u, *del6_v, *rarea;
  double *fx, *fy, *d2out;
  const int* lvl;
  int nk, accumulate;
};

constexpr int RG = 6, RPG = (H + RG - 1) / RG;  // fast path: thread = (column, row group), RPG rows per group
static_assert(W * RG <= TX * NTY, "not enough threads for the column / row-group mapping");

// Tiles whose apron holds no cube-corner cell (all but 4 per cube-corner subdomain): a thread owns one column of the
// tile and a group of rows, so that the range tests and addresses are per-thread constants / running pointers and
// the y-neighbour is carried in a register.  Same expressions as deln_tile_corner below.
__device__ __forceinline__ void deln_tile_plain(const Geo& g, const DelnFusedArgs& a, double (*d2)[W],
                                                double (*fx2)[W + 1], double (*fy2)[W], int N, int i0, int j0,
                                                int base2) {
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj;
  const int tid = threadIdx.y * TX + threadIdx.x;
  const int c = tid % W, rgp = tid / W;  // column in the tile, row group
  const bool active = rgp < RG;
  const int i = i0 - AP + c;
  const int l0 = rgp * RPG, l1 = min(l0 + RPG, H);  // my rows [l0, l1)
  auto flux_pass = [&](int nt, bool first) {
    if (active) {
      const bool okx = c >= 1 && i >= is - nt && i <= ie + 1 + nt, oky = i >= is - nt && i <= ie + nt;
      // rows of the pass: fx on j in [js-nt, je+nt], fy on [js-nt, je+1+nt] (and lj >= 1)
      const int ja = max(js - nt, j0 - AP + l0), jbx = min(je + nt, j0 - AP + l1 - 1), jby = min(je + 1 + nt, j0 - AP + l1 - 1);
      const double* pv = a.del6_v + base2 + ja * sj + i;
      const double* pu = a.del6_u + base2 + ja * sj + i;
      int lj = ja - (j0 - AP);
      double below = lj >= 1 ? d2[lj - 1][c] : 0.0;
      for (int j = ja; j <= jby; ++j, ++lj, pv += sj, pu += sj) {
        const double here = d2[lj][c];
        if (okx && j <= jbx) {
          const double diff = d2[lj][c - 1] - here;
          fx2[lj][c] = first ? *pv * diff : -*pv * diff;
        }
        if (oky && lj >= 1) {
          const double diff = below - here;
          fy2[lj][c] = first ? *pu * diff : -*pu * diff;
        }
        below = here;
      }
    }
    __syncthreads();
  };
  flux_pass(N, true);
  for (int it = 0; it < N; ++it) {
    const int nt = N - 1 - it;
    if (active && i >= is - nt - 1 && i <= ie + nt + 1) {
      const int ja = max(js - nt - 1, j0 - AP + l0), jb = min(je + nt + 1, j0 - AP + l1 - 1);
      const double* pr = a.rarea + base2 + ja * sj + i;
      int lj = ja - (j0 - AP);
      for (int j = ja; j <= jb; ++j, ++lj, pr += sj)
        d2[lj][c] = (fx2[lj][c] - fx2[lj][c + 1] + fy2[lj][c] - fy2[lj + 1][c]) * *pr;
    }
    __syncthreads();
    flux_pass(nt, false);
  }
}

// Tiles that see cube-corner cells of a level with nord > 0: d2 is read through copy_corners_x / _y.
__device__ __forceinline__ void deln_tile_corner(const Geo& g, const DelnFusedArgs& a, double (*d2)[W],
                                                 double (*fx2)[W + 1], double (*fy2)[W], const EdgeInfo& e, int N, int i0,
                                                 int j0, int base2) {
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj;
  const int tx = threadIdx.x, ty = threadIdx.y;
  auto dcx = [&](int i, int j) {
    copyx_src(e, i, j);
    return d2[j - j0 + AP][i - i0 + AP];
  };
  auto dcy = [&](int i, int j) {
    copyy_src(e, i, j);
    return d2[j - j0 + AP][i - i0 + AP];
  };
  // fluxes on the box of pass nt, restricted to what the tile holds
  auto flux_pass = [&](int nt, bool first) {
    const int ilo = max(is - nt, i0 - AP + 1), ihi_x = min(ie + 1 + nt, i0 + TX + AP - 1), ihi_y = min(ie + nt, i0 + TX + AP - 1);
    const int jlo = max(js - nt, j0 - AP + 1), jhi_x = min(je + nt, j0 + TY + AP - 1), jhi_y = min(je + 1 + nt, j0 + TY + AP - 1);
    const int ilo_y = max(is - nt, i0 - AP), jlo_x = max(js - nt, j0 - AP);
    for (int lj = ty; lj < H; lj += NTY) {
      const int j = j0 - AP + lj;
      for (int li = tx; li < W; li += TX) {
        const int i = i0 - AP + li;
        if (i >= ilo && i <= ihi_x && j >= jlo_x && j <= jhi_x) {
          const double dl = a.del6_v[base2 + j * sj + i];
          const double diff = dcx(i - 1, j) - dcx(i, j);
          fx2[lj][li] = first ? dl * diff : -dl * diff;
        }
        if (i >= ilo_y && i <= ihi_y && j >= jlo && j <= jhi_y) {
          const double dl = a.del6_u[base2 + j * sj + i];
          const double diff = dcy(i, j - 1) - dcy(i, j);
          fy2[lj][li] = first ? dl * diff : -dl * diff;
        }
      }
    }
    __syncthreads();
  };
  flux_pass(N, true);
  for (int it = 0; it < N; ++it) {
    const int nt = N - 1 - it;
    const int ilo = max(is - nt - 1, i0 - AP), ihi = min(ie + nt + 1, i0 + TX + AP - 1);
    const int jlo = max(js - nt - 1, j0 - AP), jhi = min(je + nt + 1, j0 + TY + AP - 1);
    for (int lj = ty; lj < H; lj += NTY) {
      const int j = j0 - AP + lj;
      for (int li = tx; li < W; li += TX) {
        const int i = i0 - AP + li;
        if (i >= ilo && i <= ihi && j >= jlo && j <= jhi)
          d2[lj][li] = (fx2[lj][li] - fx2[lj][li + 1] + fy2[lj][li] - fy2[lj + 1][li]) * a.rarea[base2 + j * sj + i];
      }
    }
    __syncthreads();
    flux_pass(nt, false);
  }
}

__global__ void __launch_bounds__(TX * NTY, 6) k_delnflux_fused(Geo g, DelnFusedArgs a) {
  __shared__ double d2[H][W];          // cell (i0-AP+li, j0-AP+lj)
  __shared__ double fx2[H][W + 1];     // x interface, same indexing
  __shared__ double fy2[H + 1][W];     // y interface
  const int k = blockIdx.z % a.nk, t = blockIdx.z / a.nk;
  if (a.lvl != nullptr && a.accumulate && a.lvl[t * a.nk + k] == 0) return;  // a level of zeros adds no damping flux
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj;
  const int i0 = is + blockIdx.x * TX, j0 = js + blockIdx.y * TY;
  const unsigned ed = edge_of(g, t);
  const EdgeInfo e{is, ie, js, je, (bool)(ed & EDGE_W), (bool)(ed & EDGE_E), (bool)(ed & EDGE_S), (bool)(ed & EDGE_N)};
  const int N = (int)a.nord_col[k < 3 ? k : 3];
  const bool has_mass = a.mass != nullptr;
  const long long base3 = (long long)t * g.st3 + (long long)k * g.sk;
  const int base2 = t * g.sk;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int imax = g.nx + 6, jmax = g.ny + 6;
  const double dampk = a.damp[k];

  // d2_0 on the whole tile + apron (cells outside the arrays: 0, never used).  fx2 / fy2 are not initialised: what a
  // pass leaves unwritten on the rim of the apron is never read by anything that reaches the tile's own cells.
  const int tid = ty * TX + tx;
  constexpr int NTHR = TX * NTY;
  bool nonzero = false;
  for (int n = tid; n < W * H; n += NTHR) {  // flat loops: the tile widths are not multiples of the warp size
    const int lj = n / W, li = n % W;
    const int i = i0 - AP + li, j = j0 - AP + lj;
    double v = 0.0;
    if (i >= 0 && j >= 0 && i <= imax && j <= jmax) {
      const double qv = a.q[base3 + j * sj + i];
      v = has_mass ? qv : dampk * qv;
    }
    nonzero = nonzero || v != 0.0;
    d2[lj][li] = v;
  }
  // A tile (with its apron) of zeros - q_con in dry air, a condensate outside clouds - has zero damping fluxes: when
  // they are to be added to existing fluxes there is nothing to do.
  if (!__syncthreads_or(nonzero) && a.accumulate == 1) return;
  // ask for the lines the result phase will read-modify-write (fx, fy, mass) now, so that they arrive in L2 while
  // the tile iterates in shared memory
  if (a.accumulate) {
    for (int n = tid; n < (TX + 1) * (TY + 1); n += NTHR) {
      const int lj = n / (TX + 1), li = n % (TX + 1);
      const int i = i0 + li, j = j0 + lj;
      if (i <= ie + 1 && j <= je + 1 && (li & 3) == 0) {  // one request per 32-byte sector
        const long long o = base3 + j * sj + i;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.fx + o));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(a.fy + o));
        if (has_mass) asm volatile("prefetch.global.L2 [%0];" ::"l"(a.mass + o));
      }
    }
  }

  // do cube-corner cells (both a west/east and a south/north tile edge within the apron) enter this tile?
  const bool near_w = e.W && i0 - AP < is, near_e = e.E && i0 + TX + AP - 1 > ie;
  const bool near_s = e.S && j0 - AP < js, near_n = e.N && j0 + TY + AP - 1 > je;
  if (N > 0 && (near_w || near_e) && (near_s || near_n))
    deln_tile_corner(g, a, d2, fx2, fy2, e, N, i0, j0, base2);
  else
    deln_tile_plain(g, a, d2, fx2, fy2, N, i0, j0, base2);

  // results: this tile owns the cells [i0, i0+TX) x [j0, j0+TY) and their low-side interfaces; the last tile in x / y
  // also owns the closing interface ie+1 / je+1
  for (int n = tid; n < (TX + 1) * (TY + 1); n += NTHR) {
    {
      const int lj = n / (TX + 1), li = n % (TX + 1);
      const int i = i0 + li, j = j0 + lj;
      const bool own_i = li < TX || i == ie + 1, own_j = lj < TY || j == je + 1;
      if (!own_i || !own_j) continue;
      const long long o = base3 + j * sj + i;
      if (i <= ie + 1 && j <= je) {
        const double f = fx2[lj + AP][li + AP];
        if (a.accumulate) {
          if (has_mass) a.fx[o] = a.fx[o] + 0.5 * dampk * (a.mass[o - 1] + a.mass[o]) * f;
          else a.fx[o] = a.fx[o] + f;
        } else {
          a.fx[o] = f;
        }
      }
      if (i <= ie && j <= je + 1) {
        const double f = fy2[lj + AP][li + AP];
        if (a.accumulate) {
          if (has_mass) a.fy[o] = a.fy[o] + 0.5 * dampk * (a.mass[o - sj] + a.mass[o]) * f;
          else a.fy[o] = a.fy[o] + f;
        } else {
          a.fy[o] = f;
        }
      }
    }
  }
  if (a.d2out) {
    // the last d2 on [is-1, ie+1]^2: the tiles on the low / high side also own the row / column just outside
    for (int n = tid; n < (TX + 2) * (TY + 2); n += NTHR) {
      {
        const int lj = n / (TX + 2), li = n % (TX + 2);
        const int i = i0 - 1 + li, j = j0 - 1 + lj;
        const bool own_i = (i >= i0 && i < i0 + TX && i <= ie) || (i == is - 1) || (i == ie + 1 && i <= i0 + TX);
        const bool own_j = (j >= j0 && j < j0 + TY && j <= je) || (j == js - 1) || (j == je + 1 && j <= j0 + TY);
        if (own_i && own_j) a.d2out[base3 + j * sj + i] = d2[j - j0 + AP][i - i0 + AP];
      }
    }
  }
}
}  // namespace

int delnflux_fused_launch(Core& c, const DelnArgs& a) {
  const Geo g = c.g;
  if (a.nmax > AP - 1) return -1;  // deeper damping stencils go through the unfused kernels
  DelnFusedArgs f;
  f.q = a.q; f.mass = a.mass; f.nord_col = a.nord_col; f.damp = a.damp;
  f.del6_u = c.F("damp.del6_u"); f.del6_v = c.F("damp.del6_v"); f.rarea = c.F("grid.rarea");
  f.fx = a.fx; f.fy = a.fy; f.d2out = a.d2; f.nk = a.nk; f.accumulate = a.accumulate ? (zero_paths_enabled() ? 1 : 2) : 0;  // 2: accumulate, no zero-tile shortcut
  f.lvl = a.level_nonzero;
  if (!c.err.empty()) return 1;
  dim3 blk(TX, NTY, 1);
  dim3 grd((g.nx + TX - 1) / TX, (g.ny + TY - 1) / TY, a.nk * g.nt);
  LaunchScope scope(c.stream, "delnflux_fused");
  k_delnflux_fused<<<grd, blk, 0, c.stream>>>(g, f);
  return 0;
}

#else
int delnflux_fused_launch(Core&, const DelnArgs&) { return -1; }
#endif

}  // namespace pace
