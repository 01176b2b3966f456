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
  int corner_only;  // 1: only the tiles that see cube-corner cells on levels with nord > 0 run (the rest: k_delnflux_march)
  // heat != 0 (d_sw's call for w, not accumulating): instead of storing fx2 / fy2 / d2, the kernels apply heat_diss
  // (d_sw.py:87-125) to them right away: dw = div(fx2, fy2) * rarea and heat_s = dd8 - dw * (w + 0.5 dw) on the compute
  // domain (w = q).  Saves three written and three re-read field passes and a launch.
  int heat;
  double *hdw, *hheat;
  const double *col_damp_w, *col_ke_bg;
  double absdt;
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
  // do cube-corner cells (both a west/east and a south/north tile edge within the apron) enter this tile?
  const bool near_w = e.W && i0 - AP < is, near_e = e.E && i0 + TX + AP - 1 > ie;
  const bool near_s = e.S && j0 - AP < js, near_n = e.N && j0 + TY + AP - 1 > je;
  if (a.corner_only && !(N > 0 && (near_w || near_e) && (near_s || near_n))) return;
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
        } else if (!a.heat) {
          a.fx[o] = f;
        }
      }
      if (i <= ie && j <= je + 1) {
        const double f = fy2[lj + AP][li + AP];
        if (a.accumulate) {
          if (has_mass) a.fy[o] = a.fy[o] + 0.5 * dampk * (a.mass[o - sj] + a.mass[o]) * f;
          else a.fy[o] = a.fy[o] + f;
        } else if (!a.heat) {
          a.fy[o] = f;
        }
      }
      if (a.heat && li < TX && lj < TY && i <= ie && j <= je) {
        double hs = 0.0;
        if (a.col_damp_w[k] > 1e-5) {
          const double dd8 = a.col_ke_bg[k] * a.absdt;
          const double dwv = (fx2[lj + AP][li + AP] - fx2[lj + AP][li + AP + 1] + fy2[lj + AP][li + AP] - fy2[lj + AP + 1][li + AP]) *
                             a.rarea[base2 + j * sj + i];
          a.hdw[o] = dwv;
          hs = dd8 - dwv * (a.q[o] + 0.5 * dwv);
        }
        a.hheat[o] = hs;
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
// ---- row-marching version --------------------------------------------------------------------------------------
// The tile kernel above spends ~470 thread instructions per cell, nine tenths of them on tile addressing, range tests
// and barriers (profiles/r02d_step_metrics_c384.summary.txt: issue slots 70 % busy, fp64 pipe 9 %).  Here a WARP owns
// one level, a strip of 26 columns (lane <-> i, 3-column apron each side, values exchanged by shuffles) and a chunk of
// rows, and marches over the rows: row R of q comes in, the three damping levels advance by one row each
//   d_0(R)                         fx_0(R), fy_0(R)
//   d_1(R-1) = div(fx_0, fy_0)     fx_1(R-1), fy_1(R-1)
//   d_2(R-2) = div(fx_1, fy_1)     fx_2(R-2), fy_2(R-2)
// with everything carried in registers: no shared memory, no barrier, every field read and written once.  The
// expressions are those of the tile kernel (same bits).  Cube-corner copies concern only the tiles that hold corner
// cells on levels with nord > 0: those tiles are left to the tile kernel (corner_only) and skipped here.
constexpr int MSW = 26;
#ifndef PACE_DELN_PF_ROWS
#define PACE_DELN_PF_ROWS 6
#endif
constexpr int PF_ROWS = PACE_DELN_PF_ROWS;  // L2 prefetch distance in rows beyond the register prefetch
constexpr unsigned MFULL = 0xffffffffu;

struct DelnMarchArgs {
  DelnFusedArgs f;
  int nstrips, nchunks, rows;  // column strips, row chunks, rows per chunk
  int ntx, nty;                // tiles of the tile kernel per subdomain (ownership of the corner tiles)
  long long items;
};

// ACC: accumulate into existing fluxes (their old values and the mass are extra input streams); HEAT: heat_diss epilogue.
// Separate instantiations keep the registers of the streams / epilogue a call does not use out of the others.
template <bool ACC, bool HEAT>
__global__ void __launch_bounds__(128, 4) k_delnflux_march(Geo g, DelnMarchArgs m) {
  const DelnFusedArgs& a = m.f;
  const long long item = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (item >= m.items) return;
  const int lane = threadIdx.x & 31;
  // item -> (t, k, chunk, strip), strips fastest: neighbouring warps share their column aprons in L1 / L2
  const int strip = (int)(item % m.nstrips);
  long long r1 = item / m.nstrips;
  const int chunk = (int)(r1 % m.nchunks);
  r1 /= m.nchunks;
  const int k = (int)(r1 % a.nk), t = (int)(r1 / a.nk);
  if (ACC && a.lvl != nullptr && a.lvl[t * a.nk + k] == 0) return;  // a level of zeros adds no damping flux
  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj;
  const unsigned ed = edge_of(g, t);
  const bool eW = ed & EDGE_W, eE = ed & EDGE_E, eS = ed & EDGE_S, eN = ed & EDGE_N;
  const int N = (int)a.nord_col[k < 3 ? k : 3];
  const double dampk = a.damp[k];
  const bool has_mass = a.mass != nullptr;
  const int i = is + strip * MSW - 3 + lane;
  const int ic = min(i, sj - 1);  // lanes right of the arrays read a valid address; their values are never used
  // rows this warp delivers: [j_lo, j_hi] within [js-1, je+1]
  const int j_lo = js - 1 + chunk * m.rows, j_hi = min(j_lo + m.rows - 1, je + 1);
  // 32-bit element offsets (pace_b200_create refuses fields of 2^31 or more points): the field pointers stay in the
  // constant bank and an address costs one IMAD.WIDE, instead of ten 64-bit running pointers in registers
  const int base3 = (int)((long long)t * g.st3 + (long long)k * g.sk);
  const int c3 = base3 + ic, c2 = t * g.sk + ic;
  const double* __restrict__ q = a.q;
  const double* __restrict__ dv = a.del6_v;
  const double* __restrict__ du = a.del6_u;
  const double* __restrict__ ra = a.rarea;
  // ownership / validity per lane
  const bool own_f = lane >= 3 && lane <= 28;                        // flux columns of this strip
  const bool own_d = own_f || (strip == 0 && lane == 2);             // d2out also covers i = is - 1
  const bool st_fx = own_f && i >= is && i <= ie + 1, st_fy = own_f && i >= is && i <= ie;
  const bool st_d = own_d && i >= is - 1 && i <= ie + 1;
  // tile column of this lane in the tile kernel's grid (the corner tiles are not ours when N > 0)
  const int tbx = min(max(i - is, 0) / TX, m.ntx - 1);
  const int ti0 = is + tbx * TX;
  const bool tnear_w = eW && ti0 - AP < is, tnear_e = eE && ti0 + TX + AP - 1 > ie;
  const bool lane_corner_x = N > 0 && (tnear_w || tnear_e);

  // with the heat_diss epilogue one more row is needed: dw of row j uses fy2 of row j + 1
  const int R0 = max(0, j_lo - 1 - N - 1), R1 = min(j_hi + N + (HEAT ? 1 : 0), g.ny + 6);
  double hfx_p = 0.0, hfy_p = 0.0;             // delivered fluxes of the previous step (row jo - 1)
  double qh1 = 0.0, qh2 = 0.0, qh3 = 0.0, ra3 = 0.0;  // q at rows R-1 .. R-3, rarea at row R-3
  double d0p = 0.0, fx0p = 0.0, fy0p = 0.0;   // level 0 at row R-1
  double d1p = 0.0, fx1p = 0.0, fy1p = 0.0;   // level 1 at row R-2
  double d2p = 0.0;                            // level 2 at row R-3
  double dv1 = 0.0, du1 = 0.0, ra1 = 0.0, dv2 = 0.0, du2 = 0.0, ra2 = 0.0;  // metric terms of rows R-1, R-2
  // Every stream is read two rows ahead of its use (registers): q and the metric terms of the incoming row, and — when
  // accumulating — the old fluxes and the mass of the outgoing row.  With 8 warps per scheduler that is enough distance
  // to cover the DRAM latency.
  constexpr bool acc = ACC;
  const bool accm = acc && has_mass;
  const bool ld_fx = acc && st_fx, ld_fy = acc && st_fy;
  const double* fxo_p = a.fx;
  const double* fyo_p = a.fy;
  const double* ms_p = has_mass ? a.mass : a.q;
  auto row_in = [&](int R) { return min(R, R1) * sj; };
  auto row_out = [&](int R) { return min(max(R - N, 0), je + 1) * sj; };
  double q_a = q[c3 + row_in(R0)], q_b = q[c3 + row_in(R0 + 1)];
  double dv_a = dv[c2 + row_in(R0)], dv_b = dv[c2 + row_in(R0 + 1)], du_a = du[c2 + row_in(R0)], du_b = du[c2 + row_in(R0 + 1)];
  double ra_a = ra[c2 + row_in(R0)], ra_b = ra[c2 + row_in(R0 + 1)];
  double ofx_a = 0.0, ofx_b = 0.0, ofy_a = 0.0, ofy_b = 0.0, ms_a = 0.0, ms_b = 0.0, ms_m = 0.0;  // ms_m: mass one row below
  if (ld_fx) { ofx_a = fxo_p[c3 + row_out(R0)]; ofx_b = fxo_p[c3 + row_out(R0 + 1)]; }
  if (ld_fy) { ofy_a = fyo_p[c3 + row_out(R0)]; ofy_b = fyo_p[c3 + row_out(R0 + 1)]; }
  if (accm) { ms_m = ms_p[c3 + row_out(R0 - 1)]; ms_a = ms_p[c3 + row_out(R0)]; ms_b = ms_p[c3 + row_out(R0 + 1)]; }
  for (int R = R0; R <= R1; ++R) {
    const double qv = q_a, dv0 = dv_a, du0 = du_a, ra0 = ra_a;
    const double ofx = ofx_a, ofy = ofy_a, m_0 = ms_a, m_b = ms_m;
    q_a = q_b; dv_a = dv_b; du_a = du_b; ra_a = ra_b;
    ofx_a = ofx_b; ofy_a = ofy_b; ms_m = ms_a; ms_a = ms_b;
    {
      const int rin = row_in(R + 2);
      q_b = q[c3 + rin]; dv_b = dv[c2 + rin]; du_b = du[c2 + rin]; ra_b = ra[c2 + rin];
      const int rout = row_out(R + 2);
      if (ld_fx) ofx_b = fxo_p[c3 + rout];
      if (ld_fy) ofy_b = fyo_p[c3 + rout];
      if (accm) ms_b = ms_p[c3 + rout];
      // ask L2 for the rows a few steps further ahead (no register is tied up; one request per 32-byte sector), so that
      // the two-row register prefetch above finds its data in L2 instead of DRAM
      if (PF_ROWS > 0 && (lane & 3) == 0) {
        const int pin = row_in(R + 2 + PF_ROWS), pout = row_out(R + 2 + PF_ROWS);
        asm volatile("prefetch.global.L2 [%0];" ::"l"(q + c3 + pin));
        if (ld_fx) asm volatile("prefetch.global.L2 [%0];" ::"l"(fxo_p + c3 + pout));
        if (ld_fy) asm volatile("prefetch.global.L2 [%0];" ::"l"(fyo_p + c3 + pout));
        if (accm) asm volatile("prefetch.global.L2 [%0];" ::"l"(ms_p + c3 + pout));
      }
    }
    const double d0 = has_mass ? qv : dampk * qv;
    const double fx0 = dv0 * (__shfl_up_sync(MFULL, d0, 1) - d0);
    const double fy0 = du0 * (d0p - d0);
    double fxo = fx0, fyo = fy0, dout = d0;  // what goes out at row R - N
    double d1 = 0.0, fx1 = 0.0, fy1 = 0.0, d2 = 0.0;
    if (N >= 1) {
      d1 = (fx0p - __shfl_down_sync(MFULL, fx0p, 1) + fy0p - fy0) * ra1;
      fx1 = -dv1 * (__shfl_up_sync(MFULL, d1, 1) - d1);
      fy1 = -du1 * (d1p - d1);
      fxo = fx1; fyo = fy1; dout = d1;
      if (N >= 2) {
        d2 = (fx1p - __shfl_down_sync(MFULL, fx1p, 1) + fy1p - fy1) * ra2;
        fxo = -dv2 * (__shfl_up_sync(MFULL, d2, 1) - d2);
        fyo = -du2 * (d2p - d2);
        dout = d2;
      }
    }
    const int jo = R - N;
    if (HEAT) {  // heat_diss of row jo - 1, now that fy2 of row jo is known
      const int jh = jo - 1;
      const double fx_r = __shfl_down_sync(MFULL, hfx_p, 1);
      if (jh >= max(j_lo, js) && jh <= min(j_hi, je) && own_f && i >= is && i <= ie) {
        bool skip = false;
        if (lane_corner_x) {
          const int tby = min(max(jh - js, 0) / TY, m.nty - 1);
          const int tj0 = js + tby * TY;
          skip = (eS && tj0 - AP < js) || (eN && tj0 + TY + AP - 1 > je);
        }
        if (!skip) {
          const int o = base3 + jh * sj + i;
          double hs = 0.0;
          if (a.col_damp_w[k] > 1e-5) {
            const double dd8 = a.col_ke_bg[k] * a.absdt;
            const double ra_h = N == 2 ? ra3 : N == 1 ? ra2 : ra1;   // rarea of row jh = R - N - 1
            const double w_h = N == 2 ? qh3 : N == 1 ? qh2 : qh1;    // w of row jh
            const double dwv = (hfx_p - fx_r + hfy_p - fyo) * ra_h;
            a.hdw[o] = dwv;
            hs = dd8 - dwv * (w_h + 0.5 * dwv);
          }
          a.hheat[o] = hs;
        }
      }
      hfx_p = fxo; hfy_p = fyo;
    }
    if (!HEAT && jo >= j_lo && jo <= j_hi) {  // warp-uniform
      double m_l = 0.0;
      if (accm) m_l = __shfl_up_sync(MFULL, m_0, 1);  // mass of the column to the left
      // does (i, jo) belong to a tile the tile kernel handles?
      bool skip = false;
      if (lane_corner_x) {
        const int tby = min(max(jo - js, 0) / TY, m.nty - 1);
        const int tj0 = js + tby * TY;
        skip = (eS && tj0 - AP < js) || (eN && tj0 + TY + AP - 1 > je);
      }
      if (!skip) {
        const int o = base3 + jo * sj + i;
        if (st_fx && jo >= js && jo <= je) {
          if (acc) {
            if (has_mass) a.fx[o] = ofx + 0.5 * dampk * (m_l + m_0) * fxo;
            else a.fx[o] = ofx + fxo;
          } else {
            a.fx[o] = fxo;
          }
        }
        if (st_fy && jo >= js && jo <= je + 1) {
          if (acc) {
            if (has_mass) a.fy[o] = ofy + 0.5 * dampk * (m_b + m_0) * fyo;
            else a.fy[o] = ofy + fyo;
          } else {
            a.fy[o] = fyo;
          }
        }
        if (a.d2out && st_d) a.d2out[o] = dout;
      }
    }
    d2p = d2;
    d1p = d1; fx1p = fx1; fy1p = fy1;
    d0p = d0; fx0p = fx0; fy0p = fy0;
    ra3 = ra2;
    dv2 = dv1; du2 = du1; ra2 = ra1;
    dv1 = dv0; du1 = du0; ra1 = ra0;
    qh3 = qh2; qh2 = qh1; qh1 = qv;
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
  f.corner_only = 0;
  f.heat = 0; f.hdw = nullptr; f.hheat = nullptr; f.col_damp_w = nullptr; f.col_ke_bg = nullptr; f.absdt = 0.0;
  if (a.heat && !a.accumulate) {
    f.heat = 1; f.hdw = a.heat->dw; f.hheat = a.heat->heat_s; f.col_damp_w = a.heat->col_damp_w;
    f.col_ke_bg = a.heat->col_ke_bg; f.absdt = a.heat->absdt; f.d2out = nullptr;
    a.heat->done = true;
  }
  if (!c.err.empty()) return 1;
  dim3 blk(TX, NTY, 1);
  dim3 grd((g.nx + TX - 1) / TX, (g.ny + TY - 1) / TY, a.nk * g.nt);
  static const bool tile_only = [] { const char* v = getenv("PACE_B200_DELNFLUX"); return v && std::string(v) == "tile"; }();
  static const int rows_env = [] { const char* v = getenv("PACE_B200_DELN_ROWS"); return v ? atoi(v) : 0; }();
  if (tile_only) {
    LaunchScope scope(c.stream, "delnflux_fused");
    k_delnflux_fused<<<grd, blk, 0, c.stream>>>(g, f);
    return 0;
  }
  DelnMarchArgs m;
  m.f = f;
  m.ntx = grd.x; m.nty = grd.y;
  m.nstrips = (g.nx + 1 + MSW - 1) / MSW;
  // row chunks: enough warps to fill the GPU a few times over, chunks long enough that the 3 - 4 warm-up rows stay cheap
  const int nrows = g.ny + 3;  // js-1 .. je+1
  int rows = rows_env > 0 ? rows_env : 48;
  m.nchunks = (nrows + rows - 1) / rows;
  m.rows = (nrows + m.nchunks - 1) / m.nchunks;
  m.items = (long long)m.nstrips * m.nchunks * a.nk * g.nt;
  {
    LaunchScope scope(c.stream, "delnflux_march");
    const unsigned nb = (unsigned)((m.items + 3) / 4);
    if (f.heat) k_delnflux_march<false, true><<<nb, 128, 0, c.stream>>>(g, m);
    else if (f.accumulate) k_delnflux_march<true, false><<<nb, 128, 0, c.stream>>>(g, m);
    else k_delnflux_march<false, false><<<nb, 128, 0, c.stream>>>(g, m);
  }
  // the tiles that see cube-corner cells (levels with nord > 0)
  const unsigned eb0 = (unsigned)(g.edge_bits[0] | g.edge_bits[1]), eb1 = (unsigned)((g.edge_bits[0] | g.edge_bits[1]) >> 32);
  if (a.nmax > 0 && (eb0 | eb1)) {
    f.corner_only = 1;
    LaunchScope scope(c.stream, "delnflux_fused");
    k_delnflux_fused<<<grd, blk, 0, c.stream>>>(g, f);
  }
  return 0;
}

#else
int delnflux_fused_launch(Core&, const DelnArgs&) { return -1; }
#endif

}  // namespace pace
