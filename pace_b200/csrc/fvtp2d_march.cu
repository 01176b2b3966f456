// Row-marching fvtp2d for sm_100a: one kernel per FiniteVolumeTransport call, all intermediate fields in registers,
// warp shuffles and warp-private shared-memory rows.
//
// A WARP owns one level of one subdomain, a strip of 26 columns (lane <-> i, with a 3-column apron each side) and a
// chunk of rows, and marches over the rows R = j_lo-3 .. j_hi+3 without any CTA-wide barrier.  Per step it
//   1. reads row R of q (seen through copy_corners_x and copy_corners_y where they differ: corner blocks only),
//   2. x-direction, row R:   inner x-PPM flux (edge values / slopes exchanged by shuffles), q_j(i, R) (fvtp2d.py:57-77)
//   3. y-direction, column i: the inner sweep on q and the outer sweep on q_j advance by one row; PPM edge values /
//      limited slopes are carried in registers, so each is computed once per cell                 (yppm.py)
//      -> inner y flux and final fy at interface R-2, q_i(i, R-3)                                  (fvtp2d.py:34-54)
//   4. x-direction, row R-3: outer x-PPM flux on q_i -> final fx                                   (fvtp2d.py:80-119)
// Input rows are staged global -> shared two steps ahead with cp.async.  Every field is read once and written once
// (DRAM traffic = the 7 / 9 algorithmic passes + 6 warm-up rows per chunk; the column aprons hit in L1/L2).  All expressions are those of ppm.cuh / transport.cu, evaluated per element in the same order: results
// are bit-identical to the three-kernel version and to the reference.
#include "corners.cuh"
#include "ppm.cuh"
#include "transport.hpp"
#include "tile.cuh"

namespace pace {

#ifndef PACE_HOST_EMU

namespace {

struct MarchArgs {
  const double *q, *crx, *cry, *xfx, *yfx, *xunit, *yunit, *area, *dxa, *dya;
  double *fx, *fy;
  const int* lvl;                  // optional level_nonzero flags (Fvtp2dArgs)
  int nk, rows, nstrips, nchunks;  // levels; rows per chunk; column strips; row chunks
  int zrows;                       // rows of zeros after which a step is skipped (7), or a number no chunk reaches
  long long items;                 // nstrips * nchunks * nk * subdomains
};

// State of one PPM sweep that advances along its own direction one cell per step.
template <int MORD>
struct March {
  double w[5];        // transported scalar at rows R-4 .. R; the cell being completed is w[2]
  double al_lo;       // MORD < 8: edge value at the lower interface of w[2]
  double dm_m, dm_0;  // MORD >= 8: limited slopes of w[1], w[2]
  double pbl, pbr;    // previous cell (w[1]): PPM perturbations
  bool ps;
  __device__ __forceinline__ void push(double v) {
    w[0] = w[1]; w[1] = w[2]; w[2] = w[3]; w[3] = w[4]; w[4] = v;
  }
};

// bl, br (and the smoothness flag of the mord < 8 schemes) of the cell in the middle of the window; `pos` is its
// index along the sweep, `first` = nothing carried yet, `near` = an edge formula may apply (general code path).
template <int MORD, class DF>
__device__ __forceinline__ void march_cell(March<MORD>& m, int pos, bool first, bool near, DF D, int lo, int hi,
                                           bool elo, bool ehi, double& bl, double& br, bool& s) {
  auto Q = [&](int d) { return m.w[2 + d]; };
  if (MORD < 8) {
    using namespace ppmc;
    double al_hi;
    if (near) {
      if (first) m.al_lo = ppm_al_vals(Q(-2), Q(-1), Q(0), Q(1), D(-2), D(-1), D(0), D(1), pos, lo, hi, elo, ehi);
      al_hi = ppm_al_vals(Q(-1), Q(0), Q(1), Q(2), D(-1), D(0), D(1), D(2), pos + 1, lo, hi, elo, ehi);
    } else {
      if (first) m.al_lo = p1 * (m.w[1] + m.w[2]) + p2 * (m.w[0] + m.w[3]);
      al_hi = p1 * (m.w[2] + m.w[3]) + p2 * (m.w[1] + m.w[4]);
    }
    bl = m.al_lo - m.w[2];
    br = al_hi - m.w[2];
    const double b0 = bl + br;
    s = MORD == 5 ? (bl * br < 0) : ((3.0 * fabs(b0)) < fabs(bl - br));
    m.al_lo = al_hi;
  } else {
    if (first) {
      m.dm_m = ppm_dm8_v(m.w[0], m.w[1], m.w[2]);
      m.dm_0 = ppm_dm8_v(m.w[1], m.w[2], m.w[3]);
    }
    const double dm_p = ppm_dm8_v(m.w[2], m.w[3], m.w[4]);
    if (near) {
      const BlBr r = ppm_blbr8_vals(Q(-2), Q(-1), Q(0), Q(1), Q(2), D(-2), D(-1), D(0), D(1), D(2), pos, lo, hi, elo, ehi);
      bl = r.bl;
      br = r.br;
    } else {
      const double al_lo = 0.5 * (m.w[1] + m.w[2]) + 1.0 / 3.0 * (m.dm_m - m.dm_0);
      const double al_hi = 0.5 * (m.w[2] + m.w[3]) + 1.0 / 3.0 * (m.dm_0 - dm_p);
      const double xt = 2.0 * m.dm_0;
      bl = -1.0 * fv_sign(fv_min(fabs(xt), fabs(al_lo - m.w[2])), xt);
      br = fv_sign(fv_min(fabs(xt), fabs(al_hi - m.w[2])), xt);
    }
    s = false;
    m.dm_m = m.dm_0;
    m.dm_0 = dm_p;
  }
}

// flux through the interface between the previous cell (m) and the current one (0) for Courant number c
template <int MORD>
__device__ __forceinline__ double cells_flux(double c, double qm, double blm, double brm, bool sm, double q0,
                                             double bl, double br, bool s0) {
  if (MORD < 8) {
    const double b0 = bl + br, b0m = blm + brm;
    const double mask = (sm || s0) ? 1.0 : 0.0;
    const double fx1 = c > 0.0 ? (1.0 - c) * (brm - c * b0m) : (1.0 + c) * (bl + c * b0);
    return c > 0.0 ? qm + fx1 * mask : q0 + fx1 * mask;
  }
  if (c > 0.0) {
    const double b0 = blm + brm;
    return qm + (1.0 - c) * (brm - c * b0) * 1.0;
  }
  const double b0 = bl + br;
  return q0 + (1.0 + c) * (bl + c * b0) * 1.0;
}

#ifndef PACE_MARCH_PF_ROWS
#define PACE_MARCH_PF_ROWS 6
#endif
constexpr int PF_ROWS = PACE_MARCH_PF_ROWS;  // L2 prefetch distance (rows) beyond the two-row cp.async staging
constexpr int NSLOT = 3, NSTREAM = 10;
constexpr int SW = 26;         // cells per warp strip (32 lanes - 2 x 3 apron)
constexpr int WPB = 1;         // warps per CTA (independent work items)
constexpr int PITCH = 34;      // doubles per staged row
constexpr int WARP_SMEM = (NSTREAM * NSLOT + 2 + 4) * PITCH;  // doubles per warp
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(gmem_src) : "memory");
}

// PPM flux through the left interface of lane L's cell, sweep along the lanes.  `row` holds the transported scalar of
// the 32 columns of the strip; `i` is the lane's column, D(d) the grid spacing at column i+d.  Each lane computes the
// edge value (MORD < 8) or the limited slope and bl/br (MORD >= 8) of its own interface / cell once; neighbours get
// them by shuffle.  Must be called by all 32 lanes; the result is meaningful for lanes 3..29.
template <int MORD, class DF>
__device__ __forceinline__ double xsweep_flux(const double* row, int L, int i, double c, DF D, int is, int ie, bool W,
                                              bool E) {
  // Every lane evaluates the interior formulas on its own (clamped) neighbourhood without a branch, so that the
  // independent chains of a step share basic blocks and can be interleaved by the scheduler; what the out-of-range
  // lanes 0, 1, 30, 31 produce is never used by a lane whose result is.  Lanes next to a tile edge then replace their
  // value by the edge formula.
  const int Lm1 = L > 0 ? L - 1 : 0, Lm2 = L > 1 ? L - 2 : 0;
  const double qm2 = row[Lm2], qm = row[Lm1], q0 = row[L], qp = row[L + 1];
  auto D5 = D;
  if (MORD < 8) {
    using namespace ppmc;
    double al = p1 * (qm + q0) + p2 * (qm2 + qp);
    if (L >= 2 && L <= 30 && ((W && i >= is - 1 && i <= is + 1) || (E && i >= ie && i <= ie + 2)))
      al = ppm_al_vals(qm2, qm, q0, qp, D5(-2), D5(-1), D5(0), D5(1), i, is, ie, W, E);
    const double alm = __shfl_up_sync(FULL, al, 1), alp = __shfl_down_sync(FULL, al, 1);
    const double bl = al - q0, br = alp - q0, b0 = bl + br;
    const double blm = alm - qm, brm = al - qm, b0m = blm + brm;
    bool s0, sm;
    if (MORD == 5) {
      s0 = bl * br < 0;
      sm = blm * brm < 0;
    } else {
      s0 = (3.0 * fabs(b0)) < fabs(bl - br);
      sm = (3.0 * fabs(b0m)) < fabs(blm - brm);
    }
    const double mask = (sm || s0) ? 1.0 : 0.0;
    const double fx1 = c > 0.0 ? (1.0 - c) * (brm - c * b0m) : (1.0 + c) * (bl + c * b0);
    return c > 0.0 ? qm + fx1 * mask : q0 + fx1 * mask;
  } else {
    const double dm = ppm_dm8_v(qm, q0, qp);
    const double dm_m = __shfl_up_sync(FULL, dm, 1), dm_p = __shfl_down_sync(FULL, dm, 1);
    const double al_lo = 0.5 * (qm + q0) + 1.0 / 3.0 * (dm_m - dm);
    const double al_hi = 0.5 * (q0 + qp) + 1.0 / 3.0 * (dm - dm_p);
    const double xt = 2.0 * dm;
    double bl = -1.0 * fv_sign(fv_min(fabs(xt), fabs(al_lo - q0)), xt);
    double br = fv_sign(fv_min(fabs(xt), fabs(al_hi - q0)), xt);
    if (L >= 2 && L <= 29 && ((W && i >= is - 1 && i <= is + 1) || (E && i >= ie - 1 && i <= ie + 1))) {
      const BlBr r = ppm_blbr8_vals(qm2, qm, q0, qp, row[L + 2], D5(-2), D5(-1), D5(0), D5(1), D5(2), i, is, ie, W, E);
      bl = r.bl;
      br = r.br;
    }
    const double blm = __shfl_up_sync(FULL, bl, 1), brm = __shfl_up_sync(FULL, br, 1);
    return cells_flux<8>(c, qm, blm, brm, false, q0, bl, br, false);
  }
}

template <int OIN, int OOUT>
__global__ void __launch_bounds__(WPB * 32, 20) k_fvtp2d_march(Geo g, MarchArgs a) {
  extern __shared__ double smem[];
  const int L = threadIdx.x & 31, wid = threadIdx.x >> 5;
  // work item of this warp: (strip, row chunk, level, subdomain)
  long long item = (long long)blockIdx.x * WPB + wid;
  if (item >= a.items) return;  // whole warp; no CTA-wide barrier is used anywhere below
  const int strip = (int)(item % a.nstrips);
  item /= a.nstrips;
  const int chunk = (int)(item % a.nchunks);
  item /= a.nchunks;
  const int k = (int)(item % a.nk), t = (int)(item / a.nk);
  if (a.lvl != nullptr && a.lvl[t * a.nk + k] == 0) return;  // a level of zeros: nothing to transport

  double* bufC0 = smem + wid * WARP_SMEM;  // q_i row, two copies used alternately (no barrier between the last
  double* fxring = bufC0 + 2 * PITCH;      // read of one step and the write of the next); inner x flux of the last
  double* rings = fxring + 4 * PITCH;      // 4 rows (own lane only: no synchronisation); then the staged input rows

  const int is = g.is, ie = g.ie, js = g.js, je = g.je, sj = g.sj;
  const unsigned ed = edge_of(g, t);
  const bool W = ed & EDGE_W, E = ed & EDGE_E, S = ed & EDGE_S, N = ed & EDGE_N;
  const EdgeInfo e{is, ie, js, je, W, E, S, N};
  const int i0 = is + strip * SW;                     // first cell / interface this strip writes
  const int i = i0 - 3 + L;                           // my column
  const int i_out_hi = min(i0 + SW - 1, ie);          // last cell this strip writes
  const bool col_ok = i <= ie + 3;                    // column exists (halo included)
  const bool cell_out = i >= i0 && i <= i_out_hi;     // fy, q_j owner
  const bool xface_out = i >= i0 && (i <= i_out_hi || (i == ie + 1 && i_out_hi == ie));  // fx owner
  const int j_lo = js + chunk * a.rows, j_hi = min(j_lo + a.rows - 1, je);
  const bool last_chunk = j_hi == je;

  const long long base3 = (long long)t * g.st3 + (long long)k * g.sk;
  const double *q = a.q + base3, *crx = a.crx + base3, *cry = a.cry + base3, *xfx = a.xfx + base3,
               *yfx = a.yfx + base3, *xunit = a.xunit + base3, *yunit = a.yunit + base3;
  double *fx = a.fx + base3, *fy = a.fy + base3;
  const int base2 = t * g.sk;
  const double *area = a.area + base2, *dxa = a.dxa + base2, *dya = a.dya + base2;

  constexpr int MI = OIN < 8 ? OIN : 8, MO = OOUT < 8 ? OOUT : 8;
  March<MI> mi;  // inner y sweep on q
  March<MO> mo;  // outer y sweep on q_j
#pragma unroll
  for (int n = 0; n < 5; ++n) mi.w[n] = mo.w[n] = 0.0;
  mi.al_lo = mo.al_lo = mi.dm_m = mi.dm_0 = mo.dm_m = mo.dm_0 = 0.0;
  mi.pbl = mi.pbr = mo.pbl = mo.pbr = 0.0;
  mi.ps = mo.ps = false;
  double fyin_prev = 0.0;                                 // inner y flux at interface R-3
  double yfx_prev = 0.0;                                  // yfx(i, R-3)
  int zrun = 0;  // consecutive rows, ending at R, on which q is zero in all 32 columns of the strip

  // Asynchronous staging: every input row is copied global -> shared two steps ahead of its use (cp.async, no
  // registers held), in a ring of NSLOT slots per stream.  Streams and the row they carry at step R:
  //   0 q(R)  1 crx(R)  2 xfx(R)  3 area(R)  4 cry(R-2)  5 yunit(R-2)  6 yfx(R-2)  7 area(R-3)  8 crx(R-3)  9 xunit(R-3)
  auto ring = [&](int stream, int slot) { return rings + (stream * NSLOT + slot) * PITCH; };
  auto issue = [&](int sn) {
    const int Rn = j_lo - 3 + sn;
    if (col_ok && Rn <= j_hi + 3) {
      const int slot = sn % NSLOT;
      const int on = Rn * sj + i;
      cp_async8(ring(0, slot) + L, q + on);
      cp_async8(ring(1, slot) + L, crx + on);
      cp_async8(ring(2, slot) + L, xfx + on);
      cp_async8(ring(3, slot) + L, area + on);
      if (Rn >= 2) {
        const int o2 = on - 2 * sj;
        cp_async8(ring(4, slot) + L, cry + o2);
        cp_async8(ring(5, slot) + L, yunit + o2);
        cp_async8(ring(6, slot) + L, yfx + o2);
      }
      if (Rn >= 3) {
        const int o3 = on - 3 * sj;
        cp_async8(ring(7, slot) + L, area + o3);
        cp_async8(ring(8, slot) + L, crx + o3);
        cp_async8(ring(9, slot) + L, xunit + o3);
      }
    }
    // ask L2 for the rows a few steps further ahead (one request per 32-byte sector, no register or shared-memory slot
    // tied up): the cp.async of those rows then finds them in L2 instead of waiting on DRAM
    if (PF_ROWS > 0 && col_ok && (L & 3) == 0 && Rn + PF_ROWS <= j_hi + 3) {
      const int op = (Rn + PF_ROWS) * sj + i;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(q + op));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(crx + op));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(xfx + op));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(cry + op - 2 * sj));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(yfx + op - 2 * sj));
      if (yunit != yfx) {  // mass fluxes given: two more streams
        asm volatile("prefetch.global.L2 [%0];" ::"l"(yunit + op - 2 * sj));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(xunit + op - 3 * sj));
      }
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  };
  issue(0);
  issue(1);

  for (int R = j_lo - 3, s = 0; R <= j_hi + 3; ++R, ++s) {
    const int o = R * sj + i;
    const int slot = s % NSLOT;
    asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    __syncwarp();  // row R has landed for every lane; every lane is done with the slot refilled next
    issue(s + 2);
    double* qrow = ring(0, slot);
    // 1. row R of q; the corner blocks differ between the x and the y view
    double qx = 0.0, qy = 0.0;
    if (col_ok) qx = qy = qrow[L];
    if ((S && R < js) || (N && R > je)) {  // row-uniform
      if (col_ok && ((W && i < is) || (E && i > ie))) {
        int xi = i, xj = R, yi = i, yj = R;
        copyx_src(e, xi, xj);
        copyy_src(e, yi, yj);
        qx = q[xj * sj + xi];
        qy = q[yj * sj + yi];
        qrow[L] = qx;
      }
      __syncwarp();
    }
    // Rows of zeros.  Every flux, q_i and q_j of this step is formed from rows R-6 .. R of the strip.  If q is zero on
    // all of them (condensate species outside clouds and precipitation, q_con in dry air, w at rest), every PPM
    // perturbation is a zero, the fluxes through the interfaces are +0 and so are q_i and q_j, whatever the Courant
    // numbers: the step reduces to storing the zero fluxes (times the mass fluxes: their sign is the sign of the stored
    // zero).  The carried sweep state is already the one a window of zeros leaves and stays as it is.
    zrun = __all_sync(FULL, qx == 0.0 && qy == 0.0) ? zrun + 1 : 0;
    if (zrun >= a.zrows) {
      const int jc = R - 2;
      if (cell_out && (jc <= j_hi || last_chunk)) fy[o - 2 * sj] = 0.5 * (0.0 + 0.0) * ring(5, slot)[L];
      yfx_prev = ring(6, slot)[L];
      if (xface_out) fx[o - 3 * sj] = 0.5 * (0.0 + 0.0) * ring(9, slot)[L];
      fxring[(s & 3) * PITCH + L] = 0.0;
      continue;
    }
    // 2. inner x flux and q_j on row R
    double fxin;
    {
      auto D = [&](int d) { return dxa[o + d]; };
      fxin = xsweep_flux<MI>(qrow, L, i, ring(1, slot)[L], D, is, ie, W, E);
    }
    const double fxin_p = __shfl_down_sync(FULL, fxin, 1);
    double qj = 0.0;
    {
      const double xf0 = ring(2, slot)[L], xf1 = ring(2, slot)[L + 1], ar = ring(3, slot)[L];
      const double fx10 = xf0 * fxin, fx11 = xf1 * fxin_p;
      const double area_with_x_flux = ar + xf0 - xf1;
      if (cell_out) qj = fv_div(qx * ar + fx10 - fx11, area_with_x_flux);
    }
    fxring[(s & 3) * PITCH + L] = fxin;  // needed again at step s + 3 (outer x flux of this row)
    // 3. y sweeps: cell jc = R-2 is complete
    mi.push(qy);
    mo.push(qj);
    double qi = 0.0;
    if (s >= 4) {
      const int jc = R - 2, oc = o - 2 * sj;
      const bool first = s == 4;
      const bool near = (S && jc <= js + 1) || (N && jc >= je - 1);
      auto D = [&](int d) { return dya[oc + d * sj]; };
      double bli = 0.0, bri = 0.0, blo = 0.0, bro = 0.0;
      bool si = false, so = false;
      if (near) {  // rows next to a tile edge (row-uniform): the edge formulas read the grid spacing
        if (col_ok) march_cell<MI>(mi, jc, first, true, D, js, je, S, N, bli, bri, si);
        if (cell_out) march_cell<MO>(mo, jc, first, true, D, js, je, S, N, blo, bro, so);
      } else {     // registers only: all lanes, the two sweeps in one basic block (columns that do not exist carry zeros)
        march_cell<MI>(mi, jc, first, false, D, js, je, S, N, bli, bri, si);
        march_cell<MO>(mo, jc, first, false, D, js, je, S, N, blo, bro, so);
      }
      if (s >= 5) {  // interface jc, between cells jc-1 and jc
        const double cy = ring(4, slot)[L];
        const double fyin = cells_flux<MI>(cy, mi.w[1], mi.pbl, mi.pbr, mi.ps, mi.w[2], bli, bri, si);
        const double outer = cells_flux<MO>(cy, mo.w[1], mo.pbl, mo.pbr, mo.ps, mo.w[2], blo, bro, so);
        if (cell_out && (jc <= j_hi || last_chunk)) fy[oc] = 0.5 * (outer + fyin) * ring(5, slot)[L];
        const double yf1 = ring(6, slot)[L];
        if (s >= 6 && col_ok) {  // q_i on row R-3
          const double yf0 = yfx_prev, ar = ring(7, slot)[L];
          const double fyy0 = yf0 * fyin_prev, fyy1 = yf1 * fyin;
          qi = fv_div(mi.w[1] * ar + fyy0 - fyy1, ar + yf0 - yf1);
        }
        yfx_prev = yf1;
        fyin_prev = fyin;
      }
      mi.pbl = bli; mi.pbr = bri; mi.ps = si;
      mo.pbl = blo; mo.pbr = bro; mo.ps = so;
    }
    // 4. outer x flux on row R-3
    if (s >= 6) {
      double* bufC = bufC0 + (s & 1) * PITCH;
      bufC[L] = qi;
      __syncwarp();
      const int orow = o - 3 * sj;
      auto D = [&](int d) { return dxa[orow + d]; };
      const double outer = xsweep_flux<MO>(bufC, L, i, ring(8, slot)[L], D, is, ie, W, E);
      if (xface_out) fx[orow] = 0.5 * (outer + fxring[((s + 1) & 3) * PITCH + L]) * ring(9, slot)[L];
    }
  }
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

template <int OIN, int OOUT>
void launch(Core& c, int blocks, size_t smem, cudaStream_t st, const Geo& g, const MarchArgs& f) {
  // opt in to the dynamic shared memory once per instantiation AND device (tile.cu)
  if (func_attr_once(c, (const void*)k_fvtp2d_march<OIN, OOUT>, (int)smem)) return;
  k_fvtp2d_march<OIN, OOUT><<<blocks, WPB * 32, smem, st>>>(g, f);
}
}  // namespace

// SMs of the current device (queried once per device)
static int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int fvtp2d_march_launch(Core& c, const Fvtp2dArgs& a) {
  const Geo g = c.g;
  MarchArgs f;
  f.q = a.q; f.crx = a.crx; f.cry = a.cry; f.xfx = a.xfx; f.yfx = a.yfx;
  f.xunit = a.x_mass_flux ? a.x_mass_flux : a.xfx;
  f.yunit = a.y_mass_flux ? a.y_mass_flux : a.yfx;
  f.area = c.F("grid.area"); f.dxa = c.F("grid.dxa"); f.dya = c.F("grid.dya");
  f.fx = a.fx; f.fy = a.fy;
  f.lvl = a.level_nonzero;
  f.zrows = zero_paths_enabled() ? 7 : (1 << 30);
  f.nk = a.nk;
  if (!c.err.empty()) return 1;
  const int oo = a.hord < 0 ? -a.hord : a.hord;
  const int oi = oo == 10 ? 8 : oo;
  static const int rows_env = [] { const char* v = getenv("PACE_B200_FVTP2D_ROWS"); return v ? atoi(v) : 0; }();
  f.nstrips = (g.nx + SW - 1) / SW;
  // rows per chunk: 64 (6 warm-up rows = 9 % extra work) unless that leaves fewer than ~4 waves of warps on the
  // GPU (SMs x 16 resident warps), then 32 (19 % extra) or 16 (37 %) — small per-GPU domains (multi-GPU runs) need the
  // parallelism, but at 3 subdomains per GPU the warm-up rows of 16-row chunks cost more than a shorter last wave
  f.rows = 64;
  if (rows_env > 0) {
    f.rows = rows_env;
  } else {
    const long long slots = (long long)sm_count() * 16;
    while (f.rows > 16 && (long long)f.nstrips * ((g.ny + f.rows - 1) / f.rows) * a.nk * g.nt < 4 * slots) f.rows /= 2;
  }
  if (f.rows > g.ny) f.rows = g.ny;
  f.nchunks = (g.ny + f.rows - 1) / f.rows;
  f.items = (long long)f.nstrips * f.nchunks * a.nk * g.nt;
  const int blocks = (int)((f.items + WPB - 1) / WPB);
  const size_t smem = (size_t)WPB * WARP_SMEM * sizeof(double);
  // launches that carry level flags (tracer species, q_con) are timed apart: their duration depends on how many levels
  // hold anything, so they would blur the per-launch figure of the dense launches the roofline is quoted on
  LaunchScope scope(c.stream, a.level_nonzero ? "fvtp2d_march_sparse" : "fvtp2d_march");
  if (oi == 5 && oo == 5) launch<5, 5>(c, blocks, smem, c.stream, g, f);
  else if (oi == 6 && oo == 6) launch<6, 6>(c, blocks, smem, c.stream, g, f);
  else if (oi == 7 && oo == 7) launch<7, 7>(c, blocks, smem, c.stream, g, f);
  else if (oi == 8 && oo == 8) launch<8, 8>(c, blocks, smem, c.stream, g, f);
  else if (oi == 8 && oo == 10) launch<8, 10>(c, blocks, smem, c.stream, g, f);
  else return -1;  // not instantiated: the caller falls back to the tiled kernel
  return 0;
}

#else  // host emulation keeps the straightforward three-kernel version (transport.cu)
int fvtp2d_march_launch(Core&, const Fvtp2dArgs&) { return -1; }
#endif

}  // namespace pace
