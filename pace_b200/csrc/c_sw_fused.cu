// c_sw as ONE tiled kernel: the 23 reference stencils of CGridShallowWaterDynamics.__call__ incl. DGrid2AGrid2CGridVectors
// run phase by phase on shared-memory tiles (tile.cuh); DRAM sees u, v, delp, pt, w once (TMA loads, + apron overlap)
// and the ten outputs once.  The unfused version (c_sw.cu: 13 launches, ~45 field passes through seven work arrays)
// stays as the A/B baseline (PACE_B200_CSW=unfused).
//
// Reference: pyFV3/stencils/c_sw.py:21-500 (stencils), :644-807 (__call__ order); d2a2c_vect.py:20-377, :588-699.
// Every expression tree is the reference's, as in c_sw.cu; the phases and what they need from their neighbours
// (x / y reach relative to the patch a CTA owns; 3 cells suffice, the apron along i is 4 for TMA alignment):
//   A  utmp (reads u at j-1..j+2), vtmp (v at i-1..i+2), ua / va               needed on [-3,2] x [-1,1] (+ transposed)
//   c1 corner fills utmp <- vtmp, ua <- va (x) and vtmp <- utmp, va <- ua (y)   corner patches only
//   B  uc, ut (utmp at i-2..i+1) and vc, vt (vtmp at j-2..j+1)                  needed on [-1,1] x [-1,1]
//   C  divergence_corner                                                        own cells
//   E  delpc, ptc, omga (upstream fluxes of delp, pt, w formed in registers) and ke   ke needed on [-1,0] x [-1,0]
//   F  absolute vorticity at corners                                            needed on [0,1] x [0,1]
//   G  update of uc, vc                                                         own cells
#include "tile.cuh"

namespace pace {

namespace {
constexpr double C1 = -2.0 / 14.0, C2 = 11.0 / 14.0, C3 = 5.0 / 14.0;  // d2a2c_vect.py:14-16
constexpr double A1 = 9.0 / 16.0, A2C = -1.0 / 16.0;                   // a2b_ord4.py:32-33
constexpr double BIG_NUMBER = 1e30;                                    // d2a2c_vect.py:414

#ifndef PACE_HOST_EMU
typedef TileCfg<40, 16, 4, 3, 48, 11> CswCfg;
#else
// the host emulation runs small patches so that the C48 fixture has interior patches (no edge formulas), edge patches
// and corner patches: every code path of the tiling is then checked against the reference on the CPU
typedef TileCfg<16, 8, 4, 3, 24, 2> CswCfg;
#endif
constexpr int CSW_ARRAYS = 13;
constexpr int CSW_CTAS_PER_SM = (2 * (CSW_ARRAYS * CswCfg::BYTES + 1152) <= 232448 && 2 * CswCfg::NT <= 2048) ? 2 : 1;

struct CswArgs {
  TileField u, v, delp, pt, w;                                    // loaded
  double *uc, *vc, *ua, *va, *ut, *vt, *divgd, *omga, *delpc, *ptc;  // stored
  const double *cosa_s, *rsin2, *cosa_u, *cosa_v, *rsin_u, *rsin_v, *sina_u, *sina_v, *dxa, *dya, *dx, *dy, *dxc, *dyc,
      *rdxc, *rdyc, *sin_sg1, *sin_sg2, *sin_sg3, *sin_sg4, *cos_sg1, *cos_sg2, *cos_sg3, *cos_sg4, *rarea, *rarea_c, *fC;
  double dt2;
  int nord, npt;
  int part;   // 0: every patch; 1: patches whose box lies inside the compute domain; 2: the others
  TileGrid tg;
};

// All phases of one patch.  S: CSW_ARRAYS tile arrays; the first five hold u, v, delp, pt, w.
// EDGES = false: no tile-edge formula can apply inside this patch's box (Tile::near_edge), every edge test folds away.
// wait_dpw: called once before the first read of the delp / pt / w tiles (their TMA loads run under phases A - C)
template <class C, bool EDGES, class WaitF>
__device__ __forceinline__ void c_sw_tile(const Tile<C>& T, const CswArgs& a, typename Tile<C>::Arr* S, WaitF wait_dpw) {
  constexpr int W = C::W, H = C::H;
  auto &Su = S[0], &Sv = S[1], &Sdp = S[2], &Spt = S[3], &Sw = S[4];
  auto &Sutmp = S[5], &Svtmp = S[6], &Sua = S[7], &Sva = S[8], &Suc = S[9], &Svc = S[10], &Sut = S[11], &Svt = S[12];
  auto &Ske = S[5], &Svort = S[6];  // utmp / vtmp are dead after phase B
  const Geo& g = T.g;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;
  const EdgeInfo e = EDGES ? T.e : EdgeInfo{is, ie, js, je, false, false, false, false};
  const double dt2 = a.dt2;
  const int npt = a.npt, avg_off = (npt == 0) ? -1 : 3;
  const int bi = T.bi, bj = T.bj;
#define M2(f, i, j) (a.f[T.o2((i), (j))])
  auto owned = [&](int i, int j) { return i >= T.oi0 && i < T.oi1 && j >= T.oj0 && j < T.oj1; };

  // ---- A: utmp, vtmp (set_tmps + lagrange_interpolation_{y,x}_p1 + avg_box, d2a2c_vect.py:28-95, 421-540) and
  //      ua, va (contravariant_components) in one pass over the box: utmp needs u at lj-1..lj+2, vtmp v at li-1..li+2
  T.each(0, g.nx + 5, 0, g.ny + 5, 0, W - 1, 0, H - 1, [&](int i, int j, int li, int lj) {
    const bool edge = (e.S && j < js + avg_off) || (e.N && j > je - avg_off) || (e.W && i < is + avg_off) ||
                      (e.E && i > ie - avg_off);
    const bool has_u = lj >= 1 && lj <= H - 3, has_v = li >= 1 && li <= W - 3;
    double ut_ = BIG_NUMBER, vt_ = BIG_NUMBER;
    if (has_u) {
      const int js1 = e.S ? npt + 2 : js - 1, je1 = e.N ? (je + 1) - npt : je + 1;
      const int is1 = e.W ? npt + 2 : 0, ie1 = e.E ? (ie + 1) - npt : ie + 3;
      if (i >= is1 && i <= ie1 && j >= js1 && j <= je1)
        ut_ = A2C * (Su[lj - 1][li] + Su[lj + 2][li]) + A1 * (Su[lj][li] + Su[lj + 1][li]);
      if (edge) ut_ = 0.5 * (Su[lj][li] + Su[lj + 1][li]);
      Sutmp[lj][li] = ut_;
    }
    if (has_v) {
      const int is2 = e.W ? npt + 2 : is - 1, ie2 = e.E ? (ie + 1) - npt : ie + 1;
      const int js2 = e.S ? npt + 2 : 0, je2 = e.N ? (je + 1) - npt : je + 3;
      if (i >= is2 && i <= ie2 && j >= js2 && j <= je2)
        vt_ = A2C * (Sv[lj][li - 1] + Sv[lj][li + 2]) + A1 * (Sv[lj][li] + Sv[lj][li + 1]);
      if (edge) vt_ = 0.5 * (Sv[lj][li] + Sv[lj][li + 1]);
      Svtmp[lj][li] = vt_;
    }
    // contravariant_components on compute(-2,-2)+(4,4)
    if (has_u && has_v && i >= is - 2 && i <= ie + 2 && j >= js - 2 && j <= je + 2) {
      const double cs = M2(cosa_s, i, j), rs = M2(rsin2, i, j);
      const double uav = (ut_ - vt_ * cs) * rs, vav = (vt_ - ut_ * cs) * rs;
      Sua[lj][li] = uav;
      Sva[lj][li] = vav;
      if (owned(i, j)) {
        a.ua[T.o3(i, j)] = uav;
        a.va[T.o3(i, j)] = vav;
      }
    }
  });

  // ---- c1: fill_corners_x (utmp <- vtmp 3 cells, ua <- va 2 cells), fill_corners_y (vtmp <- utmp, va <- ua);
  //      multipliers sw -1, se 1, nw 1, ne -1 (corners.py:129-304, vector fills)
  if (EDGES && T.corner_tile) {
    auto fill = [&](int id) {
      const int io = id / 12, c = (id % 12) / 3, n = id % 3 + 1;
      const int ncells = (io & 1) ? 2 : 3;
      const bool ydir = io >= 2;
      if (n > ncells) return;
      int wi, wj, ri, rj;
      double m;
      bool on;
      if (!ydir) {
        switch (c) {
          case 0: on = e.W && e.S; wi = is - n; wj = js - 1; ri = is - 1; rj = js - 1 + n; m = -1.0; break;
          case 1: on = e.E && e.S; wi = ie + n; wj = js - 1; ri = ie + 1; rj = js - 1 + n; m = 1.0; break;
          case 2: on = e.W && e.N; wi = is - n; wj = je + 1; ri = is - 1; rj = je + 1 - n; m = 1.0; break;
          default: on = e.E && e.N; wi = ie + n; wj = je + 1; ri = ie + 1; rj = je + 1 - n; m = -1.0; break;
        }
      } else {
        switch (c) {
          case 0: on = e.W && e.S; wi = is - 1; wj = js - n; ri = is - 1 + n; rj = js - 1; m = -1.0; break;
          case 1: on = e.E && e.S; wi = ie + 1; wj = js - n; ri = ie + 1 - n; rj = js - 1; m = 1.0; break;
          case 2: on = e.W && e.N; wi = is - 1; wj = je + n; ri = is - 1 + n; rj = je + 1; m = 1.0; break;
          default: on = e.E && e.N; wi = ie + 1; wj = je + n; ri = ie + 1 - n; rj = je + 1; m = -1.0; break;
        }
      }
      if (!on) return;
      const int lwi = wi - bi, lwj = wj - bj, lri = ri - bi, lrj = rj - bj;
      if (lwi < 0 || lwi >= W || lwj < 0 || lwj >= H || lri < 0 || lri >= W || lrj < 0 || lrj >= H) return;
      // io 0: utmp <- vtmp, 1: ua <- va, 2: vtmp <- utmp, 3: va <- ua
      double(*dst)[W] = io == 0 ? Sutmp : io == 1 ? Sua : io == 2 ? Svtmp : Sva;
      double(*src)[W] = io == 0 ? Svtmp : io == 1 ? Sva : io == 2 ? Sutmp : Sua;
      const double val = m * src[lrj][lri];
      dst[lwj][lwi] = val;
      if ((io & 1) && owned(wi, wj)) (io == 1 ? a.ua : a.va)[T.o3(wi, wj)] = val;
    };
#ifndef PACE_HOST_EMU
    const int id = T.ty * C::NTX + T.tx;
    if (id < 48) fill(id);
    __syncthreads();
#else
    // the sources of one op are never destinations of another (c_sw.cu), so any order gives the same result
    for (int id = 0; id < 48; ++id) fill(id);
#endif
  }

  // ---- B: uc, ut on [is-1, ie+2] x [js-1, je+1]; vc, vt on [is-1, ie+1] x [js-1, je+2];
  //      C: divergence_corner (c_sw.py:33-175) on the corner (interface x interface) compute domain, own cells.
  //      One pass: all three read only what phase A left.
  const bool do_div = a.nord > 0;
  T.each(is - 1, ie + 2, js - 1, je + 2, 1, W - 2, 1, H - 2, [&](int i, int j, int li, int lj) {
    if (li >= 2 && lj <= H - 3 && j <= je + 1) {
      double ucv, utv;
      auto contra = [&](double ucx) { return (ucx - Sv[lj][li] * M2(cosa_u, i, j)) * M2(rsin_u, i, j); };
      if ((e.W && i == is - 1) || (e.E && i == ie)) {
        ucv = C1 * Sutmp[lj][li - 2] + C2 * Sutmp[lj][li - 1] + C3 * Sutmp[lj][li];
        utv = contra(ucv);
      } else if ((e.W && i == is) || (e.E && i == ie + 1)) {  // edge_interpolate4_x (d2a2c_vect.py:360-366)
        const double t1 = M2(dxa, i - 2, j) + M2(dxa, i - 1, j);
        const double t2 = M2(dxa, i, j) + M2(dxa, i + 1, j);
        const double n1 = (t1 + M2(dxa, i - 1, j)) * Sua[lj][li - 1] - M2(dxa, i - 1, j) * Sua[lj][li - 2];
        const double n2 = (t1 + M2(dxa, i, j)) * Sua[lj][li] - M2(dxa, i, j) * Sua[lj][li + 1];
        utv = 0.5 * (n1 / t1 + n2 / t2);
        ucv = utv > 0 ? utv * M2(sin_sg3, i - 1, j) : utv * M2(sin_sg1, i, j);
      } else if ((e.W && i == is + 1) || (e.E && i == ie + 2)) {
        ucv = C1 * Sutmp[lj][li + 1] + C2 * Sutmp[lj][li] + C3 * Sutmp[lj][li - 1];
        utv = contra(ucv);
      } else {  // ut_main
        ucv = A2C * (Sutmp[lj][li - 2] + Sutmp[lj][li + 1]) + A1 * (Sutmp[lj][li - 1] + Sutmp[lj][li]);
        utv = contra(ucv);
      }
      Suc[lj][li] = ucv;
      // geoadjust_ut (c_sw.py:178-200)
      const double utf = utv > 0 ? dt2 * utv * M2(dy, i, j) * M2(sin_sg3, i - 1, j) : dt2 * utv * M2(dy, i, j) * M2(sin_sg1, i, j);
      Sut[lj][li] = utf;
      if (owned(i, j)) a.ut[T.o3(i, j)] = utf;
    }
    if (lj >= 2 && li <= W - 3 && i <= ie + 1) {
      double vcv, vtv;
      auto contra = [&](double vcx) { return (vcx - Su[lj][li] * M2(cosa_v, i, j)) * M2(rsin_v, i, j); };
      if ((e.S && j == js - 1) || (e.N && j == je)) {
        vcv = C1 * Svtmp[lj - 2][li] + C2 * Svtmp[lj - 1][li] + C3 * Svtmp[lj][li];
        vtv = contra(vcv);
      } else if ((e.S && j == js) || (e.N && j == je + 1)) {  // edge_interpolate4_y
        const double t1 = M2(dya, i, j - 2) + M2(dya, i, j - 1);
        const double t2 = M2(dya, i, j) + M2(dya, i, j + 1);
        const double n1 = (t1 + M2(dya, i, j - 1)) * Sva[lj - 1][li] - M2(dya, i, j - 1) * Sva[lj - 2][li];
        const double n2 = (t1 + M2(dya, i, j)) * Sva[lj][li] - M2(dya, i, j) * Sva[lj + 1][li];
        vtv = 0.5 * (n1 / t1 + n2 / t2);
        vcv = vtv > 0 ? vtv * M2(sin_sg4, i, j - 1) : vtv * M2(sin_sg2, i, j);
      } else if ((e.S && j == js + 1) || (e.N && j == je + 2)) {
        vcv = C1 * Svtmp[lj + 1][li] + C2 * Svtmp[lj][li] + C3 * Svtmp[lj - 1][li];
        vtv = contra(vcv);
      } else {  // vt_main / first region of north_south_edges
        vcv = A2C * (Svtmp[lj - 2][li] + Svtmp[lj + 1][li]) + A1 * (Svtmp[lj - 1][li] + Svtmp[lj][li]);
        vtv = contra(vcv);
      }
      Svc[lj][li] = vcv;
      // geoadjust_vt (c_sw.py:203-222)
      const double vtf = vtv > 0 ? dt2 * vtv * M2(dx, i, j) * M2(sin_sg4, i, j - 1) : dt2 * vtv * M2(dx, i, j) * M2(sin_sg2, i, j);
      Svt[lj][li] = vtf;
      if (owned(i, j)) a.vt[T.o3(i, j)] = vtf;
    }
    if (do_div && owned(i, j) && i >= is && j >= js && i <= ie + 1 && j <= je + 1) {
      auto uf_gen = [&](int di, int dj) {
        return (Su[lj + dj][li + di] - 0.25 * (Sva[lj + dj - 1][li + di] + Sva[lj + dj][li + di]) *
                                           (M2(cos_sg4, i + di, j + dj - 1) + M2(cos_sg2, i + di, j + dj))) *
               M2(dyc, i + di, j + dj) * 0.5 * (M2(sin_sg4, i + di, j + dj - 1) + M2(sin_sg2, i + di, j + dj));
      };
      auto vf_gen = [&](int di, int dj) {
        return (Sv[lj + dj][li + di] - 0.25 * (Sua[lj + dj][li + di - 1] + Sua[lj + dj][li + di]) *
                                           (M2(cos_sg3, i + di - 1, j + dj) + M2(cos_sg1, i + di, j + dj))) *
               M2(dxc, i + di, j + dj) * 0.5 * (M2(sin_sg3, i + di - 1, j + dj) + M2(sin_sg1, i + di, j + dj));
      };
      auto uf_edge = [&](int di, int dj) {
        return Su[lj + dj][li + di] * M2(dyc, i + di, j + dj) * 0.5 *
               (M2(sin_sg4, i + di, j + dj - 1) + M2(sin_sg2, i + di, j + dj));
      };
      auto vf_edge = [&](int di, int dj) {
        return Sv[lj + dj][li + di] * M2(dxc, i + di, j + dj) * 0.5 *
               (M2(sin_sg3, i + di - 1, j + dj) + M2(sin_sg1, i + di, j + dj));
      };
      const bool iw = e.W && i == is, iE = e.E && i == ie + 1, jS = e.S && j == js, jN = e.N && j == je + 1;
      double d;
      // later region blocks overwrite earlier ones: corners > j-edges > i-edges > general
      if ((iw || iE) && jS) {
        d = (-vf_edge(0, 0) + uf_edge(-1, 0) - uf_edge(0, 0)) * M2(rarea_c, i, j);
      } else if ((iw || iE) && jN) {
        d = (vf_edge(0, -1) + uf_edge(-1, 0) - uf_edge(0, 0)) * M2(rarea_c, i, j);
      } else if (jS || jN) {
        d = (vf_gen(0, -1) - vf_gen(0, 0) + uf_edge(-1, 0) - uf_edge(0, 0)) * M2(rarea_c, i, j);
      } else if (iw || iE) {
        d = (vf_edge(0, -1) - vf_edge(0, 0) + uf_gen(-1, 0) - uf_gen(0, 0)) * M2(rarea_c, i, j);
      } else {
        d = (vf_gen(0, -1) - vf_gen(0, 0) + uf_gen(-1, 0) - uf_gen(0, 0)) * M2(rarea_c, i, j);
      }
      a.divgd[T.o3(i, j)] = d;
    }
  });

  // ---- E: transportdelp_update_vorticity_and_kineticenergy (c_sw.py:287-391) on compute + 1 halo.  delp / pt / w are
  // read "as filled" by fill_corners_2cells_x for the x fluxes and by fill_corners_2cells_y for the y fluxes and the
  // cell's own value (the reference fills in place, x first: c_sw.py:700-760); ke is kept for the patch and one row /
  // column below it, delpc / ptc / omga are stored for own cells; zero_delpc_ptc covers the rest of compute + 3 halo.
  wait_dpw();
  const bool ct = EDGES && T.corner_tile;
  auto rdx = [&](double(*Sq)[W], int i, int j) {  // q(i, j) after fill_corners_2cells_x
    if (ct) xfill_src(e, 2, i, j);
    return Sq[j - bj][i - bi];
  };
  auto rdy = [&](double(*Sq)[W], int i, int j) {  // q(i, j) after fill_corners_2cells_y (applied after the x fill)
    if (ct) {
      const int i0 = i, j0 = j;
      yfill_src(e, 2, i, j);
      if (i == i0 && j == j0) xfill_src(e, 2, i, j);  // cells only the x fill wrote keep that value
    }
    return Sq[j - bj][i - bi];
  };
  // One pass with F (circulation_cgrid + absolute_vorticity, c_sw.py:394-436, on corners): both read phase B's uc, vc.
  T.each(0, g.nx + 5, 0, g.ny + 5, 2, W - 3, 2, H - 3, [&](int i, int j, int li, int lj) {
    const bool in_e = i >= is - 1 && i <= ie + 1 && j >= js - 1 && j <= je + 1;
    if (in_e && li <= W - 4 && lj <= H - 4) {
      if (owned(i, j)) {
        double fx1_[2], fx_[2], fx2_[2], fy1_[2], fy_[2], fy2_[2];
        for (int d = 0; d < 2; ++d) {
          // compute_nonhydrostatic_fluxes_x (c_sw.py:252-284) at i + d
          const double utc = Sut[lj][li + d];
          double f1 = utc > 0.0 ? rdx(Sdp, i + d - 1, j) : rdx(Sdp, i + d, j);
          const double f = utc > 0.0 ? rdx(Spt, i + d - 1, j) : rdx(Spt, i + d, j);
          const double f2 = utc > 0.0 ? rdx(Sw, i + d - 1, j) : rdx(Sw, i + d, j);
          f1 = utc * f1;
          fx1_[d] = f1;
          fx_[d] = f1 * f;
          fx2_[d] = f1 * f2;
        }
        for (int d = 0; d < 2; ++d) {
          const double vtc = Svt[lj + d][li];
          double f1 = vtc > 0.0 ? rdy(Sdp, i, j + d - 1) : rdy(Sdp, i, j + d);
          const double f = vtc > 0.0 ? rdy(Spt, i, j + d - 1) : rdy(Spt, i, j + d);
          const double f2 = vtc > 0.0 ? rdy(Sw, i, j + d - 1) : rdy(Sw, i, j + d);
          f1 = vtc * f1;
          fy1_[d] = f1;
          fy_[d] = f1 * f;
          fy2_[d] = f1 * f2;
        }
        const double delp0 = rdy(Sdp, i, j), pt0 = rdy(Spt, i, j), w0 = rdy(Sw, i, j), ra = M2(rarea, i, j);
        const double dpc = delp0 + (fx1_[0] - fx1_[1] + fy1_[0] - fy1_[1]) * ra;
        const long long o = T.o3(i, j);
        a.delpc[o] = dpc;
        a.ptc[o] = (pt0 * delp0 + (fx_[0] - fx_[1] + fy_[0] - fy_[1]) * ra) / dpc;
        a.omga[o] = (w0 * delp0 + (fx2_[0] - fx2_[1] + fy2_[0] - fy2_[1]) * ra) / dpc;
      }
      // vorticity / kinetic energy with upstream bias
      const double uav = Sua[lj][li], vav = Sva[lj][li];
      double kev = uav > 0.0 ? Suc[lj][li] : Suc[lj][li + 1];
      double vo = vav > 0.0 ? Svc[lj][li] : Svc[lj + 1][li];
      if ((e.S && j == js - 1) || (e.N && j == je))
        vo = (vav <= 0.0) ? vo * M2(sin_sg4, i, j) + Su[lj + 1][li] * M2(cos_sg4, i, j) : vo;
      if ((e.S && j == js) || (e.N && j == je + 1))
        vo = (vav > 0.0) ? vo * M2(sin_sg2, i, j) + Su[lj][li] * M2(cos_sg2, i, j) : vo;
      if ((e.E && i == ie) || (e.W && i == is - 1))
        kev = (uav <= 0.0) ? kev * M2(sin_sg3, i, j) + Sv[lj][li + 1] * M2(cos_sg3, i, j) : kev;
      if ((e.E && i == ie + 1) || (e.W && i == is))
        kev = (uav > 0.0) ? kev * M2(sin_sg1, i, j) + Sv[lj][li] * M2(cos_sg1, i, j) : kev;
      Ske[lj][li] = 0.5 * dt2 * (uav * kev + vav * vo);
    } else if (!in_e && owned(i, j)) {  // zero_delpc_ptc (c_sw.py:21-30) where the stencil above does not write
      a.delpc[T.o3(i, j)] = 0.0;
      a.ptc[T.o3(i, j)] = 0.0;
    }
    if (li >= 3 && lj >= 3 && i >= is && i <= ie + 1 && j >= js && j <= je + 1) {
      const double fxv = M2(dxc, i, j) * Suc[lj][li];
      const double fyv = M2(dyc, i, j) * Svc[lj][li];
      const double fx1v = M2(dxc, i, j - 1) * Suc[lj - 1][li];
      const double fy1v = M2(dyc, i - 1, j) * Svc[lj][li - 1];
      double vc_ = fx1v - fxv - fy1v + fyv;
      if (e.W && i == is && ((e.S && j == js) || (e.N && j == je + 1))) vc_ = fx1v - fxv + fyv;
      if (e.E && i == ie + 1 && ((e.S && j == js) || (e.N && j == je + 1))) vc_ = fx1v - fxv - fy1v;
      Svort[lj][li] = M2(fC, i, j) + M2(rarea_c, i, j) * vc_;
    }
  });

  // ---- G: update_y_velocity on [X, Y_INTERFACE], update_x_velocity on [X_INTERFACE, Y]; own cells ----------------
  T.each_owned(is - 1, ie + 2, js - 1, je + 2, [&](int i, int j, int li, int lj) {
    if (i <= ie + 1) {  // vc exists on [is-1, ie+1] x [js-1, je+2]
      double vcv = Svc[lj][li];
      if (i >= is && i <= ie && j >= js && j <= je + 1) {
        double tmp = dt2 * (Su[lj][li] - vcv * M2(cosa_v, i, j)) / M2(sina_v, i, j);
        if ((e.S && j == js) || (e.N && j == je + 1)) tmp = dt2 * Su[lj][li];
        const double flux = tmp > 0.0 ? Svort[lj][li] : Svort[lj][li + 1];
        vcv = vcv - tmp * flux + M2(rdyc, i, j) * (Ske[lj - 1][li] - Ske[lj][li]);
      }
      a.vc[T.o3(i, j)] = vcv;
    }
    if (j <= je + 1) {  // uc exists on [is-1, ie+2] x [js-1, je+1]
      double ucv = Suc[lj][li];
      if (i >= is && i <= ie + 1 && j >= js && j <= je) {
        double tmp = dt2 * (Sv[lj][li] - ucv * M2(cosa_u, i, j)) / M2(sina_u, i, j);
        if ((e.W && i == is) || (e.E && i == ie + 1)) tmp = dt2 * Sv[lj][li];
        const double flux = tmp > 0.0 ? Svort[lj][li] : Svort[lj + 1][li];
        ucv = ucv + tmp * flux + M2(rdxc, i, j) * (Ske[lj][li - 1] - Ske[lj][li]);
      }
      a.uc[T.o3(i, j)] = ucv;
    }
  }, false);
#undef M2
}

#ifndef PACE_HOST_EMU
__global__ void __launch_bounds__(CswCfg::NT, CSW_CTAS_PER_SM) k_c_sw_fused(Geo g, CswArgs a, const __grid_constant__ CUtensorMap m_u,
                                                              const __grid_constant__ CUtensorMap m_v,
                                                              const __grid_constant__ CUtensorMap m_dp,
                                                              const __grid_constant__ CUtensorMap m_pt,
                                                              const __grid_constant__ CUtensorMap m_w) {
  typedef CswCfg C;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long bar, bar2;   // u, v | delp, pt, w
  typename Tile<C>::Arr* S = reinterpret_cast<typename Tile<C>::Arr*>(smem_raw);
  const int k = blockIdx.z % g.nz, t = blockIdx.z / g.nz;
  Tile<C> T = make_tile<C>(g, a.tg, blockIdx.x, blockIdx.y, k, t);
  T.tx = threadIdx.x; T.ty = threadIdx.y;
  if (a.part != 0) {
    const bool inside = T.bi >= g.is && T.bi + C::W - 1 <= g.ie && T.bj >= g.js && T.bj + C::H - 1 <= g.je;
    if (inside != (a.part == 1)) return;
  }
  const bool leader = threadIdx.x == 0 && threadIdx.y == 0;
  if (leader) {
    tma::mbar_init(&bar, 1);
    tma::mbar_init(&bar2, 1);
  }
  __syncthreads();
  if (leader) {
    tma::mbar_expect_tx(&bar, 2u * C::BYTES);
    tma::load_tile(&S[0], &m_u, &bar, T.bi, T.bj, k, t);
    tma::load_tile(&S[1], &m_v, &bar, T.bi, T.bj, k, t);
    tma::mbar_expect_tx(&bar2, 3u * C::BYTES);
    tma::load_tile(&S[2], &m_dp, &bar2, T.bi, T.bj, k, t);
    tma::load_tile(&S[3], &m_pt, &bar2, T.bi, T.bj, k, t);
    tma::load_tile(&S[4], &m_w, &bar2, T.bi, T.bj, k, t);
  }
  tma::mbar_wait(&bar, 0);
  auto wait_dpw = [&]() { tma::mbar_wait(&bar2, 0); };
  if (T.near_edge) c_sw_tile<C, true>(T, a, S, wait_dpw);
  else c_sw_tile<C, false>(T, a, S, wait_dpw);
}
#endif

}  // namespace

int c_sw_unfused(Core& c, double dt2);

static bool csw_unfused_env() {
  static const bool unfused = [] { const char* v = getenv("PACE_B200_CSW"); return v && v[0] == 'u'; }();
  return unfused;
}
// the interior / rim split needs the tiled kernel (device build, patches wide enough)
bool c_sw_can_split(const Core& c) {
#ifndef PACE_HOST_EMU
  const TileGrid tg = tile_grid<CswCfg>(c.g);
  return !csw_unfused_env() && c.cfg.grid_type == 0 && !((tg.ntx > 1 && tg.tw < 10) || (tg.nty > 1 && tg.th < 10));
#else
  (void)c;
  return false;
#endif
}

int c_sw(Core& c, double dt2, int part) {
  typedef CswCfg C;
  const Geo g = c.g;
  if (csw_unfused_env()) return c_sw_unfused(c, dt2);
  if (c.cfg.grid_type != 0) {
    c.err = "unimplemented grid_type (only 0)";  // d2a2c_vect.py:394 allows 0 and 4
    return 1;
  }
  CswArgs a;
  a.u.p = c.F("u"); a.v.p = c.F("v"); a.delp.p = c.F("delp"); a.pt.p = c.F("pt"); a.w.p = c.F("w");
  a.uc = c.F("uc"); a.vc = c.F("vc"); a.ua = c.F("ua"); a.va = c.F("va"); a.ut = c.F("ad.ut"); a.vt = c.F("ad.vt");
  a.divgd = c.F("ad.divgd"); a.omga = c.F("omga"); a.delpc = c.F("csw.delpc"); a.ptc = c.F("csw.ptc");
  a.cosa_s = c.F("grid.cosa_s"); a.rsin2 = c.F("grid.rsin2"); a.cosa_u = c.F("grid.cosa_u"); a.cosa_v = c.F("grid.cosa_v");
  a.rsin_u = c.F("grid.rsin_u"); a.rsin_v = c.F("grid.rsin_v"); a.sina_u = c.F("grid.sina_u"); a.sina_v = c.F("grid.sina_v");
  a.dxa = c.F("grid.dxa"); a.dya = c.F("grid.dya"); a.dx = c.F("grid.dx"); a.dy = c.F("grid.dy"); a.dxc = c.F("grid.dxc");
  a.dyc = c.F("grid.dyc"); a.rdxc = c.F("grid.rdxc"); a.rdyc = c.F("grid.rdyc"); a.sin_sg1 = c.F("grid.sin_sg1");
  a.sin_sg2 = c.F("grid.sin_sg2"); a.sin_sg3 = c.F("grid.sin_sg3"); a.sin_sg4 = c.F("grid.sin_sg4");
  a.cos_sg1 = c.F("grid.cos_sg1"); a.cos_sg2 = c.F("grid.cos_sg2"); a.cos_sg3 = c.F("grid.cos_sg3");
  a.cos_sg4 = c.F("grid.cos_sg4"); a.rarea = c.F("grid.rarea"); a.rarea_c = c.F("grid.rarea_c"); a.fC = c.F("grid.fC");
  if (!c.err.empty()) return 1;
  a.dt2 = dt2;
  a.part = part;
  a.nord = c.cfg.nord;
  a.npt = (4 > g.nx - 1 || 4 > g.ny - 1) ? 0 : 4;
  a.tg = tile_grid<C>(g);
  // a patch narrower than the corner region would see corner cells without their fill sources
  if ((a.tg.ntx > 1 && a.tg.tw < 10) || (a.tg.nty > 1 && a.tg.th < 10)) return c_sw_unfused(c, dt2);
#ifndef PACE_HOST_EMU
  const CUtensorMap *mu = tile_tensor_map(c, a.u.p, C::W, C::H), *mv = tile_tensor_map(c, a.v.p, C::W, C::H),
                    *mdp = tile_tensor_map(c, a.delp.p, C::W, C::H), *mpt = tile_tensor_map(c, a.pt.p, C::W, C::H),
                    *mw = tile_tensor_map(c, a.w.p, C::W, C::H);
  if (!mu || !mv || !mdp || !mpt || !mw) return 1;
  constexpr int SMEM = CSW_ARRAYS * C::BYTES;
  if (func_attr_once(c, (const void*)k_c_sw_fused, SMEM)) return 1;
  {
    LaunchScope scope(c.stream, "c_sw_fused");
    dim3 blk(C::NTX, C::NTY, 1), grd(a.tg.ntx, a.tg.nty, g.nz * g.nt);
    k_c_sw_fused<<<grd, blk, SMEM, c.stream>>>(g, a, *mu, *mv, *mdp, *mpt, *mw);
  }
#else
  ++launch_counter();
  static thread_local Tile<C>::Arr S[CSW_ARRAYS];
  for (int t = 0; t < g.nt; ++t)
    for (int k = 0; k < g.nz; ++k)
      for (int by = 0; by < a.tg.nty; ++by)
        for (int bx = 0; bx < a.tg.ntx; ++bx) {
          Tile<C> T = make_tile<C>(g, a.tg, bx, by, k, t);
          host_load_tile<C>(T, S[0], a.u.p);
          host_load_tile<C>(T, S[1], a.v.p);
          host_load_tile<C>(T, S[2], a.delp.p);
          host_load_tile<C>(T, S[3], a.pt.p);
          host_load_tile<C>(T, S[4], a.w.p);
          if (T.near_edge) c_sw_tile<C, true>(T, a, S, [] {});
          else c_sw_tile<C, false>(T, a, S, [] {});
        }
#endif
  // the reference leaves delp, pt, w with their corner cells filled (x fill, then y fill)
  if (part == 1) return 0;
  for (int ydir = 0; ydir < 2; ++ydir) {
    CornerOps ops;
    ops.n = 3; ops.nk = g.nz;
    ops.op[0] = CornerOp{a.delp.p, a.delp.p, 2, ydir, 1.0, 1.0, 1.0, 1.0};
    ops.op[1] = CornerOp{a.pt.p, a.pt.p, 2, ydir, 1.0, 1.0, 1.0, 1.0};
    ops.op[2] = CornerOp{a.w.p, a.w.p, 2, ydir, 1.0, 1.0, 1.0, 1.0};
    run_corner_ops(c.stream, g, ops);
  }
  return 0;
}

}  // namespace pace
