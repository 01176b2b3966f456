// riem_solver_c / riem_solver3 with the tridiagonal work arrays on chip.
//
// The straightforward column kernels (riem_solver.cu) keep ~20 per-level work arrays in global scratch: measured
// 4.9 GB of DRAM traffic per riem_solver_c launch at C128 against 0.5 GB of compulsory traffic, 35 long-scoreboard
// stall cycles per issued instruction (profiles/r01_columns_c128.txt).  Here a warp owns 32 adjacent columns (lane
// <-> column, coalesced rows); of the two arrays that must survive between the k-sweeps, Y[k] lives on chip - in
// tensor memory for its first 64 levels and in shared memory for the rest (tmem.cuh; 16 warps per SM), or wholly in
// shared memory (10 warps per SM) - and X in one global work array that every sweep prefetches a chunk ahead (these
// sweeps are chains of dependent fp64 divisions / exp / log and need the occupancy):
//   A  forward : pressures (pem, peg cumulative sums), pm, first guess pe, tridiagonal elimination   X=gam  Y=pp
//   B  backward: back substitution                                                                    Y=pp
//   C  forward : aa (pem summed again), w elimination                                                 X=gam' Y=w behind the sweep
//   D  backward: w back substitution                                                                  Y=w
//   E  forward : pe cumulative sum, output pressures (pem summed again)                               Y=pe behind the sweep
//   F  backward: p1 recurrence, new dz, heights
// Everything else is recomputed from the inputs (identical expressions give identical bits) or carried in
// registers; pm and X are the only global work arrays.
// Statement order and expression trees per level are those of riem_solver.cu / the reference (sim1_solver.py:20-141,
// riem_solver_c.py:21-123, riem_solver3.py:26-146), so results are bit-identical.
#include <cstdlib>
#include <string>

#include "tile.cuh"
#include "tmem.cuh"

namespace pace {

#ifndef PACE_HOST_EMU

namespace {

struct RiemArgs {
  const double *delp, *cappa, *ptv, *q_con, *w_in, *zsurf, *ws;
  double* zh;                       // heights: ad.gz (C) / ad.zh (D), read then rewritten
  double *pm, *xwork;  // global work arrays
  double *pef;                      // ad.pkc
  double *w_out, *delz, *pk3, *peln, *pk, *pe;  // D-grid solver only
  double ptop, p_fac, dt, t1g, rdt, peln1, ptk;
  int km, last_call, beta_neg;
  Box b;
};

// Y lives in shared memory (k_riem, 20.7 KB per warp at L79: 10 warps per SM) or, for its first 64 levels, in tensor
// memory (k_riem_tmem: 16 warps per SM, 1.6 x the residency for these latency-bound sweeps): tmem.cuh.

// The six sweeps of one column per lane.  `active` = this lane's column exists; a lane without one (tail of the last
// warp) works on a copy of the last column and stores nothing, because the tensor-memory accesses are warp-collective.
template <bool D3, bool TM, int UU>
__device__ __forceinline__ void riem_column(const Geo& g, const RiemArgs& a, const YAcc<TM>& Y, int col, int t, bool active) {
  const int km = a.km;
#define yld(k) Y.ld(k)
#define yst(k, v) Y.st((k), (v))
  const int i = a.b.i0 + col % a.b.ni, j = a.b.j0 + col / a.b.ni;
  const long long o0 = (long long)t * g.st3 + j * g.sj + i;  // level 0
  const int o2 = t * g.sk + j * g.sj + i;
  const long long sk = g.sk;
#define G(f, k) (f)[o0 + (long long)(k) * sk]
#define GST(f, k, v) do { const double gst_v_ = (v); if (active) G(f, k) = gst_v_; } while (0)
  // all arrays are distinct (w_in == w_out in the D-grid solver: neither is restrict-qualified), so loads of the
  // following levels may be hoisted above the stores of the current one
  const double* __restrict__ delp = a.delp;
  const double* __restrict__ cappa = a.cappa;
  const double* __restrict__ ptv = a.ptv;
  const double* __restrict__ q_con = a.q_con;
  const double* w_in = a.w_in;
  double* w_out = a.w_out;
  double* __restrict__ zh = a.zh;
  double* __restrict__ pm = a.pm;
  double* __restrict__ xg = a.xwork;
  double* __restrict__ pef = a.pef;
  double* __restrict__ delz = a.delz;
  double* __restrict__ pk3 = a.pk3;
  double* __restrict__ peln = a.peln;
  double* __restrict__ pk = a.pk;
  double* __restrict__ pe = a.pe;
  constexpr double RGRAV = 1.0 / GRAV;
  auto DM = [&](int k) { return D3 ? G(delp, k) * RGRAV : G(delp, k) / GRAV; };
  auto GM = [&](int k) { return 1.0 / (1.0 - G(cappa, k)); };
  const double ptop = a.ptop, dt = a.dt, t1g = a.t1g, rdt = a.rdt, p_fac = a.p_fac;
  const double ws = a.ws[o2];

  // Every sweep walks the column in chunks of U levels: the global inputs of the next chunk are requested before the
  // current chunk is computed, so that ~5 x U loads per warp are in flight behind the (latency-bound) recurrences.
  // Measured at C384 (ms per step, U = 2 / 3 / 4): riem_solver_c 32.5 / 35.5 / 37.5, riem_solver3 37.2 / 36.8 / 31.8.
  constexpr int U = UU;
  // ---- A: forward ---------------------------------------------------------------------------------------------
  {
    double pem_k = ptop, peg_k = ptop, logpg_k = a.peln1;
    if (D3) {
      GST(pk3, 0, a.ptk);
      if (a.last_call) { GST(peln, 0, a.peln1); GST(pk, 0, a.ptk); GST(pe, 0, ptop); }
    }
    double zh_k = G(zh, 0);
    double dm_j = 0.0, pe0_j = 0.0;            // level k-1
    double bet_prev = 0.0, grat_prev = 0.0, pp_prev = 0.0;
    double n_delp[U], n_q[U], n_cap[U], n_zh[U], n_ptv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int kk = min(u, km - 1);
      n_delp[u] = G(delp, kk); n_q[u] = G(q_con, kk); n_cap[u] = G(cappa, kk); n_zh[u] = G(zh, kk + 1);
      n_ptv[u] = G(ptv, kk);
    }
    for (int k0 = 0; k0 < km; k0 += U) {
      double c_delp[U], c_q[U], c_cap[U], c_zh[U], c_ptv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        c_delp[u] = n_delp[u]; c_q[u] = n_q[u]; c_cap[u] = n_cap[u]; c_zh[u] = n_zh[u]; c_ptv[u] = n_ptv[u];
      }
      if (k0 + U < km) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int kk = min(k0 + U + u, km - 1);
          n_delp[u] = G(delp, kk); n_q[u] = G(q_con, kk); n_cap[u] = G(cappa, kk); n_zh[u] = G(zh, kk + 1);
          n_ptv[u] = G(ptv, kk);
        }
      }
      // The U levels of the chunk: first the cheap sequential sums, then the exp / log of all U levels as independent
      // chains (libm::log_n / exp_n), then the sequential elimination steps.  Levels >= km of the last chunk compute
      // on clamped inputs and are never stored.
      double pemn[U], pegn[U], pegk[U], dmv[U], gmv[U], dzv[U], pmv[U], pe0v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        pegk[u] = peg_k;
        pemn[u] = pem_k + c_delp[u];
        pegn[u] = peg_k + c_delp[u] * (1.0 - c_q[u]);
        pem_k = pemn[u];
        peg_k = pegn[u];
        dmv[u] = D3 ? c_delp[u] * RGRAV : c_delp[u] / GRAV;
        gmv[u] = 1.0 / (1.0 - c_cap[u]);
        dzv[u] = c_zh[u] - zh_k;
        zh_k = c_zh[u];
      }
      if (D3) {
        double lp[U], lg[U], ka[U], pk3n[U];
        libm::log_n(pemn, lp);
        libm::log_n(pegn, lg);
#pragma unroll
        for (int u = 0; u < U; ++u) ka[u] = KAPPA * lp[u];
        libm::exp_n(ka, pk3n);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          pmv[u] = (pegn[u] - pegk[u]) / (lg[u] - logpg_k);
          logpg_k = lg[u];
          const int k = k0 + u;
          if (k < km) {
            GST(pk3, k + 1, pk3n[u]);
            if (a.last_call) { GST(peln, k + 1, lp[u]); GST(pk, k + 1, pk3n[u]); GST(pe, k + 1, pemn[u]); }
          }
        }
      } else {
        double ratio[U], lr[U];
#pragma unroll
        for (int u = 0; u < U; ++u) ratio[u] = pegn[u] / pegk[u];
        libm::log_n(ratio, lr);
#pragma unroll
        for (int u = 0; u < U; ++u) pmv[u] = (pegn[u] - pegk[u]) / lr[u];
      }
      {
        double targ[U], tl[U], earg[U], ev[U];
#pragma unroll
        for (int u = 0; u < U; ++u) targ[u] = ((-dmv[u]) / dzv[u] * RDGAS) * c_ptv[u];
        libm::log_n(targ, tl);
#pragma unroll
        for (int u = 0; u < U; ++u) earg[u] = gmv[u] * tl[u];
        libm::exp_n(earg, ev);
#pragma unroll
        for (int u = 0; u < U; ++u) pe0v[u] = ev[u] - pmv[u];
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = k0 + u;
        if (k < km) {
          GST(pm, k, pmv[u]);
          const double dm_k = dmv[u];
          const double pe0_k = pe0v[u];
          if (k >= 1) {  // elimination step of level j = k-1
            const int jl = k - 1;
            const double gr = dm_j / dm_k;
            const double bb = 2.0 * (1.0 + gr);
            const double dd = 3.0 * (pe0_j + gr * pe0_k);
            if (jl == 0) {
              yst(0, 0.0);
              pp_prev = dd / bb;
              yst(1, pp_prev);
              bet_prev = bb;
            } else {
              const double gam = grat_prev / bet_prev;
              const double bet = bb - gam;
              GST(xg, jl, gam);
              pp_prev = (dd - pp_prev) / bet;
              yst(jl + 1, pp_prev);
              bet_prev = bet;
            }
            grat_prev = gr;
          }
          dm_j = dm_k;
          pe0_j = pe0_k;
        }
      }
    }
    {  // level j = km-1: bb = 2, dd = 3 pe
      const int jl = km - 1;
      const double bb = 2.0;
      const double dd = 3.0 * pe0_j;
      const double gam = grat_prev / bet_prev;
      const double bet = bb - gam;
      GST(xg, jl, gam);
      yst(jl + 1, (dd - pp_prev) / bet);
    }
    Y.fence_st();
  }
  // ---- B: backward, pp (k = km-1 .. 1).  aa, which the reference forms here, only needs the final pp of its own
  // level: it is formed in sweep C, where the pressure sum runs forward again (no aa / pem arrays in global memory)
  {
    double pp_next = yld(km);
    double n_x[U];
#pragma unroll
    for (int u = 0; u < U; ++u) n_x[u] = G(xg, max(km - 1 - u, 1));
    for (int k0 = km - 1; k0 >= 1; k0 -= U) {
      double c_x[U];
#pragma unroll
      for (int u = 0; u < U; ++u) c_x[u] = n_x[u];
      if (k0 - U >= 1) {
#pragma unroll
        for (int u = 0; u < U; ++u) n_x[u] = G(xg, max(k0 - U - u, 1));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = k0 - u;
        if (k >= 1) {
          const double ppk = yld(k) - c_x[u] * pp_next;
          yst(k, ppk);
          pp_next = ppk;
        }
      }
    }
    Y.fence_st();
  }
  // ---- C: forward, aa and the w elimination ---------------------------------------------------------------------
  // aa(k) = t1g * 0.5 * (gm(k-1) + gm(k)) / (dz(k-1) + dz(k)) * (pem(k) + pp(k)), 1 <= k <= km-1 (sim1_solver.py), with
  // pem(k) = pem(k-1) + delp(k-1) summed in the order of sweep A, gm = 1 / (1 - cappa), dz(k) = zh(k+1) - zh(k)
  {
    const double delp0 = G(delp, 0);
    const double dm0 = D3 ? delp0 * RGRAV : delp0 / GRAV;
    double pem_k = ptop + delp0;                       // pem(1)
    double zh_lo = G(zh, 1);
    double dz_m = zh_lo - G(zh, 0);                    // dz(0)
    double gm_m = GM(0);
    double pp_k = yld(1);
    double aa_k;                                       // aa(k) of the level in work
    {
      const double zh_hi = G(zh, 2);
      const double dz_k = zh_hi - zh_lo;               // dz(1)
      const double gm_k = GM(1);
      aa_k = t1g * 0.5 * (gm_m + gm_k) / (dz_m + dz_k) * (pem_k + pp_k);
      zh_lo = zh_hi; dz_m = dz_k; gm_m = gm_k;         // now dz(1), gm(1), zh(2)
    }
    double betk = dm0 - aa_k;
    double wprev = (dm0 * G(w_in, 0) + dt * pp_k) / betk;
    yst(0, wprev);   // pp(0) is not read: from here on Y(k) holds the eliminated w of the levels this sweep has passed
    double n_delp[U], n_w[U], n_cap[U], n_zh[U];       // delp(k), w(k), cappa(k+1), zh(k+2)
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int kk = min(1 + u, km - 1);
      n_delp[u] = G(delp, kk); n_w[u] = G(w_in, kk); n_cap[u] = G(cappa, min(kk + 1, km - 1)); n_zh[u] = G(zh, min(kk + 2, km));
    }
    for (int k0 = 1; k0 < km - 1; k0 += U) {
      double c_delp[U], c_w[U], c_cap[U], c_zh[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { c_delp[u] = n_delp[u]; c_w[u] = n_w[u]; c_cap[u] = n_cap[u]; c_zh[u] = n_zh[u]; }
      if (k0 + U < km - 1) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int kk = min(k0 + U + u, km - 1);
          n_delp[u] = G(delp, kk); n_w[u] = G(w_in, kk); n_cap[u] = G(cappa, min(kk + 1, km - 1)); n_zh[u] = G(zh, min(kk + 2, km));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = k0 + u;
        if (k < km - 1) {
          // aa(k + 1)
          const double pp_n = yld(k + 1);
          pem_k = pem_k + c_delp[u];                   // pem(k + 1)
          const double dz_k = c_zh[u] - zh_lo;         // dz(k + 1)
          const double gm_k = 1.0 / (1.0 - c_cap[u]);  // gm(k + 1)
          const double aa_n = t1g * 0.5 * (gm_m + gm_k) / (dz_m + dz_k) * (pem_k + pp_n);
          zh_lo = c_zh[u]; dz_m = dz_k; gm_m = gm_k;
          // elimination step of level k
          const double aak = aa_k;
          const double gm = aak / betk;
          GST(xg, k, gm);
          const double dm_k = D3 ? c_delp[u] * RGRAV : c_delp[u] / GRAV;
          betk = dm_k - (aak + aa_n + aak * gm);
          wprev = (dm_k * c_w[u] + dt * (pp_n - pp_k) - aak * wprev) / betk;
          yst(k, wprev);
          pp_k = pp_n;
          aa_k = aa_n;
        }
      }
    }
    {
      // level km-1: dz_m = dz(km-1), gm_m = gm(km-1), pem_k = pem(km-1) at this point
      const int k = km - 1;
      const double aak = aa_k;
      const double delp_k = G(delp, k);
      const double pp_n = yld(k + 1);
      const double p1 = t1g * gm_m / dz_m * ((pem_k + delp_k) + pp_n);
      const double gm = aak / betk;
      GST(xg, k, gm);
      const double dm_k = D3 ? delp_k * RGRAV : delp_k / GRAV;
      betk = dm_k - (aak + p1 + aak * gm);
      wprev = (dm_k * G(w_in, k) + dt * (pp_n - pp_k) - p1 * ws - aak * wprev) / betk;
      yst(k, wprev);
    }
    Y.fence_st();
  }
  // ---- D: backward, w back substitution, in place in Y (pp is dead once sweep C has passed a level) ---------------
  {
    double wn = yld(km - 1);
    double n_x[U];
#pragma unroll
    for (int u = 0; u < U; ++u) n_x[u] = G(xg, max(km - 2 - u, 0) + 1);
    for (int k0 = km - 2; k0 >= 0; k0 -= U) {
      double c_x[U];
#pragma unroll
      for (int u = 0; u < U; ++u) c_x[u] = n_x[u];
      if (k0 - U >= 0) {
#pragma unroll
        for (int u = 0; u < U; ++u) n_x[u] = G(xg, max(k0 - U - u, 0) + 1);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = k0 - u;
        if (k >= 0) {
          wn = yld(k) - c_x[u] * wn;
          yst(k, wn);
        }
      }
    }
    Y.fence_st();
  }
  // ---- E: forward, pe and the output pressure (pe, or pe + pem with the pressure sum of sweep A run again); the
  // D-grid solver publishes w here (its old value is needed up to this point)
  {
    double pe_k = 0.0;
    double pem_k = ptop;
    double w_cur = yld(0);   // Y(k) holds the new w of level k until this sweep replaces it by pe(k): read one level ahead
    yst(0, 0.0);
    if (D3) GST(pef, 0, a.beta_neg ? 0.0 + ptop : 0.0);   // pe(0) = 0, pem(0) = ptop
    else GST(pef, 0, ptop);
    double n_delp[U], n_w1[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int kk = min(u, km - 1);
      n_delp[u] = G(delp, kk); n_w1[u] = G(w_in, kk);
    }
    for (int k0 = 1; k0 <= km; k0 += U) {
      double c_delp[U], c_w1[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { c_delp[u] = n_delp[u]; c_w1[u] = n_w1[u]; }
      if (k0 + U <= km) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int kk = min(k0 + U + u - 1, km - 1);
          n_delp[u] = G(delp, kk); n_w1[u] = G(w_in, kk);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = k0 + u;
        if (k <= km) {
          const double wk = w_cur;                     // w(k-1)
          if (k < km) w_cur = yld(k);                  // w(k), before pe(k) takes its place
          const double dm_m = D3 ? c_delp[u] * RGRAV : c_delp[u] / GRAV;
          pe_k = pe_k + dm_m * (wk - c_w1[u]) * rdt;
          yst(k, pe_k);
          pem_k = pem_k + c_delp[u];                   // pem(k)
          if (D3) {
            GST(w_out, k - 1, wk);
            GST(pef, k, a.beta_neg ? pe_k + pem_k : pe_k);
          } else {
            GST(pef, k, pe_k + pem_k);
          }
        }
      }
    }
    Y.fence_st();
  }
  // ---- F: backward, p1, new dz, heights, output pressures ----------------------------------------------------
  {
    double zcur = a.zsurf[o2];
    GST(zh, km, zcur);
    double p1_next = 0.0;
    double dm_n = 0.0;  // dm(k+1)
    double y_p1 = yld(km), y_p2 = 0.0;  // Y[k+1], Y[k+2] of the level in work: every level of Y is read once
    double n_delp[U], n_pm[U], n_ptv[U], n_cap[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int kk = max(km - 1 - u, 0);
      n_delp[u] = G(delp, kk); n_pm[u] = G(pm, kk); n_ptv[u] = G(ptv, kk); n_cap[u] = G(cappa, kk);
    }
    for (int k0 = km - 1; k0 >= 0; k0 -= U) {
      double c_delp[U], c_pm[U], c_ptv[U], c_cap[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        c_delp[u] = n_delp[u]; c_pm[u] = n_pm[u]; c_ptv[u] = n_ptv[u]; c_cap[u] = n_cap[u];
      }
      if (k0 - U >= 0) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int kk = max(k0 - U - u, 0);
          n_delp[u] = G(delp, kk); n_pm[u] = G(pm, kk); n_ptv[u] = G(ptv, kk); n_cap[u] = G(cappa, kk);
        }
      }
      // sequential p1 recurrence of the U levels, then their exp / log as independent chains, then the stores
      double dmv[U], maxp[U], lm[U], earg[U], ev[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = max(k0 - u, 0);
        const double dm_k = D3 ? c_delp[u] * RGRAV : c_delp[u] / GRAV;
        const double y_k = yld(k);
        double p1;
        if (k == km - 1) {
          p1 = (y_k + 2.0 * y_p1) * 1.0 / 3.0;
        } else {
          const double gr = dm_k / dm_n;
          const double bb = 2.0 * (1.0 + gr);
          p1 = (y_k + bb * y_p1 + gr * y_p2) * 1.0 / 3.0 - gr * p1_next;
        }
        y_p2 = y_p1;
        y_p1 = y_k;
        const double pm_k = c_pm[u];
        const double aa = p1 + pm_k;
        maxp[u] = (p_fac * dm_k > aa) ? p_fac * pm_k : aa;
        dmv[u] = dm_k;
        p1_next = p1;
        dm_n = dm_k;
      }
      libm::log_n(maxp, lm);
#pragma unroll
      for (int u = 0; u < U; ++u) earg[u] = (c_cap[u] - 1.0) * lm[u];
      libm::exp_n(earg, ev);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int k = k0 - u;
        if (k >= 0) {
          const double dzn = (-dmv[u]) * RDGAS * c_ptv[u] * ev[u];
          if (D3) {
            GST(delz, k, dzn);
            zcur = zcur - dzn;
          } else {
            zcur = zcur - dzn * GRAV;
          }
          GST(zh, k, zcur);
        }
      }
    }
  }
#undef G
#undef GST
#undef yld
#undef yst
}

// D3 = false: riem_solver_c, true: riem_solver3.  One warp per CTA, Y in shared memory.
template <bool D3>
__global__ void __launch_bounds__(32) k_riem(Geo g, RiemArgs a) {
  extern __shared__ double sm[];
  const int col = blockIdx.x * 32 + threadIdx.x;
  // columns of the box are numbered row by row and dealt to lanes in that order (no idle lanes at row ends)
  if (col >= a.b.ni * a.b.nj) return;  // no barrier is used anywhere below
  YAcc<false> Y{sm + threadIdx.x, 0u};
  riem_column<D3, false, D3 ? 4 : 2>(g, a, Y, col, blockIdx.y, true);
}

// Four warps per CTA, the first 64 levels of Y in tensor memory.
template <bool D3, int UU>
__global__ void __launch_bounds__(128, 4) k_riem_tmem(Geo g, RiemArgs a) {
  extern __shared__ double sm[];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = tmem_alloc_cta(&tmem_slot);
  const int ncol = a.b.ni * a.b.nj;
  const int col0 = blockIdx.x * 128 + warp * 32;       // first column of this warp
  if (col0 < ncol) {                                   // warp-uniform: the TMEM accesses below are warp-collective
    const int nsm = a.km + 1 > TMEM_LEVELS ? a.km + 1 - TMEM_LEVELS : 0;   // levels kept in shared memory
    YAcc<true> Y{sm + (size_t)warp * nsm * 32 + lane, tmem_warp_addr(base)};
    const int col = col0 + lane;
    riem_column<D3, true, UU>(g, a, Y, col < ncol ? col : ncol - 1, blockIdx.y, col < ncol);
  }
  tmem_free_cta(base);
}

// PACE_B200_RIEM=smem selects the shared-memory layout (A/B runs)
bool riem_use_tmem() {
  static const bool v = [] {
    const char* e = std::getenv("PACE_B200_RIEM");
    return !(e != nullptr && std::string(e) == "smem");
  }();
  return v;
}

template <bool D3>
int launch_riem(Core& c, RiemArgs& a, const char* label) {
  const Geo g = c.g;
  a.pm = c.S("sim1.pm"); a.xwork = c.S("sim1.x");
  if (!c.err.empty()) return 1;
  const int ncol = a.b.ni * a.b.nj;
  LaunchScope scope(c.stream, label);
  if (riem_use_tmem()) {
    const int nsm = a.km + 1 > TMEM_LEVELS ? a.km + 1 - TMEM_LEVELS : 0;
    const size_t smem = (size_t)nsm * 32 * 4 * sizeof(double);
    if (smem > 48 * 1024) { c.err = "riem_solver: too many levels for the on-chip kernel"; return 1; }
    dim3 grd((ncol + 127) / 128, g.nt, 1);
    // levels per chunk (PACE_B200_RIEM_U = 2 / 3): at C384 riem_solver_c 20.5 / 21.2, riem_solver3 23.3 / 24.7 ms per step
    // (three levels of exp / log chains spill 8 / 80 bytes at the 128 registers that 16 warps per SM allow)
    static const int u = [] { const char* e = std::getenv("PACE_B200_RIEM_U"); return e ? std::atoi(e) : 2; }();
    if (u == 3) k_riem_tmem<D3, 3><<<grd, 128, smem, c.stream>>>(g, a);
    else k_riem_tmem<D3, 2><<<grd, 128, smem, c.stream>>>(g, a);
    return 0;
  }
  const size_t smem = (size_t)(a.km + 1) * 32 * sizeof(double);
  if (func_attr_once(c, (const void*)k_riem<D3>, 96 * 1024)) return 1;  // once per instantiation and device (tile.cu)
  if (smem > 96 * 1024) { c.err = "riem_solver: too many levels for the on-chip kernel"; return 1; }
  dim3 grd((ncol + 31) / 32, g.nt, 1);
  k_riem<D3><<<grd, 32, smem, c.stream>>>(g, a);
  return 0;
}

}  // namespace

int riem_solver_c_onchip(Core& c, double dt2) {
  const Geo g = c.g;
  RiemArgs a{};
  a.delp = c.F("csw.delpc"); a.cappa = c.F("ad.cappa"); a.w_in = c.F("omga"); a.ptv = c.F("csw.ptc");
  a.q_con = c.F("q_con"); a.zsurf = c.F("phis"); a.ws = c.F("ad.ws3"); a.zh = c.F("ad.gz"); a.pef = c.F("ad.pkc");
  if (!c.err.empty()) return 1;
  a.ptop = c.cfg.ptop; a.p_fac = c.cfg.p_fac; a.dt = dt2; a.t1g = 2.0 * dt2 * dt2; a.rdt = 1.0 / dt2;
  a.peln1 = 0.0; a.ptk = 0.0; a.km = g.nz; a.last_call = 0; a.beta_neg = 0;
  a.b = box_compute(g, -1, -1, 2, 2, 1);
  return launch_riem<false>(c, a, "riem_solver_c");
}

int riem_solver3_onchip(Core& c, double dt, int last_call) {
  const Geo g = c.g;
  RiemArgs a{};
  a.delp = c.F("delp"); a.cappa = c.F("ad.cappa"); a.q_con = c.F("q_con"); a.ptv = c.F("pt"); a.zsurf = c.F("ad.zs");
  a.ws = c.F("ad.wsd"); a.pe = c.F("pe"); a.zh = c.F("ad.zh"); a.w_out = c.F("w"); a.w_in = a.w_out;
  a.delz = c.F("delz"); a.pef = c.F("ad.pkc"); a.pk3 = c.F("ad.pk3"); a.pk = c.F("pk"); a.peln = c.F("peln");
  if (!c.err.empty()) return 1;
  a.ptop = c.cfg.ptop; a.p_fac = c.cfg.p_fac; a.dt = dt; a.t1g = 2.0 * dt * dt; a.rdt = 1.0 / dt;
  a.peln1 = std::log(a.ptop);           // math.log(ptop)           (riem_solver3.py:287)
  a.ptk = std::exp(KAPPA * a.peln1);    // math.exp(KAPPA * peln1)  (riem_solver3.py:288)
  a.km = g.nz; a.last_call = last_call; a.beta_neg = c.cfg.beta < -0.1;
  a.b = box_compute(g, 0, 0, 0, 0, 1);
  return launch_riem<true>(c, a, "riem_solver3");
}

#else
int riem_solver_c_onchip(Core&, double) { return -1; }
int riem_solver3_onchip(Core&, double, int) { return -1; }
#endif

}  // namespace pace
