// Semi-implicit nonhydrostatic vertical solves: riem_solver_c (C-grid half step) and riem_solver3
// (D-grid full step), both around the SIM1 tridiagonal solver.
//
// Reference: pyFV3/stencils/sim1_solver.py:20-141 (sim1_solver), riem_solver_c.py:21-123,200-278
// (precompute / finalize / __call__), riem_solver3.py:26-146,225-339.
// The three reference stencils (precompute, sim1_solver, finalize) are column-local, so each solver
// is ONE kernel here: one thread per column, i fastest across the warp so every k-level access is a
// coalesced row segment.  Expression trees and k-sweep order follow the reference statement by
// statement (comments give the reference interval).
#include "core.hpp"

namespace pace {

struct Sim1Work {  // per-column work arrays, stored as scratch 3-D fields (coalesced like the inputs)
  double *w1, *g_rat, *bb, *dd, *bet, *gam, *pp, *aa, *p1;
};

#define K_(f, kk) (f)[p.o + (off3_t)(kk) * p.sk]

// sim1_solver.py:20-141.  km = number of layers; arrays indexed [0,km] (interfaces) or [0,km).
__device__ __forceinline__ void sim1_column(const Pt& p, int km, double* w, const double* dm,
                                            const double* gamma, double* dz, const double* ptv,
                                            const double* pm, double* pe, const double* pem, double ws,
                                            const double* cp3, double dt, double t1g, double rdt,
                                            double p_fac, const Sim1Work& s) {
  // PARALLEL interval(0,-1)
  for (int k = 0; k < km; ++k) {
    K_(pe, k) = fv_exp(K_(gamma, k) * fv_log(((-K_(dm, k)) / K_(dz, k) * RDGAS) * K_(ptv, k))) - K_(pm, k);
    K_(s.w1, k) = K_(w, k);
  }
  // FORWARD interval(0,-2) / (-2,-1)
  for (int k = 0; k < km - 1; ++k) {
    double gr = K_(dm, k) / K_(dm, k + 1);
    K_(s.g_rat, k) = gr;
    K_(s.bb, k) = 2.0 * (1.0 + gr);
    K_(s.dd, k) = 3.0 * (K_(pe, k) + gr * K_(pe, k + 1));
  }
  K_(s.bb, km - 1) = 2.0;
  K_(s.dd, km - 1) = 3.0 * K_(pe, km - 1);
  // bet = bb[k=0] everywhere
  double bb0 = K_(s.bb, 0);
  K_(s.bet, 0) = bb0;
  // pp[0] = 0, pp[1] = dd[0] / bet[1] (bet[1] still bb[0] here)
  K_(s.pp, 0) = 0.0;
  K_(s.pp, 1) = K_(s.dd, 0) / bb0;
  // FORWARD interval(1,-1): gam = g_rat[k-1]/bet[k-1]; bet = bb - gam
  for (int k = 1; k < km; ++k) {
    double gm = K_(s.g_rat, k - 1) / K_(s.bet, k - 1);
    K_(s.gam, k) = gm;
    K_(s.bet, k) = K_(s.bb, k) - gm;
  }
  // FORWARD interval(2,None): pp = (dd[k-1] - pp[k-1]) / bet[k-1]
  for (int k = 2; k <= km; ++k) K_(s.pp, k) = (K_(s.dd, k - 1) - K_(s.pp, k - 1)) / K_(s.bet, k - 1);
  // BACKWARD interval(1,-1)
  for (int k = km - 1; k >= 1; --k) {
    double ppk = K_(s.pp, k) - K_(s.gam, k) * K_(s.pp, k + 1);
    K_(s.pp, k) = ppk;
    K_(s.aa, k) = t1g * 0.5 * (K_(gamma, k - 1) + K_(gamma, k)) / (K_(dz, k - 1) + K_(dz, k)) * (K_(pem, k) + ppk);
  }
  // bet = delta_mass[0] - aa[1] everywhere, then the w sweep
  double betk = K_(dm, 0) - K_(s.aa, 1);
  // FORWARD interval(0,1)
  K_(w, 0) = (K_(dm, 0) * K_(s.w1, 0) + dt * K_(s.pp, 1)) / betk;
  // FORWARD interval(1,-2)
  for (int k = 1; k < km - 1; ++k) {
    double aak = K_(s.aa, k);
    double gm = aak / betk;
    K_(s.gam, k) = gm;
    betk = K_(dm, k) - (aak + K_(s.aa, k + 1) + aak * gm);
    K_(w, k) = (K_(dm, k) * K_(s.w1, k) + dt * (K_(s.pp, k + 1) - K_(s.pp, k)) - aak * K_(w, k - 1)) / betk;
  }
  // FORWARD interval(-2,-1)
  {
    int k = km - 1;
    double aak = K_(s.aa, k);
    double p1 = t1g * K_(gamma, k) / K_(dz, k) * (K_(pem, k + 1) + K_(s.pp, k + 1));
    double gm = aak / betk;
    K_(s.gam, k) = gm;
    betk = K_(dm, k) - (aak + p1 + aak * gm);
    K_(w, k) = (K_(dm, k) * K_(s.w1, k) + dt * (K_(s.pp, k + 1) - K_(s.pp, k)) - p1 * ws - aak * K_(w, k - 1)) / betk;
  }
  // BACKWARD interval(0,-2)
  for (int k = km - 2; k >= 0; --k) K_(w, k) = K_(w, k) - K_(s.gam, k + 1) * K_(w, k + 1);
  // FORWARD: pe
  K_(pe, 0) = 0.0;
  for (int k = 1; k <= km; ++k)
    K_(pe, k) = K_(pe, k - 1) + K_(dm, k - 1) * (K_(w, k - 1) - K_(s.w1, k - 1)) * rdt;
  // BACKWARD: p1
  K_(s.p1, km - 1) = (K_(pe, km - 1) + 2.0 * K_(pe, km)) * 1.0 / 3.0;
  for (int k = km - 2; k >= 0; --k)
    K_(s.p1, k) = (K_(pe, k) + K_(s.bb, k) * K_(pe, k + 1) + K_(s.g_rat, k) * K_(pe, k + 2)) * 1.0 / 3.0 -
                  K_(s.g_rat, k) * K_(s.p1, k + 1);
  // PARALLEL interval(0,-1)
  for (int k = 0; k < km; ++k) {
    double a = K_(s.p1, k) + K_(pm, k);
    double maxp = (p_fac * K_(dm, k) > a) ? p_fac * K_(pm, k) : a;
    K_(dz, k) = (-K_(dm, k)) * RDGAS * K_(ptv, k) * fv_exp((K_(cp3, k) - 1.0) * fv_log(maxp));
  }
}

static Sim1Work sim1_work(Core& c) {
  Sim1Work s;
  s.w1 = c.S("sim1.w1"); s.g_rat = c.S("sim1.g_rat"); s.bb = c.S("sim1.bb"); s.dd = c.S("sim1.dd");
  s.bet = c.S("sim1.bet"); s.gam = c.S("sim1.gam"); s.pp = c.S("sim1.pp"); s.aa = c.S("sim1.aa");
  s.p1 = c.S("sim1.p1");
  return s;
}

int riem_solver_c_onchip(Core& c, double dt2);             // riem_solver_onchip.cu
int riem_solver3_onchip(Core& c, double dt, int last_call);

// PACE_B200_RIEM=global selects the straightforward kernels of this file on the GPU too (A/B measurements)
static bool riem_use_onchip() {
#ifdef PACE_HOST_EMU
  return false;
#else
  static const bool on = [] { const char* v = getenv("PACE_B200_RIEM"); return !(v && std::string(v) == "global"); }();
  return on;
#endif
}

// riem_solver_c.py:200-278.  in: delpc, cappa, w3(omga), ptc, q_con, gz, ws3, phis -> gz, pkc(pef)
int riem_solver_c(Core& c, double dt2) {
  if (riem_use_onchip()) return riem_solver_c_onchip(c, dt2);
  const Geo g = c.g;
  const int km = g.nz;
  const double* delpc = c.F("csw.delpc");
  const double* cappa = c.F("ad.cappa");
  const double* w3 = c.F("omga");
  const double* ptc = c.F("csw.ptc");
  const double* q_con = c.F("q_con");
  const double* hs = c.F("phis");
  const double* ws = c.F("ad.ws3");
  double* gz = c.F("ad.gz");
  double* pef = c.F("ad.pkc");
  double *dm = c.S("riemc.dm"), *w = c.S("riemc.w"), *pem = c.S("riemc.pem"), *pe = c.S("riemc.pe"),
         *gm = c.S("riemc.gm"), *dz = c.S("riemc.dz"), *pm = c.S("riemc.pm"), *peg = c.S("riemc.peg");
  Sim1Work s = sim1_work(c);
  if (!c.err.empty()) return 1;
  const double ptop = c.cfg.ptop, p_fac = c.cfg.p_fac;
  const double dt = dt2, t1g = 2.0 * dt * dt, rdt = 1.0 / dt;
  // origin_compute(add=(-1,-1,0)), domain_compute(add=(2,2,1))
  Box b = box_compute(g, -1, -1, 2, 2, 1);
  launch_ij(c.stream, g, b, [=] __device__(const Pt& p) {
    // precompute (riem_solver_c.py:21-88)
    for (int k = 0; k <= km; ++k) {
      K_(dm, k) = K_(delpc, k);
      K_(w, k) = K_(w3, k);
    }
    K_(pem, 0) = ptop;
    K_(peg, 0) = ptop;
    for (int k = 1; k <= km; ++k) {
      K_(pem, k) = K_(pem, k - 1) + K_(dm, k - 1);
      K_(peg, k) = K_(peg, k - 1) + K_(dm, k - 1) * (1.0 - K_(q_con, k - 1));
    }
    for (int k = 0; k < km; ++k) K_(dz, k) = K_(gz, k + 1) - K_(gz, k);
    for (int k = 0; k <= km; ++k) {
      K_(gm, k) = 1.0 / (1.0 - K_(cappa, k));
      K_(dm, k) = K_(dm, k) / GRAV;
    }
    for (int k = 0; k < km; ++k) K_(pm, k) = (K_(peg, k + 1) - K_(peg, k)) / fv_log(K_(peg, k + 1) / K_(peg, k));
    sim1_column(p, km, w, dm, gm, dz, ptc, pm, pe, pem, A20(ws), cappa, dt, t1g, rdt, p_fac, s);
    // finalize (riem_solver_c.py:91-123)
    K_(pef, 0) = ptop;
    for (int k = 1; k <= km; ++k) K_(pef, k) = K_(pe, k) + K_(pem, k);
    K_(gz, km) = A20(hs);
    for (int k = km - 1; k >= 0; --k) K_(gz, k) = K_(gz, k + 1) - K_(dz, k) * GRAV;
  }, "riem_solver_c");
  return 0;
}

}  // namespace pace

namespace pace {

// NonhydrostaticVerticalSolver.__call__ (riem_solver3.py:225-339): precompute + sim1_solver + finalize
// in one column kernel.  in: delp cappa pe ad.zh q_con pt w delz ad.wsd ad.zs
// out: w delz ad.zh pe ad.pkc(ppe) ad.pk3 (+ pk, peln when last_call)
int riem_solver3(Core& c, double dt, int last_call) {
  const Geo g = c.g;
  const int km = g.nz;
  if (c.cfg.a_imp <= 0.999) { c.err = "a_imp <= 0.999 is not implemented"; return 1; }
  if (riem_use_onchip()) return riem_solver3_onchip(c, dt, last_call);
  const double *delp = c.F("delp"), *cappa = c.F("ad.cappa"), *q_con = c.F("q_con"), *pt = c.F("pt"),
               *zs = c.F("ad.zs"), *ws = c.F("ad.wsd");
  double *pe = c.F("pe"), *zh = c.F("ad.zh"), *w = c.F("w"), *delz = c.F("delz"), *ppe = c.F("ad.pkc"),
         *pk3 = c.F("ad.pk3"), *pk = c.F("pk"), *peln = c.F("peln");
  double *dm = c.S("riem3.dm"), *pe_init = c.S("riem3.pe_init"), *p_int = c.S("riem3.p_interface"),
         *logp = c.S("riem3.log_p_interface"), *gm = c.S("riem3.gamma"), *p_gas = c.S("riem3.p_gas"),
         *pg = c.S("riem3.pg"), *logpg = c.S("riem3.logpg");
  Sim1Work s = sim1_work(c);
  if (!c.err.empty()) return 1;
  const double ptop = c.cfg.ptop, p_fac = c.cfg.p_fac;
  const double t1g = 2.0 * dt * dt, rdt = 1.0 / dt;
  const double peln1 = std::log(ptop);            // math.log(ptop)           (riem_solver3.py:287)
  const double ptk = std::exp(KAPPA * peln1);     // math.exp(KAPPA * peln1)  (riem_solver3.py:288)
  constexpr double RGRAV = 1.0 / GRAV;
  const bool beta_neg = c.cfg.beta < -0.1;
  launch_ij(c.stream, g, box_compute(g, 0, 0, 0, 0, 1), [=] __device__(const Pt& p) {
    // precompute (riem_solver3.py:26-90)
    for (int k = 0; k <= km; ++k) {
      K_(dm, k) = K_(delp, k);
      K_(pe_init, k) = K_(pe, k);
    }
    K_(p_int, 0) = ptop; K_(logp, 0) = peln1; K_(pk3, 0) = ptk; K_(pg, 0) = ptop; K_(logpg, 0) = peln1;
    for (int k = 1; k <= km; ++k) {
      double pi = K_(p_int, k - 1) + K_(dm, k - 1);
      K_(p_int, k) = pi;
      double lp = fv_log(pi);
      K_(logp, k) = lp;
      double pgk = K_(pg, k - 1) + K_(dm, k - 1) * (1.0 - K_(q_con, k - 1));
      K_(pg, k) = pgk;
      K_(logpg, k) = fv_log(pgk);
      K_(pk3, k) = fv_exp(KAPPA * lp);
    }
    for (int k = 0; k <= km; ++k) {
      K_(gm, k) = 1.0 / (1.0 - K_(cappa, k));
      K_(dm, k) = K_(dm, k) * RGRAV;
    }
    for (int k = 0; k < km; ++k) {
      K_(p_gas, k) = (K_(pg, k + 1) - K_(pg, k)) / (K_(logpg, k + 1) - K_(logpg, k));
      K_(delz, k) = K_(zh, k + 1) - K_(zh, k);
    }
    sim1_column(p, km, w, dm, gm, delz, pt, p_gas, pe, p_int, A20(ws), cappa, dt, t1g, rdt, p_fac, s);
    // finalize (riem_solver3.py:93-146)
    for (int k = 0; k <= km; ++k) {
      K_(ppe, k) = beta_neg ? K_(pe, k) + K_(p_int, k) : K_(pe, k);
      if (last_call) {
        K_(peln, k) = K_(logp, k);
        K_(pk, k) = K_(pk3, k);
        K_(pe, k) = K_(p_int, k);
      } else {
        K_(pe, k) = K_(pe_init, k);
      }
    }
    K_(zh, km) = A20(zs);
    for (int k = km - 1; k >= 0; --k) K_(zh, k) = K_(zh, k + 1) - K_(delz, k);
  }, "riem_solver3");
  return 0;
}

}  // namespace pace
