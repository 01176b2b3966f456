// Argument bundles of the transport operators shared by d_sw, updatedzd and tracer advection.
#pragma once
#include "core.hpp"

namespace pace {

// FiniteVolumeTransport.__call__ (fvtp2d.py:251-368)
struct Fvtp2dArgs {
  double* q = nullptr;                       // transported scalar (its halo corners get x-copied, as in the reference)
  const double *crx = nullptr, *cry = nullptr, *xfx = nullptr, *yfx = nullptr;
  double *fx = nullptr, *fy = nullptr;       // out: fluxes on x / y interfaces
  const double *x_mass_flux = nullptr, *y_mass_flux = nullptr, *mass = nullptr;
  int hord = 6;
  int nk = 0;
  // optional DelnFlux (fvtp2d.py:168-180): per-level nord and damp = (damp_c*da_min)**(nord+1)
  const double *nord_col = nullptr, *damp = nullptr;
  int nmax = 0;
};
int fvtp2d(Core& c, const Fvtp2dArgs& a);

// DelnFluxNoSG.__call__ (+ DelnFlux accumulation when accumulate=true)
struct DelnArgs {
  const double* q = nullptr;
  double *fx = nullptr, *fy = nullptr;       // accumulate: fluxes updated in place; else: receive fx2, fy2
  const double* mass = nullptr;
  const double *nord_col = nullptr, *damp = nullptr;
  double* d2 = nullptr;                      // optional work/output array
  int nmax = 0, nk = 0;
  bool accumulate = false;
};
int delnflux(Core& c, const DelnArgs& a);

}  // namespace pace
