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
  // optional, device: level_nonzero[t * nk + k] == 0 says that q is zero on the whole level k of subdomain t (halo
  // included); the fluxes of such a level are zeros that nobody needs to form: fx / fy are then left untouched there
  const int* level_nonzero = nullptr;
};
int fvtp2d(Core& c, const Fvtp2dArgs& a);
// level_nonzero flags of a field (see Fvtp2dArgs); returns a library-owned device array of nt * nk ints, or nullptr
// when the build has no such kernel (host emulation: everything is computed)
const int* level_nonzero_flags(Core& c, const double* q, int nk, int slot = 9, bool frame_only = false);
const int* tracer_zero_fields(Core& c, int nq);

// heat_diss (d_sw.py:87-125) applied to the damping fluxes of w inside the DelnFlux kernels (optional)
struct DelnHeat {
  double *dw, *heat_s;
  const double *col_damp_w, *col_ke_bg;
  double absdt;
  bool done = false;   // set by delnflux when its kernels evaluated heat_diss (the fluxes are then not stored)
};

// DelnFluxNoSG.__call__ (+ DelnFlux accumulation when accumulate=true)
struct DelnArgs {
  const double* q = nullptr;
  double *fx = nullptr, *fy = nullptr;       // accumulate: fluxes updated in place; else: receive fx2, fy2
  const double* mass = nullptr;
  const double *nord_col = nullptr, *damp = nullptr;
  double* d2 = nullptr;                      // optional work/output array
  int nmax = 0, nk = 0;
  bool accumulate = false;
  const int* level_nonzero = nullptr;        // accumulate only: levels of zeros (see Fvtp2dArgs) add nothing
  DelnHeat* heat = nullptr;                  // not accumulating: see DelnHeat
};
int delnflux(Core& c, const DelnArgs& a);

}  // namespace pace
