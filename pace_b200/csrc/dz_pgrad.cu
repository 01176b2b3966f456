// updatedzc (C-grid geopotential height advection) and the C-grid pressure-gradient update.
//
// Reference: pyFV3/stencils/updatedzc.py:18-118 (functions + update_dz_c), :185-226 (__call__);
//            pyFV3/stencils/dyn_core.py:122-173 (p_grad_c_stencil).
#include "core.hpp"
#include "corners.cuh"
#include "update_dz_c_body.cuh"

namespace pace {

// UpdateGeopotentialHeightOnCGrid.__call__(zs, ut, vt, gz, ws, dt): kernels in update_dz_c_body.cuh
int update_dz_c(Core& c, double dt) {
  double* gz0 = c.S("dzc.gz0");
  if (!c.err.empty()) return 1;
  return update_dz_c_with(c, dt, gz0);
}

// The same inside AcousticDynamics.__call__ (step.cu), where zh holds the pre-update heights: see update_dz_c_with.
int update_dz_c_from_zh(Core& c, double dt, bool gz_equals_zh) {
  double* zh = c.F("ad.zh");
  if (!c.err.empty()) return 1;
  return update_dz_c_with(c, dt, zh, gz_equals_zh ? 1 : 2);
}

// p_grad_c_stencil (dyn_core.py:122-173), nonhydrostatic branch (wk = delpc)
int p_grad_c(Core& c, double dt2) {
  const Geo g = c.g;
  const double *rdxc = c.F("grid.rdxc"), *rdyc = c.F("grid.rdyc"), *delpc = c.F("csw.delpc"),
               *pkc = c.F("ad.pkc"), *gz = c.F("ad.gz");
  double *uc = c.F("uc"), *vc = c.F("vc");
  if (!c.err.empty()) return 1;
  if (c.cfg.hydrostatic) {
    c.err = "Hydrostatic is not implemented";
    return 1;
  }
  launch_ijk(c.stream, g, box_compute(g, 0, 0, 1, 1, 0), [=] __device__(const Pt& p) {
    A0(uc) = A0(uc) + dt2 * A20(rdxc) / (A(delpc, -1, 0, 0) + A0(delpc)) *
                          ((A(gz, -1, 0, 1) - A0(gz)) * (A(pkc, 0, 0, 1) - A(pkc, -1, 0, 0)) +
                           (A(gz, -1, 0, 0) - A(gz, 0, 0, 1)) * (A(pkc, -1, 0, 1) - A0(pkc)));
    A0(vc) = A0(vc) + dt2 * A20(rdyc) / (A(delpc, 0, -1, 0) + A0(delpc)) *
                          ((A(gz, 0, -1, 1) - A0(gz)) * (A(pkc, 0, 0, 1) - A(pkc, 0, -1, 0)) +
                           (A(gz, 0, -1, 0) - A(gz, 0, 0, 1)) * (A(pkc, 0, -1, 1) - A0(pkc)));
  });
  return 0;
}

}  // namespace pace
