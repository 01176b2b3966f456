// Pointwise stencils of DynamicalCore._compute around the acoustic / tracer / remap sections.
//
// Reference: pyFV3/stencils/moist_cv.py:15-234 (fv_setup and helpers); pyFV3/stencils/fv_dynamics.py:33-58
// (pt_to_potential_density_pt, omega_from_w); basic_operations.py:7-16 (copy_defn).
#include "core.hpp"

namespace pace {

// fv_setup (moist_cv.py:201-234), moist_phys=True, nwat=6, compute domain
int fv_setup(Core& c) {
  const Geo g = c.g;
  if (!c.cfg.nwat || c.cfg.nwat != 6) { c.err = "Dynamical core (fv_dynamics): nwat != 6 is not implemented. Only nwat=6 has been implemented."; return 1; }
  const double *qvapor = c.F("qvapor"), *qliquid = c.F("qliquid"), *qrain = c.F("qrain"), *qsnow = c.F("qsnow"),
               *qice = c.F("qice"), *qgraupel = c.F("qgraupel"), *pt = c.F("pt"), *delp = c.F("delp"), *delz = c.F("delz");
  double *q_con = c.F("q_con"), *cvm = c.F("dyn.cvm"), *pkz = c.F("pkz"), *cappa = c.F("ad.cappa"), *dp1 = c.F("dyn.dp1");
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_compute(g), [=] __device__(const Pt& p) {
    double ql = A0(qliquid) + A0(qrain);
    double qs = A0(qice) + A0(qsnow) + A0(qgraupel);
    double gz = ql + qs;
    double qv = A0(qvapor);
    double cv = (1.0 - (qv + gz)) * CV_AIR + qv * CV_VAP + ql * C_LIQ + qs * C_ICE;
    A0(cvm) = cv;
    A0(q_con) = gz;
    double d1 = ZVIR * qv;
    A0(dp1) = d1;
    double cap = RDGAS / (RDGAS + cv / (1.0 + d1));
    A0(cappa) = cap;
    A0(pkz) = fv_exp(cap * fv_log(RDG * A0(delp) * A0(pt) * (1.0 + d1) * (1.0 - gz) / A0(delz)));
  });
  return 0;
}

// pt_to_potential_density_pt (fv_dynamics.py:33-46)
int pt_to_potential_density_pt(Core& c) {
  const Geo g = c.g;
  const double *pkz = c.F("pkz"), *dp1 = c.F("dyn.dp1"), *q_con = c.F("q_con");
  double* pt = c.F("pt");
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_compute(g), [=] __device__(const Pt& p) {
    A0(pt) = A0(pt) * (1.0 + A0(dp1)) * (1.0 - A0(q_con)) / A0(pkz);
  });
  return 0;
}

// omega_from_w (fv_dynamics.py:49-58)
int omega_from_w(Core& c) {
  const Geo g = c.g;
  const double *delp = c.F("delp"), *delz = c.F("delz"), *w = c.F("w");
  double* omga = c.F("omga");
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_compute(g), [=] __device__(const Pt& p) { A0(omga) = A0(delp) / A0(delz) * A0(w); });
  return 0;
}

// copy_defn on the full domain, npz levels (fv_dynamics.py:275-279, 524-527)
int copy_levels(Core& c, const char* src_name, const char* dst_name) {
  const Geo g = c.g;
  const double* src = c.F(src_name);
  double* dst = c.F(dst_name);
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_full(g), [=] __device__(const Pt& p) { A0(dst) = A0(src); });
  return 0;
}

}  // namespace pace

namespace pace {
// Test hook: evaluate the transcendentals used by the kernels (csrc/libm_port.cuh) elementwise on device
// arrays, so tests can compare them bit-for-bit with the host libm.  op 0: exp(x), 1: log(x), 2: pow(x, y).
int math_eval(Core& c, int op, int n, const double* x, const double* y, double* out) {
  Geo g = c.g;
  g.nt = 1;
  launch_ijk(c.stream, g, Box{0, 0, 0, n, 1, 1}, [=] __device__(const Pt& p) {
    const int i = p.i;
    out[i] = op == 0 ? fv_exp(x[i]) : (op == 1 ? fv_log(x[i]) : fv_pow(x[i], y[i]));
  });
  return 0;
}
}  // namespace pace
