// Dycore <-> physics coupling: the two calls the reference's driver makes around its physics in the
// time loop (driver/pace/driver/driver.py:645-661), with the state staying on the device.
//
// Reference: pySHiELD/update/update_atmos_state.py:22-37 (fill_gfs_delp), :40-98 (prepare_tendencies_and_update_tracers),
//            :101-147, 184-231 (DycoreToPhysics: copy_dycore_to_physics), :234-349 (UpdateAtmosphereState);
//            pySHiELD/update/fv_update_phys.py:21-76 (moist_cv, update_pressure_and_surface_winds), :152-195
//            (ApplyPhysicsToDycore.__call__); pySHiELD/update/update_dwind_phys.py:21-150, 490-701 (AGrid2DGridPhysics).
//
// Field names: "phy.<name>" is the reference's PhysicsState.<name>, "tend.u_dt / v_dt / pt_dt" its TendencyState,
// "dgrid.<name>" its DriverGridData (edge_vect_s / edge_vect_n, 1-D along i in the reference, are registered as 2-D
// fields constant along j).  The physics parameterisations themselves are not part of this library: the caller fills
// phy.prsi and phy.physics_updated_* between the two calls.
//
// AGrid2DGridPhysics is 23 reference stencils through 12 work fields (the 3-D Cartesian tendency v3, its edge sums ue /
// ve and their tile-edge corrections ut / vt).  Every one of them is a function of u_dt / v_dt at the point and its
// neighbours, so here ONE kernel evaluates, per D-grid wind, the expression tree the reference's stencils compose
// (same operations in the same order: same bits) and a second one zeroes the tendencies once nothing reads them.
#include "core.hpp"
#include "corners.cuh"

namespace pace {

namespace {
struct Ptr16 { const double* src[16]; double* dst[16]; };
const char* const COPIED[16] = {"qvapor", "qliquid", "qrain", "qsnow", "qice", "qgraupel", "qo3mr", "qsgs_tke",
                                "qcld", "pt", "delp", "delz", "ua", "va", "w", "omga"};
constexpr int MAX_SUBDOMAINS = 96;
struct Origins { int i[MAX_SUBDOMAINS], j[MAX_SUBDOMAINS]; };
}  // namespace

// DycoreToPhysics.__call__ without the dry convective adjustment (update_atmos_state.py:184-231): sixteen copies on
// [is, ie+1] x [js, je+1] x [0, npz)
int dycore_to_physics(Core& c) {
  const Geo g = c.g;
  Ptr16 p16;
  for (int n = 0; n < 16; ++n) {
    p16.src[n] = c.F(COPIED[n]);
    p16.dst[n] = c.F((std::string("phy.") + COPIED[n]).c_str());
  }
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
#pragma unroll
    for (int n = 0; n < 16; ++n) p16.dst[n][p.o] = p16.src[n][p.o];
  });
  return 0;
}

// fill_gfs_delp + prepare_tendencies_and_update_tracers (UpdateAtmosphereState.__call__, update_atmos_state.py:296-339)
int coupling_tendencies(Core& c, double dt_atmos, int dycore_only) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const int km = g.nz;
  double* delp = c.F("delp");
  double* qfill = c.F(dycore_only ? "qvapor" : "phy.physics_updated_specific_humidity");
  if (!c.err.empty()) return 1;
  // fill_gfs_delp(delp, q, q_min = 1e-9).  The reference runs it on the whole array, halo included, with npz + 1
  // levels (the intervals below are those of its npz + 1 level domain); the halo columns of a physics field hold
  // nothing anyone reads, so the compute domain is enough.
  const double q_min = 1.0e-9;
  launch_ij(st, g, box_compute(g), [=] __device__(const Pt& p) {
    // BACKWARD, interval(0, -2): levels km-2 .. 0 read the level below, already updated
    for (int k = km - 2; k >= 0; --k) {
      const double qb = A(qfill, 0, 0, k + 1);
      if (qb < q_min) A(qfill, 0, 0, k) = A(qfill, 0, 0, k) + (qb - q_min) * A(delp, 0, 0, k + 1) / A(delp, 0, 0, k);
    }
    // PARALLEL, interval(1, -1)
    for (int k = 1; k <= km - 1; ++k)
      if (A(qfill, 0, 0, k) < q_min) A(qfill, 0, 0, k) = q_min;
    // FORWARD, interval(1, -1): reads the level above, already updated
    for (int k = 1; k <= km - 1; ++k) {
      const double qa = A(qfill, 0, 0, k - 1);
      if (qa < 0.0) A(qfill, 0, 0, k) = A(qfill, 0, 0, k) + qa * A(delp, 0, 0, k - 1) / A(delp, 0, 0, k);
    }
    // FORWARD, interval(0, -1)
    for (int k = 0; k <= km - 1; ++k)
      if (A(qfill, 0, 0, k) < 0.0) A(qfill, 0, 0, k) = 0.0;
  });
  if (dycore_only) return 0;
  double *u_dt = c.F("tend.u_dt"), *v_dt = c.F("tend.v_dt"), *pt_dt = c.F("tend.pt_dt");
  const double *u_t1 = c.F("phy.physics_updated_ua"), *v_t1 = c.F("phy.physics_updated_va"),
               *pt_t1 = c.F("phy.physics_updated_pt"), *qv1 = c.F("phy.physics_updated_specific_humidity"),
               *ql1 = c.F("phy.physics_updated_qliquid"), *qr1 = c.F("phy.physics_updated_qrain"),
               *qs1 = c.F("phy.physics_updated_qsnow"), *qi1 = c.F("phy.physics_updated_qice"),
               *qg1 = c.F("phy.physics_updated_qgraupel"), *u_t0 = c.F("phy.ua"), *v_t0 = c.F("phy.va"),
               *pt_t0 = c.F("phy.pt"), *prsi = c.F("phy.prsi");
  double *qv0 = c.F("qvapor"), *ql0 = c.F("qliquid"), *qr0 = c.F("qrain"), *qs0 = c.F("qsnow"), *qi0 = c.F("qice"),
         *qg0 = c.F("qgraupel");
  if (!c.err.empty()) return 1;
  const double rdt = 1.0 / dt_atmos;
  launch_ijk(st, g, box_compute(g), [=] __device__(const Pt& p) {
    A0(u_dt) = A0(u_dt) + (A0(u_t1) - A0(u_t0)) * rdt;
    A0(v_dt) = A0(v_dt) + (A0(v_t1) - A0(v_t0)) * rdt;
    A0(pt_dt) = A0(pt_dt) + (A0(pt_t1) - A0(pt_t0)) * rdt;
    const double dp = A(prsi, 0, 0, 1) - A0(prsi);
    const double qwat_qv = dp * A0(qv1), qwat_ql = dp * A0(ql1), qwat_qr = dp * A0(qr1), qwat_qs = dp * A0(qs1),
                 qwat_qi = dp * A0(qi1), qwat_qg = dp * A0(qg1);
    const double qt = qwat_qv + qwat_ql + qwat_qr + qwat_qs + qwat_qi + qwat_qg;
    const double q_sum = A0(qv0) + A0(ql0) + A0(qr0) + A0(qs0) + A0(qi0) + A0(qg0);
    const double q0 = A0(delp) * (1.0 - q_sum) + qt;
    A0(delp) = q0;
    A0(qv0) = qwat_qv / q0;
    A0(ql0) = qwat_ql / q0;
    A0(qr0) = qwat_qr / q0;
    A0(qs0) = qwat_qs / q0;
    A0(qi0) = qwat_qi / q0;
    A0(qg0) = qwat_qg / q0;
  });
  return 0;
}

// moist_cv + update_pressure_and_surface_winds (ApplyPhysicsToDycore.__call__, fv_update_phys.py:159-186); the surface
// winds the reference also extracts there go to two fields nothing reads (its own TODO) and are left out
int coupling_thermo(Core& c, double dt) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const int km = g.nz;
  const double *qv = c.F("qvapor"), *ql_ = c.F("qliquid"), *qr = c.F("qrain"), *qs_ = c.F("qsnow"), *qi = c.F("qice"),
               *qg = c.F("qgraupel"), *delp = c.F("delp");
  double *pt = c.F("pt"), *t_dt = c.F("tend.pt_dt"), *pe = c.F("pe"), *peln = c.F("peln"), *pk = c.F("pk"), *ps = c.F("ps");
  if (!c.err.empty()) return 1;
  const double con_cp = CP_AIR;
  launch_ijk(st, g, box_compute(g), [=] __device__(const Pt& p) {
    const double ql = A0(ql_) + A0(qr);
    const double qs = A0(qi) + A0(qs_) + A0(qg);
    const double gz = ql + qs;
    const double cvm = (1.0 - (A0(qv) + gz)) * CV_AIR + A0(qv) * CV_VAP + ql * C_LIQ + qs * C_ICE;
    A0(pt) = A0(pt) + A0(t_dt) * dt * con_cp / cvm;
    A0(t_dt) = 0.0;
  });
  launch_ij(st, g, box_compute(g), [=] __device__(const Pt& p) {
    double pe_k = A(pe, 0, 0, 0);
    for (int k = 1; k <= km; ++k) {
      pe_k = pe_k + A(delp, 0, 0, k - 1);
      A(pe, 0, 0, k) = pe_k;
      const double ln = fv_log(pe_k);
      A(peln, 0, 0, k) = ln;
      A(pk, 0, 0, k) = fv_exp(KAPPA * ln);
    }
    A20(ps) = pe_k;
  });
  return 0;
}

// AGrid2DGridPhysics.__call__ (update_dwind_phys.py:490-701), grid_type < 4.  origin_i / origin_j: what has to be added
// to a local (halo-inclusive) index of subdomain t to get the reference's "global" index on the tile
// (local_to_global_1d: subtile_index * subtile_width).
int update_dwind_phys(Core& c, double dt_atmos, const int* origin_i, const int* origin_j, int npx, int npy) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  if (c.cfg.grid_type > 3) { c.err = "update_dwind_phys: doubly periodic grids (grid_type > 3) are not implemented"; return 1; }
  if (g.nt > MAX_SUBDOMAINS) { c.err = "update_dwind_phys: more subdomains in one handle than supported (96)"; return 1; }
  double *u = c.F("u"), *v = c.F("v"), *u_dt = c.F("tend.u_dt"), *v_dt = c.F("tend.v_dt");
  const double *vlon1 = c.F("dgrid.vlon1"), *vlon2 = c.F("dgrid.vlon2"), *vlon3 = c.F("dgrid.vlon3"),
               *vlat1 = c.F("dgrid.vlat1"), *vlat2 = c.F("dgrid.vlat2"), *vlat3 = c.F("dgrid.vlat3"),
               *ev_w = c.F("dgrid.edge_vect_w"), *ev_e = c.F("dgrid.edge_vect_e"), *ev_s = c.F("dgrid.edge_vect_s"),
               *ev_n = c.F("dgrid.edge_vect_n"), *es1_1 = c.F("dgrid.es1_1"), *es1_2 = c.F("dgrid.es1_2"),
               *es1_3 = c.F("dgrid.es1_3"), *ew2_1 = c.F("dgrid.ew2_1"), *ew2_2 = c.F("dgrid.ew2_2"),
               *ew2_3 = c.F("dgrid.ew2_3");
  if (!c.err.empty()) return 1;
  Origins org;
  for (int t = 0; t < g.nt; ++t) { org.i[t] = origin_i[t]; org.j[t] = origin_j[t]; }
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;
  const int im2 = (npx - 1) / 2 + 2, jm2 = (npy - 1) / 2 + 2;
  const double dt5 = 0.5 * dt_atmos;
  launch_ijk(st, g, box_dims(g, 1, 1, 0), [=] __device__(const Pt& p) {
    // v3_n at offset (di, dj): the physics tendency as a 3-D Cartesian vector (update_dwind_prep_stencil)
    auto v3 = [&](int n, int di, int dj) {
      const double* lon = n == 0 ? vlon1 : n == 1 ? vlon2 : vlon3;
      const double* lat = n == 0 ? vlat1 : n == 1 ? vlat2 : vlat3;
      return A(u_dt, di, dj, 0) * A2(lon, di, dj) + A(v_dt, di, dj, 0) * A2(lat, di, dj);
    };
    auto ue = [&](int n, int di) { return v3(n, di, -1) + v3(n, di, 0); };   // ue_n(i + di, j)
    auto ve = [&](int n, int dj) { return v3(n, -1, dj) + v3(n, 0, dj); };   // ve_n(i, j + dj)
    const int gi = p.i + org.i[p.t], gj = p.j + org.j[p.t];
    const int g_is = is + org.i[p.t];
    if (p.i <= ie) {   // u on [is, ie] x [js, je+1]
      double e[3];
#pragma unroll
      for (int n = 0; n < 3; ++n) e[n] = ue(n, 0);
      // tile edges: south row js, north row je+1; west half of the edge leans on the point to the east, east half on
      // the point to the west (update_dwind_x_edge_west / east stencils)
      const bool south = p.S && p.j == js, north = p.N && p.j == je + 1;
      if (south || north) {
        const double w = south ? A20(ev_s) : A20(ev_n);
        const bool lower = gi <= im2 && (south || g_is < im2);   // the north edge's lower part: global_is < im2 (:399)
        const bool upper = gi > im2;
        if (lower || upper) {
          const int di = lower ? 1 : -1;
#pragma unroll
          for (int n = 0; n < 3; ++n) e[n] = w * ue(n, di) + (1.0 - w) * e[n];
        }
      }
      A0(u) = A0(u) + dt5 * (e[0] * A20(es1_1) + e[1] * A20(es1_2) + e[2] * A20(es1_3));
    }
    if (p.j <= je) {   // v on [is, ie+1] x [js, je]
      double e[3];
#pragma unroll
      for (int n = 0; n < 3; ++n) e[n] = ve(n, 0);
      const bool west = p.W && p.i == is, east = p.E && p.i == ie + 1;
      if (west || east) {
        const double w = west ? A20(ev_w) : A20(ev_e);
        const int dj = gj <= jm2 ? 1 : -1;
#pragma unroll
        for (int n = 0; n < 3; ++n) e[n] = w * ve(n, dj) + (1.0 - w) * e[n];
      }
      A0(v) = A0(v) + dt5 * (e[0] * A20(ew2_1) + e[1] * A20(ew2_2) + e[2] * A20(ew2_3));
    }
  });
  // set_winds_zero on [is-1, ie+1] x [js-1, je+1] (the reference does it right after forming ue / ve)
  launch_ijk(st, g, box_compute(g, -1, -1, 2, 2, 0), [=] __device__(const Pt& p) {
    A0(u_dt) = 0.0;
    A0(v_dt) = 0.0;
  });
  return 0;
}

}  // namespace pace
