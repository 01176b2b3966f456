// Whole-phase entry points: the call order of the reference's Python driver classes behind the C ABI, so that a
// caller in any language steps the core with one call per phase (or one per dycore step).
//
// Reference: AcousticDynamics.__call__ (pyFV3/stencils/dyn_core.py:714-1013), TracerAdvection.__call__
// (tracer_2d_1l.py:289-417), LagrangianToEulerian.__call__ (remapping.py:522-731), DynamicalCore.step_dynamics /
// _compute (fv_dynamics.py:445-655).  Where the reference starts a halo update early and waits later
// (HaloUpdater.start / wait) the fields are not modified in between, so the update is issued at the wait point.
// Host code only: it orders module calls (each a few kernel launches on the handle's stream) and halo updates.
#include <cstdlib>

#include "../../include/pace_b200.h"
#include "core.hpp"

namespace pace {

namespace {
enum Plan { A3 = 0, A2 = 1, B3 = 2, D3 = 3, C3 = 4, SYNC = 5 };  // plan ids of pace_b200.halo / INTEGRATION.md
const char* const TRACERS[8] = {"qvapor", "qliquid", "qrain", "qice", "qsnow", "qgraupel", "qo3mr", "qsgs_tke"};

struct Names {
  const char* v[8];
  int n = 0;
  Names() {}
  Names(const char* const* names, int count) {
    for (int i = 0; i < count; ++i) v[n++] = names[i];
  }
  Names(std::initializer_list<const char*> l) {
    for (const char* s : l) v[n++] = s;
  }
};

#define RUN(call)            \
  do {                       \
    if ((call) != 0) return 1; \
    if (!c.err.empty()) return 1; \
  } while (0)

// get_nk_heat_dissipation (dyn_core.py:176-191)
int nk_heat_dissipation(const Config& cfg) {
  if (cfg.convert_ke || cfg.vtdm4 > 1.0e-4) return cfg.npz;
  if (cfg.d2_bg_k1 < 1.0e-3) return 0;
  if (cfg.d2_bg_k2 < 1.0e-3) return 1;
  return 2;
}
}  // namespace

// One HaloUpdater.update(): fields whose neighbours all live on this GPU are filled by the gather kernel; a
// registered exchange callback (other GPUs / processes hold neighbours) takes the whole update instead.
static int halo(Core& c, int plan, const Names& x, const Names* y = nullptr, int nk = -1) {
  if (nk < 0) nk = c.g.nz;
  if (plan < 16 && c.peer_plans[plan].set && !c.arena_ptrs.empty())
    return peer_halo_update(c, plan, x.v, y ? y->v : nullptr, x.n, nk);
  if (c.halo_cb) {
    if (c.halo_cb(c.halo_user, plan, x.v, y ? y->v : nullptr, x.n, nk) != 0) {
      if (c.err.empty()) c.err = "halo exchange callback failed";
      return 1;
    }
    return 0;
  }
  return halo_update(c, plan, x.v, y ? y->v : nullptr, x.n, nk);
}

// two updates the reference waits for at the same point
static int halo2(Core& c, int plan_a, const Names& xa, const Names* ya, int nka, int plan_b, const Names& xb, const Names* yb,
                 int nkb) {
  if (nka < 0) nka = c.g.nz;
  if (nkb < 0) nkb = c.g.nz;
  if (plan_a < 16 && plan_b < 16 && c.peer_plans[plan_a].set && c.peer_plans[plan_b].set && !c.arena_ptrs.empty())
    return peer_halo_update2(c, plan_a, xa.v, ya ? ya->v : nullptr, xa.n, nka, plan_b, xb.v, yb ? yb->v : nullptr, xb.n, nkb);
  if (halo(c, plan_a, xa, ya, nka)) return 1;
  return halo(c, plan_b, xb, yb, nkb);
}

// HaloUpdater.start() ... wait() with compute in between (dyn_core.py:763/810: divgd across update_dz_c / riem_solver_c /
// p_grad_c; the delp / pt / q_con update is not needed before pk3_halo either): with peer arenas the whole exchange is
// queued on a side stream behind an event, the main stream goes on with kernels that touch neither the halos being
// filled nor the arenas, and halo_end makes the main stream wait for it.  Every rank opens and closes the same
// exchanges in the same order, and none is issued in between, so the barriers stay in step across the GPUs.
template <class Issue>
static int halo_async(Core& c, bool peer, Issue issue) {
#ifndef PACE_HOST_EMU
  static const bool no_overlap = [] { const char* v = getenv("PACE_B200_HALO_OVERLAP"); return v && v[0] == '0'; }();
  if (peer && !no_overlap && !c.async_open) {
    if (!c.side_stream) {
      cudaEvent_t a, b;
      if (cudaStreamCreateWithFlags(&c.side_stream, cudaStreamNonBlocking) != cudaSuccess ||
          cudaEventCreateWithFlags(&a, cudaEventDisableTiming) != cudaSuccess ||
          cudaEventCreateWithFlags(&b, cudaEventDisableTiming) != cudaSuccess) {
        c.err = "cannot create the halo side stream";
        return 1;
      }
      c.ev_fork = a; c.ev_join = b;
    }
    cudaEventRecord((cudaEvent_t)c.ev_fork, c.stream);
    cudaStreamWaitEvent(c.side_stream, (cudaEvent_t)c.ev_fork, 0);
    cudaStream_t main_stream = c.stream;
    c.stream = c.side_stream;
    const int rc = issue();
    c.stream = main_stream;
    if (rc) return 1;
    cudaEventRecord((cudaEvent_t)c.ev_join, c.side_stream);
    c.async_open = true;
    return 0;
  }
#endif
  (void)peer;
  return issue();
}
static bool peer_plan(const Core& c, int plan) { return plan < 16 && c.peer_plans[plan].set && !c.arena_ptrs.empty(); }
static int halo_begin(Core& c, int plan, const Names& x, const Names* y = nullptr, int nk = -1) {
  return halo_async(c, peer_plan(c, plan), [&] { return halo(c, plan, x, y, nk); });
}
static int halo2_begin(Core& c, int plan_a, const Names& xa, const Names* ya, int nka, int plan_b, const Names& xb,
                       const Names* yb, int nkb) {
  return halo_async(c, peer_plan(c, plan_a) && peer_plan(c, plan_b),
                    [&] { return halo2(c, plan_a, xa, ya, nka, plan_b, xb, yb, nkb); });
}
static int halo_end(Core& c) {
#ifndef PACE_HOST_EMU
  if (c.async_open) {
    cudaStreamWaitEvent(c.stream, (cudaEvent_t)c.ev_join, 0);
    c.async_open = false;
  }
#endif
  return 0;
}

static int stage(Core& c, const char* savepoint) {
  if (!c.stage_cb) return 0;
  if (c.stage_cb(c.stage_user, savepoint) != 0) {
    if (c.err.empty()) c.err = std::string("stage callback failed at ") + savepoint;
    return 1;
  }
  return 0;
}

// ---- AcousticDynamics.__call__ (dyn_core.py:714-1013) -----------------------------------------------------
int acoustic(Core& c, double timestep, int n_map) {
  const Config& cfg = c.cfg;
  const int npz = c.g.nz, n_split = cfg.n_split;
  const bool end_step = n_map == cfg.k_split;
  const double dt = timestep / n_split, dt2 = 0.5 * dt;
  const Names u{"u"}, v{"v"}, uc{"uc"}, vc{"vc"};
  struct InAcoustic {   // d_sw may reuse q_con's level flags between the substeps of this call only
    Core& c;
    explicit InAcoustic(Core& c_) : c(c_) { c.in_acoustic = true; c.qcon_flags_fresh = false; }
    ~InAcoustic() { c.in_acoustic = false; c.qcon_flags_fresh = false; }
  } in_acoustic(c);
  RUN(halo(c, A3, Names{"q_con", "ad.cappa"}));
  RUN(zero_data(c, n_map == 1));
  for (int it = 0; it < n_split; ++it) {
    const bool remap_step = it == n_split - 1;
    if (it == 0) {
      RUN(gz_from_zs_delz(c));
      RUN(halo(c, A3, Names{"delp", "pt"}));
    }
    if (it == n_split - 1 && end_step && cfg.use_old_omega) RUN(pem_from_delp(c));
    // c_sw needs the halos of u, v, w only in the patches on the rim of a subdomain: with peer arenas (and no
    // savepoint callback, which expects the halos before C_SW-In) the exchange runs on the side stream under the
    // interior patches, and the rim patches follow it
    if (!c.stage_cb && peer_plan(c, D3) && peer_plan(c, A3) && c_sw_can_split(c)) {
      RUN(halo2_begin(c, D3, u, &v, -1, A3, Names{"w"}, nullptr, -1));
      RUN(c_sw(c, dt2, 1));
      RUN(halo_end(c));
      RUN(c_sw(c, dt2, 2));
    } else {
      RUN(halo2(c, D3, u, &v, -1, A3, Names{"w"}, nullptr, -1));
      RUN(stage(c, "C_SW-In"));
      RUN(c_sw(c, dt2));
      RUN(stage(c, "C_SW-Out"));
    }
    // divgd is next read by d_sw: its exchange overlaps the C-grid height update, Riemann solver and pressure gradient
    // (first substep: the gz update follows at once, so it is done in line)
    if (cfg.nord > 0) {
      if (it == 0) RUN(halo(c, B3, Names{"ad.divgd"}));
      else RUN(halo_begin(c, B3, Names{"ad.divgd"}));
    }
    // the reference copies gz -> zh (first substep) or zh -> gz (later ones) and then updates gz from two corner-filled
    // copies of it; zh IS that pre-update copy, so the update reads zh and (later substeps) writes all of gz
    if (it == 0) {
      RUN(halo(c, A3, Names{"ad.gz"}, nullptr, npz + 1));
      RUN(copy_full(c, "ad.gz", "ad.zh"));
    }
    RUN(update_dz_c_from_zh(c, dt2, it == 0));
    RUN(riem_solver_c(c, dt2));
    RUN(p_grad_c(c, dt2));
    RUN(halo_end(c));
    RUN(halo(c, C3, uc, &vc));
    RUN(stage(c, "D_SW-In"));
    RUN(d_sw(c, dt));
    RUN(stage(c, "D_SW-Out"));
    // update_dz_d and riem_solver3 read delp, pt, q_con on the compute domain only: their halo update overlaps both
    RUN(halo_begin(c, A3, Names{"delp", "pt", "q_con"}));
    RUN(update_dz_d(c, dt));
    RUN(riem_solver3(c, dt, remap_step));
    RUN(halo_end(c));
    RUN(halo2(c, A3, Names{"ad.zh"}, nullptr, npz + 1, A2, Names{"ad.pkc"}, nullptr, npz + 1));
    if (remap_step) RUN(edge_pe(c));
    RUN(pk3_halo(c));
    RUN(compute_geopotential(c));
    RUN(nh_p_grad(c, dt, /*keep_agrid=*/it != n_split - 1));
    if (cfg.rf_fast) RUN(ray_fast(c, dt));
    if (it == n_split - 1 && cfg.grid_type < 4) RUN(halo(c, SYNC, u, &v));
  }
  const int nk_heat = nk_heat_dissipation(cfg);
  if (nk_heat != 0 && cfg.d_con > 1.0e-5) {
    RUN(halo(c, A3, Names{"ad.heat_source"}));
    double* hs = c.F("ad.heat_source");
    if (!hs) return 1;
    const int nmax = 3 < cfg.nord + 1 ? 3 : cfg.nord + 1;
    RUN(del2cubed(c, hs, 0.2 * cfg.da_min, nmax));
    RUN(apply_diffusive_heating(c, fabs(dt * cfg.delt_max), nk_heat));
  }
  return 0;
}

// ---- TracerAdvection.__call__ (tracer_2d_1l.py:289-417) ---------------------------------------------------
int tracer_advection(Core& c) {
  const double cmax_max_all_ranks = 2.0;  // hard-coded in the reference in place of an allreduce (:337-364)
  const int n_split = (int)floor(1.0 + cmax_max_all_ranks);
  const Names q(TRACERS, 8);
  RUN(tracer_flux_prep(c, n_split));
  RUN(halo(c, A3, q));
  for (int it = 0; it < n_split; ++it) {
    RUN(tracer_substep(c, 8, it == 0));
    if (it != n_split - 1) {
      RUN(halo(c, A3, q));
      RUN(tracer_swap_dp(c));
    }
  }
  return 0;
}

// ---- LagrangianToEulerian.__call__ (remapping.py:522-731) -------------------------------------------------
int remapping(Core& c, int last_step, double mdt) {
  RUN(remap_main(c, 8));
  if (c.cfg.do_sat_adj) {
    if (!c.sat_cfg) {
      c.err = "do_sat_adj is set but no saturation-adjustment configuration was given (pace_b200_set_satadj_config)";
      return 1;
    }
    // first level with pfull > 10 hPa (remapping.py:347-351)
    const std::vector<double>* p = c.H("grid.p");
    if (!p) return 1;
    int kmp = c.g.nz - 1;
    for (int k = 0; k < c.g.nz && k < (int)p->size(); ++k)
      if ((*p)[k] > 10.0e2) { kmp = k; break; }
    RUN(sat_adjust(c, *c.sat_cfg, mdt, last_step, kmp));
  }
  RUN(remap_finish(c, last_step));
  return 0;
}

// ---- DynamicalCore.step_dynamics / _compute (fv_dynamics.py:445-655) --------------------------------------
int step(Core& c) {
  const Config& cfg = c.cfg;
  RUN(stage(c, "FVDynamics-In"));
  RUN(fv_setup(c));                     // compute_preamble (fv_dynamics.py:461-508)
  RUN(pt_to_potential_density_pt(c));
  for (int k = 0; k < cfg.k_split; ++k) {
    const int n_map = k + 1;
    const bool last_step = k == cfg.k_split - 1;
    RUN(copy_levels(c, "delp", "dyn.dp1"));
    RUN(acoustic(c, cfg.dt_atmos / cfg.k_split, n_map));
    RUN(stage(c, "Tracer2D1L-In"));
    RUN(tracer_advection(c));
    RUN(stage(c, "Tracer2D1L-Out"));
    if (c.g.nz > 4) {
      RUN(stage(c, "Remapping-In"));
      // TracerAdvection has just run in this call and nothing but (optionally) the caller's savepoint hook came in
      // between: without a hook its level flags describe the tracers the remap is about to read
      c.tracer_flags_fresh = c.stage_cb == nullptr;
      const int remap_rc = remapping(c, last_step, cfg.dt_atmos / cfg.k_split);
      c.tracer_flags_fresh = false;
      if (remap_rc) return remap_rc;
      RUN(stage(c, "Remapping-Out"));
      if (last_step) {
        RUN(omega_from_w(c));
        if (cfg.nf_omega > 0) {
          RUN(halo(c, A3, Names{"omga"}));
          double* om = c.F("omga");
          if (!om) return 1;
          RUN(del2cubed(c, om, 0.18 * cfg.da_min, cfg.nf_omega));
        }
      }
    }
  }
  RUN(neg_adj3(c));
  const Names u{"u"}, v{"v"};
  RUN(halo(c, D3, u, &v));
  RUN(c2l_ord4(c));
  // keep the number of halo updates per step even, so that every step starts on the same message half (a captured
  // launch sequence can then be replayed) — the extra barrier also keeps the two-half protocol valid across steps
  if (!c.arena_ptrs.empty() && (c.peer_updates & 1)) {
    ++c.peer_updates;
    RUN(peer_barrier(c));
  }
  RUN(stage(c, "FVDynamics-Out"));
  return 0;
}

// ---- UpdateAtmosphereState.__call__ (pySHiELD/update/update_atmos_state.py:296-349) ---------------------------
// with ApplyPhysicsToDycore.__call__ (fv_update_phys.py:152-195): fill and tracer update, moist_cv, pressures, the halo
// update of the wind tendencies, their transfer to the D-grid winds, and the A-grid winds of the new state.  The two
// halo updates (the reference's three: u_dt and v_dt travel together here) keep the number of updates per step even.
int update_atmos_state(Core& c, double dt, int dycore_only, int apply_tendencies, const int* origin_i,
                       const int* origin_j, int npx, int npy) {
  RUN(coupling_tendencies(c, c.cfg.dt_atmos, dycore_only));
  if (!apply_tendencies) return 0;
  RUN(coupling_thermo(c, dt));
  RUN(halo(c, A3, Names{"tend.u_dt", "tend.v_dt"}));
  RUN(update_dwind_phys(c, c.cfg.dt_atmos, origin_i, origin_j, npx, npy));
  const Names u{"u"}, v{"v"};
  RUN(halo(c, D3, u, &v));
  RUN(c2l_ord4(c));
  return 0;
}

}  // namespace pace
