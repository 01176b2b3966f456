// pace_b200 — host-side state of one dycore instance (all subdomains resident on one GPU).
#pragma once
#include <map>
#include <memory>
#include <tuple>
#include <string>
#include <vector>

#include "common.cuh"

namespace pace {

// Mirrors the subset of pyFV3.DynamicalCoreConfig (_config.py:143-281) the hot path reads.
struct Config {
  int npx, npy, npz;       // tile size + 1, levels
  int nx, ny;              // subdomain compute size
  int n_tiles_local;       // subdomains on this GPU
  int n_split, k_split;
  int hord_dp, hord_mt, hord_tm, hord_tr, hord_vt;
  int kord_mt, kord_tm, kord_tr, kord_wz;
  int nord, n_sponge, nf_omega;
  int grid_type;
  int do_sat_adj, do_vort_damp, rf_fast, hydrostatic, convert_ke, fill, adiabatic, do_qa;
  int use_old_omega, do_f3d, inline_q, nwat;
  double dt_atmos;
  double a_imp, beta, p_fac, d_con, d2_bg, d2_bg_k1, d2_bg_k2, d4_bg, dddmp, delt_max, ke_bg;
  double rf_cutoff, tau, vtdm4, d_ext, consv_te;
  double ptop, da_min, da_min_c;
  int ks;
};

// mirror of pace_b200_satadj_config (SatAdjustConfig, pyFV3/_config.py:15-66)
struct SatCfg {
  int rad_snow, rad_rain, rad_graupel, tintqs, icloud_f, do_qa;
  double sat_adj0, ql_gen, qs_mlt, ql0_max, t_sub, qi_gen, qi_lim, qi0_max, dw_ocean, dw_land, cld_min;
  double tau_i2s, tau_v2l, tau_r2g, tau_l2r, tau_l2v, tau_imlt, tau_smlt;
};

typedef int (*halo_callback_t)(void* user, int plan_id, const char* const* x_names, const char* const* y_names, int n,
                               int nk);
typedef int (*stage_callback_t)(void* user, const char* savepoint);

struct Core {
  Config cfg;
  Geo g;
  cudaStream_t stream = 0;
  std::map<std::string, double*> fields;    // caller-owned device arrays (state, grid, temporaries)
  std::map<std::string, double*> scratch;   // library-owned 3-D scratch, allocated on first use
  std::map<std::string, std::vector<double>> hcol;  // host copies of K-columns set via pace_b200_set_column
  std::map<std::string, int> colsize;
  void* halo_ops[256] = {nullptr};               // device copies of halo plans (pace_b200_halo_plan_set)
  int halo_nops[256] = {0};
  std::vector<void*> owned;
  std::string err;
  // whole-phase entry points (step.cu): exchange callback for halo updates whose neighbours are not all on this GPU,
  // savepoint callback at the reference's checkpointer hooks, SatAdjust3d configuration
  halo_callback_t halo_cb = nullptr;
  void* halo_user = nullptr;
  stage_callback_t stage_cb = nullptr;
  void* stage_user = nullptr;
  // tensor maps of the tiled kernels (tile.cu), keyed by (field pointer, box width, box height); storage is 64-byte
  // aligned inside the allocation
  std::map<std::tuple<const void*, int, int>, std::unique_ptr<unsigned char[]>> tmaps;
  SatCfg sat_store;
  const SatCfg* sat_cfg = nullptr;
  // halo exchange with subdomains on other GPUs by stores into the peers' receive arenas (halo.cu, peer_*):
  // arena_ptrs[r] = rank r's arena as mapped into this process (two message halves of half_doubles doubles, then one
  // 8-byte flag slot per rank); a plan's messages: one pack plan per peer GPU, one unpack plan
  struct PeerPlan {
    std::vector<int> peer_ranks, pack_ids;
    int unpack_id = -1;
    long long max_recv_cells = 0;   // largest number of cells any rank receives in one update of this plan
    bool set = false;
    // all pack plans and the local gather plan in one device list (one launch per update): ops + the peer slot each op
    // stores to (-1: same-GPU gather)
    void* merged_ops = nullptr;
    int* merged_slot = nullptr;
    int n_merged = 0;
  };
  std::vector<char*> arena_ptrs;
  int peer_rank = 0;
  long long half_doubles = 0;
  unsigned long long peer_updates = 0;    // halo updates issued so far: its parity selects the message half
  void* peer_dev = nullptr;               // device block: barrier sequence number, error word, flag-slot pointers
  // an exchange whose halos are not needed right away runs on a side stream while the main stream computes
  // (step.cu: halo_begin / halo_end); at most one is open, and no other exchange is issued while it is
  cudaStream_t side_stream = 0;
  void *ev_fork = nullptr, *ev_join = nullptr;   // cudaEvent_t
  bool async_open = false;
  // inside AcousticDynamics q_con only changes by transport between two d_sw calls: after the first substep its level
  // flags are kept up to date from the halo frame (tracer.cu: level_nonzero_flags)
  bool in_acoustic = false, qcon_flags_fresh = false;
  // the level flags of the tracers (tracer.cu) are those of a TracerAdvection call that has just run: set by step()
  // around its remapping call only, so that a caller who changes the tracers between the two phases is never affected
  bool tracer_flags_fresh = false;
  PeerPlan peer_plans[16];

  double* F(const char* name);              // caller-registered field, error if missing
  double* S(const char* name);              // zero-initialised library scratch field [nt][nz+1][ny+7][sj]
  double* FS(const char* name) { return fields.count(name) ? fields[name] : S(name); }  // registered, else scratch
  bool has(const char* name) const { return fields.count(name) != 0; }
  const std::vector<double>* H(const char* name);  // host copy of a K-column, error if missing
  int col_max(const char* name);                    // int(max) of a K-column (e.g. nmax of a nord column)
  size_t field_elems() const { return (size_t)g.st3 * g.nt; }
  ~Core();
};

// ---- module entry points (one per reference class / stencil group) -------------------------------
int riem_solver_c(Core& c, double dt2);
// part 0: the whole call; 1: only the patches that lie inside the compute domain (they read no halo cell); 2: the rest
int c_sw(Core& c, double dt2, int part = 0);
bool c_sw_can_split(const Core& c);
int update_dz_c(Core& c, double dt2);
int update_dz_c_from_zh(Core& c, double dt2, bool gz_equals_zh);
int p_grad_c(Core& c, double dt2);
int set_column(Core& c, const char* name, const double* host_values, int n);
int halo_plan_set(Core& c, int plan_id, const void* host_ops, int n_ops);
int halo_update(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk);
int halo_pack(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk, double* buf);
int halo_unpack(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                const double* buf);
int peer_arena_set(Core& c, int my_rank, int n_ranks, void* const* arena_ptrs, long long half_doubles);
int peer_plan_set(Core& c, int plan_id, int n_peers, const int* peer_ranks, const int* pack_plan_ids, int unpack_plan_id,
                  long long max_recv_cells);
int peer_halo_update(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk);
// two updates that share a wait point as one exchange (one barrier): plan A's message first, plan B's behind it
int peer_halo_update2(Core& c, int plan_a, const char* const* xa, const char* const* ya, int na, int nka, int plan_b,
                      const char* const* xb, const char* const* yb, int nb, int nkb);
int peer_barrier(Core& c);
int peer_status(Core& c, unsigned long long* seq, unsigned long long* error);
int tracer_flux_prep(Core& c, int n_split);
int tracer_substep(Core& c, int nq, bool first_subcycle = true);
int tracer_swap_dp(Core& c);
int math_eval(Core& c, int op, int n, const double* x, const double* y, double* out);
int map_single(Core& c, double* q, const double* pe1, const double* pe2, const double* qs, int iv, int kord, bool xi,
               bool yi, double qmin);
int map_single_f32(Core& c, float* q, const float* pe1, const float* pe2, const float* qs, int iv, int kord, bool xi,
                   bool yi, float qmin);
int remap_main(Core& c, int nq);
int remap_finish(Core& c, int last_step);
int sat_adjust(Core& c, const SatCfg& sc, double mdt, int last_step, int kmp);
int neg_adj3(Core& c);
int c2l_ord4(Core& c);
int fv_setup(Core& c);
int pt_to_potential_density_pt(Core& c);
int omega_from_w(Core& c);
int copy_levels(Core& c, const char* src_name, const char* dst_name);
int fv_prep(Core& c, double dt, bool accumulate_courant = false);
int riem_solver3(Core& c, double dt, int last_call);
int update_dz_d(Core& c, double dt);
int zero_data(Core& c, int first_timestep);
int gz_from_zs_delz(Core& c);
int pem_from_delp(Core& c);
int copy_full(Core& c, const char* src_name, const char* dst_name);
int edge_pe(Core& c);
int pk3_halo(Core& c);
int compute_geopotential(Core& c);
int nh_p_grad(Core& c, double dt, bool keep_agrid = false);
int ray_fast(Core& c, double dt);
int del2cubed(Core& c, double* q, double cd, int nmax);
int apply_diffusive_heating(Core& c, double delt_time_factor, int nk);
int d_sw(Core& c, double dt);
int a2b_ord4(Core& c, const double* qin, double* qout, int k0, int nk, bool replace);
int acoustic(Core& c, double timestep, int n_map);
int tracer_advection(Core& c);
int remapping(Core& c, int last_step, double mdt);
int step(Core& c);
// diagnostics egress (egress.cu)
int egress_pack(Core& c, const char* name, int kind, int level, int x_interface, int y_interface, int nk, double* dst);
int column_integral(Core& c, const char* q_name, const char* delp_name, const char* out_name);
// dycore <-> physics coupling (coupling.cu; sequencing with the halo updates in step.cu)
int dycore_to_physics(Core& c);
int coupling_tendencies(Core& c, double dt_atmos, int dycore_only);
int coupling_thermo(Core& c, double dt);
int update_dwind_phys(Core& c, double dt_atmos, const int* origin_i, const int* origin_j, int npx, int npy);
int update_atmos_state(Core& c, double dt, int dycore_only, int apply_tendencies, const int* origin_i,
                       const int* origin_j, int npx, int npy);
int divergence_damping(Core& c, double* vort_b, double* delpc, double* ke, const double* vort_a, double dt);

}  // namespace pace
