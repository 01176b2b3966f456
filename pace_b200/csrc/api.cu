// C ABI of libpace_b200 (declared in include/pace_b200.h) and the Core plumbing behind it.
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "../../include/pace_b200.h"
#include "core.hpp"
#include "transport.hpp"

namespace pace {

static thread_local std::string g_err;

// ---- launch accounting ------------------------------------------------------------------------------------
// Process-wide by design (pace_b200_launch_count / pace_b200_profile_* take no handle: bench.py reads the totals of
// everything the library launched); the event record is guarded so that handles driven from several host threads
// cannot corrupt it, and the counter is per thread-of-execution exact only when one thread launches at a time.
static long long g_launches = 0;
long long& launch_counter() { return g_launches; }
#ifndef PACE_HOST_EMU
struct ProfEntry { const char* label; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfEntry> g_prof;
static std::vector<cudaEvent_t> g_event_pool;
static std::mutex g_prof_mutex;
static thread_local cudaEvent_t t_open_end = nullptr;   // end event of the scope this thread has open
static cudaEvent_t get_event() {
  if (!g_event_pool.empty()) { cudaEvent_t e = g_event_pool.back(); g_event_pool.pop_back(); return e; }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}
bool prof_enabled() { return g_prof_on; }
void prof_begin(cudaStream_t s, const char* label) {
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  ProfEntry p{label, get_event(), get_event()};
  cudaEventRecord(p.a, s);
  g_prof.push_back(p);
  t_open_end = p.b;
}
void prof_end(cudaStream_t s) {
  if (t_open_end) cudaEventRecord(t_open_end, s);
  t_open_end = nullptr;
}
#else
bool prof_enabled() { return false; }
void prof_begin(cudaStream_t, const char*) {}
void prof_end(cudaStream_t) {}
#endif

bool zero_paths_enabled() {
  static const bool on = [] { const char* v = getenv("PACE_B200_ZERO_PATHS"); return !(v && v[0] == '0'); }();
  return on;
}

void set_error(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
}

double* Core::F(const char* name) {
  auto it = fields.find(name);
  if (it == fields.end()) {
    if (err.empty()) err = std::string("field not registered: ") + name;
    return nullptr;
  }
  return it->second;
}

double* Core::S(const char* name) {
  auto it = scratch.find(name);
  if (it != scratch.end()) return it->second;
  void* ptr = nullptr;
  size_t bytes = field_elems() * sizeof(double);
  cudaError_t e = cudaMalloc(&ptr, bytes);
  if (e != cudaSuccess) {
    if (err.empty()) err = std::string("cudaMalloc failed for scratch ") + name + ": " + cudaGetErrorString(e);
    return nullptr;
  }
  cudaMemsetAsync(ptr, 0, bytes, stream);
  owned.push_back(ptr);
  scratch[name] = (double*)ptr;
  return (double*)ptr;
}

const std::vector<double>* Core::H(const char* name) {
  auto it = hcol.find(name);
  if (it == hcol.end()) {
    if (err.empty()) err = std::string("column not set: ") + name;
    return nullptr;
  }
  return &it->second;
}

int Core::col_max(const char* name) {
  const std::vector<double>* v = H(name);
  if (!v) return 0;
  double m = 0;
  for (int k = 0; k < g.nz && k < (int)v->size(); ++k) m = (*v)[k] > m ? (*v)[k] : m;
  return (int)m;
}

int set_column(Core& c, const char* name, const double* host_values, int n) {
  c.hcol[name] = std::vector<double>(host_values, host_values + n);
  double* dev = nullptr;
  auto it = c.fields.find(name);
  if (it != c.fields.end() && c.colsize.count(name) && c.colsize[name] >= n) {
    dev = it->second;
  } else {
    void* ptr = nullptr;
    if (cudaMalloc(&ptr, sizeof(double) * n) != cudaSuccess) {
      c.err = "cudaMalloc failed for column";
      return 1;
    }
    c.owned.push_back(ptr);
    dev = (double*)ptr;
    c.fields[name] = dev;
    c.colsize[name] = n;
  }
  // stream-ordered so a column rewritten between calls (e.g. col.rf) cannot overtake queued kernels
  if (cudaMemcpyAsync(dev, c.hcol[name].data(), sizeof(double) * n, cudaMemcpyHostToDevice, c.stream) != cudaSuccess) {
    c.err = "cudaMemcpy failed for column";
    return 1;
  }
  return 0;
}

Core::~Core() {
  for (void* p : owned) cudaFree(p);
#ifndef PACE_HOST_EMU
  if (ev_fork) cudaEventDestroy((cudaEvent_t)ev_fork);
  if (ev_join) cudaEventDestroy((cudaEvent_t)ev_join);
  if (side_stream) cudaStreamDestroy(side_stream);
#endif
}

static int finish(Core* c, int rc) {
  if (rc == 0 && c->err.empty()) {
    cudaError_t e = cudaPeekAtLastError();
    if (e == cudaSuccess) return 0;
    c->err = std::string("CUDA launch error: ") + cudaGetErrorString(e);
  }
  if (c->err.empty()) c->err = g_err.empty() ? "unknown error" : g_err;
  return 1;
}

}  // namespace pace

using namespace pace;

extern "C" {

int pace_b200_row_stride(int nx) { return (nx + 2 * HALO + 1 + 7) / 8 * 8; }

int pace_b200_create(const pace_b200_config* cfg, const unsigned char* edge_flags, void** handle) {
  static_assert(sizeof(pace_b200_config) == sizeof(Config), "config mirror out of sync");
  if (!cfg || !handle) return 1;
  if (cfg->n_tiles_local < 1 || cfg->n_tiles_local > MAX_TILES) {
    set_error("n_tiles_local must be in [1,%d]", MAX_TILES);
    return 1;
  }
  Core* c = new Core();
  memcpy(&c->cfg, cfg, sizeof(Config));
  Geo& g = c->g;
  g.nx = cfg->nx; g.ny = cfg->ny; g.nz = cfg->npz; g.nt = cfg->n_tiles_local;
  g.sj = pace_b200_row_stride(cfg->nx);
  g.sk = g.sj * (cfg->ny + 2 * HALO + 1);
  g.st3 = (long long)g.sk * (cfg->npz + 1);
  if (g.st3 * (long long)cfg->n_tiles_local >= (1LL << 31)) {
    g_err = "pace_b200_create: 2^31 or more points per field on one GPU (32-bit field offsets); use more GPUs";
    delete c;
    return 1;
  }
  g.is = HALO; g.ie = HALO + g.nx - 1; g.js = HALO; g.je = HALO + g.ny - 1;
  g.edge_bits[0] = g.edge_bits[1] = 0;
  for (int t = 0; t < g.nt; ++t)
    g.edge_bits[t >> 4] |= (unsigned long long)((edge_flags ? edge_flags[t] : 15) & 15) << ((t & 15) * 4);
  *handle = c;
  return 0;
}

int pace_b200_destroy(void* h) {
  delete (Core*)h;
  return 0;
}

const char* pace_b200_last_error(void* h) {
  if (!h) return g_err.c_str();
  return ((Core*)h)->err.c_str();
}

int pace_b200_set_stream(void* h, void* s) {
  ((Core*)h)->stream = (cudaStream_t)s;
  return 0;
}

int pace_b200_set_field(void* h, const char* name, double* p) {
  ((Core*)h)->fields[name] = p;
  return 0;
}

int pace_b200_set_column(void* h, const char* name, const double* host_values, int n) {
  Core* c = (Core*)h;
  c->err.clear();
  return set_column(*c, name, host_values, n);
}

static double* opt_field(Core* c, const char* name) { return (name && name[0]) ? c->F(name) : nullptr; }

long long pace_b200_launch_count(void) { return launch_counter(); }

int pace_b200_profile_enable(int on) {
#ifndef PACE_HOST_EMU
  g_prof_on = on != 0;
#endif
  return 0;
}

// Writes "label count total_ms\n" lines for every labelled kernel timed since the last report; clears the record.
int pace_b200_profile_report(char* buf, int n) {
  std::string out;
#ifndef PACE_HOST_EMU
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lock(g_prof_mutex);
  std::map<std::string, std::pair<long long, double>> acc;
  for (auto& p : g_prof) {
    float ms = 0;
    cudaEventElapsedTime(&ms, p.a, p.b);
    auto& e = acc[p.label];
    e.first += 1;
    e.second += ms;
    g_event_pool.push_back(p.a);
    g_event_pool.push_back(p.b);
  }
  g_prof.clear();
  char line[256];
  for (auto& kv : acc) {
    snprintf(line, sizeof line, "%s %lld %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
    out += line;
  }
#endif
  if (n > 0) {
    strncpy(buf, out.c_str(), n - 1);
    buf[n - 1] = 0;
  }
  return 0;
}

#define ENTRY(call)          \
  Core* c = (Core*)h;        \
  c->err.clear();            \
  return finish(c, (call));

int pace_b200_riem_solver_c(void* h, double dt2) { ENTRY(riem_solver_c(*c, dt2)) }
int pace_b200_c_sw(void* h, double dt2) { ENTRY(c_sw(*c, dt2)) }
int pace_b200_update_dz_c(void* h, double dt2) { ENTRY(update_dz_c(*c, dt2)) }
int pace_b200_p_grad_c(void* h, double dt2) { ENTRY(p_grad_c(*c, dt2)) }
int pace_b200_fv_prep(void* h, double dt) { ENTRY(fv_prep(*c, dt)) }
int pace_b200_fv_setup(void* h) { ENTRY(fv_setup(*c)) }
int pace_b200_neg_adj3(void* h) { ENTRY(neg_adj3(*c)) }
int pace_b200_c2l_ord4(void* h) { ENTRY(c2l_ord4(*c)) }
int pace_b200_sat_adjust(void* h, const pace_b200_satadj_config* cfg, double mdt, int last_step, int kmp) {
  static_assert(sizeof(pace_b200_satadj_config) == sizeof(SatCfg), "satadj config mirror out of sync");
  ENTRY(sat_adjust(*c, *(const SatCfg*)cfg, mdt, last_step, kmp))
}
int pace_b200_set_satadj_config(void* h, const pace_b200_satadj_config* cfg) {
  Core* c = (Core*)h;
  if (!cfg) { c->sat_cfg = nullptr; return 0; }
  memcpy(&c->sat_store, cfg, sizeof(SatCfg));
  c->sat_cfg = &c->sat_store;
  return 0;
}
int pace_b200_set_halo_callback(void* h, pace_b200_halo_callback cb, void* user) {
  Core* c = (Core*)h;
  c->halo_cb = cb; c->halo_user = user;
  return 0;
}
int pace_b200_set_stage_callback(void* h, pace_b200_stage_callback cb, void* user) {
  Core* c = (Core*)h;
  c->stage_cb = cb; c->stage_user = user;
  return 0;
}
int pace_b200_halo_peer_arena(void* h, int my_rank, int n_ranks, void* const* arena_ptrs, long long half_doubles) {
  ENTRY(peer_arena_set(*c, my_rank, n_ranks, arena_ptrs, half_doubles))
}
int pace_b200_halo_peer_plan(void* h, int plan_id, int n_peers, const int* peer_ranks, const int* pack_plan_ids,
                             int unpack_plan_id, long long max_recv_cells) {
  ENTRY(peer_plan_set(*c, plan_id, n_peers, peer_ranks, pack_plan_ids, unpack_plan_id, max_recv_cells))
}
int pace_b200_halo_peer_status(void* h, unsigned long long* barriers_done, unsigned long long* timed_out) {
  ENTRY(peer_status(*c, barriers_done, timed_out))
}
int pace_b200_halo_peer_update(void* h, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk) {
  Core* c = (Core*)h;
  c->err.clear();
  if (plan_id < 0 || plan_id >= 16 || !c->peer_plans[plan_id].set || c->arena_ptrs.empty()) {
    c->err = "halo_peer_update: no peer plan / arena set for this plan id";
    return 1;
  }
  return finish(c, peer_halo_update(*c, plan_id, x_names, y_names, n, nk));
}
int pace_b200_acoustic(void* h, double timestep, int n_map) { ENTRY(acoustic(*c, timestep, n_map)) }
int pace_b200_tracer_advection(void* h) { ENTRY(tracer_advection(*c)) }
int pace_b200_remap(void* h, int last_step, double mdt) { ENTRY(remapping(*c, last_step, mdt)) }
int pace_b200_step(void* h) { ENTRY(step(*c)) }
int pace_b200_dycore_to_physics(void* h) { ENTRY(dycore_to_physics(*c)) }
int pace_b200_egress_pack(void* h, const char* name, int kind, int level, int x_interface, int y_interface, int nk, double* dst) {
  ENTRY(egress_pack(*c, name, kind, level, x_interface, y_interface, nk, dst))
}
int pace_b200_column_integral(void* h, const char* q, const char* delp, const char* out2d) { ENTRY(column_integral(*c, q, delp, out2d)) }
int pace_b200_update_atmos_state(void* h, double dt, int dycore_only, int apply_tendencies, const int* origin_i,
                                 const int* origin_j, int npx, int npy) {
  ENTRY(update_atmos_state(*c, dt, dycore_only, apply_tendencies, origin_i, origin_j, npx, npy))
}
int pace_b200_remap_main(void* h, int nq) { ENTRY(remap_main(*c, nq)) }
int pace_b200_remap_finish(void* h, int last_step) { ENTRY(remap_finish(*c, last_step)) }
int pace_b200_map_single(void* h, const char* q, const char* pe1, const char* pe2, const char* qs, int iv, int kord,
                         int x_interface, int y_interface, double qmin) {
  Core* c = (Core*)h;
  c->err.clear();
  double *qp = c->F(q), *p1 = c->F(pe1), *p2 = c->F(pe2), *qsp = opt_field(c, qs);
  if (!c->err.empty()) return 1;
  return finish(c, map_single(*c, qp, p1, p2, qsp, iv, kord, x_interface != 0, y_interface != 0, qmin));
}
int pace_b200_map_single_f32(void* h, float* q, const float* pe1, const float* pe2, const float* qs, int iv, int kord,
                             int x_interface, int y_interface, float qmin) {
  Core* c = (Core*)h;
  c->err.clear();
  if (!q || !pe1 || !pe2) { c->err = "map_single_f32: null field pointer"; return 1; }
  return finish(c, map_single_f32(*c, q, pe1, pe2, qs, iv, kord, x_interface != 0, y_interface != 0, qmin));
}
int pace_b200_math_eval(void* h, int op, int n, const double* x, const double* y, double* out) {
  ENTRY(math_eval(*c, op, n, x, y, out))
}
int pace_b200_tracer_flux_prep(void* h, int n_split) { ENTRY(tracer_flux_prep(*c, n_split)) }
int pace_b200_tracer_substep(void* h, int nq) { ENTRY(tracer_substep(*c, nq)) }
int pace_b200_tracer_swap_dp(void* h) { ENTRY(tracer_swap_dp(*c)) }
int pace_b200_pt_to_potential_density_pt(void* h) { ENTRY(pt_to_potential_density_pt(*c)) }
int pace_b200_omega_from_w(void* h) { ENTRY(omega_from_w(*c)) }
int pace_b200_copy_levels(void* h, const char* src, const char* dst) { ENTRY(copy_levels(*c, src, dst)) }
int pace_b200_halo_plan_set(void* h, int plan_id, const void* ops, int n_ops) { ENTRY(halo_plan_set(*c, plan_id, ops, n_ops)) }
int pace_b200_halo_update(void* h, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk) {
  ENTRY(halo_update(*c, plan_id, x_names, y_names, n, nk))
}
int pace_b200_halo_pack(void* h, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                        void* message) {
  ENTRY(halo_pack(*c, plan_id, x_names, y_names, n, nk, (double*)message))
}
int pace_b200_halo_unpack(void* h, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                          const void* message) {
  ENTRY(halo_unpack(*c, plan_id, x_names, y_names, n, nk, (const double*)message))
}
int pace_b200_riem_solver3(void* h, double dt, int last_call) { ENTRY(riem_solver3(*c, dt, last_call)) }
int pace_b200_update_dz_d(void* h, double dt) { ENTRY(update_dz_d(*c, dt)) }
int pace_b200_zero_data(void* h, int first_timestep) { ENTRY(zero_data(*c, first_timestep)) }
int pace_b200_gz_from_zs_delz(void* h) { ENTRY(gz_from_zs_delz(*c)) }
int pace_b200_pem_from_delp(void* h) { ENTRY(pem_from_delp(*c)) }
int pace_b200_copy_full(void* h, const char* src, const char* dst) { ENTRY(copy_full(*c, src, dst)) }
int pace_b200_edge_pe(void* h) { ENTRY(edge_pe(*c)) }
int pace_b200_pk3_halo(void* h) { ENTRY(pk3_halo(*c)) }
int pace_b200_compute_geopotential(void* h) { ENTRY(compute_geopotential(*c)) }
int pace_b200_nh_p_grad(void* h, double dt) { ENTRY(nh_p_grad(*c, dt)) }
int pace_b200_ray_fast(void* h, double dt) { ENTRY(ray_fast(*c, dt)) }
int pace_b200_del2cubed(void* h, const char* q, double cd, int nmax) {
  Core* c = (Core*)h;
  c->err.clear();
  double* qp = c->F(q);
  if (!c->err.empty()) return 1;
  return finish(c, del2cubed(*c, qp, cd, nmax));
}
int pace_b200_apply_diffusive_heating(void* h, double delt_time_factor, int nk) {
  ENTRY(apply_diffusive_heating(*c, delt_time_factor, nk))
}
int pace_b200_d_sw(void* h, double dt) { ENTRY(d_sw(*c, dt)) }

int pace_b200_fvtp2d(void* h, const char* q, const char* crx, const char* cry, const char* xfx, const char* yfx,
                     const char* fx, const char* fy, const char* x_mass_flux, const char* y_mass_flux,
                     const char* mass, int hord, const char* nord_col, const char* damp, int nk) {
  Core* c = (Core*)h;
  c->err.clear();
  Fvtp2dArgs a;
  a.q = c->F(q); a.crx = c->F(crx); a.cry = c->F(cry); a.xfx = c->F(xfx); a.yfx = c->F(yfx);
  a.fx = c->F(fx); a.fy = c->F(fy);
  a.x_mass_flux = opt_field(c, x_mass_flux); a.y_mass_flux = opt_field(c, y_mass_flux); a.mass = opt_field(c, mass);
  a.hord = hord; a.nk = nk;
  if (nord_col && nord_col[0]) {
    a.nord_col = c->F(nord_col); a.damp = c->F(damp); a.nmax = c->col_max(nord_col);
  }
  if (!c->err.empty()) return 1;
  return finish(c, fvtp2d(*c, a));
}

int pace_b200_delnflux_nosg(void* h, const char* q, const char* fx2, const char* fy2, const char* damp,
                            const char* d2, const char* nord_col, int nk) {
  Core* c = (Core*)h;
  c->err.clear();
  DelnArgs d;
  d.q = c->F(q); d.fx = c->F(fx2); d.fy = c->F(fy2); d.damp = c->F(damp); d.d2 = opt_field(c, d2);
  d.nord_col = c->F(nord_col); d.nmax = c->col_max(nord_col); d.nk = nk; d.accumulate = false;
  if (!c->err.empty()) return 1;
  return finish(c, delnflux(*c, d));
}

int pace_b200_a2b_ord4(void* h, const char* qin, const char* qout, int k0, int nk, int replace) {
  Core* c = (Core*)h;
  c->err.clear();
  double *a = c->F(qin), *b = c->F(qout);
  if (!c->err.empty()) return 1;
  return finish(c, a2b_ord4(*c, a, b, k0, nk, replace != 0));
}

int pace_b200_divergence_damping(void* h, const char* vort_b, const char* delpc, const char* ke,
                                 const char* vort_a, double dt) {
  Core* c = (Core*)h;
  c->err.clear();
  double *a = c->F(vort_b), *b = c->F(delpc), *k = c->F(ke), *va = c->F(vort_a);
  if (!c->err.empty()) return 1;
  return finish(c, divergence_damping(*c, a, b, k, va, dt));
}

}  // extern "C"
