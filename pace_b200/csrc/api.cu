// C ABI of libpace_b200 (declared in include/pace_b200.h) and the Core plumbing behind it.
#include <cstdarg>
#include <cstring>

#include "../../include/pace_b200.h"
#include "core.hpp"

namespace pace {

static thread_local std::string g_err;

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

Core::~Core() {
  for (void* p : owned) cudaFree(p);
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
  g.is = HALO; g.ie = HALO + g.nx - 1; g.js = HALO; g.je = HALO + g.ny - 1;
  memset(g.edge, 0, sizeof g.edge);
  for (int t = 0; t < g.nt; ++t) g.edge[t] = edge_flags ? edge_flags[t] : 15;
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

#define ENTRY(call)          \
  Core* c = (Core*)h;        \
  c->err.clear();            \
  return finish(c, (call));

int pace_b200_riem_solver_c(void* h, double dt2) { ENTRY(riem_solver_c(*c, dt2)) }
int pace_b200_c_sw(void* h, double dt2) { ENTRY(c_sw(*c, dt2)) }
int pace_b200_update_dz_c(void* h, double dt2) { ENTRY(update_dz_c(*c, dt2)) }
int pace_b200_p_grad_c(void* h, double dt2) { ENTRY(p_grad_c(*c, dt2)) }

}  // extern "C"
