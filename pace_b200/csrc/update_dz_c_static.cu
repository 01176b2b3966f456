// pace_b200_update_dz_c_static: the C-grid height advection (update_dz_c, dz_pgrad.cu) for callers that must not
// allocate device memory.
//
// Who that is: a host model that hands the whole GPU to its own memory pool when it starts (a coupled-model driver
// with a pre-sized arena, a framework's caching allocator configured to take everything), so that a cudaMalloc made
// later by a library either fails or eats into memory the pool believes it owns; and callers inside a time-critical
// loop, where the device-wide synchronisation that cudaMalloc implies is not acceptable.  pace_b200_update_dz_c gets
// its one work array (the pre-update copy of gz) from cudaMalloc on first use.  The entry point below keeps that copy
// in a fixed array that is part of the library image instead, and never calls cudaMalloc, cudaFree or any other
// allocating routine.
//
// What to know before using it:
//  * The array holds PACE_B200_STATIC_SCRATCH_DOUBLES doubles (16 MiB).  Every GPU on which this library is loaded
//    carries those 16 MiB from load until the process exits, used or not.
//  * A handle whose fields are larger than that (subdomains x (npz+1) x (ny+7) x row stride doubles) is refused with
//    an error message that names both numbers; nothing is launched in that case.
//  * The first call on each device looks up the device address of the array (cudaGetSymbolAddress).  Make that
//    first call during start-up, before any stream capture is begun on that device.
//  * All handles and all calls on a device share the one array, and it keeps whatever the previous call left in
//    it.  Nothing in it has to start at zero: each call first overwrites every element it will later read
//    (all (npz+1) levels of every subdomain, halo included), so no clearing is done.  Two calls whose kernels can
//    run at the same time on one GPU (different streams, different handles, different host threads) race on it;
//    order them on one stream or with events.
//  * Inputs and outputs are the caller's registered fields, exactly as for pace_b200_update_dz_c, and the results
//    are the same bit for bit.
#include "core.hpp"
#include "corners.cuh"
#include "update_dz_c_body.cuh"

#include "../../include/pace_b200.h"  // PACE_B200_STATIC_SCRATCH_DOUBLES (2097152 doubles = 16 MiB)

namespace pace {

#ifndef PACE_HOST_EMU

namespace {

// In device code this name is the array itself.  In host code it is only a stand-in that identifies the symbol to the
// runtime: its host-side address is NOT a device pointer and must never be passed to a kernel.
static __device__ __align__(256) double g_dzc_static_scratch[PACE_B200_STATIC_SCRATCH_DOUBLES];

constexpr int MAX_DEVICES = 64;
double* g_scratch_addr[MAX_DEVICES] = {nullptr};  // device address of the array, per device ordinal

// Device address of the array on the current device; looked up once per device.
double* static_scratch(Core& c) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
    c.err = std::string("update_dz_c_static: cannot identify the current device: ") + cudaGetErrorString(e);
    return nullptr;
  }
  if (!g_scratch_addr[dev]) {
    double* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_dzc_static_scratch);
    if (e != cudaSuccess || !ptr) {
      c.err = std::string("update_dz_c_static: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e);
      return nullptr;
    }
    g_scratch_addr[dev] = ptr;
  }
  return g_scratch_addr[dev];
}

}  // namespace

// Same three kernels as update_dz_c (update_dz_c_body.cuh); only the home of gz0 differs.
static int update_dz_c_static(Core& c, double dt) {
  const size_t need = c.field_elems();
  if (need > (size_t)PACE_B200_STATIC_SCRATCH_DOUBLES) {
    c.err = "update_dz_c_static: this handle needs " + std::to_string(need) +
            " doubles of work space, the built-in array holds " + std::to_string(PACE_B200_STATIC_SCRATCH_DOUBLES) +
            "; use pace_b200_update_dz_c (which allocates) or smaller / fewer subdomains per handle";
    return 1;
  }
  double* gz0 = static_scratch(c);
  if (!gz0) return 1;
  if (update_dz_c_with(c, dt, gz0) != 0) return 1;
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    c.err = std::string("update_dz_c_static: CUDA launch error: ") + cudaGetErrorString(e);
    return 1;
  }
  return 0;
}

#else  // host emulation has no device image to keep an array in

static int update_dz_c_static(Core& c, double) {
  c.err = "update_dz_c_static exists only in the CUDA library";
  return 1;
}

#endif

}  // namespace pace

extern "C" int pace_b200_update_dz_c_static(void* handle, double dt2) {
  pace::Core* c = (pace::Core*)handle;
  c->err.clear();
  return pace::update_dz_c_static(*c, dt2);
}
