// pace_b200 — shared device/host helpers for the FV3 dycore kernels (sm_100a).
//
// Memory layout (all 3-D fields of all subdomains resident on this GPU):
//     f[t][k][j][i]   i fastest, row stride sj (>= nx+7, padded), plane stride sk = sj*(ny+7),
//                     subdomain ("tile") stride st3 = sk*(nz+1).
// Index convention equals the reference's storage indices (NDSL sizer.py:131-154): 3-cell halo,
// compute cells i in [3, 3+nx), interface dims one more.  2-D fields are f[t][j][i] (stride sk).
//
// Parity rule: every kernel keeps the reference's per-element expression tree and evaluation
// order (GT4Py numpy semantics); the library is compiled with -fmad=false so nothing is contracted.
#pragma once
#ifdef PACE_HOST_EMU
// Test-only host emulation (tests/hostemu.py): the SAME kernel source compiled by g++ with the launch
// helpers turned into loops, so kernel logic can be debugged against the golden fixtures on a machine
// without a GPU.  Never built or loaded by the product package.
#include "host_emu.hpp"
#else
#include <cuda_runtime.h>
#endif
#include <cmath>
#include <cstdint>
#include <cstdio>

namespace pace {

constexpr int MAX_TILES = 32;
constexpr int HALO = 3;

enum EdgeBits : unsigned char { EDGE_W = 1, EDGE_E = 2, EDGE_S = 4, EDGE_N = 8 };

// GFS constant set (reference NDSL/ndsl/constants.py:91-153, PACE_CONSTANTS default)
constexpr double GRAV = 9.80665;
constexpr double RDGAS = 287.05;
constexpr double RVGAS = 461.50;
constexpr double CP_AIR = 1004.6;
constexpr double KAPPA = RDGAS / CP_AIR;
constexpr double CV_AIR = CP_AIR - RDGAS;
constexpr double RDG = -RDGAS / GRAV;
constexpr double K1K = RDGAS / CV_AIR;
constexpr double ZVIR = RVGAS / RDGAS - 1;
constexpr double CV_VAP = 3.0 * RVGAS;
constexpr double CP_VAP = 4.0 * RVGAS;
constexpr double C_ICE = 1972.0;
constexpr double C_LIQ = 4.1855e3;
constexpr double TICE = 273.16;
constexpr double DZ_MIN = 2.0;
constexpr double HUGE_R = 1.0e40;

struct Geo {
  int nx, ny, nz, nt;   // compute cells per subdomain, number of subdomains on this GPU
  int sj, sk;           // row / plane strides (elements)
  long long st3;        // subdomain stride of 3-D fields
  int is, ie, js, je;   // first/last compute cell (3, nx+2, 3, ny+2)
  // EdgeBits per subdomain (does it touch the tile edge?), 4 bits each, subdomains 0-15 / 16-31.  Packed words and not
  // an array: a kernel parameter array indexed by the subdomain makes the compiler copy the whole struct to local
  // memory in every thread.
  unsigned long long edge_bits[2];
};
__host__ __device__ __forceinline__ unsigned edge_of(const Geo& g, int t) {
  return (unsigned)((t < 16 ? g.edge_bits[0] : g.edge_bits[1]) >> ((t & 15) * 4)) & 15u;
}

// One grid point handed to a stencil body.
// Linear offsets into 3-D fields: 32-bit on the device (one IMAD.WIDE per access instead of 64-bit index chains;
// pace_b200_create refuses configurations with 2^31 or more points per field), 64-bit in the host emulation.
#ifdef PACE_HOST_EMU
typedef long long off3_t;
#else
typedef int off3_t;
#endif

struct Pt {
  off3_t o;      // linear offset of (t,k,j,i) in 3-D fields
  int o2;        // linear offset of (t,j,i) in 2-D fields
  int i, j, k, t;
  int sj, sk;
  bool W, E, S, N;  // subdomain touches west/east/south/north tile edge
};

__device__ __forceinline__ Pt make_pt(const Geo& g, int t, int i, int j, int k) {
  Pt p;
  p.i = i; p.j = j; p.k = k; p.t = t;
  p.sj = g.sj; p.sk = g.sk;
  p.o2 = t * g.sk + j * g.sj + i;
  p.o = (off3_t)((off3_t)t * (off3_t)g.st3 + (off3_t)k * g.sk + j * g.sj + i);
  const unsigned e = edge_of(g, t);
  p.W = e & EDGE_W; p.E = e & EDGE_E; p.S = e & EDGE_S; p.N = e & EDGE_N;
  return p;
}

// field access relative to the current point `p`
#define A(f, di, dj, dk) (f)[p.o + (di) + (dj) * p.sj + (off3_t)(dk) * p.sk]
#define A0(f) (f)[p.o]
#define A2(f, di, dj) (f)[p.o2 + (di) + (dj) * p.sj]
#define A20(f) (f)[p.o2]

struct Box {
  int i0, j0, k0, ni, nj, nk;
  unsigned magic = 0;  // set by launch_ijk: floor(2^32 / ni) + 1 when n / ni == umulhi(n, magic) for every point n of a level
};

// ---- launch accounting (api.cu): every kernel launch is counted; launches given a label are additionally
// bracketed by CUDA events on their stream while profiling is enabled (pace_b200_profile_*), which is how
// bench.py measures the dominant kernel live inside its timed region.
long long& launch_counter();
bool prof_enabled();
void prof_begin(cudaStream_t s, const char* label);
void prof_end(cudaStream_t s);
struct LaunchScope {
  cudaStream_t s;
  bool on;
  LaunchScope(cudaStream_t s_, const char* label) : s(s_), on(label != nullptr && prof_enabled()) {
    ++launch_counter();
    if (on) prof_begin(s, label);
  }
  ~LaunchScope() {
    if (on) prof_end(s);
  }
};

#ifndef PACE_HOST_EMU
// The points of one level are numbered row by row and dealt to the threads in that order: no idle lanes at the row
// ends (box widths such as 199 would leave 11 % of a 32-wide block column idle).
// Grid: x = blocks of 256 points of one level, y = level, z = subdomain (no division for k / t); the row of a point is
// a multiply-high by the launch's magic number (launch_ijk).  The index preamble is ~12 instructions: for the many
// light stencils it used to be more than half of the executed instructions.
template <class F>
__global__ void __launch_bounds__(256) k_ijk(Geo g, Box b, F f) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= b.ni * b.nj) return;
  const int j = b.magic ? (int)__umulhi((unsigned)n, b.magic) : n / b.ni, i = n - j * b.ni;
  const int k = blockIdx.y, t = blockIdx.z;
  Pt p = make_pt(g, t, b.i0 + i, b.j0 + j, b.k0 + k);
  f(p);
}

// the same kernel compiled for at least MINB resident CTAs per SM (stencils with long per-point code)
template <int MINB, class F>
__global__ void __launch_bounds__(256, MINB) k_ijk_occ(Geo g, Box b, F f) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= b.ni * b.nj) return;
  const int j = b.magic ? (int)__umulhi((unsigned)n, b.magic) : n / b.ni, i = n - j * b.ni;
  const int k = blockIdx.y, t = blockIdx.z;
  Pt p = make_pt(g, t, b.i0 + i, b.j0 + j, b.k0 + k);
  f(p);
}

// column kernels: one thread per (t,j,i); the body loops over k itself (p.k = b.k0)
template <class F>
__global__ void __launch_bounds__(128) k_ij(Geo g, Box b, F f) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= b.ni || j >= b.nj) return;
  int t = blockIdx.z;
  Pt p = make_pt(g, t, b.i0 + i, b.j0 + j, b.k0);
  f(p);
}

// column kernel compiled for at least MINB resident CTAs per SM (register cap 65536 / (128 MINB)): the long column
// kernels are bound by the latency of their thread-local arrays and scale with the number of resident warps
template <int MINB, class F>
__global__ void __launch_bounds__(128, MINB) k_ij_occ(Geo g, Box b, F f) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int j = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= b.ni || j >= b.nj) return;
  int t = blockIdx.z;
  Pt p = make_pt(g, t, b.i0 + i, b.j0 + j, b.k0);
  f(p);
}
template <int MINB, class F>
inline void launch_ij_occ(cudaStream_t s, const Geo& g, Box b, F f, const char* label = nullptr) {
  if (b.ni <= 0 || b.nj <= 0) return;
  LaunchScope scope(s, label);
  dim3 blk(32, 4, 1);
  dim3 grd((b.ni + 31) / 32, (b.nj + 3) / 4, g.nt);
  k_ij_occ<MINB><<<grd, blk, 0, s>>>(g, b, f);
}

template <class F>
inline void launch_ijk(cudaStream_t s, const Geo& g, Box b, F f, const char* label = nullptr) {
  if (b.ni <= 0 || b.nj <= 0 || b.nk <= 0) return;
  LaunchScope scope(s, label);
  dim3 blk(256, 1, 1);
  dim3 grd((b.ni * b.nj + 255) / 256, b.nk, g.nt);
  // row index by multiplication: exact while n * ni < 2^32 for all n < ni * nj (n * (magic * ni - 2^32) < 2^32)
  if (b.ni > 1 && (unsigned long long)b.ni * b.nj * b.ni < (1ull << 32)) b.magic = (unsigned)((1ull << 32) / (unsigned)b.ni) + 1u;
  k_ijk<<<grd, blk, 0, s>>>(g, b, f);
}

// pad_smem: dynamic shared memory the kernel does not use; it only lowers the number of resident CTAs per SM for
// column kernels whose thread-local arrays should stay within the L2
template <int MINB, class F>
inline void launch_ijk_occ(cudaStream_t s, const Geo& g, Box b, F f, const char* label = nullptr) {
  if (b.ni <= 0 || b.nj <= 0 || b.nk <= 0) return;
  LaunchScope scope(s, label);
  dim3 blk(256, 1, 1);
  dim3 grd((b.ni * b.nj + 255) / 256, b.nk, g.nt);
  if (b.ni > 1 && (unsigned long long)b.ni * b.nj * b.ni < (1ull << 32)) b.magic = (unsigned)((1ull << 32) / (unsigned)b.ni) + 1u;
  k_ijk_occ<MINB><<<grd, blk, 0, s>>>(g, b, f);
}

template <class F>
inline void launch_ij(cudaStream_t s, const Geo& g, Box b, F f, const char* label = nullptr, int pad_smem = 0) {
  if (b.ni <= 0 || b.nj <= 0) return;
  LaunchScope scope(s, label);
  dim3 blk(32, 4, 1);
  dim3 grd((b.ni + 31) / 32, (b.nj + 3) / 4, g.nt);
  if (pad_smem > 48 * 1024) cudaFuncSetAttribute(k_ij<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, pad_smem);
  k_ij<<<grd, blk, pad_smem, s>>>(g, b, f);
}

#else   // PACE_HOST_EMU: plain loops, same iteration space
template <class F>
inline void launch_ijk(cudaStream_t, const Geo& g, Box b, F f, const char* = nullptr) {
  ++launch_counter();
#ifdef PACE_HOST_EMU_OMP
#pragma omp parallel for collapse(2) schedule(static)
#endif
  for (int t = 0; t < g.nt; ++t)
    for (int k = 0; k < b.nk; ++k)
      for (int j = 0; j < b.nj; ++j)
        for (int i = 0; i < b.ni; ++i) f(make_pt(g, t, b.i0 + i, b.j0 + j, b.k0 + k));
}
template <int MINB, class F>
inline void launch_ij_occ(cudaStream_t s, const Geo& g, Box b, F f, const char* label = nullptr) {
  launch_ij(s, g, b, f, label);
}
template <int MINB, class F>
inline void launch_ijk_occ(cudaStream_t s, const Geo& g, Box b, F f, const char* label = nullptr) {
  launch_ijk(s, g, b, f, label);
}
template <class F>
inline void launch_ij(cudaStream_t, const Geo& g, Box b, F f, const char* = nullptr) {
  ++launch_counter();
#ifdef PACE_HOST_EMU_OMP
#pragma omp parallel for collapse(2) schedule(static)
#endif
  for (int t = 0; t < g.nt; ++t)
    for (int j = 0; j < b.nj; ++j)
      for (int i = 0; i < b.ni; ++i) f(make_pt(g, t, b.i0 + i, b.j0 + j, b.k0));
}
#endif

// ---- boxes in the reference's vocabulary (GridIndexing.get_origin_domain, stencil.py:768-788) -----
// compute domain of a staggering, grown by halos; xi/yi/zi = interface dims (+1 extent)
inline Box box_dims(const Geo& g, bool xi, bool yi, bool zi, int hx = 0, int hy = 0) {
  Box b;
  b.i0 = HALO - hx; b.j0 = HALO - hy; b.k0 = 0;
  b.ni = g.nx + (xi ? 1 : 0) + 2 * hx;
  b.nj = g.ny + (yi ? 1 : 0) + 2 * hy;
  b.nk = g.nz + (zi ? 1 : 0);
  return b;
}
inline Box box_full(const Geo& g, int addk = 0) {  // origin_full / domain_full(add=(0,0,addk))
  return Box{0, 0, 0, g.nx + 2 * HALO, g.ny + 2 * HALO, g.nz + addk};
}
inline Box box_compute(const Geo& g, int oi = 0, int oj = 0, int ai = 0, int aj = 0, int ak = 0) {
  return Box{HALO + oi, HALO + oj, 0, g.nx + ai, g.ny + aj, g.nz + ak};
}

// ---- math: single point of definition for the transcendentals (parity) -------------
}  // namespace pace
#include "libm_port.cuh"
namespace pace {
__device__ __forceinline__ double fv_exp(double x) { return libm::exp_(x); }
__device__ __forceinline__ double fv_log(double x) { return libm::log_(x); }
__device__ __forceinline__ double fv_pow(double x, double y) { return libm::pow_(x, y); }
__device__ __forceinline__ double fv_sqrt(double x) { return sqrt(x); }
__device__ __forceinline__ double fv_abs(double x) { return fabs(x); }
// gtscript min/max/sign as the numpy backend evaluates them (np.minimum/np.maximum semantics on
// non-NaN data; sign(a,b) = |a| if b>=0 else -|a|, basic_operations.py:32-45)
__device__ __forceinline__ double fv_min(double a, double b) { return a < b ? a : b; }
__device__ __forceinline__ double fv_max(double a, double b) { return a > b ? a : b; }
// n / d.  A zero numerator over a positive denominator is n itself (same bits, signed zero included); answering that
// without dividing keeps fields that are zero over most of the domain (condensates, q_con) off the out-of-line
// special-operand path of the fp64 division, which a warp enters as soon as one lane has a zero numerator.
template <class T>
__device__ __forceinline__ T fv_div(T n, T d) {
#ifndef PACE_HOST_EMU
  if (n == T(0.0) && d > T(0.0)) return n;
#endif
  return n / d;
}

#define CUDA_CHECK_RET(call)                                                              \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      ::pace::set_error("%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return 1;                                                                           \
    }                                                                                     \
  } while (0)

void set_error(const char* fmt, ...);
// The shortcuts for data that is zero (rows / levels / columns / tiles of zeros: fvtp2d_march.cu, tracer.cu, remap.cu,
// delnflux_fused.cu) are exact and on by default; PACE_B200_ZERO_PATHS=0 turns them off so that the cost of the dense
// computation can be measured on the same inputs (bench.py reports both).
bool zero_paths_enabled();

}  // namespace pace
