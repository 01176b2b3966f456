// Tiled phase kernels: one CTA owns a TXM x TYM patch of one level of one subdomain, stages the fields it reads —
// patch plus an apron of APX / APY cells — in shared memory with TMA (cp.async.bulk.tensor, one elected thread, an mbarrier counts
// the bytes), runs a chain of stencil phases on the shared-memory tiles with a CTA barrier between phases, and writes
// only final results to HBM.  This is how a halo-delimited phase of the acoustic substep (c_sw: 23 reference stencils)
// becomes one kernel whose DRAM traffic is its inputs (x apron overlap) plus its outputs.
//
// The patches partition the WHOLE array of a level (halo rows / columns included, since the reference's stencils write
// into the halo), so patch (0, 0) starts at array index (0, 0); tensor-map coordinates left of / below the array, and
// right of / above it, are filled with zeros by the TMA unit and never enter a result.
// The tensor maps cover this GPU's own field allocations only; no multicast.
//
// The same phase code runs in the host emulation (PACE_HOST_EMU: loops over the CTA's points, plain copies for the loads),
// which is how the fused kernels are checked against the reference fixtures on a machine without a GPU.
#pragma once
#include "core.hpp"
#include "corners.cuh"

#ifndef PACE_HOST_EMU
#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through the runtime, no libcuda link)
#endif

namespace pace {

// APX must be even and the patch width is kept even: a TMA box has to start on a 16-byte boundary along i (an odd fp64
// start coordinate raises an illegal-instruction fault; measured with tools/probe/tma_probe.cu), so the apron along i
// is 4 cells where the stencils need 3.
template <int TXM_, int TYM_, int APX_, int APY_, int NTX_, int NTY_>
struct TileCfg {
  static constexpr int TXM = TXM_, TYM = TYM_, APX = APX_, APY = APY_, NTX = NTX_, NTY = NTY_;
  static constexpr int W = TXM + 2 * APX, H = TYM + 2 * APY;
  static_assert(APX % 2 == 0 && TXM % 2 == 0, "TMA: box start coordinates along i must be even");
  static constexpr int NT = NTX * NTY;
  static constexpr int BYTES = W * H * 8;
  static_assert(W % 2 == 0, "TMA: the inner box extent must be a multiple of 16 bytes");
  static_assert(NTX >= W, "one thread column per tile column");
  static_assert(BYTES % 128 == 0, "tile arrays sit back to back and each is a 128-byte aligned TMA destination");
};

// One field as a tiled kernel sees it: the plain pointer (stores, host emulation) and its tensor map (TMA loads).
struct TileField {
  double* p = nullptr;
#ifndef PACE_HOST_EMU
  const CUtensorMap* map = nullptr;  // device copy is passed by value inside the kernel's argument struct
#endif
};

#ifndef PACE_HOST_EMU
// Tensor map over one registered 3-D field: dims (i, j, k, t) = (sj, ny+7, nz+1, nt), box (W, H, 1, 1).
// Cached per (pointer, box) in the Core; HostPipeline alternates two copies of the state, both stay cached.
const CUtensorMap* tile_tensor_map(Core& c, const double* field, int box_w, int box_h);
// cudaFuncAttributeMaxDynamicSharedMemorySize, set once per (kernel, device)
int func_attr_once(Core& c, const void* kernel, int smem_bytes);

namespace tma {
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
// box (W, H, 1, 1) at (i, j, k, t) -> dense shared-memory tile
__device__ __forceinline__ void load_tile(void* dst, const CUtensorMap* map, unsigned long long* bar, int i, int j, int k,
                                          int t) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(i), "r"(j), "r"(k), "r"(t), "r"(smem_u32(bar))
      : "memory");
}
}  // namespace tma
#endif

// What a phase body sees.  Coordinates: (i, j) global array indices, (li, lj) = (i - bi, j - bj) tile-local.
template <class C>
struct Tile {
  typedef double Arr[C::H][C::W];
  Geo g;
  int t, k;
  int bi, bj;              // global index of tile-local (0, 0)  (= first owned cell - apron)
  int oi0, oi1, oj0, oj1;  // owned cells [oi0, oi1) x [oj0, oj1)
  EdgeInfo e;
  bool corner_tile;        // the tile's box holds cube-corner halo cells (corner fills apply)
  bool near_edge;          // some cell of the box lies within 4 cells of a cube-tile edge (edge formulas may apply)
  long long base3;         // offset of (t, k, 0, 0) in 3-D fields
  int base2;               // offset of (t, 0, 0) in 2-D fields
#ifndef PACE_HOST_EMU
  int tx, ty;
#endif

  __device__ __forceinline__ long long o3(int i, int j) const { return base3 + (long long)j * g.sj + i; }
  __device__ __forceinline__ int o2(int i, int j) const { return base2 + j * g.sj + i; }

  // f(i, j, li, lj) for every point of the global rectangle [ia, ib] x [ja, jb] that lies inside the tile-local
  // window [wi0, wi1] x [wj0, wj1] (all inclusive); a CTA barrier follows unless sync is false.
  template <class F>
  __device__ __forceinline__ void each(int ia, int ib, int ja, int jb, int wi0, int wi1, int wj0, int wj1, F f,
                                       bool sync = true) const {
    const int la = max(ia - bi, wi0), lb = min(ib - bi, wi1), lc = max(ja - bj, wj0), ld = min(jb - bj, wj1);
#ifndef PACE_HOST_EMU
    const int li = tx;
    if (li >= la && li <= lb)
      for (int lj = lc + ty; lj <= ld; lj += C::NTY) f(bi + li, bj + lj, li, lj);
    if (sync) __syncthreads();
#else
    (void)sync;
    for (int lj = lc; lj <= ld; ++lj)
      for (int li = la; li <= lb; ++li) f(bi + li, bj + lj, li, lj);
#endif
  }
  // the same over the owned cells only (result stores)
  template <class F>
  __device__ __forceinline__ void each_owned(int ia, int ib, int ja, int jb, F f, bool sync = true) const {
    each(max(ia, oi0), min(ib, oi1 - 1), max(ja, oj0), min(jb, oj1 - 1), 0, C::W - 1, 0, C::H - 1, f, sync);
  }
};

// Patch widths: the array extent (nx + 7) is split into the smallest number of patches that fit TXM, of equal width.
struct TileGrid {
  int ntx, nty, tw, th;
};
template <class C>
inline TileGrid tile_grid(const Geo& g) {
  TileGrid tg;
  const int ni = g.nx + 2 * HALO + 1, nj = g.ny + 2 * HALO + 1;
  tg.ntx = (ni + C::TXM - 1) / C::TXM;
  tg.nty = (nj + C::TYM - 1) / C::TYM;
  tg.tw = ((ni + tg.ntx - 1) / tg.ntx + 1) & ~1;  // even: every patch starts at an even i
  tg.th = (nj + tg.nty - 1) / tg.nty;
  return tg;
}

template <class C>
__device__ __forceinline__ Tile<C> make_tile(const Geo& g, const TileGrid& tg, int bx, int by, int k, int t) {
  Tile<C> T;
  T.g = g; T.t = t; T.k = k;
  const int ni = g.nx + 2 * HALO + 1, nj = g.ny + 2 * HALO + 1;
  T.oi0 = bx * tg.tw; T.oi1 = min(T.oi0 + tg.tw, ni);
  T.oj0 = by * tg.th; T.oj1 = min(T.oj0 + tg.th, nj);
  T.bi = T.oi0 - C::APX; T.bj = T.oj0 - C::APY;
  const unsigned ed = edge_of(g, t);
  T.e = EdgeInfo{g.is, g.ie, g.js, g.je, (bool)(ed & EDGE_W), (bool)(ed & EDGE_E), (bool)(ed & EDGE_S), (bool)(ed & EDGE_N)};
  const bool nw = T.e.W && T.bi < g.is, ne = T.e.E && T.bi + C::W - 1 > g.ie;
  const bool ns = T.e.S && T.bj < g.js, nn = T.e.N && T.bj + C::H - 1 > g.je;
  T.corner_tile = (nw || ne) && (ns || nn);
  T.near_edge = (T.e.W && T.bi <= g.is + 4) || (T.e.E && T.bi + C::W - 1 >= g.ie - 4) ||
                (T.e.S && T.bj <= g.js + 4) || (T.e.N && T.bj + C::H - 1 >= g.je - 4);
  T.base3 = (long long)t * g.st3 + (long long)k * g.sk;
  T.base2 = t * g.sk;
  return T;
}

// Host emulation of a tile load: the box, zeros outside the array (what the TMA unit does).
template <class C>
inline void host_load_tile(const Tile<C>& T, double (*dst)[C::W], const double* field) {
  const int ni = T.g.sj, nj = T.g.ny + 2 * HALO + 1;
  for (int lj = 0; lj < C::H; ++lj)
    for (int li = 0; li < C::W; ++li) {
      const int i = T.bi + li, j = T.bj + lj;
      dst[lj][li] = (i >= 0 && i < ni && j >= 0 && j < nj) ? field[T.o3(i, j)] : 0.0;
    }
}

}  // namespace pace
