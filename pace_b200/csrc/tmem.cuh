// Tensor memory (TMEM, 512 columns x 128 lanes x 32 bit per SM) as per-thread scratch for the column solvers.
//
// The column kernels of this path (Riemann solvers, the interface spline of update_dz_d) are sequential in k and need
// one value per level to survive between a forward and a backward sweep.  In shared memory that array caps the SM at
// 10 warps; tensor memory is otherwise idle here (there is no tcgen05.mma on this path).  With the 32x32b access shape
// lane l of warp w of a CTA owns TMEM lane 32 (w % 4) + l, so every thread has a private row of 32-bit words and an
// fp64 level is two adjacent columns.  A CTA of four warps allocates 128 columns = 64 levels; four such CTAs fill the
// 512 columns: 16 warps per SM.  Levels beyond the first 64 live in shared memory.
#pragma once
#include <cstdint>

namespace pace {

#ifndef PACE_HOST_EMU

constexpr int TMEM_LEVELS = 64;
constexpr int TMEM_COLS = 2 * TMEM_LEVELS;   // a power of two >= 32, as tcgen05.alloc requires

// One fp64 per level and thread.  TM = false: all levels in shared memory (sm[k * 32], lane-interleaved).
template <bool TM>
struct YAcc {
  double* sm;        // this lane's first shared-memory level (stride 32 doubles)
  uint32_t taddr;    // TMEM address of this warp's lanes, first allocated column
  __device__ __forceinline__ double ld(int k) const {
    if (TM && k < TMEM_LEVELS) {
      uint32_t lo, hi;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(taddr + 2 * k));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      return __hiloint2double((int)hi, (int)lo);
    }
    return sm[(TM ? k - TMEM_LEVELS : k) * 32];
  }
  __device__ __forceinline__ void st(int k, double v) const {
    if (TM && k < TMEM_LEVELS) {
      asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};"
                   :: "r"(taddr + 2 * k), "r"((uint32_t)__double2loint(v)), "r"((uint32_t)__double2hiint(v)) : "memory");
    } else {
      sm[(TM ? k - TMEM_LEVELS : k) * 32] = v;
    }
  }
  // all earlier st() of this thread have landed (needed before a later ld() of the same level)
  __device__ __forceinline__ void fence_st() const {
    if (TM) asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
};

// Called by every thread of a CTA of four warps: warp 0 allocates TMEM_COLS columns; returns the base address.
// `slot` is a __shared__ word.  Contains a CTA barrier.
__device__ __forceinline__ uint32_t tmem_alloc_cta(uint32_t* slot) {
  if ((threadIdx.x >> 5) == 0) {
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(slot);
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(dst), "r"((uint32_t)TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *slot;
}
// Called by every thread once no warp of the CTA touches its tensor memory any more.  Contains a CTA barrier.
__device__ __forceinline__ void tmem_free_cta(uint32_t base) {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if ((threadIdx.x >> 5) == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(base), "r"((uint32_t)TMEM_COLS) : "memory");
  }
}
// TMEM address of the calling warp's 32 lanes
__device__ __forceinline__ uint32_t tmem_warp_addr(uint32_t base) { return base + ((uint32_t)((threadIdx.x >> 5) * 32) << 16); }

#endif  // PACE_HOST_EMU

}  // namespace pace
