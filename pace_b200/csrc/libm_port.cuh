// exp / log / pow (double) bit-identical to the host libm the parity oracle uses.
//
// The oracle (reference numpy backend with NumPy's SIMD transcendentals pinned off) evaluates exp, log and
// power through glibc 2.39's __exp_fma / __log_fma / __pow_fma on an FMA-capable x86-64 host.  CUDA's libdevice
// versions differ from those in the last bit for a few percent of arguments, and the FV3 Riemann solves amplify
// that (p' = p_full - p_hydrostatic cancels ~4 digits), so these routines restate the exact instruction sequence
// of the glibc build — including where GCC contracted a*b+c into FMA (read off the disassembly; each statement
// notes the contraction) — over the same tables (libm_tables.inc, extracted by tools/gen_libm_tables.py).
// Algorithm: S. Nagy's table-driven exp/log/pow (glibc sysdeps/ieee754/dbl-64/e_exp.c, e_log.c, e_pow.c).
//
// Only the main paths are restated (normal positive arguments, |x| < 512 for exp, results far from
// overflow/underflow) — the only ones the dycore can reach with finite fields; anything else falls back to the
// CUDA / host library routine.
#pragma once
#include "common.cuh"

namespace pace {
namespace libm {

#ifdef PACE_HOST_EMU
#define PACE_LIBM_TABLE static const
#else
#define PACE_LIBM_TABLE static __device__ const
#endif
#include "libm_tables.inc"

__device__ __forceinline__ double as_d(unsigned long long b) {
#ifdef PACE_HOST_EMU
  double d;
  memcpy(&d, &b, 8);
  return d;
#else
  return __longlong_as_double((long long)b);
#endif
}
__device__ __forceinline__ unsigned long long as_u(double d) {
#ifdef PACE_HOST_EMU
  unsigned long long b;
  memcpy(&b, &d, 8);
  return b;
#else
  return (unsigned long long)__double_as_longlong(d);
#endif
}
__device__ __forceinline__ double fma_(double a, double b, double c) { return fma(a, b, c); }

// exp core shared by exp() and pow(): x = xhi (+ xtail), result scaled by 2^(sign_bias bit) (pow only)
template <bool CHECK = true>
__device__ __forceinline__ bool exp_core(double x, double xtail, bool has_tail, double& out) {
  if (CHECK) {
    const unsigned abstop = (unsigned)(as_u(x) >> 52) & 0x7ff;
    if (abstop - 0x3c9u > 0x3eu) return false;  // |x| < 2^-54 or |x| >= 512: library path
  }
  const double InvLn2N = as_d(EXP_HEAD[0]), Shift = as_d(EXP_HEAD[1]), NegLn2hiN = as_d(EXP_HEAD[2]),
               NegLn2loN = as_d(EXP_HEAD[3]), C2 = as_d(EXP_HEAD[4]), C3 = as_d(EXP_HEAD[5]), C4 = as_d(EXP_HEAD[6]),
               C5 = as_d(EXP_HEAD[7]);
  double kd = fma_(x, InvLn2N, Shift);          // z + Shift, contracted
  const unsigned long long ki = as_u(kd);
  kd = kd - Shift;
  double r = fma_(kd, NegLn2loN, fma_(kd, NegLn2hiN, x));  // x + kd*NegLn2hiN + kd*NegLn2loN, both contracted
  if (has_tail) r = xtail + r;
  const unsigned idx = 2 * (unsigned)(ki & 127);
  const unsigned long long top = ki << 45;
  const double tail = as_d(EXP_TAB[idx]);
  const unsigned long long sbits = EXP_TAB[idx + 1] + top;
  const double r2 = r * r;
  // tmp = tail + r + r2*(C2 + r*C3) + r2*r2*(C4 + r*C5)
  const double tmp = fma_(r2 * r2, fma_(r, C5, C4), fma_(fma_(r, C3, C2), r2, tail + r));
  const double scale = as_d(sbits);
  out = fma_(scale, tmp, scale);                 // scale + scale*tmp, contracted
  return true;
}

__device__ __forceinline__ double exp_(double x) {
  double y;
  if (exp_core(x, 0.0, false, y)) return y;
  return ::exp(x);
}

__device__ __forceinline__ double log_(double x) {
  const unsigned long long ix = as_u(x);
  const unsigned top = (unsigned)(ix >> 48);
  if (top - 0x0010u >= 0x7ff0u - 0x0010u) return ::log(x);  // zero, subnormal, negative, inf, nan
  const double* H = nullptr;
  (void)H;
  if (ix - 0x3fee000000000000ULL < 0x3090000000000ULL) {
    // 1 - 0x1p-4 <= x < 1 + 0x1.09p-4: polynomial in r = x - 1 with a double-double head
    if (ix == 0x3ff0000000000000ULL) return 0.0;
    const double B0 = as_d(LOG_HEAD[7]), B1 = as_d(LOG_HEAD[8]), B2 = as_d(LOG_HEAD[9]), B3 = as_d(LOG_HEAD[10]),
                 B4 = as_d(LOG_HEAD[11]), B5 = as_d(LOG_HEAD[12]), B6 = as_d(LOG_HEAD[13]), B7 = as_d(LOG_HEAD[14]),
                 B8 = as_d(LOG_HEAD[15]), B9 = as_d(LOG_HEAD[16]), B10 = as_d(LOG_HEAD[17]);
    const double r = x - 1.0;
    const double r2 = r * r;
    const double r3 = r * r2;
    const double p1 = fma_(r2, B3, fma_(r, B2, B1));
    const double p2 = fma_(r2, B6, fma_(r, B5, B4));
    const double p3 = fma_(r3, B10, fma_(r2, B9, fma_(r, B8, B7)));
    const double poly = fma_(fma_(p3, r3, p2), r3, p1);
    // rhi = r + w - w with w = r*2^27, the sum and the difference both contracted
    const double t = fma_(r, 134217728.0, r);
    const double rhi = fma_(-134217728.0, r, t);
    const double rlo = r - rhi;
    const double rhi2 = rhi * rhi;
    const double hi = fma_(rhi2, B0, r);
    double lo = fma_(rhi2, B0, r - hi);
    lo = fma_(B0 * rlo, rhi + r, lo);
    const double y = fma_(poly, r3, lo);
    return y + hi;
  }
  const double Ln2hi = as_d(LOG_HEAD[0]), Ln2lo = as_d(LOG_HEAD[1]), A0 = as_d(LOG_HEAD[2]), A1 = as_d(LOG_HEAD[3]),
               A2 = as_d(LOG_HEAD[4]), A3 = as_d(LOG_HEAD[5]), A4 = as_d(LOG_HEAD[6]);
  const unsigned long long tmp = ix - 0x3fe6000000000000ULL;
  const unsigned i = (unsigned)(tmp >> 45) & 127;
  const int k = (int)((long long)tmp >> 52);
  const unsigned long long iz = ix - (tmp & 0xfff0000000000000ULL);
  const double invc = as_d(LOG_TAB[2 * i]), logc = as_d(LOG_TAB[2 * i + 1]);
  const double z = as_d(iz);
  const double kd = (double)k;
  const double r = fma_(z, invc, -1.0);
  const double w = fma_(kd, Ln2hi, logc);
  const double hi = r + w;
  const double lo = fma_(kd, Ln2lo, (w - hi) + r);
  const double r2 = r * r;
  const double r3 = r * r2;
  const double q = fma_(fma_(r, A4, A3), r2, fma_(r, A2, A1));
  const double y = fma_(r3, q, fma_(r2, A0, lo));
  return y + hi;
}

// ---- N independent evaluations at once ----------------------------------------------------------------------
// exp_ / log_ return early on their rare paths, which keeps the compiler from interleaving several calls.  These
// evaluate the main path of N arguments as straight-line code (N independent dependency chains for the scheduler)
// when all of them are main-path arguments, and fall back to the scalar routines otherwise.  Same operations per
// argument, hence the same bits.
__device__ __forceinline__ bool exp_is_main(double x) {
  const unsigned abstop = (unsigned)(as_u(x) >> 52) & 0x7ff;
  return abstop - 0x3c9u <= 0x3eu;
}
__device__ __forceinline__ bool log_is_main(double x) {
  const unsigned long long ix = as_u(x);
  const unsigned top = (unsigned)(ix >> 48);
  return top - 0x0010u < 0x7ff0u - 0x0010u && !(ix - 0x3fee000000000000ULL < 0x3090000000000ULL);
}
__device__ __forceinline__ double log_main(double x) {  // the table-driven path of log_ (x normal, not near 1)
  const unsigned long long ix = as_u(x);
  const double Ln2hi = as_d(LOG_HEAD[0]), Ln2lo = as_d(LOG_HEAD[1]), A0 = as_d(LOG_HEAD[2]), A1 = as_d(LOG_HEAD[3]),
               A2 = as_d(LOG_HEAD[4]), A3 = as_d(LOG_HEAD[5]), A4 = as_d(LOG_HEAD[6]);
  const unsigned long long tmp = ix - 0x3fe6000000000000ULL;
  const unsigned i = (unsigned)(tmp >> 45) & 127;
  const int k = (int)((long long)tmp >> 52);
  const unsigned long long iz = ix - (tmp & 0xfff0000000000000ULL);
  const double invc = as_d(LOG_TAB[2 * i]), logc = as_d(LOG_TAB[2 * i + 1]);
  const double z = as_d(iz);
  const double kd = (double)k;
  const double r = fma_(z, invc, -1.0);
  const double w = fma_(kd, Ln2hi, logc);
  const double hi = r + w;
  const double lo = fma_(kd, Ln2lo, (w - hi) + r);
  const double r2 = r * r;
  const double r3 = r * r2;
  const double q = fma_(fma_(r, A4, A3), r2, fma_(r, A2, A1));
  const double y = fma_(r3, q, fma_(r2, A0, lo));
  return y + hi;
}
__device__ __forceinline__ bool log_is_near1(double x) {  // 1 - 0x1p-4 <= x < 1 + 0x1.09p-4 and x != 1
  const unsigned long long ix = as_u(x);
  return ix - 0x3fee000000000000ULL < 0x3090000000000ULL && ix != 0x3ff0000000000000ULL;
}
__device__ __forceinline__ double log_near1(double x) {  // the polynomial path of log_ around 1
  const double B0 = as_d(LOG_HEAD[7]), B1 = as_d(LOG_HEAD[8]), B2 = as_d(LOG_HEAD[9]), B3 = as_d(LOG_HEAD[10]),
               B4 = as_d(LOG_HEAD[11]), B5 = as_d(LOG_HEAD[12]), B6 = as_d(LOG_HEAD[13]), B7 = as_d(LOG_HEAD[14]),
               B8 = as_d(LOG_HEAD[15]), B9 = as_d(LOG_HEAD[16]), B10 = as_d(LOG_HEAD[17]);
  const double r = x - 1.0;
  const double r2 = r * r;
  const double r3 = r * r2;
  const double p1 = fma_(r2, B3, fma_(r, B2, B1));
  const double p2 = fma_(r2, B6, fma_(r, B5, B4));
  const double p3 = fma_(r3, B10, fma_(r2, B9, fma_(r, B8, B7)));
  const double poly = fma_(fma_(p3, r3, p2), r3, p1);
  const double t = fma_(r, 134217728.0, r);
  const double rhi = fma_(-134217728.0, r, t);
  const double rlo = r - rhi;
  const double rhi2 = rhi * rhi;
  const double hi = fma_(rhi2, B0, r);
  double lo = fma_(rhi2, B0, r - hi);
  lo = fma_(B0 * rlo, rhi + r, lo);
  const double y = fma_(poly, r3, lo);
  return y + hi;
}
template <int N>
__device__ __forceinline__ void log_n(const double (&x)[N], double (&y)[N]) {
  bool all_main = true, all_near1 = true;
#pragma unroll
  for (int u = 0; u < N; ++u) {
    all_main = all_main && log_is_main(x[u]);
    all_near1 = all_near1 && log_is_near1(x[u]);
  }
  if (all_main) {
#pragma unroll
    for (int u = 0; u < N; ++u) y[u] = log_main(x[u]);
  } else if (all_near1) {
#pragma unroll
    for (int u = 0; u < N; ++u) y[u] = log_near1(x[u]);
  } else {
#pragma unroll
    for (int u = 0; u < N; ++u) y[u] = log_(x[u]);
  }
}
template <int N>
__device__ __forceinline__ void exp_n(const double (&x)[N], double (&y)[N]) {
  bool all_main = true;
#pragma unroll
  for (int u = 0; u < N; ++u) all_main = all_main && exp_is_main(x[u]);
  if (all_main) {
#pragma unroll
    for (int u = 0; u < N; ++u) {
      double out;
      exp_core<false>(x[u], 0.0, false, out);
      y[u] = out;
    }
  } else {
#pragma unroll
    for (int u = 0; u < N; ++u) y[u] = exp_(x[u]);
  }
}

__device__ __forceinline__ double pow_(double x, double y) {
  const unsigned long long ix = as_u(x), iy = as_u(y);
  const unsigned topx = (unsigned)(ix >> 52), topy = (unsigned)(iy >> 52);
  // main path only: x positive normal, 2^-65 <= |y| < 2^63
  if (topx - 1u > 0x7fdu || (topy & 0x7ff) - 0x3beu > 0x7fu) return ::pow(x, y);
  const double Ln2hi = as_d(POWLOG_HEAD[0]), Ln2lo = as_d(POWLOG_HEAD[1]), A0 = as_d(POWLOG_HEAD[2]),
               A1 = as_d(POWLOG_HEAD[3]), A2 = as_d(POWLOG_HEAD[4]), A3 = as_d(POWLOG_HEAD[5]), A4 = as_d(POWLOG_HEAD[6]),
               A5 = as_d(POWLOG_HEAD[7]), A6 = as_d(POWLOG_HEAD[8]);
  // log_inline (FMA variant)
  const unsigned long long tmp = ix - 0x3fe6955500000000ULL;
  const unsigned i = (unsigned)(tmp >> 45) & 127;
  const int k = (int)((long long)tmp >> 52);
  const unsigned long long iz = ix - (tmp & 0xfff0000000000000ULL);
  const double z = as_d(iz), kd = (double)k;
  const double invc = as_d(POWLOG_TAB[4 * i]), logc = as_d(POWLOG_TAB[4 * i + 2]), logctail = as_d(POWLOG_TAB[4 * i + 3]);
  const double r = fma_(z, invc, -1.0);
  const double t1 = fma_(kd, Ln2hi, logc);
  const double t2 = t1 + r;
  const double lo1 = fma_(kd, Ln2lo, logctail);
  const double lo2 = (t1 - t2) + r;
  const double ar = A0 * r;
  const double ar2 = r * ar;
  const double ar3 = r * ar2;
  const double hi = t2 + ar2;
  const double lo3 = fma_(ar, r, -ar2);
  const double lo4 = (t2 - hi) + ar2;
  const double p = fma_(ar2, fma_(fma_(r, A6, A5), ar2, fma_(r, A4, A3)), fma_(r, A2, A1));
  const double lo = fma_(ar3, p, ((lo1 + lo2) + lo3) + lo4);
  const double lhi = hi + lo;
  const double llo = (hi - lhi) + lo;
  // exp_inline(y*log(x))
  const double ehi = y * lhi;
  const double elo = fma_(y, llo, fma_(lhi, y, -ehi));
  double out;
  if (exp_core(ehi, elo, true, out)) return out;
  return ::pow(x, y);
}

}  // namespace libm
}  // namespace pace
