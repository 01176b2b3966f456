// UpdateHeightOnDGrid (Fortran update_dz_d): advect the interface heights zh with the D-grid fluxes.
//
// Reference: pyFV3/stencils/updatedzd.py:28-109 (apply_height_fluxes), :130-209 (cubic spline),
//            :316-389 (__call__ order).  29 reference launches -> 1 + fvtp2d(4) + delnflux(5) + 2.
#include <cstdlib>

#include "tmem.cuh"
#include "transport.hpp"

namespace pace {

#ifndef PACE_HOST_EMU
namespace {
// cubic_spline_interpolation_from_layer_center_to_interfaces (updatedzd.py:130-209) with the forward sweep's values in
// tensor memory (tmem.cuh) instead of the output field: a thread owns one column of ONE of the four fields, reads its
// km levels once and writes its km+1 interface values once (8 field passes for the four fields instead of 16: the
// kernel is bound by DRAM bandwidth).  Same expressions, level by level, as the four-field kernel in update_dz_d.
struct SplineArgs {
  const double* in[4];
  double* out[4];
  const double *gk, *beta, *gamma;
  int km, ni, nj, i0, j0;
};
constexpr int SPL_U = 8;   // levels requested ahead of the forward recurrence

__global__ void __launch_bounds__(128, 4) k_dzd_spline_tmem(Geo g, SplineArgs a) {
  extern __shared__ double sm[];          // [3][km] level constants, then the levels >= TMEM_LEVELS of Y per warp
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int km = a.km;
  double *s_gk = sm, *s_beta = sm + km, *s_gamma = sm + 2 * km;
  for (int k = threadIdx.x; k < km; k += blockDim.x) {
    s_gk[k] = a.gk[k]; s_beta[k] = a.beta[k]; s_gamma[k] = a.gamma[k];
  }
  const uint32_t base = tmem_alloc_cta(&tmem_slot);   // contains the CTA barrier that publishes the constants
  const int ncol = a.ni * a.nj;
  const int col0 = blockIdx.x * 128 + warp * 32;
  if (col0 < ncol) {   // warp-uniform: the TMEM accesses are warp-collective
    const int nsm = km > TMEM_LEVELS ? km - TMEM_LEVELS : 0;
    const YAcc<true> Y{sm + 3 * km + (size_t)warp * nsm * 32 + lane, tmem_warp_addr(base)};
    const bool active = col0 + lane < ncol;           // lanes past the end work on a copy of the last column, store nothing
    const int col = active ? col0 + lane : ncol - 1;
    const int i = a.i0 + col % a.ni, j = a.j0 + col / a.ni, t = blockIdx.y, f = blockIdx.z;
    const long long o0 = (long long)t * g.st3 + j * g.sj + i;
    const long long sk = g.sk;
    const double* __restrict__ in = (f == 0 ? a.in[0] : f == 1 ? a.in[1] : f == 2 ? a.in[2] : a.in[3]) + o0;
    double* __restrict__ out = (f == 0 ? a.out[0] : f == 1 ? a.out[1] : f == 2 ? a.out[2] : a.out[3]) + o0;
    // forward sweep
    double nxt[SPL_U];
#pragma unroll
    for (int u = 0; u < SPL_U; ++u) nxt[u] = in[(long long)min(1 + u, km - 1) * sk];
    const double c_top = in[0];
    double p = fv_div(2.0 * s_gk[0] * (s_gk[0] + 1.0) * c_top + nxt[0], s_beta[0]);
    Y.st(0, p);
    double m = c_top, m2 = c_top;                     // levels k-1, k-2
    for (int k0 = 1; k0 < km; k0 += SPL_U) {
      double cur[SPL_U];
#pragma unroll
      for (int u = 0; u < SPL_U; ++u) cur[u] = nxt[u];
      if (k0 + SPL_U < km) {
#pragma unroll
        for (int u = 0; u < SPL_U; ++u) nxt[u] = in[(long long)min(k0 + SPL_U + u, km - 1) * sk];
      }
#pragma unroll
      for (int u = 0; u < SPL_U; ++u) {
        const int k = k0 + u;
        if (k < km) {
          p = fv_div(3.0 * (m + s_gk[k] * cur[u]) - p, s_beta[k]);
          Y.st(k, p);
          m2 = m;
          m = cur[u];
        }
      }
    }
    {
      const double gl = s_gk[km - 1];
      const double a_bot = 1.0 + gl * (gl + 1.5);
      const double xt1b = 2.0 * gl * (gl + 1.0);
      const double xt2 = gl * (gl + 0.5) - a_bot * s_gamma[km - 1];
      p = (xt1b * m + m2 - a_bot * p) / xt2;          // m: level km-1, m2: level km-2
      if (active) out[(long long)km * sk] = p;
    }
    Y.fence_st();
    // backward sweep
    for (int k = km - 1; k >= 0; --k) {
      p = Y.ld(k) - s_gamma[k] * p;
      if (active) out[(long long)k * sk] = p;
    }
  }
  tmem_free_cta(base);
}
}  // namespace
#endif

// levels ahead of the sweeps for which L2 is asked (one request per 32-byte sector); measured per update_dz_d call at
// C384: 6.58 ms without, 6.12 ms with 6 levels, 6.43 ms with 12
constexpr int DZD_PF = 6;

int update_dz_d(Core& c, double dt) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  const int km = g.nz;
  const double *zs = c.F("ad.zs"), *crx = c.F("ad.crx"), *cry = c.F("ad.cry"), *xfx = c.F("ad.xfx"), *yfx = c.F("ad.yfx"),
               *area = c.F("grid.area"), *gk = c.F("col.dzd_gk"), *beta = c.F("col.dzd_beta"), *gamma = c.F("col.dzd_gamma"),
               *nord_v = c.F("col.nord_v"), *damp_vt = c.F("col.damp_vt");
  double *zh = c.F("ad.zh"), *ws = c.F("ad.wsd");
  double *crx_i = c.FS("dzd.crx_interface"), *cry_i = c.FS("dzd.cry_interface"), *xfx_i = c.FS("dzd.x_area_flux_interface"),
         *yfx_i = c.FS("dzd.y_area_flux_interface"), *wk = c.FS("dzd.wk"), *hx = c.FS("dzd.height_x_diffusive_flux"),
         *hy = c.FS("dzd.height_y_diffusive_flux"), *fx = c.FS("dzd.fx"), *fy = c.FS("dzd.fy");
  if (!c.err.empty()) return 1;
  const int nmax_v = c.col_max("col.nord_v");
  {
    const std::vector<double>* dv = c.H("col.damp_vt");
    if (!dv) return 1;
    for (int k = 0; k < km; ++k)
      if ((*dv)[k] <= 1e-5) { c.err = "damp <= 1e-5 in column_namelist is untested"; return 1; }
  }

#ifndef PACE_HOST_EMU
  // PACE_B200_DZD_SPLINE=field selects the four-field kernel below (A/B runs; the host emulation always uses it)
  static const bool spline_tmem = [] { const char* e = getenv("PACE_B200_DZD_SPLINE"); return !(e != nullptr && e[0] == 'f'); }();
#else
  const bool spline_tmem = false;
#endif
  // cubic_spline_interpolation_from_layer_center_to_interfaces x4, full horizontal domain
  // The four fields are independent: their recurrences advance together, level by level, so that four division
  // chains are in flight per thread instead of one (the kernel is bound by the latency of the dependent divisions).
  if (!spline_tmem) launch_ij(st, g, box_full(g), [=] __device__(const Pt& p) {
    double p0, p1, p2, p3;  // crx, xfx, cry, yfx
    {
      const double xt1 = 2.0 * gk[0] * (gk[0] + 1.0), b0 = beta[0];
      p0 = fv_div(xt1 * A(crx, 0, 0, 0) + A(crx, 0, 0, 1), b0);
      p1 = fv_div(xt1 * A(xfx, 0, 0, 0) + A(xfx, 0, 0, 1), b0);
      p2 = fv_div(xt1 * A(cry, 0, 0, 0) + A(cry, 0, 0, 1), b0);
      p3 = fv_div(xt1 * A(yfx, 0, 0, 0) + A(yfx, 0, 0, 1), b0);
      A(crx_i, 0, 0, 0) = p0; A(xfx_i, 0, 0, 0) = p1; A(cry_i, 0, 0, 0) = p2; A(yfx_i, 0, 0, 0) = p3;
    }
    double m0 = A(crx, 0, 0, 0), m1 = A(xfx, 0, 0, 0), m2 = A(cry, 0, 0, 0), m3 = A(yfx, 0, 0, 0);  // level k-1
    for (int k = 1; k < km; ++k) {
      const double g_k = gk[k], b_k = beta[k];
#ifndef PACE_HOST_EMU
      if (DZD_PF > 0 && (p.i & 3) == 0 && k + DZD_PF < km) {  // L2 prefetch of the rows a few levels ahead (one per sector)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(crx, 0, 0, k + DZD_PF)));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(xfx, 0, 0, k + DZD_PF)));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(cry, 0, 0, k + DZD_PF)));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(yfx, 0, 0, k + DZD_PF)));
      }
#endif
      const double c0 = A(crx, 0, 0, k), c1 = A(xfx, 0, 0, k), c2 = A(cry, 0, 0, k), c3 = A(yfx, 0, 0, k);
      p0 = fv_div(3.0 * (m0 + g_k * c0) - p0, b_k);
      p1 = fv_div(3.0 * (m1 + g_k * c1) - p1, b_k);
      p2 = fv_div(3.0 * (m2 + g_k * c2) - p2, b_k);
      p3 = fv_div(3.0 * (m3 + g_k * c3) - p3, b_k);
      A(crx_i, 0, 0, k) = p0; A(xfx_i, 0, 0, k) = p1; A(cry_i, 0, 0, k) = p2; A(yfx_i, 0, 0, k) = p3;
      m0 = c0; m1 = c1; m2 = c2; m3 = c3;
    }
    {
      const double gl = gk[km - 1];
      const double a_bot = 1.0 + gl * (gl + 1.5);
      const double xt1b = 2.0 * gl * (gl + 1.0);
      const double xt2 = gl * (gl + 0.5) - a_bot * gamma[km - 1];
      // m* hold level km-1; level km-2 is read again
      p0 = (xt1b * m0 + A(crx, 0, 0, km - 2) - a_bot * p0) / xt2;
      p1 = (xt1b * m1 + A(xfx, 0, 0, km - 2) - a_bot * p1) / xt2;
      p2 = (xt1b * m2 + A(cry, 0, 0, km - 2) - a_bot * p2) / xt2;
      p3 = (xt1b * m3 + A(yfx, 0, 0, km - 2) - a_bot * p3) / xt2;
      A(crx_i, 0, 0, km) = p0; A(xfx_i, 0, 0, km) = p1; A(cry_i, 0, 0, km) = p2; A(yfx_i, 0, 0, km) = p3;
    }
    for (int k = km - 1; k >= 0; --k) {
      const double gm = gamma[k];
#ifndef PACE_HOST_EMU
      if (DZD_PF > 0 && (p.i & 3) == 0 && k - DZD_PF >= 0) {
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(crx_i, 0, 0, k - DZD_PF)));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(xfx_i, 0, 0, k - DZD_PF)));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(cry_i, 0, 0, k - DZD_PF)));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(&A(yfx_i, 0, 0, k - DZD_PF)));
      }
#endif
      p0 = A(crx_i, 0, 0, k) - gm * p0;
      p1 = A(xfx_i, 0, 0, k) - gm * p1;
      p2 = A(cry_i, 0, 0, k) - gm * p2;
      p3 = A(yfx_i, 0, 0, k) - gm * p3;
      A(crx_i, 0, 0, k) = p0; A(xfx_i, 0, 0, k) = p1; A(cry_i, 0, 0, k) = p2; A(yfx_i, 0, 0, k) = p3;
    }
  }, "dzd_spline");
#ifndef PACE_HOST_EMU
  if (spline_tmem) {
    const Box b = box_full(g);
    SplineArgs sa;
    sa.in[0] = crx; sa.in[1] = xfx; sa.in[2] = cry; sa.in[3] = yfx;
    sa.out[0] = crx_i; sa.out[1] = xfx_i; sa.out[2] = cry_i; sa.out[3] = yfx_i;
    sa.gk = gk; sa.beta = beta; sa.gamma = gamma; sa.km = km; sa.ni = b.ni; sa.nj = b.nj; sa.i0 = b.i0; sa.j0 = b.j0;
    const int nsm = km > TMEM_LEVELS ? km - TMEM_LEVELS : 0;
    const size_t smem = ((size_t)3 * km + (size_t)nsm * 32 * 4) * sizeof(double);
    if (km < 3 || smem > 48 * 1024) { c.err = "update_dz_d: level count outside the range of the spline kernel"; return 1; }
    LaunchScope scope(st, "dzd_spline");
    k_dzd_spline_tmem<<<dim3((b.ni * b.nj + 127) / 128, g.nt, 4), 128, smem, st>>>(g, sa);
  }
#endif
  Fvtp2dArgs a;
  a.q = zh; a.crx = crx_i; a.cry = cry_i; a.xfx = xfx_i; a.yfx = yfx_i; a.fx = fx; a.fy = fy; a.hord = c.cfg.hord_tm; a.nk = km + 1;
  if (fvtp2d(c, a)) return 1;
  DelnArgs d;
  d.q = zh; d.fx = hx; d.fy = hy; d.nord_col = nord_v; d.damp = damp_vt; d.d2 = wk; d.nmax = nmax_v; d.nk = km + 1;
  if (delnflux(c, d)) return 1;
  // apply_height_fluxes: PARALLEL part, then the column part (ws, monotone zh)
  launch_ijk(st, g, box_compute(g, 0, 0, 0, 0, 1), [=] __device__(const Pt& p) {
    double ar = A20(area);
    double area_after_flux = ((ar + A0(xfx_i) - A(xfx_i, 1, 0, 0)) + (ar + A0(yfx_i) - A(yfx_i, 0, 1, 0))) - ar;
    A0(zh) = (A0(zh) * ar + A0(fx) - A(fx, 1, 0, 0) + A0(fy) - A(fy, 0, 1, 0)) / area_after_flux +
             (A0(hx) - A(hx, 1, 0, 0) + A0(hy) - A(hy, 0, 1, 0)) / ar;
  });
  launch_ij(st, g, box_compute(g), [=] __device__(const Pt& p) {
    double below = A(zh, 0, 0, km);
    A20(ws) = (A20(zs) - below) / dt;
    for (int k = km - 1; k >= 0; --k) {
      double other = below + DZ_MIN;
      double h = A(zh, 0, 0, k);
      h = h > other ? h : other;
      A(zh, 0, 0, k) = h;
      below = h;
    }
  });
  return 0;
}

}  // namespace pace
