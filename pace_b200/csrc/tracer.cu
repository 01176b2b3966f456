// Tracer advection (Fortran tracer_2d_1L): sub-cycled fvtp2d transport of the NQ tracers with the mass
// fluxes / Courant numbers accumulated over the acoustic substeps.
//
// Reference: pyFV3/stencils/tracer_2d_1l.py:25-176 (stencils), :289-417 (__call__).  The halo updates between
// sub-cycles are issued by the caller (engine.py), exactly where the reference calls its HaloUpdater.
#include "transport.hpp"

namespace pace {

static const char* TRACER_NAMES[9] = {"qvapor", "qliquid", "qrain", "qice", "qsnow", "qgraupel", "qo3mr", "qsgs_tke", "qcld"};
const char* tracer_name(int n) { return TRACER_NAMES[n]; }  // ndsl/dsl/gt4py_utils.py:28-38

// flux_compute + divide_fluxes_by_n_substeps (tracer_2d_1l.py:47-108)
int tracer_flux_prep(Core& c, int n_split) {
  const Geo g = c.g;
  double *cx = c.F("cxd"), *cy = c.F("cyd"), *mfx = c.F("mfxd"), *mfy = c.F("mfyd");
  double *xfx = c.FS("tr.x_area_flux"), *yfx = c.FS("tr.y_area_flux");
  const double *dxa = c.F("grid.dxa"), *dya = c.F("grid.dya"), *dx = c.F("grid.dx"), *dy = c.F("grid.dy"),
               *sin_sg1 = c.F("grid.sin_sg1"), *sin_sg2 = c.F("grid.sin_sg2"), *sin_sg3 = c.F("grid.sin_sg3"),
               *sin_sg4 = c.F("grid.sin_sg4");
  if (!c.err.empty()) return 1;
  const int is = g.is, ie = g.ie, js = g.js, je = g.je;
  const double frac = 1.0 / n_split;
  // origin_full, domain_full(add=(1,1,0)): the whole (nx+7) x (ny+7) allocation
  launch_ijk(c.stream, g, Box{0, 0, 0, g.nx + 7, g.ny + 7, g.nz}, [=] __device__(const Pt& p) {
    double xv = A0(xfx), yv = A0(yfx);
    if (p.i >= is && p.i <= ie + 1 && p.j >= js - 3 && p.j <= je + 3) {
      double cxv = A0(cx);
      xv = cxv > 0 ? cxv * A2(dxa, -1, 0) * A20(dy) * A2(sin_sg3, -1, 0) : cxv * A20(dxa) * A20(dy) * A20(sin_sg1);
    }
    if (p.i >= is - 3 && p.i <= ie + 3 && p.j >= js && p.j <= je + 1) {
      double cyv = A0(cy);
      yv = cyv > 0 ? cyv * A2(dya, 0, -1) * A20(dx) * A2(sin_sg4, 0, -1) : cyv * A20(dya) * A20(dx) * A20(sin_sg2);
    }
    if (n_split > 1) {
      A0(cx) = A0(cx) * frac;
      xv = xv * frac;
      A0(mfx) = A0(mfx) * frac;
      A0(cy) = A0(cy) * frac;
      yv = yv * frac;
      A0(mfy) = A0(mfy) * frac;
    }
    A0(xfx) = xv;
    A0(yfx) = yv;
  });
  return 0;
}

#ifndef PACE_HOST_EMU
namespace {
// flags[t * nk + k] = does level k of subdomain t (the whole padded level, halo included) hold a non-zero value?
__global__ void __launch_bounds__(256) k_level_nonzero(Geo g, const double* q, int nk, int* flags) {
  const int k = blockIdx.x, t = blockIdx.y;
  const double* lev = q + (long long)t * g.st3 + (long long)k * g.sk;
  bool nz = false;
  for (int n = threadIdx.x; n < g.sk; n += blockDim.x) nz = nz || lev[n] != 0.0;
  nz = __syncthreads_or(nz);
  if (threadIdx.x == 0) flags[t * nk + k] = nz ? 1 : 0;
}
// Later sub-cycles of TracerAdvection: a level that was zero on the whole padded array before a sub-cycle is still zero
// on its compute cells afterwards (transport of zeros; the kernels skip it), so only the halo update in between can
// have brought something in: flag |= "the 3-cell frame of the level holds a non-zero value".  Levels already flagged
// stay flagged (a flag only permits skipping).  Reads 6 % of a level instead of all of it.
__global__ void __launch_bounds__(256) k_level_frame_or(Geo g, const double* q, int nk, int* flags) {
  const int k = blockIdx.x, t = blockIdx.y;
  if (flags[t * nk + k] != 0) return;
  const double* lev = q + (long long)t * g.st3 + (long long)k * g.sk;
  const int nj = g.ny + 2 * HALO + 1, sj = g.sj;
  bool nz = false;
  // rows below / above the compute rows: whole rows; compute rows: the columns left / right of the compute cells
  const int nrow_full = 2 * HALO + 1;                    // rows 0..2 and ny+3..ny+6
  for (int n = threadIdx.x; n < nrow_full * sj; n += blockDim.x) {
    const int r = n / sj, i = n - r * sj;
    const int j = r < HALO ? r : g.ny + r;               // r = 3.. -> ny+3..
    nz = nz || lev[j * sj + i] != 0.0;
  }
  const int nside = sj - g.nx;                            // columns 0..2 and nx+3..sj-1
  for (int n = threadIdx.x; n < g.ny * nside; n += blockDim.x) {
    const int r = n / nside, c = n - r * nside;
    const int i = c < HALO ? c : g.nx + c;
    nz = nz || lev[(HALO + r) * sj + i] != 0.0;
  }
  (void)nj;
  nz = __syncthreads_or(nz);
  if (threadIdx.x == 0 && nz) flags[t * nk + k] = 1;
}
}  // namespace
#endif

// Condensate species are zero on whole levels (above the clouds; everywhere in a dry test case).  One read pass finds
// those levels; their transport (fvtp2d + apply_tracer_flux) is skipped: a level of zeros stays a level of zeros.
// slot: one of 10 flag arrays (0 - 7: the tracers of TracerAdvection, 8: q_con inside AcousticDynamics, 9: any other); frame_only: update the
// slot's flags from the halo frame alone (see k_level_frame_or; the slot must hold the flags of the same field from
// before the last transport + halo update)
const int* level_nonzero_flags(Core& c, const double* q, int nk, int slot, bool frame_only) {
#ifndef PACE_HOST_EMU
  if (!zero_paths_enabled()) return nullptr;
  const Geo g = c.g;
  auto it = c.scratch.find("tr.level_nonzero");
  int* flags;
  const size_t per_slot = (size_t)g.nt * (g.nz + 1);
  if (it == c.scratch.end()) {
    void* ptr = nullptr;
    if (cudaMalloc(&ptr, sizeof(int) * per_slot * 10) != cudaSuccess) {
      c.err = "cudaMalloc failed for tr.level_nonzero";
      return nullptr;
    }
    c.owned.push_back(ptr);
    c.scratch["tr.level_nonzero"] = (double*)ptr;
    flags = (int*)ptr;
  } else {
    flags = (int*)it->second;
  }
  flags += per_slot * (size_t)slot;
  ++launch_counter();
  if (frame_only) k_level_frame_or<<<dim3(nk, g.nt), 256, 0, c.stream>>>(g, q, nk, flags);
  else k_level_nonzero<<<dim3(nk, g.nt), 256, 0, c.stream>>>(g, q, nk, flags);
  return flags;
#else
  (void)c; (void)q; (void)nk; (void)slot; (void)frame_only;
  return nullptr;
#endif
}

#ifndef PACE_HOST_EMU
namespace {
// zero[t * 9 + n] = 1 when every level flag of tracer n on subdomain t is 0 (slots 0 - 7 of level_nonzero_flags)
__global__ void k_fields_all_zero(const int* flags, int per_slot, int nk, int nq, int* zero) {
  const int t = blockIdx.x, n = blockIdx.y;
  bool any = false;
  if (n < 8) {
    const int* f = flags + (size_t)per_slot * n + t * nk;
    for (int k = threadIdx.x; k < nk; k += blockDim.x) any = any || f[k] != 0;
  } else {
    any = true;   // no flags are kept for a ninth tracer
  }
  any = __syncthreads_or(any);
  if (threadIdx.x == 0) zero[t * 9 + n] = any ? 0 : 1;
}
}  // namespace
#endif

// After TracerAdvection: which tracers are zero on a whole subdomain (all their level flags of the last sub-cycle are
// 0: the levels, halo included, held no non-zero value before the last transport, so the compute cells hold none
// after it).  The vertical remap leaves such a field as it is.  nullptr when the flags are not those of a
// TracerAdvection call that has just run (Core::tracer_flags_fresh, set by step()) or the shortcuts are off.
const int* tracer_zero_fields(Core& c, int nq) {
#ifndef PACE_HOST_EMU
  if (!zero_paths_enabled() || !c.tracer_flags_fresh) return nullptr;
  auto it = c.scratch.find("tr.level_nonzero");
  if (it == c.scratch.end()) return nullptr;
  const Geo g = c.g;
  int* zero;
  auto iz = c.scratch.find("tr.field_zero");
  if (iz == c.scratch.end()) {
    void* ptr = nullptr;
    if (cudaMalloc(&ptr, sizeof(int) * (size_t)g.nt * 9) != cudaSuccess) { c.err = "cudaMalloc failed for tr.field_zero"; return nullptr; }
    c.owned.push_back(ptr);
    c.scratch["tr.field_zero"] = (double*)ptr;
    zero = (int*)ptr;
  } else {
    zero = (int*)iz->second;
  }
  ++launch_counter();
  k_fields_all_zero<<<dim3(g.nt, 9), 128, 0, c.stream>>>((const int*)it->second, g.nt * (g.nz + 1), g.nz, nq, zero);
  return zero;
#else
  (void)c; (void)nq;
  return nullptr;
#endif
}

// one sub-cycle: apply_mass_flux, then fvtp2d + apply_tracer_flux for each tracer (tracer_2d_1l.py:381-411)
// first_subcycle = false: the level flags of the previous sub-cycle are brought up to date from the halo frame
int tracer_substep(Core& c, int nq, bool first_subcycle) {
  const Geo g = c.g;
  cudaStream_t st = c.stream;
  double *dp1 = c.F("dyn.dp1"), *dp2 = c.FS("tr.tmp_dp");
  const double *cx = c.F("cxd"), *cy = c.F("cyd"), *mfx = c.F("mfxd"), *mfy = c.F("mfyd"), *rarea = c.F("grid.rarea");
  double *xfx = c.FS("tr.x_area_flux"), *yfx = c.FS("tr.y_area_flux"), *fx = c.FS("tr.x_flux"), *fy = c.FS("tr.y_flux");
  if (!c.err.empty()) return 1;
  launch_ijk(st, g, box_compute(g), [=] __device__(const Pt& p) {
    A0(dp2) = A0(dp1) + (A0(mfx) - A(mfx, 1, 0, 0) + A0(mfy) - A(mfy, 0, 1, 0)) * A20(rarea);
  });
  for (int n = 0; n < nq; ++n) {
    double* q = c.F(TRACER_NAMES[n]);
    if (!c.err.empty()) return 1;
    Fvtp2dArgs a;
    a.q = q; a.crx = cx; a.cry = cy; a.xfx = xfx; a.yfx = yfx; a.fx = fx; a.fy = fy;
    a.x_mass_flux = mfx; a.y_mass_flux = mfy; a.hord = c.cfg.hord_tr; a.nk = g.nz;
    const int* lvl = level_nonzero_flags(c, q, g.nz, n < 8 ? n : 9, !first_subcycle && n < 8);
    if (!c.err.empty()) return 1;
    a.level_nonzero = lvl;
    if (fvtp2d(c, a)) return 1;
    const int nkf = g.nz;
    launch_ijk(st, g, box_compute(g), [=] __device__(const Pt& p) {
      if (lvl != nullptr && lvl[p.t * nkf + p.k] == 0) return;  // q = (0 * dp1 + 0) / dp2
      A0(q) = fv_div(A0(q) * A0(dp1) + (A0(fx) - A(fx, 1, 0, 0) + A0(fy) - A(fy, 0, 1, 0)) * A20(rarea), A0(dp2));
    });
  }
  return 0;
}

// swap_dp (tracer_2d_1l.py:171-176), compute domain
int tracer_swap_dp(Core& c) {
  const Geo g = c.g;
  double *dp1 = c.F("dyn.dp1"), *dp2 = c.FS("tr.tmp_dp");
  if (!c.err.empty()) return 1;
  launch_ijk(c.stream, g, box_compute(g), [=] __device__(const Pt& p) {
    double tmp = A0(dp1);
    A0(dp1) = A0(dp2);
    A0(dp2) = tmp;
  });
  return 0;
}

}  // namespace pace
