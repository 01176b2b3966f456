// Device-side halo exchange between subdomains resident on the same GPU.
//
// Replaces NDSL's HaloUpdater pack / Isend / Irecv / unpack (NDSL/ndsl/halo/updater.py:217-303,
// halo/data_transformer.py:404-560, halo/cuda_kernels.py:5-236).  The plan (which source cell fills which
// halo cell, with rotation folded into an affine index map, pace_b200/halo.py) is uploaded once; an update is
// ONE kernel per field group: no index arrays, no staging buffers, rotations applied in flight.
#include "core.hpp"

namespace pace {

struct HaloOpDev {
  int dst_t, src_t, comp, src_comp;
  double sign;
  int i0, j0, ni, nj;
  int aii, aij, bi, aji, ajj, bj;
};

struct HaloFields {
  double* x[8];
  double* y[8];
  int n;
};

__device__ __forceinline__ void halo_cell(const Geo& g, const HaloOpDev& o, const HaloFields& f, int q, int k, int c) {
  int li = c % o.ni, lj = c / o.ni;
  int i = o.i0 + li, j = o.j0 + lj;
  int si = o.aii * i + o.aij * j + o.bi, sj = o.aji * i + o.ajj * j + o.bj;
  const double* src = (o.src_comp == 0 ? f.x[q] : f.y[q]);
  double* dst = (o.comp == 0 ? f.x[q] : f.y[q]);
  double v = src[(long long)o.src_t * g.st3 + (long long)k * g.sk + sj * g.sj + si];
  dst[(long long)o.dst_t * g.st3 + (long long)k * g.sk + j * g.sj + i] = (o.sign == 1.0) ? v : o.sign * v;
}

#ifndef PACE_HOST_EMU
__global__ void k_halo(Geo g, const HaloOpDev* ops, HaloFields f, int nk) {
  const HaloOpDev o = ops[blockIdx.x];
  const int k = blockIdx.y;
  const int cells = o.ni * o.nj;
  for (int q = 0; q < f.n; ++q)
    for (int c = threadIdx.x; c < cells; c += blockDim.x) halo_cell(g, o, f, q, k, c);
}
#endif

int halo_plan_set(Core& c, int plan_id, const void* host_ops, int n_ops) {
  if (plan_id < 0 || plan_id >= 16) { c.err = "halo plan id out of range"; return 1; }
  static_assert(sizeof(HaloOpDev) == 64, "HaloOpDev layout is part of the C ABI (pace_b200_halo_op)");
  void* dev = nullptr;
  size_t bytes = sizeof(HaloOpDev) * (size_t)n_ops;
  if (cudaMalloc(&dev, bytes ? bytes : 8) != cudaSuccess) { c.err = "cudaMalloc failed for halo plan"; return 1; }
  if (cudaMemcpy(dev, host_ops, bytes, cudaMemcpyHostToDevice) != cudaSuccess) { c.err = "cudaMemcpy failed for halo plan"; return 1; }
  c.owned.push_back(dev);
  c.halo_ops[plan_id] = dev;
  c.halo_nops[plan_id] = n_ops;
  return 0;
}

// x_names / y_names: arrays of n registered field names (y_names NULL for scalar plans)
int halo_update(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk) {
  if (plan_id < 0 || plan_id >= 16 || !c.halo_ops[plan_id]) { c.err = "halo plan not set"; return 1; }
  if (n < 1 || n > 8) { c.err = "halo_update: 1..8 fields per call"; return 1; }
  HaloFields f;
  f.n = n;
  for (int q = 0; q < n; ++q) {
    f.x[q] = c.F(x_names[q]);
    f.y[q] = y_names ? c.F(y_names[q]) : nullptr;
  }
  if (!c.err.empty()) return 1;
  const int n_ops = c.halo_nops[plan_id];
  if (n_ops == 0) return 0;
  const HaloOpDev* ops = (const HaloOpDev*)c.halo_ops[plan_id];
#ifndef PACE_HOST_EMU
  LaunchScope scope(c.stream, "halo_update");
  k_halo<<<dim3(n_ops, nk), 128, 0, c.stream>>>(c.g, ops, f, nk);
#else
  for (int o = 0; o < n_ops; ++o)
    for (int k = 0; k < nk; ++k)
      for (int q = 0; q < n; ++q)
        for (int cc = 0; cc < ops[o].ni * ops[o].nj; ++cc) halo_cell(c.g, ops[o], f, q, k, cc);
#endif
  return 0;
}

}  // namespace pace
