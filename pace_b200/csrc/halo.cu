// Device-side halo exchange: between subdomains resident on the same GPU in one gather kernel, and, for
// neighbours on another GPU, pack / unpack kernels: the pack kernel stores its message straight into the receiver's
// buffer over NVLink when it is given a peer-mapped address (pace_b200/distributed.py PeerArena, the default on GPUs),
// or fills a local send buffer for an isend / irecv transfer (PACE_B200_HALO=p2p; gloo in the CPU tests).
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

// Message layout shared by sender and receiver: the ops of one plan in plan order; op number m owns the cells
// [prefix_m, prefix_m + ni*nj) of every (field q, level k) slab: element = (prefix_m * nq + q * cells) * nk + k * cells + c.
// Pack ops carry the prefix in dst_t (the destination subdomain is remote), unpack ops in src_t.
__device__ __forceinline__ long long msg_index(int prefix, int cells, int nq, int nk, int q, int k, int c) {
  return ((long long)prefix * nq + (long long)q * cells) * nk + (long long)k * cells + c;
}
__device__ __forceinline__ void halo_pack_cell(const Geo& g, const HaloOpDev& o, const HaloFields& f, double* buf, int nk,
                                               int q, int k, int c) {
  int li = c % o.ni, lj = c / o.ni;
  int i = o.i0 + li, j = o.j0 + lj;
  int si = o.aii * i + o.aij * j + o.bi, sj = o.aji * i + o.ajj * j + o.bj;
  const double* src = (o.src_comp == 0 ? f.x[q] : f.y[q]);
  double v = src[(long long)o.src_t * g.st3 + (long long)k * g.sk + sj * g.sj + si];
  buf[msg_index(o.dst_t, o.ni * o.nj, f.n, nk, q, k, c)] = (o.sign == 1.0) ? v : o.sign * v;
}
__device__ __forceinline__ void halo_unpack_cell(const Geo& g, const HaloOpDev& o, const HaloFields& f, const double* buf,
                                                 int nk, int q, int k, int c) {
  int li = c % o.ni, lj = c / o.ni;
  int i = o.i0 + li, j = o.j0 + lj;
  double* dst = (o.comp == 0 ? f.x[q] : f.y[q]);
  dst[(long long)o.dst_t * g.st3 + (long long)k * g.sk + j * g.sj + i] =
      buf[msg_index(o.src_t, o.ni * o.nj, f.n, nk, q, k, c)];
}

#ifndef PACE_HOST_EMU
// mode 0: subdomain -> subdomain, 1: subdomain -> message, 2: message -> subdomain
__global__ void k_halo(Geo g, const HaloOpDev* ops, HaloFields f, int nk, int mode, double* buf) {
  const HaloOpDev o = ops[blockIdx.x];
  const int k = blockIdx.y;
  const int cells = o.ni * o.nj;
  for (int q = 0; q < f.n; ++q)
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
      if (mode == 0) halo_cell(g, o, f, q, k, c);
      else if (mode == 1) halo_pack_cell(g, o, f, buf, nk, q, k, c);
      else halo_unpack_cell(g, o, f, buf, nk, q, k, c);
    }
}
#endif

int halo_plan_set(Core& c, int plan_id, const void* host_ops, int n_ops) {
  if (plan_id < 0 || plan_id >= 256) { c.err = "halo plan id out of range"; return 1; }
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
// mode 0: local update; 1: pack the ops of `plan_id` into buf; 2: unpack them from buf
static int halo_run(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                    int mode, double* buf) {
  if (plan_id < 0 || plan_id >= 256 || !c.halo_ops[plan_id]) { c.err = "halo plan not set"; return 1; }
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
  LaunchScope scope(c.stream, mode == 0 ? "halo_update" : mode == 1 ? "halo_pack" : "halo_unpack");
  k_halo<<<dim3(n_ops, nk), 128, 0, c.stream>>>(c.g, ops, f, nk, mode, buf);
#else
  for (int o = 0; o < n_ops; ++o)
    for (int k = 0; k < nk; ++k)
      for (int q = 0; q < n; ++q)
        for (int cc = 0; cc < ops[o].ni * ops[o].nj; ++cc) {
          if (mode == 0) halo_cell(c.g, ops[o], f, q, k, cc);
          else if (mode == 1) halo_pack_cell(c.g, ops[o], f, buf, nk, q, k, cc);
          else halo_unpack_cell(c.g, ops[o], f, buf, nk, q, k, cc);
        }
#endif
  return 0;
}

int halo_update(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk) {
  return halo_run(c, plan_id, x_names, y_names, n, nk, 0, nullptr);
}
int halo_pack(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk, double* buf) {
  if (!buf) { c.err = "halo_pack: null message buffer"; return 1; }
  return halo_run(c, plan_id, x_names, y_names, n, nk, 1, buf);
}
int halo_unpack(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                const double* buf) {
  if (!buf) { c.err = "halo_unpack: null message buffer"; return 1; }
  return halo_run(c, plan_id, x_names, y_names, n, nk, 2, const_cast<double*>(buf));
}

}  // namespace pace
