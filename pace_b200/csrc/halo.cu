// Device-side halo exchange: between subdomains resident on the same GPU in one gather kernel, and, for
// neighbours on another GPU, pack / unpack kernels: the pack kernel stores its message straight into the receiver's
// buffer over NVLink when it is given a peer-mapped address (pace_b200/distributed.py PeerArena, the default on GPUs),
// or fills a local send buffer for an isend / irecv transfer (PACE_B200_HALO=p2p; gloo in the CPU tests).
//
// Replaces NDSL's HaloUpdater pack / Isend / Irecv / unpack (NDSL/ndsl/halo/updater.py:217-303,
// halo/data_transformer.py:404-560, halo/cuda_kernels.py:5-236).  The plan (which source cell fills which
// halo cell, with rotation folded into an affine index map, pace_b200/halo.py) is uploaded once; an update is
// ONE kernel per field group: no index arrays, no staging buffers, rotations applied in flight.
#include <cstring>

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

#ifndef PACE_HOST_EMU
// One launch for the sending half of an update: the ops of every peer's pack plan (slot >= 0: stored into that peer's
// arena) and of the same-GPU gather plan (slot -1).
struct PeerBufs { double* b[16]; };
__global__ void k_halo_send(Geo g, const HaloOpDev* ops, const int* slot, HaloFields f, int nk, PeerBufs bufs) {
  const HaloOpDev o = ops[blockIdx.x];
  const int s = slot[blockIdx.x];
  const int k = blockIdx.y;
  const int cells = o.ni * o.nj;
  for (int q = 0; q < f.n; ++q)
    for (int c = threadIdx.x; c < cells; c += blockDim.x) {
      if (s < 0) halo_cell(g, o, f, q, k, c);
      else halo_pack_cell(g, o, f, bufs.b[s], nk, q, k, c);
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


// ---- neighbours on other GPUs: peer stores + flag barrier, all on the handle's stream -----------------------------
// One halo update = for every peer GPU a pack kernel that stores its message straight into the peer's arena (NVLink
// peer stores: plain st.global on a peer-mapped address), the local gather for same-GPU neighbours, a barrier kernel
// (every rank writes the update's sequence number into its slot of every peer's flag array with release semantics at
// system scope, then waits until all slots of its own array carry that number), and the unpack kernel.  Two message
// halves alternate so that update n+1 cannot overwrite what a peer still unpacks from update n: the barrier of update
// n+1 separates them.  No host code runs between the kernels, so a step is a pure launch sequence on one stream.
// The arenas are mapped by the caller (torch symmetric memory in pace_b200/distributed.py); no TMA is involved.
#ifndef PACE_HOST_EMU
namespace {
struct PeerDev {
  unsigned long long seq;        // barriers executed so far
  unsigned long long error;      // set to the sequence number of a barrier that timed out
  unsigned long long* slot[64];  // slot[r]: this rank's flag slot inside rank r's arena
  unsigned long long* mine;      // this rank's own flag array
  int n, me;
};

__global__ void k_peer_barrier(PeerDev* d) {
  __shared__ unsigned long long seq_s;
  if (threadIdx.x == 0) {
    seq_s = d->seq + 1;
    d->seq = seq_s;
  }
  __syncthreads();
  const unsigned long long seq = seq_s;
  const int r = threadIdx.x;
  if (r < d->n && r != d->me) {
    __threadfence_system();  // the pack kernels before this one have completed; order their peer stores before the flag
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(d->slot[r]), "l"(seq) : "memory");
    unsigned long long t0, t1, v;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    const unsigned long long* p = d->mine + r;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
      if (v >= seq) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 30000000000ull) {  // 30 s: a peer has failed; give up instead of hanging the GPU
        d->error = seq;
        break;
      }
    }
  }
}
}  // namespace
#endif

int peer_arena_set(Core& c, int my_rank, int n_ranks, void* const* arena_ptrs, long long half_doubles) {
#ifndef PACE_HOST_EMU
  if (n_ranks < 2 || n_ranks > 64 || my_rank < 0 || my_rank >= n_ranks) { c.err = "peer_arena_set: bad rank / world size"; return 1; }
  c.arena_ptrs.assign(n_ranks, nullptr);
  for (int r = 0; r < n_ranks; ++r) c.arena_ptrs[r] = (char*)arena_ptrs[r];
  c.peer_rank = my_rank;
  c.half_doubles = half_doubles;
  c.peer_updates = 0;
  PeerDev h;
  memset(&h, 0, sizeof h);
  h.n = n_ranks; h.me = my_rank;
  const size_t flag_off = (size_t)2 * half_doubles * sizeof(double);
  for (int r = 0; r < n_ranks; ++r) h.slot[r] = (unsigned long long*)(c.arena_ptrs[r] + flag_off) + my_rank;
  h.mine = (unsigned long long*)(c.arena_ptrs[my_rank] + flag_off);
  if (!c.peer_dev) {
    if (cudaMalloc(&c.peer_dev, sizeof(PeerDev)) != cudaSuccess) { c.err = "cudaMalloc failed for the peer barrier block"; return 1; }
    c.owned.push_back(c.peer_dev);
  }
  CUDA_CHECK_RET(cudaMemcpy(c.peer_dev, &h, sizeof h, cudaMemcpyHostToDevice));
  return 0;
#else
  (void)my_rank; (void)n_ranks; (void)arena_ptrs; (void)half_doubles;
  c.err = "peer-store halo exchange exists only in the CUDA library";
  return 1;
#endif
}

int peer_status(Core& c, unsigned long long* seq, unsigned long long* error) {
#ifndef PACE_HOST_EMU
  if (!c.peer_dev) { c.err = "peer_status: no arena set"; return 1; }
  unsigned long long v[2];
  CUDA_CHECK_RET(cudaMemcpy(v, c.peer_dev, sizeof v, cudaMemcpyDeviceToHost));  // synchronises with the stream's work
  if (seq) *seq = v[0];
  if (error) *error = v[1];
  return 0;
#else
  (void)seq; (void)error;
  c.err = "peer-store halo exchange exists only in the CUDA library";
  return 1;
#endif
}

int peer_plan_set(Core& c, int plan_id, int n_peers, const int* peer_ranks, const int* pack_plan_ids, int unpack_plan_id,
                  long long max_recv_cells) {
  if (plan_id < 0 || plan_id >= 16) { c.err = "peer_plan_set: plan id out of range"; return 1; }
  Core::PeerPlan& p = c.peer_plans[plan_id];
  p.peer_ranks.assign(peer_ranks, peer_ranks + n_peers);
  p.pack_ids.assign(pack_plan_ids, pack_plan_ids + n_peers);
  p.unpack_id = unpack_plan_id;
  p.max_recv_cells = max_recv_cells;
  p.set = true;
#ifndef PACE_HOST_EMU
  // merged send list: the pack plans (in peer order) followed by the local gather plan
  if (n_peers > 16) { c.err = "peer_plan_set: more than 16 peer GPUs"; return 1; }
  int total = c.halo_nops[plan_id];
  for (int m = 0; m < n_peers; ++m) total += c.halo_nops[pack_plan_ids[m]];
  p.n_merged = total;
  if (total > 0) {
    void *ops = nullptr, *slots = nullptr;
    if (cudaMalloc(&ops, sizeof(HaloOpDev) * total) != cudaSuccess || cudaMalloc(&slots, sizeof(int) * total) != cudaSuccess) {
      c.err = "cudaMalloc failed for the merged halo plan";
      return 1;
    }
    c.owned.push_back(ops);
    c.owned.push_back(slots);
    std::vector<int> hs(total);
    int at = 0;
    for (int m = 0; m <= n_peers; ++m) {
      const int id = m < n_peers ? pack_plan_ids[m] : plan_id;
      const int n = c.halo_nops[id];
      if (n > 0)
        CUDA_CHECK_RET(cudaMemcpy((HaloOpDev*)ops + at, c.halo_ops[id], sizeof(HaloOpDev) * n, cudaMemcpyDeviceToDevice));
      for (int q = 0; q < n; ++q) hs[at + q] = m < n_peers ? m : -1;
      at += n;
    }
    CUDA_CHECK_RET(cudaMemcpy(slots, hs.data(), sizeof(int) * total, cudaMemcpyHostToDevice));
    p.merged_ops = ops;
    p.merged_slot = (int*)slots;
  }
#endif
  return 0;
}

int peer_barrier(Core& c) {
#ifndef PACE_HOST_EMU
  if (!c.peer_dev) { c.err = "peer_barrier: no arena set"; return 1; }
  LaunchScope scope(c.stream, "halo_barrier");
  k_peer_barrier<<<1, 64, 0, c.stream>>>((PeerDev*)c.peer_dev);
  return 0;
#else
  c.err = "peer-store halo exchange exists only in the CUDA library";
  return 1;
#endif
}

// the sending half of one update of `plan_id`: pack into every peer's arena at byte offset `off`, gather locally
static int peer_send(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk,
                     long long off_doubles) {
#ifndef PACE_HOST_EMU
  const Core::PeerPlan& p = c.peer_plans[plan_id];
  if (n < 1 || n > 8) { c.err = "halo_update: 1..8 fields per call"; return 1; }
  if (p.n_merged == 0) return 0;
  HaloFields f;
  f.n = n;
  for (int q = 0; q < n; ++q) {
    f.x[q] = c.F(x_names[q]);
    f.y[q] = y_names ? c.F(y_names[q]) : nullptr;
  }
  if (!c.err.empty()) return 1;
  PeerBufs bufs;
  for (size_t m = 0; m < p.peer_ranks.size(); ++m) bufs.b[m] = (double*)c.arena_ptrs[p.peer_ranks[m]] + off_doubles;
  LaunchScope scope(c.stream, "halo_send");
  k_halo_send<<<dim3(p.n_merged, nk), 128, 0, c.stream>>>(c.g, (const HaloOpDev*)p.merged_ops, p.merged_slot, f, nk, bufs);
  return 0;
#else
  (void)plan_id; (void)x_names; (void)y_names; (void)n; (void)nk; (void)off_doubles;
  c.err = "peer-store halo exchange exists only in the CUDA library";
  return 1;
#endif
}

int peer_halo_update(Core& c, int plan_id, const char* const* x_names, const char* const* y_names, int n, int nk) {
  const Core::PeerPlan& p = c.peer_plans[plan_id];
  const long long half = (long long)(c.peer_updates & 1) * c.half_doubles;
  ++c.peer_updates;
  if (p.max_recv_cells * n * nk > c.half_doubles) { c.err = "halo update larger than the peer arena"; return 1; }
  if (peer_send(c, plan_id, x_names, y_names, n, nk, half)) return 1;
  if (peer_barrier(c)) return 1;
  if (p.unpack_id >= 0 && halo_unpack(c, p.unpack_id, x_names, y_names, n, nk, (double*)c.arena_ptrs[c.peer_rank] + half))
    return 1;
  return 0;
}

int peer_halo_update2(Core& c, int plan_a, const char* const* xa, const char* const* ya, int na, int nka, int plan_b,
                      const char* const* xb, const char* const* yb, int nb, int nkb) {
  const Core::PeerPlan &pa = c.peer_plans[plan_a], &pb = c.peer_plans[plan_b];
  const long long half = (long long)(c.peer_updates & 1) * c.half_doubles;
  ++c.peer_updates;
  // every rank places B's message behind the largest A message any rank receives: the same offset everywhere
  const long long off_b = pa.max_recv_cells * na * nka;
  if (off_b + pb.max_recv_cells * nb * nkb > c.half_doubles) { c.err = "merged halo update larger than the peer arena"; return 1; }
  if (peer_send(c, plan_a, xa, ya, na, nka, half)) return 1;
  if (peer_send(c, plan_b, xb, yb, nb, nkb, half + off_b)) return 1;
  if (peer_barrier(c)) return 1;
  double* mine = (double*)c.arena_ptrs[c.peer_rank] + half;
  if (pa.unpack_id >= 0 && halo_unpack(c, pa.unpack_id, xa, ya, na, nka, mine)) return 1;
  if (pb.unpack_id >= 0 && halo_unpack(c, pb.unpack_id, xb, yb, nb, nkb, mine + off_b)) return 1;
  return 0;
}

}  // namespace pace
