// Diagnostics egress: what the reference's MonitorDiagnostics.store assembles for its monitor
// (driver/pace/driver/diagnostics.py:164-197) leaves the device as compact arrays in the reference's own order, so that
// the host side of an output step is a copy and a hand-over, not a transposition of gigabytes.
//
//  * egress_pack: the compute domain of a field (with its closing interface row / column / level where the field has
//    one: `quantity.view[:]`) from the library's padded [t][k][j][i] layout to a dense [t][i][j][k] buffer (the
//    reference's (i, j, k) per rank), through a shared-memory transpose so that both sides are coalesced; also one
//    level of a 3-D field as a 2-D array (ZSelect.select_data, diagnostics.py:44-64) and 2-D fields.
//  * column_integral: _compute_column_integral (diagnostics.py:226-249), RGRAV * np.sum(q * delp, axis=2), with NumPy's
//    summation order for a contiguous axis (pairwise_sum in numpy/_core/src/umath/loops_utils.h.src: eight running
//    sums over the first n - n % 8 elements, combined as a balanced tree, then the remainder one by one) so that the
//    result has the bits the reference's monitor would have been handed.
#include "core.hpp"

namespace pace {

#ifndef PACE_HOST_EMU
namespace {
// one (k, i) plane of row j of subdomain t -> (i, k): 32 x 32 tiles
__global__ void __launch_bounds__(256) k_pack3(Geo g, const double* src, double* dst, int ni, int nj, int nk) {
  __shared__ double tile[32][33];
  const int t = blockIdx.z / nj, j = blockIdx.z % nj;
  const int i0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const double* s = src + (long long)t * g.st3 + (long long)(g.js + j) * g.sj + g.is;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int k = k0 + r, i = i0 + threadIdx.x;
    if (k < nk && i < ni) tile[r][threadIdx.x] = s[(long long)k * g.sk + i];
  }
  __syncthreads();
  double* d = dst + ((long long)t * ni * nj + j) * nk;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int i = i0 + r, k = k0 + threadIdx.x;
    if (i < ni && k < nk) d[(long long)i * nj * nk + k] = tile[threadIdx.x][r];
  }
}
// a horizontal plane (2-D field, or one level of a 3-D field) -> [t][i][j]
__global__ void __launch_bounds__(256) k_pack2(Geo g, const double* src, long long stride_t, double* dst, int ni, int nj) {
  __shared__ double tile[32][33];
  const int t = blockIdx.z;
  const int i0 = blockIdx.x * 32, j0 = blockIdx.y * 32;
  const double* s = src + (long long)t * stride_t + (long long)g.js * g.sj + g.is;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int j = j0 + r, i = i0 + threadIdx.x;
    if (j < nj && i < ni) tile[r][threadIdx.x] = s[(long long)j * g.sj + i];
  }
  __syncthreads();
  double* d = dst + (long long)t * ni * nj;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int i = i0 + r, j = j0 + threadIdx.x;
    if (i < ni && j < nj) d[(long long)i * nj + j] = tile[threadIdx.x][r];
  }
}
}  // namespace
#endif

// kind 0: 3-D field, levels [0, nk); kind 1: 2-D field; kind 2: level `level` of a 3-D field
int egress_pack(Core& c, const char* name, int kind, int level, int x_interface, int y_interface, int nk, double* dst) {
  const Geo g = c.g;
  const double* src = c.F(name);
  if (!c.err.empty()) return 1;
  if (dst == nullptr) { c.err = "egress_pack: no destination"; return 1; }
  const int ni = g.nx + (x_interface ? 1 : 0), nj = g.ny + (y_interface ? 1 : 0);
  if (kind == 0 && (nk < 1 || nk > g.nz + 1)) { c.err = "egress_pack: level count out of range"; return 1; }
  if (kind == 2 && (level < 0 || level > g.nz)) { c.err = "egress_pack: level out of range"; return 1; }
  ++launch_counter();
#ifndef PACE_HOST_EMU
  if (kind == 0) {
    k_pack3<<<dim3((ni + 31) / 32, (nk + 31) / 32, g.nt * nj), dim3(32, 8), 0, c.stream>>>(g, src, dst, ni, nj, nk);
  } else {
    const long long stride_t = kind == 1 ? (long long)g.sk : g.st3;
    const double* s = kind == 1 ? src : src + (long long)level * g.sk;
    k_pack2<<<dim3((ni + 31) / 32, (nj + 31) / 32, g.nt), dim3(32, 8), 0, c.stream>>>(g, s, stride_t, dst, ni, nj);
  }
#else
  for (int t = 0; t < g.nt; ++t)
    for (int i = 0; i < ni; ++i)
      for (int j = 0; j < nj; ++j) {
        if (kind == 0) {
          for (int k = 0; k < nk; ++k)
            dst[(((long long)t * ni + i) * nj + j) * nk + k] =
                src[(long long)t * g.st3 + (long long)k * g.sk + (long long)(g.js + j) * g.sj + g.is + i];
        } else {
          const long long base = kind == 1 ? (long long)t * g.sk : (long long)t * g.st3 + (long long)level * g.sk;
          dst[((long long)t * ni + i) * nj + j] = src[base + (long long)(g.js + j) * g.sj + g.is + i];
        }
      }
#endif
  return 0;
}

// out(i, j) = RGRAV * sum_k q(i, j, k) * delp(i, j, k) over the npz layers, compute domain; out is a 2-D field
int column_integral(Core& c, const char* q_name, const char* delp_name, const char* out_name) {
  const Geo g = c.g;
  const double *q = c.F(q_name), *delp = c.F(delp_name);
  double* out = c.F(out_name);
  if (!c.err.empty()) return 1;
  const int n = g.nz;
  if (n > 128) { c.err = "column_integral: more than 128 levels (NumPy's blocked pairwise sum) is not implemented"; return 1; }
  constexpr double RGRAV = 1.0 / GRAV;
  launch_ij(c.stream, g, box_compute(g), [=] __device__(const Pt& p) {
    auto a = [&](int k) { return A(q, 0, 0, k) * A(delp, 0, 0, k); };
    double res;
    if (n < 8) {
      res = 0.0;
      for (int k = 0; k < n; ++k) res = res + a(k);
    } else {
      double r[8];
#pragma unroll
      for (int m = 0; m < 8; ++m) r[m] = a(m);
      int k = 8;
      for (; k < n - (n % 8); k += 8) {
#pragma unroll
        for (int m = 0; m < 8; ++m) r[m] = r[m] + a(k + m);
      }
      res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
      for (; k < n; ++k) res = res + a(k);
    }
    A20(out) = RGRAV * (0.0 + res);   // add.reduce starts from the identity
  });
  return 0;
}

}  // namespace pace
