// Stand-alone probe of the 4-D fp64 TMA tile load used by pace_b200/csrc/tile.cuh: loads one box and checks it.
//   tma_probe <sj> <nj> <nk> <nt> <boxw> <boxh> <ci> <cj>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__global__ void k(const __grid_constant__ CUtensorMap m, double* out, int n, int ci, int cj, int ck, int ct) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(n * 8) : "memory");
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
        ::"r"(smem_u32(smem)), "l"(&m), "r"(ci), "r"(cj), "r"(ck), "r"(ct), "r"(smem_u32(&bar))
        : "memory");
  }
  asm volatile(
      "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(
          smem_u32(&bar)),
      "r"(0)
      : "memory");
  const double* s = (const double*)smem;
  for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = s[i];
}

int main(int argc, char** argv) {
  if (argc < 9) return 2;
  int sj = atoi(argv[1]), nj = atoi(argv[2]), nk = atoi(argv[3]), nt = atoi(argv[4]), bw = atoi(argv[5]), bh = atoi(argv[6]),
      ci = atoi(argv[7]), cj = atoi(argv[8]);
  size_t n = (size_t)sj * nj * nk * nt;
  std::vector<double> h(n);
  for (size_t i = 0; i < n; ++i) h[i] = (double)i + 0.25;
  double *d, *o;
  cudaMalloc(&d, n * 8);
  cudaMemcpy(d, h.data(), n * 8, cudaMemcpyHostToDevice);
  cudaMalloc(&o, (size_t)bw * bh * 8);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  typedef CUresult (*enc_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                            const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap m;
  cuuint64_t dims[4] = {(cuuint64_t)sj, (cuuint64_t)nj, (cuuint64_t)nk, (cuuint64_t)nt};
  cuuint64_t str[3] = {(cuuint64_t)sj * 8, (cuuint64_t)sj * nj * 8, (cuuint64_t)sj * nj * nk * 8};
  cuuint32_t box[4] = {(cuuint32_t)bw, (cuuint32_t)bh, 1, 1}, es[4] = {1, 1, 1, 1};
  int dt = argc > 9 ? atoi(argv[9]) : (int)CU_TENSOR_MAP_DATA_TYPE_FLOAT64;
  CUresult r = ((enc_t)fn)(&m, (CUtensorMapDataType)dt, 4, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode rc=%d\n", (int)r);
  if (r) return 1;
  int ck = nk - 1, ct = nt - 1;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bw * bh * 8);
  k<<<1, 128, bw * bh * 8>>>(m, o, bw * bh, ci, cj, ck, ct);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  if (e) return 1;
  std::vector<double> got((size_t)bw * bh);
  cudaMemcpy(got.data(), o, got.size() * 8, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int lj = 0; lj < bh; ++lj)
    for (int li = 0; li < bw; ++li) {
      int i = ci + li, j = cj + lj;
      double want = (i >= 0 && i < sj && j >= 0 && j < nj) ? h[(((size_t)ct * nk + ck) * nj + j) * sj + i] : 0.0;
      if (got[(size_t)lj * bw + li] != want) ++bad;
    }
  printf("mismatches=%d\n", bad);
  return bad != 0;
}
