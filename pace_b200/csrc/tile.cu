// Host side of the tiled kernels (tile.cuh): tensor maps over the caller's field allocations, per-device kernel
// attributes.
#include "tile.cuh"

#include <mutex>
#include <set>
#include <tuple>

namespace pace {

#ifndef PACE_HOST_EMU
namespace {
typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_fn_t encoder() {
  static encode_fn_t fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (encode_fn_t)p;
  }();
  return fn;
}
std::mutex g_attr_mutex;
std::set<std::pair<const void*, int>> g_attr_done;  // (kernel, device)
}  // namespace

const CUtensorMap* tile_tensor_map(Core& c, const double* field, int box_w, int box_h) {
  const auto key = std::make_tuple((const void*)field, box_w, box_h);
  auto it = c.tmaps.find(key);
  if (it != c.tmaps.end()) return (const CUtensorMap*)(((uintptr_t)it->second.get() + 63) & ~(uintptr_t)63);
  encode_fn_t enc = encoder();
  if (!enc) {
    c.err = "cuTensorMapEncodeTiled is not available from this driver";
    return nullptr;
  }
  const Geo& g = c.g;
  std::unique_ptr<unsigned char[]> mem(new unsigned char[sizeof(CUtensorMap) + 64]);
  CUtensorMap* m = (CUtensorMap*)(((uintptr_t)mem.get() + 63) & ~(uintptr_t)63);
  const cuuint64_t dims[4] = {(cuuint64_t)g.sj, (cuuint64_t)(g.ny + 2 * HALO + 1), (cuuint64_t)(g.nz + 1), (cuuint64_t)g.nt};
  const cuuint64_t strides[3] = {(cuuint64_t)g.sj * 8, (cuuint64_t)g.sk * 8, (cuuint64_t)g.st3 * 8};  // bytes, dims 1..3
  const cuuint32_t box[4] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, (void*)field, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    c.err = "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ") for a " + std::to_string(box_w) + " x " +
            std::to_string(box_h) + " box";
    return nullptr;
  }
  const CUtensorMap* out = m;
  c.tmaps[key] = std::move(mem);
  return out;
}

int func_attr_once(Core& c, const void* kernel, int smem_bytes) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { c.err = "cudaGetDevice failed"; return 1; }
  std::lock_guard<std::mutex> lock(g_attr_mutex);
  if (g_attr_done.count({kernel, dev})) return 0;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e != cudaSuccess) {
    c.err = std::string("cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed: ") + cudaGetErrorString(e);
    return 1;
  }
  g_attr_done.insert({kernel, dev});
  return 0;
}
#endif

}  // namespace pace
