#include "host_common.h"

#include <cudaTypedefs.h>

#include "../include/dtf_b200.h"

namespace dtf {

static thread_local std::string g_last_error;
static std::string g_last_error_global;
static std::mutex g_err_mu;

void set_last_error(const std::string& s) {
  g_last_error = s;
  std::lock_guard<std::mutex> lk(g_err_mu);
  g_last_error_global = s;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

struct TmapKey {
  const void* ptr;
  int rows, cols, box_rows, dev;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && box_rows == o.box_rows && dev == o.dev;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    h ^= (size_t)k.rows * 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
    h ^= (size_t)k.cols * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
    h ^= (size_t)(k.box_rows * 64 + k.dev) * 0x165667B19E3779F9ull + (h << 6) + (h >> 2);
    return h;
  }
};

const CUtensorMap* get_tmap_2d_f32_sw128(const float* ptr, int rows, int cols, int box_rows) {
  static std::unordered_map<TmapKey, CUtensorMap*, TmapKeyHash> cache;
  static std::mutex mu;
  int dev = 0;
  cudaGetDevice(&dev);
  TmapKey key{ptr, rows, cols, box_rows, dev};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled entry point not available");
    return nullptr;
  }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (cols % 4) != 0) {
    set_last_error("TMA operand must be 16-byte aligned with a row pitch that is a multiple of 16 bytes");
    return nullptr;
  }
  // CUtensorMap must be 64-byte aligned.
  void* mem = nullptr;
  if (posix_memalign(&mem, 64, sizeof(CUtensorMap)) != 0) return nullptr;
  CUtensorMap* tm = reinterpret_cast<CUtensorMap*>(mem);
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * sizeof(float)};
  cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[160];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed (%d) rows=%d cols=%d box_rows=%d", (int)r, rows, cols,
             box_rows);
    set_last_error(buf);
    free(mem);
    return nullptr;
  }
  cache.emplace(key, tm);
  return tm;
}

}  // namespace dtf

extern "C" int dtf_version(void) { return 100; }

extern "C" const char* dtf_last_error(void) {
  static thread_local std::string copy;
  std::lock_guard<std::mutex> lk(dtf::g_err_mu);
  copy = dtf::g_last_error_global;
  return copy.c_str();
}
