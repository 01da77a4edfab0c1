#include "host_common.h"

#include <cudaTypedefs.h>

#include <vector>

#include "../include/dtf_b200.h"

namespace dtf {

int g_pdl_enabled = 1;

static thread_local std::string g_last_error;
static std::string g_last_error_global;
static std::mutex g_err_mu;

void set_last_error(const std::string& s) {
  g_last_error = s;
  std::lock_guard<std::mutex> lk(g_err_mu);
  g_last_error_global = s;
}

int hw_num_sms() {
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

// SMs the compute kernels size their grids for.  While a parameter-server push runs beside the backward pass it owns
// a few SMs outright (its blocks take a whole register file, see ps_update.cu), and the persistent convolution
// kernels are launched with correspondingly fewer CTAs instead of queueing a second wave behind the push.
static int g_sm_reserved = 0;      // process-wide: the backward pass launches from the autograd thread
int num_sms() {
  const int n = hw_num_sms() - g_sm_reserved;
  return n > 8 ? n : 8;
}
}  // namespace dtf

extern "C" int dtf_reserve_sms(int n) {
  const int prev = dtf::g_sm_reserved;
  dtf::g_sm_reserved = n > 0 ? n : 0;
  return prev;
}

namespace dtf {

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

// The caches are keyed by address and shape.  Under a captured graph the addresses repeat and the caches stay small;
// an eager run with a churning allocator would grow them without bound, so a cache that passes TMAP_CACHE_CAP entries
// is emptied.  Descriptors are handed to kernels by value (__grid_constant__), but a launcher may still hold the
// pointer it was just given, so the memory of an emptied cache is released one overflow later, not at once.
constexpr size_t TMAP_CACHE_CAP = 8192;
template <typename Map>
static void bound_cache(Map& cache, std::vector<CUtensorMap*>& retired) {
  if (cache.size() < TMAP_CACHE_CAP) return;
  for (CUtensorMap* p : retired) free(p);
  retired.clear();
  for (auto& kv : cache)
    if (kv.second != nullptr) retired.push_back(kv.second);
  cache.clear();
}

const CUtensorMap* get_tmap_2d_f32_sw128(const float* ptr, int rows, int cols, int box_rows) {
  static std::unordered_map<TmapKey, CUtensorMap*, TmapKeyHash> cache;
  static std::mutex mu;
  int dev = 0;
  cudaGetDevice(&dev);
  TmapKey key{ptr, rows, cols, box_rows, dev};
  static std::vector<CUtensorMap*> retired;
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  bound_cache(cache, retired);
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

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeIm2colFn get_encode_im2col_fn() {
  static EncodeIm2colFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeIm2colFn>(p);
  }
  return fn;
}

struct Im2colKey {
  Im2colSpec s;
  int dev;
  bool operator==(const Im2colKey& o) const { return dev == o.dev && memcmp(&s, &o.s, sizeof(Im2colSpec)) == 0; }
};
struct Im2colKeyHash {
  size_t operator()(const Im2colKey& k) const {
    const unsigned char* b = reinterpret_cast<const unsigned char*>(&k.s);
    size_t h = 1469598103934665603ull ^ (size_t)k.dev;
    for (size_t i = 0; i < sizeof(Im2colSpec); ++i) h = (h ^ b[i]) * 1099511628211ull;
    return h;
  }
};

const CUtensorMap* get_tmap_im2col_f32(const Im2colSpec& spec) {
  static std::unordered_map<Im2colKey, CUtensorMap*, Im2colKeyHash> cache;
  static std::mutex mu;
  Im2colKey key;
  memset(&key, 0, sizeof(key));          // padding bytes take part in the comparison
  key.s.ptr = spec.ptr;
  key.s.N = spec.N; key.s.H = spec.H; key.s.W = spec.W; key.s.C = spec.C; key.s.ld = spec.ld;
  key.s.lower_w = spec.lower_w; key.s.lower_h = spec.lower_h; key.s.upper_w = spec.upper_w; key.s.upper_h = spec.upper_h;
  key.s.tw = spec.tw; key.s.th = spec.th; key.s.channels = spec.channels; key.s.nan_fill = spec.nan_fill;
  cudaGetDevice(&key.dev);
  static std::vector<CUtensorMap*> retired;
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  bound_cache(cache, retired);
  EncodeIm2colFn enc = get_encode_im2col_fn();
  if (!enc) return nullptr;
  const Im2colSpec& s = key.s;
  if ((reinterpret_cast<uintptr_t>(s.ptr) & 15) != 0 || (s.ld % 4) != 0) return nullptr;
  if (s.channels != 16 && s.channels != 32) return nullptr;
  if (s.lower_w < -128 || s.lower_w > 127 || s.lower_h < -128 || s.lower_h > 127) return nullptr;
  if (s.upper_w < -128 || s.upper_w > 127 || s.upper_h < -128 || s.upper_h > 127) return nullptr;
  if (s.tw < 1 || s.tw > 8 || s.th < 1 || s.th > 8) return nullptr;
  void* mem = nullptr;
  if (posix_memalign(&mem, 64, sizeof(CUtensorMap)) != 0) return nullptr;
  CUtensorMap* tm = reinterpret_cast<CUtensorMap*>(mem);
  cuuint64_t gdim[4] = {(cuuint64_t)s.C, (cuuint64_t)s.W, (cuuint64_t)s.H, (cuuint64_t)s.N};
  cuuint64_t gstride[3] = {(cuuint64_t)s.ld * 4ull, (cuuint64_t)s.ld * 4ull * (cuuint64_t)s.W,
                           (cuuint64_t)s.ld * 4ull * (cuuint64_t)s.W * (cuuint64_t)s.H};
  int lower[2] = {s.lower_w, s.lower_h};
  int upper[2] = {s.upper_w, s.upper_h};
  cuuint32_t estr[4] = {1u, (cuuint32_t)s.tw, (cuuint32_t)s.th, 1u};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(s.ptr), gdim, gstride, lower, upper,
                   (cuuint32_t)s.channels, 128u, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   s.channels == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   s.nan_fill ? CU_TENSOR_MAP_FLOAT_OOB_FILL_NAN_REQUEST_ZERO_FMA : CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    free(mem);
    cache.emplace(key, nullptr);         // remember: this shape uses the gather path
    return nullptr;
  }
  // Drivers up to CUDA 13.1 set a descriptor bit that breaks im2col loads of tensors smaller than 128 KB; the
  // vendored CUTLASS tree (cute/atom/copy_traits_sm90_im2col.hpp) clears it the same way.
  int drv = 0;
  if (cudaDriverGetVersion(&drv) == cudaSuccess && drv <= 13010) {
    const unsigned long long bytes = (unsigned long long)s.N * s.H * s.W * s.ld * 4ull;
    if (bytes < 131072ull) reinterpret_cast<uint64_t*>(tm)[1] &= ~(1ull << 21);
  }
  cache.emplace(key, tm);
  return tm;
}

}  // namespace dtf

extern "C" int dtf_version(void) { return 100; }

extern "C" int dtf_debug_set_pdl(int enabled) {
  dtf::g_pdl_enabled = enabled;
  return 0;
}

extern "C" const char* dtf_last_error(void) {
  static thread_local std::string copy;
  std::lock_guard<std::mutex> lk(dtf::g_err_mu);
  copy = dtf::g_last_error_global;
  return copy.c_str();
}

// CRC-32C (Castagnoli), slice-by-8, host only: checkpoint bundles and TFRecord framing checksum every block /
// record / tensor with it, and the interpreted byte loop manages ~10 MB/s.  crc = dtf_crc32c(data, n, 0) for a
// fresh checksum; pass a previous result as `crc` to continue over the next chunk.
namespace {
struct Crc32cTables {
  uint32_t t[8][256];
  Crc32cTables() {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
      t[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
      for (int s = 1; s < 8; ++s) t[s][i] = (t[s - 1][i] >> 8) ^ t[0][t[s - 1][i] & 0xff];
  }
};
}  // namespace

extern "C" uint32_t dtf_crc32c(const void* data, unsigned long long n, uint32_t crc) {
  static const Crc32cTables T;
  const unsigned char* p = static_cast<const unsigned char*>(data);
  uint32_t c = ~crc;
  while (n && (reinterpret_cast<uintptr_t>(p) & 7)) { c = T.t[0][(c ^ *p++) & 0xff] ^ (c >> 8); --n; }
  while (n >= 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    w ^= c;                       // little-endian hosts only (x86-64 / aarch64 as deployed)
    c = T.t[7][w & 0xff] ^ T.t[6][(w >> 8) & 0xff] ^ T.t[5][(w >> 16) & 0xff] ^ T.t[4][(w >> 24) & 0xff] ^
        T.t[3][(w >> 32) & 0xff] ^ T.t[2][(w >> 40) & 0xff] ^ T.t[1][(w >> 48) & 0xff] ^ T.t[0][(w >> 56) & 0xff];
    p += 8; n -= 8;
  }
  while (n--) c = T.t[0][(c ^ *p++) & 0xff] ^ (c >> 8);
  return ~c;
}
