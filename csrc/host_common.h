// Host-side helpers shared by the launchers: error plumbing, device properties and the
// TMA tensor-map cache (cuTensorMapEncodeTiled is resolved through the runtime's driver
// entry-point API so the library never links libcuda directly and still loads on a CPU box).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <unordered_map>

namespace dtf {

void set_last_error(const std::string& s);

#define DTF_CHECK(expr)                                                                     \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      ::dtf::set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e));            \
      return static_cast<int>(_e);                                                          \
    }                                                                                       \
  } while (0)

#define DTF_REQUIRE(cond, msg)                                                              \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      ::dtf::set_last_error(std::string("requirement failed: ") + #cond + " — " + (msg));   \
      return -1;                                                                            \
    }                                                                                       \
  } while (0)

int num_sms();      // SMs available to compute kernels (hardware count minus dtf_reserve_sms)
int hw_num_sms();

// [rows][cols] fp32 row-major matrix, box = {32 cols (128 B), box_rows}, 128-byte swizzle.
// Returns nullptr (and sets last error) on failure.  Cached per (ptr, rows, cols, box_rows).
const CUtensorMap* get_tmap_2d_f32_sw128(const float* ptr, int rows, int cols, int box_rows);

// NHWC fp32 activation [N][H][W][ld >= C] as an im2col-mode tensor map: one load brings `channels` (16 -> 64-byte
// swizzle, 32 -> 128-byte swizzle) channels of 128 consecutive output pixels for one filter tap; the hardware
// walks the base pixels along W, H, N with traversal strides (tw, th), adds the tap offset given with the copy,
// and fills out-of-range pixels with zeros (nan_fill = false) or the NaN marker (true).
// lower/upper are the bounding-box corner offsets in [W, H] order.  nullptr when the shape is not encodable.
struct Im2colSpec {
  const float* ptr;
  int N, H, W, C, ld;
  int lower_w, lower_h, upper_w, upper_h;
  int tw, th;
  int channels;
  int nan_fill;
};
const CUtensorMap* get_tmap_im2col_f32(const Im2colSpec& s);

// Launch with the programmatic-dependent-launch attribute (see pdl_wait in umma.cuh); dtf_debug_set_pdl(0) turns
// the attribute off (plain stream order).
extern int g_pdl_enabled;
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl_enabled ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Shared-memory budget of the persistent tensor-core kernels (the most a CTA may take).
constexpr int SMEM_BUDGET = 227 * 1024;

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

}  // namespace dtf
