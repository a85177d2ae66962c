// Host-side helpers shared by every .cu translation unit of libppb200: error reporting for the C ABI,
// TMA tensor-map construction through the driver entry point (no link-time libcuda dependency).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace ppb {

void set_error(const char* fmt, ...);
const char* get_error();
int sm_count();  // SM count of the current device (cached per device)
// device pointer to a step-varying 64-bit value that every dropout-using kernel adds to its seed (may be null);
// lets CUDA graphs replay with fresh masks without re-capturing.
const unsigned long long* seed_offset_ptr();

// inner-most dimension first, like the driver API. strides are in bytes for dims 1..rank-1.
bool make_tmap(CUtensorMap* out, CUtensorMapDataType dtype, int rank, const void* ptr, const uint64_t* dims,
               const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz);

inline bool make_tmap_2d(CUtensorMap* out, CUtensorMapDataType dtype, const void* ptr, uint64_t d0, uint64_t d1,
                         uint64_t stride1_bytes, uint32_t b0, uint32_t b1,
                         CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  uint64_t dims[2] = {d0, d1};
  uint64_t strides[1] = {stride1_bytes};
  uint32_t box[2] = {b0, b1};
  return make_tmap(out, dtype, 2, ptr, dims, strides, box, swz);
}

}  // namespace ppb

#define PPB_CUDA_OK(expr)                                                                  \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ppb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return -1;                                                                           \
    }                                                                                      \
  } while (0)

#define PPB_REQUIRE(cond, ...)     \
  do {                             \
    if (!(cond)) {                 \
      ppb::set_error(__VA_ARGS__); \
      return -2;                   \
    }                              \
  } while (0)
