// Error channel + TMA tensor-map encoder for libppb200 (see host_common.h).
#include "host_common.h"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

namespace ppb {

static thread_local char g_err[1024] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

static int g_reserved_sms = -1;

// SMs the persistent kernels may occupy: the device's SM count minus a reserve left free for concurrently running
// communication kernels (NCCL P2P CTAs that spin while a posted receive waits for its sender would otherwise push
// one persistent CTA into a second wave).  The reserve comes from ppb_set_reserved_sms() or $PPB_RESERVED_SMS.
int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  if (g_reserved_sms < 0) {
    const char* e = getenv("PPB_RESERVED_SMS");
    g_reserved_sms = e ? atoi(e) : 0;
    if (g_reserved_sms < 0) g_reserved_sms = 0;
  }
  const int n = cached[dev] - g_reserved_sms;
  return n > 8 ? n : 8;
}

static const unsigned long long* g_seed_off[64] = {nullptr};
const unsigned long long* seed_offset_ptr() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  return g_seed_off[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

bool make_tmap(CUtensorMap* out, CUtensorMapDataType dtype, int rank, const void* ptr, const uint64_t* dims,
               const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled driver entry point not available");
    return false;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bdim[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = fn(out, dtype, (cuuint32_t)rank, const_cast<void*>(ptr), gdim, gstr, bdim, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (CUresult %d): rank=%d ptr=%p dims=[%llu,%llu,..] stride1=%llu box=[%u,%u,..]",
              (int)r, rank, ptr, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
              (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), box[0], rank > 1 ? box[1] : 0);
    return false;
  }
  return true;
}

}  // namespace ppb

extern "C" const char* ppb_last_error() { return ppb::get_error(); }
extern "C" int ppb_version() { return 100; }
extern "C" int ppb_set_reserved_sms(int n) {
  ppb::g_reserved_sms = n < 0 ? 0 : n;
  return 0;
}
extern "C" int ppb_set_dropout_offset_ptr(const unsigned long long* p) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    ppb::set_error("ppb_set_dropout_offset_ptr: no current device");
    return -1;
  }
  ppb::g_seed_off[dev] = p;
  return 0;
}
