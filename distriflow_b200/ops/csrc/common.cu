#include "common.cuh"

#include <stdarg.h>

#include <mutex>

namespace dfb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int g_sms[64];
static int g_smem[64];
static bool g_have[64];

static void query_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!g_have[dev]) {
    cudaDeviceGetAttribute(&g_sms[dev], cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&g_smem[dev], cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    g_have[dev] = true;
  }
}
int num_sms() {
  query_device();
  int dev = 0;
  cudaGetDevice(&dev);
  return g_sms[dev] > 0 ? g_sms[dev] : 148;
}
int max_smem_optin() {
  query_device();
  int dev = 0;
  cudaGetDevice(&dev);
  return g_smem[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

static void load_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) g_encode = reinterpret_cast<EncodeTiledFn>(fn);
}

// cuTensorMapEncodeTiled is a DRIVER call and needs a current context on the calling thread.  A fresh thread (the
// autograd backward thread of device 0 is the usual case: torch skips cudaSetDevice when the target equals the
// default device) has none until some runtime call binds one.  Bind the context of the device that owns `ptr`.
static void ensure_thread_context(const void* ptr) {
  static thread_local bool bound = false;
  if (bound) return;
  cudaPointerAttributes attr;
  int dev = 0;
  if (cudaPointerGetAttributes(&attr, ptr) == cudaSuccess && (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged))
    dev = attr.device;
  else {
    cudaGetLastError();
    cudaGetDevice(&dev);
  }
  cudaSetDevice(dev);
  bound = true;
}

int encode_tmap(CUtensorMap* out, const void* base, int elem, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box, const uint32_t* elem_strides, int swizzle_128b, int l2_promo_256) {
  std::call_once(g_encode_once, load_encode);
  ensure_thread_context(base);
  if (!g_encode) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return 3;
  }
  cuuint64_t gdims[5], gstr[5];
  cuuint32_t gbox[5], gel[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    gel[i] = elem_strides ? elem_strides[i] : 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = g_encode(out, elem == 0 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, rank,
                        const_cast<void*>(base), gdims, gstr, gbox, gel, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        swizzle_128b ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        l2_promo_256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rank %d dims %llu,%llu box %u,%u stride0 %llu base %p)", (int)r,
              rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0), box[0], rank > 1 ? box[1] : 0,
              (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), base);
    return 4;
  }
  return 0;
}

}  // namespace dfb

DFB_API const char* dfb_last_error(void) { return dfb::g_err; }

DFB_API int dfb_version(void) { return 100; }
