// Host-side helpers shared by every translation unit of libdistriflow_b200.so: last-error string,
// CUDA error checks, device properties cache, and TMA descriptor encoding through the driver entry
// point (so the library does not link against libcuda and builds on a machine without a driver).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define DFB_API extern "C" __attribute__((visibility("default")))

namespace dfb {

void set_error(const char* fmt, ...);
int num_sms();                       // SM count of the current device (cached per device)
int max_smem_optin();                // max dynamic shared memory per block (opt-in)

enum DType : int { DT_F32 = 0, DT_BF16 = 1, DT_F16 = 2 };
enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_GELU = 2 };

#define DFB_CUDA_CHECK(expr)                                                                    \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      dfb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));     \
      return 1;                                                                                 \
    }                                                                                           \
  } while (0)

#define DFB_REQUIRE(cond, ...)             \
  do {                                     \
    if (!(cond)) {                         \
      dfb::set_error(__VA_ARGS__);         \
      return 2;                            \
    }                                      \
  } while (0)

// Encode a tiled TMA descriptor.  dims/strides innermost first; strides in bytes for dims 1..rank-1.
// elem: 0 = bf16, 1 = fp32.  swizzle: 0 = none, 3 = 128B.  Returns 0 on success.
int encode_tmap(CUtensorMap* out, const void* base, int elem, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box, const uint32_t* elem_strides, int swizzle_128b, int l2_promo_256 = 1);

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

}  // namespace dfb
