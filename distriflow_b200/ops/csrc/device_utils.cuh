// Small device helpers shared by the elementwise / reduction kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace dfb {

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffff, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffff, v, o));
  return v;
}

// 16 bytes of T
template <typename T>
struct Vec16;
template <>
struct Vec16<__nv_bfloat16> {
  static constexpr int kN = 8;
  uint4 raw;
  __device__ float get(int i) const { return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(&raw)[i]); }
  __device__ void set(int i, float v) { reinterpret_cast<__nv_bfloat16*>(&raw)[i] = __float2bfloat16(v); }
};
template <>
struct Vec16<float> {
  static constexpr int kN = 4;
  uint4 raw;
  __device__ float get(int i) const { return reinterpret_cast<const float*>(&raw)[i]; }
  __device__ void set(int i, float v) { reinterpret_cast<float*>(&raw)[i] = v; }
};

}  // namespace dfb
