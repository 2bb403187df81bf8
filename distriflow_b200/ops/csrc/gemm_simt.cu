// CUDA-core GEMM for the shapes the tensor-core path does not take: fp32 models (the MNIST experiment of the
// reference runs fp32), tiny N (Dense(10), Dense(5): /root/reference/experiment/mnist/mnist_server.ts:16-22,
// model.json) where a 128x128 tcgen05 tile would be almost all padding, and unaligned pitches.
//
//   D[m,n] = act(sum_k A(m,k) * B(n,k) + bias[n]),  A(m,k) = A[m*sa_m + k*sa_k],  B(n,k) = B[n*sb_n + k*sb_k]
//
// 64x64 tile, BK = 16, 256 threads, 4x4 outputs per thread, fp32 accumulate; optional split over K with
// atomicAdd into a zeroed fp32 output (used for weight gradients where K = batch can be long).
#include "common.cuh"
#include "device_utils.cuh"

namespace dfb {

__device__ __forceinline__ float act_apply(float x, int act) {
  if (act == ACT_RELU) return fmaxf(x, 0.f);
  if (act == ACT_GELU) {
    float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    return 0.5f * x * (1.f + tanhf(u));
  }
  return x;
}

template <typename TA, typename TB, typename TD>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const TA* __restrict__ A, const TB* __restrict__ B, TD* __restrict__ D, const void* __restrict__ bias,
                 int bias_dtype, int M, int N, int K, long sa_m, long sa_k, long sb_n, long sb_k, long ldd, int act,
                 int ksplit, int atomic_out) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tx = tid & 15, ty = tid >> 4;
  const int kchunk = ceil_div(ceil_div(K, BK), ksplit) * BK;
  const int kbeg = blockIdx.z * kchunk, kend = min(K, kbeg + kchunk);
  float acc[4][4] = {};
  const bool a_kcontig = (sa_k == 1), b_kcontig = (sb_k == 1);
  for (int k0 = kbeg; k0 < kend; k0 += BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int idx = tid + i * 256;
      int kk = a_kcontig ? (idx & 15) : (idx >> 6), mm = a_kcontig ? (idx >> 4) : (idx & 63);
      int gm = m0 + mm, gk = k0 + kk;
      As[kk][mm] = (gm < M && gk < kend) ? to_f32<TA>(A[gm * sa_m + gk * sa_k]) : 0.f;
      kk = b_kcontig ? (idx & 15) : (idx >> 6);
      int nn = b_kcontig ? (idx >> 4) : (idx & 63);
      int gn = n0 + nn;
      gk = k0 + kk;
      Bs[kk][nn] = (gn < N && gk < kend) ? to_f32<TB>(B[gn * sb_n + gk * sb_k]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float v = acc[i][j];
      if (atomic_out) {
        if constexpr (sizeof(TD) == 4) atomicAdd(reinterpret_cast<float*>(D) + gm * ldd + gn, v);
      } else {
        if (bias != nullptr)
          v += bias_dtype == DT_F32 ? reinterpret_cast<const float*>(bias)[gn]
                                    : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias)[gn]);
        D[gm * ldd + gn] = from_f32<TD>(act_apply(v, act));
      }
    }
  }
}

template <typename TA, typename TB, typename TD>
static int launch_simt(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K, long sa_m,
                       long sa_k, long sb_n, long sb_k, long ldd, int act, int ksplit, cudaStream_t stream) {
  dim3 grid(ceil_div(N, 64), ceil_div(M, 64), ksplit);
  gemm_simt_kernel<TA, TB, TD><<<grid, 256, 0, stream>>>((const TA*)A, (const TB*)B, (TD*)D, bias, bias_dtype, M, N, K, sa_m,
                                                         sa_k, sb_n, sb_k, ldd, act, ksplit, ksplit > 1 ? 1 : 0);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace dfb

using namespace dfb;

// General-stride GEMM on CUDA cores.  With ksplit > 1 the output must be fp32 and zero-filled by the caller
// (partials are combined with atomicAdd; bias/activation are then not applied).
DFB_API int dfb_gemm_simt(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K,
                          long sa_m, long sa_k, long sb_n, long sb_k, long ldd, int a_dtype, int b_dtype, int out_dtype,
                          int act, int ksplit, cudaStream_t stream) {
  DFB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_simt: empty problem");
  if (ksplit < 1) ksplit = 1;
  DFB_REQUIRE(ksplit == 1 || (out_dtype == DT_F32 && bias == nullptr && act == ACT_NONE),
              "gemm_simt: split-K needs a zeroed fp32 output and no epilogue");
#define DFB_SIMT_CASE(ta, tb, td, TA, TB, TD)                                                                            \
  if (a_dtype == ta && b_dtype == tb && out_dtype == td)                                                                 \
    return launch_simt<TA, TB, TD>(A, B, D, bias, bias_dtype, M, N, K, sa_m, sa_k, sb_n, sb_k, ldd, act, ksplit, stream);
  DFB_SIMT_CASE(DT_F32, DT_F32, DT_F32, float, float, float)
  DFB_SIMT_CASE(DT_BF16, DT_BF16, DT_BF16, __nv_bfloat16, __nv_bfloat16, __nv_bfloat16)
  DFB_SIMT_CASE(DT_BF16, DT_BF16, DT_F32, __nv_bfloat16, __nv_bfloat16, float)
  DFB_SIMT_CASE(DT_F32, DT_F32, DT_BF16, float, float, __nv_bfloat16)
#undef DFB_SIMT_CASE
  set_error("gemm_simt: unsupported dtype combination a=%d b=%d out=%d", a_dtype, b_dtype, out_dtype);
  return 2;
}
