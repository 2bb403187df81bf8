// Fused softmax-cross-entropy, forward + backward in one launch.
//
// Replaces the reference's loss site `tf.losses.softmaxCrossEntropy(...).mean()` and its tape backward
// (/root/reference/src/common/models.ts:139).  One CTA per row:
//   pass 1  online max / sum-exp over the row (16-byte loads); the label logit is captured on the way;
//   pass 2  dlogits = (softmax - onehot) * scale, in place or into a separate buffer (the row was just
//           read, so at 100 KB per GPT-2 row pass 2 is an L2 hit on B200's 126 MB L2).
// Index labels (`int64/int32 [B]`, one-hot folded in) are the perf path; dense labels (`[B,C]` one-hot or
// soft) serve the MNIST configs; MODE_REFCOMPAT_* is the reference's swapped-argument loss with
// both gradient variants.
#include <stdlib.h>

#include "common.cuh"
#include "device_utils.cuh"

namespace dfb {

__device__ __forceinline__ void online_combine(float& m, float& s, float m2, float s2) {
  float mx = fmaxf(m, m2);
  s = (m == -INFINITY ? 0.f : s * __expf(m - mx)) + (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - mx));
  m = mx;
}

template <int kThreads>
__device__ __forceinline__ void block_online_reduce(float& m, float& s, float* sh_m, float* sh_s) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float m2 = __shfl_xor_sync(0xffffffff, m, o), s2 = __shfl_xor_sync(0xffffffff, s, o);
    online_combine(m, s, m2, s2);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sh_m[w] = m; sh_s[w] = s; }
  __syncthreads();
  if (w == 0) {
    m = l < kThreads / 32 ? sh_m[l] : -INFINITY;
    s = l < kThreads / 32 ? sh_s[l] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float m2 = __shfl_xor_sync(0xffffffff, m, o), s2 = __shfl_xor_sync(0xffffffff, s, o);
      online_combine(m, s, m2, s2);
    }
    if (l == 0) { sh_m[0] = m; sh_s[0] = s; }
  }
  __syncthreads();
  m = sh_m[0];
  s = sh_s[0];
  __syncthreads();
}

// ------------------------------------------------------------------ index labels
// `logits` and `dlogits` may alias (in-place gradient).
// kCache: the row (up to ~110 KB: a 50304-wide bf16 vocabulary is 100.6 KB) is parked in shared memory by the first pass, so
// the gradient pass does not go back to L2 / HBM for it.  Kept as an option; measured slower than the L2-served second pass.
template <typename T, typename LabelT, int kThreads, bool kCache>
__global__ void __launch_bounds__(kThreads)
softmax_ce_index_kernel(const T* logits, const LabelT* __restrict__ labels, float* __restrict__ loss_rows, T* dlogits, int C,
                        long ld, long ldg, float scale, long ignore_index) {
  extern __shared__ uint4 row_cache[];
  __shared__ float sh_m[32], sh_s[32];
  __shared__ float sh_label;
  const long row = blockIdx.x;
  const T* x = logits + row * ld;
  const long label = (long)labels[row];
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  const bool vec_ok = ((ld % VN) == 0) && ((ldg % VN) == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0) &&
                      ((reinterpret_cast<uintptr_t>(dlogits) & 15) == 0);
  const int cvec = vec_ok ? (C / VN) * VN : 0;
  float m = -INFINITY, s = 0.f;
  auto consume = [&](const V& v, int c) {
    if constexpr (kCache) row_cache[c / VN] = v.raw;
    float lm = v.get(0);
#pragma unroll
    for (int i = 1; i < VN; ++i) lm = fmaxf(lm, v.get(i));
    float ls = 0.f;
#pragma unroll
    for (int i = 0; i < VN; ++i) ls += __expf(v.get(i) - lm);
    online_combine(m, s, lm, ls);
    if (label >= c && label < c + VN) sh_label = to_f32<T>(x[label]);   // scalar reload: keeps v in registers
  };
  {
    constexpr int U = 4;                            // 4 x 16-byte loads in flight per thread: the row pass is latency bound
    constexpr int kStep = kThreads * VN;
    int c = threadIdx.x * VN;
    for (; c + (U - 1) * kStep < cvec; c += U * kStep) {
      V v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) v[u].raw = *reinterpret_cast<const uint4*>(x + c + u * kStep);
#pragma unroll
      for (int u = 0; u < U; ++u) consume(v[u], c + u * kStep);
    }
    for (; c < cvec; c += kStep) {
      V v;
      v.raw = *reinterpret_cast<const uint4*>(x + c);
      consume(v, c);
    }
  }
  for (int c = cvec + threadIdx.x; c < C; c += kThreads) {
    float xv = to_f32<T>(x[c]);
    online_combine(m, s, xv, 1.f);
    if (c == label) sh_label = xv;
  }
  block_online_reduce<kThreads>(m, s, sh_m, sh_s);       // contains the barriers that publish sh_label
  const float lse = m + __logf(s);
  const bool ignored = (label == ignore_index) || label < 0 || label >= C;
  if (threadIdx.x == 0) loss_rows[row] = ignored ? 0.f : (lse - sh_label);
  if (dlogits == nullptr) return;
  T* g = dlogits + row * ldg;
  const float sc = ignored ? 0.f : scale;
#pragma unroll 4
  for (int c = threadIdx.x * VN; c < cvec; c += kThreads * VN) {
    V v, o;
    if constexpr (kCache) v.raw = row_cache[c / VN];       // written by this same thread in the first pass
    else v.raw = *reinterpret_cast<const uint4*>(x + c);
#pragma unroll
    for (int i = 0; i < VN; ++i) {
      float p = __expf(v.get(i) - lse);
      o.set(i, (p - ((c + i) == label ? 1.f : 0.f)) * sc);
    }
    *reinterpret_cast<uint4*>(g + c) = o.raw;
  }
  for (int c = cvec + threadIdx.x; c < C; c += kThreads) {
    float p = __expf(to_f32<T>(x[c]) - lse);
    g[c] = from_f32<T>((p - (c == label ? 1.f : 0.f)) * sc);
  }
}

// ------------------------------------------------------------------ dense labels / reference-compat
enum CeMode : int { MODE_STANDARD = 0, MODE_REFCOMPAT_CUSTOM = 1, MODE_REFCOMPAT_TRUE = 2 };

// MODE_STANDARD : z = logits (gradient target), y = labels:
//                 loss = -sum_c y*log_softmax(z);  dz = (softmax(z)*sum(y) - y) * scale
// MODE_REFCOMPAT: z = dataset one-hot (plays "logits"), y = model output p (gradient target):
//                 loss = -sum_c p*log_softmax(z);  dp = (p - softmax(z))*scale (custom) | -log_softmax(z)*scale (true)
template <typename T, int kThreads>
__global__ void __launch_bounds__(kThreads)
softmax_ce_dense_kernel(const T* z, const T* y, float* __restrict__ loss_rows, T* grad, int C, long ldz, long ldy, long ldg,
                        float scale, int mode) {
  __shared__ float sh_m[32], sh_s[32];
  __shared__ float sh_r[32];
  const long row = blockIdx.x;
  const T* zr = z + row * ldz;
  const T* yr = y + row * ldy;
  float m = -INFINITY, s = 0.f;
  for (int c = threadIdx.x; c < C; c += kThreads) online_combine(m, s, to_f32<T>(zr[c]), 1.f);
  block_online_reduce<kThreads>(m, s, sh_m, sh_s);
  const float lse = m + __logf(s);
  float part = 0.f, ysum = 0.f;
  for (int c = threadIdx.x; c < C; c += kThreads) {
    float yv = to_f32<T>(yr[c]);
    part += -yv * (to_f32<T>(zr[c]) - lse);
    ysum += yv;
  }
  part = warp_sum(part);
  ysum = warp_sum(ysum);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sh_r[w] = part; sh_m[w] = ysum; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f, ys = 0.f;
    for (int i = 0; i < kThreads / 32; ++i) { t += sh_r[i]; ys += sh_m[i]; }
    loss_rows[row] = t;
    sh_s[0] = ys;
  }
  __syncthreads();
  if (grad == nullptr) return;
  ysum = sh_s[0];
  T* g = grad + row * ldg;
  for (int c = threadIdx.x; c < C; c += kThreads) {
    float zl = to_f32<T>(zr[c]) - lse, yv = to_f32<T>(yr[c]);
    float gv;
    if (mode == MODE_STANDARD) gv = (__expf(zl) * ysum - yv) * scale;
    else if (mode == MODE_REFCOMPAT_CUSTOM) gv = (yv - __expf(zl)) * scale;
    else gv = -zl * scale;
    g[c] = from_f32<T>(gv);
  }
}

}  // namespace dfb

using namespace dfb;

// Index-label softmax-CE.  logits [rows, C] with pitch ld; labels int64 (label_dtype 0) or int32 (1);
// loss_rows fp32 [rows]; dlogits may be null (forward only) or alias logits.  scale multiplies the gradient
// (pass upstream_grad / rows for a mean loss).
DFB_API int dfb_softmax_ce_index(const void* logits, const void* labels, int label_dtype, float* loss_rows, void* dlogits,
                                 int rows, int C, long ld, long ldg, int dtype, float scale, long ignore_index,
                                 cudaStream_t stream) {
  DFB_REQUIRE(rows > 0 && C > 0, "softmax_ce: empty problem");
  constexpr int T = 512;
  // row cache: only when a gradient is wanted, the vector path applies and two CTAs still fit on an SM
  const size_t esz = dtype == DT_BF16 ? 2 : 4;
  const int vn = (int)(16 / esz);
  const bool vec_ok = (ld % vn) == 0 && (ldg % vn) == 0 && ((uintptr_t)logits & 15) == 0 && ((uintptr_t)dlogits & 15) == 0;
  const size_t cache_bytes = (size_t)(C / vn) * 16;
  // Measured on B200 (scripts/bench_ce.py, 8192 x 50304 bf16): 426.8 us with the row cache (two 100 KB CTAs per SM leave too
  // few loads in flight) against 325.6 us without it, where the second pass is served by L2: off unless DFB_CE_CACHE is set.
  static const bool want_cache = getenv("DFB_CE_CACHE") != nullptr;
  const bool cache = want_cache && dlogits != nullptr && vec_ok && cache_bytes >= 16 * 1024 && cache_bytes <= 110 * 1024;
#define DFB_CE_CASE(dt, lt, TT, LT)                                                                                         \
  if (dtype == dt && label_dtype == lt) {                                                                                   \
    if (cache) {                                                                                                            \
      auto kern = softmax_ce_index_kernel<TT, LT, T, true>;                                                                 \
      static bool configured = false;                                                                                       \
      if (!configured) {                                                                                                    \
        DFB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));                \
        configured = true;                                                                                                  \
      }                                                                                                                     \
      kern<<<rows, T, cache_bytes, stream>>>((const TT*)logits, (const LT*)labels, loss_rows, (TT*)dlogits, C, ld, ldg,     \
                                             scale, ignore_index);                                                          \
    } else {                                                                                                                \
      softmax_ce_index_kernel<TT, LT, T, false><<<rows, T, 0, stream>>>((const TT*)logits, (const LT*)labels, loss_rows,    \
                                                                        (TT*)dlogits, C, ld, ldg, scale, ignore_index);    \
    }                                                                                                                       \
    DFB_CUDA_CHECK(cudaGetLastError());                                                                                     \
    return 0;                                                                                                               \
  }
  DFB_CE_CASE(DT_BF16, 0, __nv_bfloat16, long long)
  DFB_CE_CASE(DT_BF16, 1, __nv_bfloat16, int)
  DFB_CE_CASE(DT_F32, 0, float, long long)
  DFB_CE_CASE(DT_F32, 1, float, int)
#undef DFB_CE_CASE
  set_error("softmax_ce_index: unsupported dtype %d / label dtype %d", dtype, label_dtype);
  return 2;
}

// Dense-label softmax-CE and the reference-compat loss (mode: 0 standard, 1 refcompat custom grad, 2 refcompat true grad).
DFB_API int dfb_softmax_ce_dense(const void* z, const void* y, float* loss_rows, void* grad, int rows, int C, long ldz,
                                 long ldy, long ldg, int dtype, float scale, int mode, cudaStream_t stream) {
  DFB_REQUIRE(rows > 0 && C > 0, "softmax_ce: empty problem");
  DFB_REQUIRE(mode >= 0 && mode <= 2, "softmax_ce_dense: mode");
  if (dtype == DT_F32) {
    if (C <= 1024)
      softmax_ce_dense_kernel<float, 128><<<rows, 128, 0, stream>>>((const float*)z, (const float*)y, loss_rows, (float*)grad,
                                                                    C, ldz, ldy, ldg, scale, mode);
    else
      softmax_ce_dense_kernel<float, 512><<<rows, 512, 0, stream>>>((const float*)z, (const float*)y, loss_rows, (float*)grad,
                                                                    C, ldz, ldy, ldg, scale, mode);
  } else if (dtype == DT_BF16) {
    softmax_ce_dense_kernel<__nv_bfloat16, 256><<<rows, 256, 0, stream>>>(
        (const __nv_bfloat16*)z, (const __nv_bfloat16*)y, loss_rows, (__nv_bfloat16*)grad, C, ldz, ldy, ldg, scale, mode);
  } else {
    set_error("softmax_ce_dense: unsupported dtype %d", dtype);
    return 2;
  }
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
