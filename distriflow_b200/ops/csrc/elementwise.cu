// Memory-bound helpers around the GEMM/conv kernels: bias-gradient column sums, activation backward,
// NHWC max-pool forward/backward (reference layer MaxPooling2D in /root/reference/experiment/mnist/model.json),
// and the metrics kernel that replaces `model.evaluate(...).dataSync()`
// (/root/reference/src/common/models.ts:106-115) with a write into a device ring.
#include "common.cuh"
#include "device_utils.cuh"

namespace dfb {

// ---------------------------------------------------------------- column sums (bias gradient)
// out[n] += sum_m x[m, n]; block = 32 x 8 threads, each thread 16 bytes wide; rows are split over blockIdx.y.
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ x, float* __restrict__ out, int M, int N, long ld,
                                                     int rows_per_block) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  __shared__ float sh[8][32 * VN + 1];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int n = (blockIdx.x * 32 + tx) * VN;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float acc[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) acc[i] = 0.f;
  if (n + VN <= N) {
    for (int r = r0 + ty; r < r1; r += 8) {
      V v;
      v.raw = *reinterpret_cast<const uint4*>(x + (long)r * ld + n);
#pragma unroll
      for (int i = 0; i < VN; ++i) acc[i] += v.get(i);
    }
  } else {
    for (int r = r0 + ty; r < r1; r += 8)
      for (int i = 0; i < VN; ++i)
        if (n + i < N) acc[i] += to_f32<T>(x[(long)r * ld + n + i]);
  }
#pragma unroll
  for (int i = 0; i < VN; ++i) sh[ty][tx * VN + i] = acc[i];
  __syncthreads();
  for (int c = threadIdx.x; c < 32 * VN; c += 256) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += sh[j][c];
    int gn = blockIdx.x * 32 * VN + c;
    if (gn < N) atomicAdd(out + gn, s);
  }
}

// ---------------------------------------------------------------- activations
__device__ __forceinline__ float gelu_f(float x) {
  float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_grad_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t = tanhf(u);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k0 * (1.f + 3.f * k1 * x * x);
}

// mode 0: y = gelu(x)   mode 1: dx = dy * gelu'(x)   mode 2: dx = dy * (y > 0)   mode 3: y = relu(x)
template <typename T>
__global__ void __launch_bounds__(256) act_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out,
                                                  long n, int mode) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  const long nv = n / VN;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < nv; i += (long)gridDim.x * 256) {
    V va, vb, vo;
    va.raw = reinterpret_cast<const uint4*>(a)[i];
    if (mode == 1 || mode == 2) vb.raw = reinterpret_cast<const uint4*>(b)[i];
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      float x = va.get(j), r;
      if (mode == 0) r = gelu_f(x);
      else if (mode == 1) r = x * gelu_grad_f(vb.get(j));
      else if (mode == 2) r = vb.get(j) > 0.f ? x : 0.f;
      else r = fmaxf(x, 0.f);
      vo.set(j, r);
    }
    reinterpret_cast<uint4*>(out)[i] = vo.raw;
  }
  if (blockIdx.x == 0 && threadIdx.x < n - nv * VN) {
    long i = nv * VN + threadIdx.x;
    float x = to_f32<T>(a[i]), r;
    if (mode == 0) r = gelu_f(x);
    else if (mode == 1) r = x * gelu_grad_f(to_f32<T>(b[i]));
    else if (mode == 2) r = to_f32<T>(b[i]) > 0.f ? x : 0.f;
    else r = fmaxf(x, 0.f);
    out[i] = from_f32<T>(r);
  }
}

// ---------------------------------------------------------------- NHWC max-pool
template <typename T>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int H, int W, int C,
                                                          int OH, int OW, int kh, int kw, int sh, int sw, int pt, int pl) {
  const long total = (long)N * OH * OW * C;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    int c = (int)(i % C);
    long t = i / C;
    int ow = (int)(t % OW); t /= OW;
    int oh = (int)(t % OH);
    int n = (int)(t / OH);
    float m = -INFINITY;
    for (int r = 0; r < kh; ++r) {
      int ih = oh * sh - pt + r;
      if (ih < 0 || ih >= H) continue;
      for (int s = 0; s < kw; ++s) {
        int iw = ow * sw - pl + s;
        if (iw < 0 || iw >= W) continue;
        m = fmaxf(m, to_f32<T>(x[(((long)n * H + ih) * W + iw) * C + c]));
      }
    }
    y[i] = from_f32<T>(m);
  }
}

// dx must be zero-filled (fp32 accumulation buffer when windows overlap; same dtype otherwise).
template <typename T>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx,
                                                          int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh,
                                                          int sw, int pt, int pl) {
  const long total = (long)N * OH * OW * C;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    int c = (int)(i % C);
    long t = i / C;
    int ow = (int)(t % OW); t /= OW;
    int oh = (int)(t % OH);
    int n = (int)(t / OH);
    float m = -INFINITY;
    long arg = -1;
    for (int r = 0; r < kh; ++r) {
      int ih = oh * sh - pt + r;
      if (ih < 0 || ih >= H) continue;
      for (int s = 0; s < kw; ++s) {
        int iw = ow * sw - pl + s;
        if (iw < 0 || iw >= W) continue;
        long idx = (((long)n * H + ih) * W + iw) * C + c;
        float v = to_f32<T>(x[idx]);
        if (v > m) { m = v; arg = idx; }
      }
    }
    if (arg >= 0) {
      if constexpr (sizeof(T) == 4) atomicAdd(reinterpret_cast<float*>(dx) + arg, to_f32<T>(dy[i]));
      else dx[arg] = from_f32<T>(to_f32<T>(dx[arg]) + to_f32<T>(dy[i]));   // non-overlapping windows only (checked on host)
    }
  }
}

// ---------------------------------------------------------------- metrics: arg-max accuracy into a device ring
// slot[0] += number of rows whose arg-max matches the label; slot[1] += rows.  One warp per row.
template <typename T>
__global__ void __launch_bounds__(256) accuracy_kernel(const T* __restrict__ pred, const void* __restrict__ labels,
                                                       int label_kind, float* __restrict__ slot, int rows, int C, long ldp,
                                                       long ldl) {
  const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const T* p = pred + (long)warp * ldp;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int c = lane; c < C; c += 32) {
    float v = to_f32<T>(p[c]);
    if (v > best) { best = v; bi = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xffffffff, best, o);
    int oi = __shfl_xor_sync(0xffffffff, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  int lab;
  if (label_kind == 0) lab = (int)reinterpret_cast<const long long*>(labels)[warp];
  else if (label_kind == 1) lab = reinterpret_cast<const int*>(labels)[warp];
  else {   // dense labels of type T: arg-max of the label row
    const T* l = reinterpret_cast<const T*>(labels) + (long)warp * ldl;
    float lb = -INFINITY;
    int li = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      float v = to_f32<T>(l[c]);
      if (v > lb) { lb = v; li = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ob = __shfl_xor_sync(0xffffffff, lb, o);
      int oi = __shfl_xor_sync(0xffffffff, li, o);
      if (ob > lb || (ob == lb && oi < li)) { lb = ob; li = oi; }
    }
    lab = li;
  }
  if (lane == 0) {
    if (bi == lab) atomicAdd(slot, 1.f);
    atomicAdd(slot + 1, 1.f);
  }
}

}  // namespace dfb

using namespace dfb;

DFB_API int dfb_colsum(const void* x, float* out, int M, int N, long ld, int dtype, cudaStream_t stream) {
  DFB_REQUIRE(M > 0 && N > 0, "colsum: empty problem");
  DFB_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)N * sizeof(float), stream));
  const int vn = dtype == DT_BF16 ? 8 : 4;
  DFB_REQUIRE(dtype == DT_BF16 || dtype == DT_F32, "colsum: dtype");
  DFB_REQUIRE((ld % vn) == 0 && ((uintptr_t)x % 16) == 0, "colsum: pitch/alignment");
  const int bx = ceil_div(N, 32 * vn);
  int by = ceil_div(num_sms() * 4, bx);
  if (by > ceil_div(M, 64)) by = ceil_div(M, 64);
  if (by < 1) by = 1;
  const int rpb = ceil_div(M, by);
  dim3 grid(bx, ceil_div(M, rpb));
  if (dtype == DT_BF16) colsum_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)x, out, M, N, ld, rpb);
  else colsum_kernel<float><<<grid, 256, 0, stream>>>((const float*)x, out, M, N, ld, rpb);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// mode 0: out = gelu(a); 1: out = a * gelu'(b); 2: out = a * (b > 0); 3: out = relu(a).  Contiguous tensors.
DFB_API int dfb_act(const void* a, const void* b, void* out, long n, int dtype, int mode, cudaStream_t stream) {
  DFB_REQUIRE(n > 0, "act: empty");
  DFB_REQUIRE(((uintptr_t)a % 16) == 0 && ((uintptr_t)out % 16) == 0 && ((uintptr_t)b % 16) == 0, "act: alignment");
  long blocks = (n / 8 + 255) / 256;
  if (blocks > num_sms() * 16) blocks = num_sms() * 16;
  if (blocks < 1) blocks = 1;
  if (dtype == DT_BF16)
    act_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, stream>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b,
                                                              (__nv_bfloat16*)out, n, mode);
  else if (dtype == DT_F32)
    act_kernel<float><<<(int)blocks, 256, 0, stream>>>((const float*)a, (const float*)b, (float*)out, n, mode);
  else {
    set_error("act: dtype");
    return 2;
  }
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_maxpool_nhwc(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh, int sw,
                             int pt, int pl, int dtype, cudaStream_t stream) {
  long total = (long)N * OH * OW * C;
  DFB_REQUIRE(total > 0, "maxpool: empty");
  long blocks = (total + 255) / 256;
  if (blocks > num_sms() * 32) blocks = num_sms() * 32;
  if (dtype == DT_BF16)
    maxpool_fwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, H, W, C, OH,
                                                                      OW, kh, kw, sh, sw, pt, pl);
  else
    maxpool_fwd_kernel<float><<<(int)blocks, 256, 0, stream>>>((const float*)x, (float*)y, N, H, W, C, OH, OW, kh, kw, sh, sw, pt,
                                                              pl);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// dx is zero-filled here.  bf16 requires non-overlapping windows (stride >= kernel); fp32 handles any.
DFB_API int dfb_maxpool_nhwc_bwd(const void* x, const void* dy, void* dx, int N, int H, int W, int C, int OH, int OW, int kh,
                                 int kw, int sh, int sw, int pt, int pl, int dtype, cudaStream_t stream) {
  long total = (long)N * OH * OW * C;
  DFB_REQUIRE(total > 0, "maxpool_bwd: empty");
  DFB_REQUIRE(dtype == DT_F32 || (sh >= kh && sw >= kw), "maxpool_bwd: bf16 needs non-overlapping windows");
  DFB_CUDA_CHECK(cudaMemsetAsync(dx, 0, (size_t)N * H * W * C * (dtype == DT_BF16 ? 2 : 4), stream));
  long blocks = (total + 255) / 256;
  if (blocks > num_sms() * 32) blocks = num_sms() * 32;
  if (dtype == DT_BF16)
    maxpool_bwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                                      (__nv_bfloat16*)dx, N, H, W, C, OH, OW, kh, kw, sh, sw, pt,
                                                                      pl);
  else
    maxpool_bwd_kernel<float><<<(int)blocks, 256, 0, stream>>>((const float*)x, (const float*)dy, (float*)dx, N, H, W, C, OH, OW,
                                                              kh, kw, sh, sw, pt, pl);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// label_kind: 0 = int64 indices, 1 = int32 indices, 2 = dense rows of the prediction dtype.
// Accumulates into slot[0] (matches) and slot[1] (rows); the caller zeroes or rotates ring slots.
// ---------------------------------------------------------------------------------------------- embedding backward
// table_grad[idx[t], :] += dy[t, :] for every token t, added IN PLACE with vector reductions (`red.global.add.v4.bf16x2`,
// 16 bytes per operation) to a gradient buffer that already holds another contribution: the tied LM head's weight gradient
// is written first into the same bucket slice, the token embedding's gradient lands on top of it.  No dense [V, D]
// temporary, no sort (PyTorch's embedding_dense_backward + accumulate + copy into the bucket cost 0.6 ms of the GPT-2 step
// and moved 0.5 GB).  Repeated tokens add in an arbitrary order: the sum is rounded per addition in the table's dtype.
// One warp per token row.  Reference call site: the Flatten/Dense input of /root/reference/experiment/mnist/model.json has
// no embedding; GPT-2 (BASELINE.json config 5) does.
__device__ __forceinline__ void red_add_v4_bf16x2(void* p, uint4 v) {
  asm volatile("red.relaxed.gpu.global.add.noftz.v4.bf16x2 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void red_add_v4_f32(float* p, float4 v) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(256) embedding_bwd_scatter_kernel(const T* __restrict__ dy, const long long* __restrict__ idx,
                                                                    T* __restrict__ table_grad, long n_tok, int D, long V) {
  const int lane = threadIdx.x & 31;
  const long warp = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), n_warps = (long)gridDim.x * (blockDim.x >> 5);
  constexpr int VN = 16 / (int)sizeof(T);
  for (long t = warp; t < n_tok; t += n_warps) {
    const long row = idx[t];
    if (row < 0 || row >= V) continue;
    const T* src = dy + t * D;
    T* dst = table_grad + row * D;
    for (int c = lane * VN; c < D; c += 32 * VN) {
      if constexpr (sizeof(T) == 2) red_add_v4_bf16x2(dst + c, *reinterpret_cast<const uint4*>(src + c));
      else red_add_v4_f32(reinterpret_cast<float*>(dst + c), *reinterpret_cast<const float4*>(src + c));
    }
  }
}

// D a multiple of 8 (bf16) / 4 (fp32); dy rows contiguous (pitch D); idx int64.
DFB_API int dfb_embedding_bwd_scatter(const void* dy, const long long* idx, void* table_grad, long n_tok, int D, long V, int dtype,
                                      cudaStream_t stream) {
  DFB_REQUIRE(dtype == DT_BF16 || dtype == DT_F32, "embedding_bwd: dtype");
  DFB_REQUIRE(n_tok > 0 && D > 0 && D % (dtype == DT_BF16 ? 8 : 4) == 0, "embedding_bwd: D=%d must be a multiple of %d", D,
              dtype == DT_BF16 ? 8 : 4);
  DFB_REQUIRE(((uintptr_t)dy % 16) == 0 && ((uintptr_t)table_grad % 16) == 0, "embedding_bwd: 16-byte alignment");
  long blocks = (n_tok + 7) / 8;
  if (blocks > num_sms() * 16L) blocks = num_sms() * 16L;
  if (dtype == DT_BF16)
    embedding_bwd_scatter_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, stream>>>((const __nv_bfloat16*)dy, idx, (__nv_bfloat16*)table_grad,
                                                                                    n_tok, D, V);
  else
    embedding_bwd_scatter_kernel<float><<<(unsigned)blocks, 256, 0, stream>>>((const float*)dy, idx, (float*)table_grad, n_tok, D, V);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_accuracy(const void* pred, const void* labels, int label_kind, float* slot, int rows, int C, long ldp, long ldl,
                         int dtype, cudaStream_t stream) {
  DFB_REQUIRE(rows > 0 && C > 0, "accuracy: empty");
  int blocks = ceil_div(rows * 32, 256);
  if (dtype == DT_BF16)
    accuracy_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>((const __nv_bfloat16*)pred, labels, label_kind, slot, rows, C, ldp,
                                                              ldl);
  else if (dtype == DT_F32)
    accuracy_kernel<float><<<blocks, 256, 0, stream>>>((const float*)pred, labels, label_kind, slot, rows, C, ldp, ldl);
  else {
    set_error("accuracy: dtype");
    return 2;
  }
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------- im2col for thin-channel convolutions (ResNet stem: C_in = 3)
// patches[p, k] = X[n, oh*sh-pt+r, ow*sw-pl+s, ci] with k = (r*kw+s)*C+ci, zero for padding and for k >= kh*kw*C (row pitch
// Kpad, a multiple of 8).  With C_in < 64 the implicit-GEMM TMA path has no 128-byte channel rows to fetch, so the patch
// matrix is materialised once (bf16) and the tcgen05 Dense GEMM does forward and filter gradient on it.
namespace dfb {
// Division by a runtime constant as multiply-high + shift (host-computed magic; exact for dividends below 2^31): the patch
// index decomposition (pixel / tap / channel) was the whole cost of this kernel when done with hardware-emulated division.
struct FastDiv {
  uint32_t d, magic, shift;
  __device__ __forceinline__ uint32_t div(uint32_t n) const { return (__umulhi(n, magic) + n) >> shift; }
  __device__ __forceinline__ void divmod(uint32_t n, uint32_t& q, uint32_t& r) const {
    q = div(n);
    r = n - q * d;
  }
};
static FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  uint32_t shift = 0;
  while ((1ull << shift) < d) ++shift;
  f.shift = shift;
  f.magic = (uint32_t)((((1ull << 32) * ((1ull << shift) - d)) / d) + 1);
  return f;
}

__global__ void __launch_bounds__(256) im2col_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, int N, int H,
                                                     int W, int C, int OH, int OW, int kh, int kw, int sh, int sw, int pt, int pl,
                                                     int Kpad, FastDiv d_k8n, FastDiv d_ow, FastDiv d_oh, long total) {
  const int Kreal = kh * kw * C;
  __shared__ int tab[1024];                        // k -> (r << 24) | (s << 16) | ci, built once per block
  const bool vec = (C & 7) == 0;
  const int entries = vec ? (Kreal >> 3) : Kreal;  // vector path: one entry per 8 consecutive channels of a tap
  for (int e = threadIdx.x; e < entries; e += 256) {
    const int k = vec ? e * 8 : e;
    const int ci = k % C, rs = k / C;
    tab[e] = ((rs / kw) << 24) | ((rs % kw) << 16) | ci;
  }
  __syncthreads();
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    uint32_t pix, k8, t, ow, n, oh;
    d_k8n.divmod((uint32_t)i, pix, k8);
    d_ow.divmod(pix, t, ow);
    d_oh.divmod(t, n, oh);
    const int ih0 = (int)oh * sh - pt, iw0 = (int)ow * sw - pl;
    const __nv_bfloat16* xb = x + (((long)n * H + ih0) * W + iw0) * C;
    if (vec) {                    // 8 consecutive k = 8 consecutive channels of one tap: a single 16-byte load
      uint4 q = make_uint4(0, 0, 0, 0);
      if ((int)k8 * 8 < Kreal) {
        const int e = tab[k8];
        const int r = e >> 24, s2 = (e >> 16) & 0xff;
        const int ih = ih0 + r, iw = iw0 + s2;
        if (ih >= 0 && ih < H && iw >= 0 && iw < W) q = *reinterpret_cast<const uint4*>(xb + (r * W + s2) * C + (e & 0xffff));
      }
      *reinterpret_cast<uint4*>(out + (long)pix * Kpad + k8 * 8) = q;
      continue;
    }
    __align__(16) __nv_bfloat16 v[8];              // thin inputs (the 3-channel stem): element-wise gather
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = (int)k8 * 8 + j;
      __nv_bfloat16 val = __float2bfloat16(0.f);
      if (k < Kreal) {
        const int e = tab[k];
        const int r = e >> 24, s2 = (e >> 16) & 0xff;
        const int ih = ih0 + r, iw = iw0 + s2;
        if (ih >= 0 && ih < H && iw >= 0 && iw < W) val = xb[(r * W + s2) * C + (e & 0xffff)];
      }
      v[j] = val;
    }
    *reinterpret_cast<uint4*>(out + (long)pix * Kpad + k8 * 8) = *reinterpret_cast<const uint4*>(v);
  }
}
}  // namespace dfb

namespace dfb {
// dX[n, ih, iw, c..c+8) = sum over taps (r, s) with (ih+pt-r) % sh == 0, (iw+pl-s) % sw == 0 of
// dP[(n, (ih+pt-r)/sh, (iw+pl-s)/sw), (r*kw+s)*C + c]   — the gather form of col2im (no atomics), C % 8 == 0.
__global__ void __launch_bounds__(256) col2im_kernel(const __nv_bfloat16* __restrict__ dp, __nv_bfloat16* __restrict__ dx, int N, int H,
                                                     int W, int C, int OH, int OW, int kh, int kw, int sh, int sw, int pt, int pl,
                                                     int Kpad) {
  const int c8n = C >> 3;
  const long total = (long)N * H * W * c8n;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    const int c = (int)(i % c8n) * 8;
    long t = i / c8n;
    const int iw = (int)(t % W); t /= W;
    const int ih = (int)(t % H);
    const int n = (int)(t / H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < kh; ++r) {
      const int th = ih + pt - r;
      if (th < 0 || th % sh) continue;
      const int oh = th / sh;
      if (oh >= OH) continue;
      for (int s = 0; s < kw; ++s) {
        const int tw = iw + pl - s;
        if (tw < 0 || tw % sw) continue;
        const int ow = tw / sw;
        if (ow >= OW) continue;
        const uint4 q = *reinterpret_cast<const uint4*>(dp + (((long)n * OH + oh) * OW + ow) * Kpad + (r * kw + s) * C + c);
        const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&q);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += __bfloat162float(h[j]);
      }
    }
    __align__(16) __nv_bfloat16 o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = __float2bfloat16(acc[j]);
    *reinterpret_cast<uint4*>(dx + i * 8) = *reinterpret_cast<const uint4*>(o);
  }
}
}  // namespace dfb

DFB_API int dfb_col2im_nhwc(const void* dp, void* dx, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh, int sw,
                            int pt, int pl, int Kpad, cudaStream_t stream) {
  DFB_REQUIRE(C % 8 == 0 && Kpad % 8 == 0 && Kpad >= kh * kw * C, "col2im: C and Kpad must be multiples of 8");
  const long total = (long)N * H * W * (C / 8);
  long blocks = (total + 255) / 256;
  if (blocks > dfb::num_sms() * 32) blocks = dfb::num_sms() * 32;
  if (blocks < 1) blocks = 1;
  dfb::col2im_kernel<<<(int)blocks, 256, 0, stream>>>((const __nv_bfloat16*)dp, (__nv_bfloat16*)dx, N, H, W, C, OH, OW, kh, kw, sh,
                                                   sw, pt, pl, Kpad);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_im2col_nhwc(const void* x, void* out, int N, int H, int W, int C, int OH, int OW, int kh, int kw, int sh, int sw,
                            int pt, int pl, int Kpad, cudaStream_t stream) {
  DFB_REQUIRE(Kpad % 8 == 0 && Kpad >= kh * kw * C, "im2col: Kpad must be a multiple of 8 covering kh*kw*C");
  const long total = (long)N * OH * OW * (Kpad / 8);
  DFB_REQUIRE(total > 0 && total < (1L << 31), "im2col: empty or more than 2^31 patch vectors");
  DFB_REQUIRE(kh < 128 && kw < 256 && C < 65536 && (((C & 7) == 0 ? kh * kw * C / 8 : kh * kw * C) <= 1024),
              "im2col: filter too large for the tap table (kh*kw*C <= 1024 scalars or 8192 with C %% 8 == 0)");
  long blocks = (total + 255) / 256;
  if (blocks > dfb::num_sms() * 16) blocks = dfb::num_sms() * 16;
  dfb::im2col_kernel<<<(int)blocks, 256, 0, stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, N, H, W, C, OH, OW, kh, kw, sh,
                                                   sw, pt, pl, Kpad, dfb::make_fastdiv((uint32_t)(Kpad / 8)),
                                                   dfb::make_fastdiv((uint32_t)OW), dfb::make_fastdiv((uint32_t)OH), total);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
