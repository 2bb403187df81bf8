// Channels-last BatchNorm for training, fused with the residual add and ReLU that follow it in ResNet blocks.
//
//   forward : y = act(gamma * (x - mean) * invstd + beta [+ residual])      x, y, residual: [rows, C] (NHWC flattened)
//   backward: dz = dy * (y > 0) [ReLU]; dgamma, dbeta; dx = gamma*invstd*(dz - mean(dz) - xhat*mean(dz*xhat)); dres = dz
//
// The reference has no BatchNorm (its models are Dense / Conv2D / MaxPool); this serves the ResNet
// configs of BASELINE.json.  The step profile of ResNet-50 (profiles/README.md) had PyTorch's four channels-last
// batch-norm kernels at 54 % of the step and the separate add / ReLU kernels at another 13 %.  Here the activation
// tensor is read once for the statistics and once for the transform, and the add + ReLU ride on the transform.
//
// Work split: a block is `tx` channel-vectors (16 bytes each) wide and 256/tx rows deep, so that a warp reads whole
// 128..512-byte row segments; per-block partial sums go to a scratch buffer and a second tiny kernel merges them in a
// fixed order (one warp per channel, fp64 accumulation of the partial sums), so results are deterministic.
#include "common.cuh"
#include "device_utils.cuh"

namespace dfb {

constexpr int kBnThreads = 256;

struct BnGeom {
  int vn;      // elements per 16-byte vector
  int cvec;    // C / vn
  int tx, ty;  // threads across channel vectors / across rows
  int gx, gy;  // grid
  long rpb;    // rows per block
};

static BnGeom bn_geom(long rows, int C, int dtype, int blocks_per_sm) {
  BnGeom g;
  g.vn = dtype == DT_BF16 ? 8 : 4;
  g.cvec = C / g.vn;
  g.tx = 1;
  while (g.tx * 2 <= g.cvec && g.tx < 32) g.tx *= 2;
  g.ty = kBnThreads / g.tx;
  g.gx = ceil_div(g.cvec, g.tx);
  long want = (long)num_sms() * blocks_per_sm / g.gx;
  if (want < 1) want = 1;
  long max_gy = ceil_div((int)((rows + g.ty - 1) / g.ty), 4);    // at least ~4 rows per thread
  if (max_gy < 1) max_gy = 1;
  g.gy = (int)(want < max_gy ? want : max_gy);
  if (g.gy > 65535) g.gy = 65535;
  g.rpb = (rows + g.gy - 1) / g.gy;
  g.rpb = (g.rpb + g.ty - 1) / g.ty * g.ty;
  g.gy = (int)((rows + g.rpb - 1) / g.rpb);
  return g;
}

__device__ __forceinline__ float param_at(const void* p, int dtype, int i) {
  return dtype == DT_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]) : reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void param_store(void* p, int dtype, int i, float v) {
  if (dtype == DT_BF16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(v);
  else reinterpret_cast<float*>(p)[i] = v;
}

// Block-level reduction over the row lanes: `vals` holds K floats per thread; the K*tx column results are written by
// `emit(channel_vector, k, sum)`.
template <int K, typename Emit>
__device__ __forceinline__ void bn_block_reduce(const float (&vals)[K], int tx, int ty, float* sh, Emit emit) {
#pragma unroll
  for (int k = 0; k < K; ++k) sh[k * kBnThreads + threadIdx.x] = vals[k];
  __syncthreads();
  for (int t = threadIdx.x; t < tx * K; t += kBnThreads) {
    const int c = t % tx, k = t / tx;
    float acc = 0.f;
    for (int r = 0; r < ty; ++r) acc += sh[k * kBnThreads + r * tx + c];
    emit(c, k, acc);
  }
}

// ---------------------------------------------------------------- forward statistics
template <typename T>
__global__ void __launch_bounds__(kBnThreads) bn_stats_kernel(const T* __restrict__ x, float* __restrict__ part, long rows, int C,
                                                              int tx, long rpb) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  __shared__ float sh[2 * VN * kBnThreads];
  const int ty = kBnThreads / tx;
  const int lc = threadIdx.x % tx, lr = threadIdx.x / tx;
  const int cv = blockIdx.x * tx + lc;
  const bool active = cv < C / VN;
  const long r0 = (long)blockIdx.y * rpb, r1 = min(rows, r0 + rpb);
  float acc[2 * VN];
#pragma unroll
  for (int i = 0; i < 2 * VN; ++i) acc[i] = 0.f;
  if (active) {
    const T* p = x + (long)cv * VN;
#pragma unroll 4
    for (long r = r0 + lr; r < r1; r += ty) {
      V v;
      v.raw = *reinterpret_cast<const uint4*>(p + r * C);
#pragma unroll
      for (int i = 0; i < VN; ++i) {
        const float f = v.get(i);
        acc[i] += f;
        acc[VN + i] += f * f;
      }
    }
  }
  float* out = part + (long)blockIdx.y * 2 * C;
  const int cbase = blockIdx.x * tx;
  bn_block_reduce<2 * VN>(acc, tx, ty, sh, [&](int c, int k, float s) {
    const int ch = (cbase + c) * VN + (k % VN);
    if (ch < C) out[(k / VN) * C + ch] = s;
  });
}

// One WARP per channel: the lanes stride over the per-block (sum, sum of squares) partials, accumulate in fp64 and
// butterfly-reduce (fixed order, deterministic); lane 0 derives mean / biased variance, updates the running statistics
// and emits the affine coefficients of the transform (a = gamma*invstd, b = beta - mean*a).
__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// A lane's share of the per-block partials of channel c; four independent loads per round (the loop is latency bound).
__device__ __forceinline__ void bn_gather_partials(const float* __restrict__ part, int gy, int C, int c, int lane, double& s0,
                                                   double& s1) {
  const long pitch = 2L * C;
  int b = lane;
  for (; b + 96 < gy; b += 128) {
    const float a0 = part[b * pitch + c], a1 = part[(b + 32) * pitch + c], a2 = part[(b + 64) * pitch + c],
                a3 = part[(b + 96) * pitch + c];
    const float q0 = part[b * pitch + C + c], q1 = part[(b + 32) * pitch + C + c], q2 = part[(b + 64) * pitch + C + c],
                q3 = part[(b + 96) * pitch + C + c];
    s0 += ((double)a0 + (double)a1) + ((double)a2 + (double)a3);
    s1 += ((double)q0 + (double)q1) + ((double)q2 + (double)q3);
  }
  for (; b < gy; b += 32) {
    s0 += (double)part[b * pitch + c];
    s1 += (double)part[b * pitch + C + c];
  }
}

__global__ void __launch_bounds__(256) bn_stats_finalize_kernel(const float* __restrict__ part, int gy, long rows, long rpb, int C,
                                                                const void* gamma, const void* beta, int pdt, float eps,
                                                                float momentum, float* running_mean, float* running_var,
                                                                float* save_mean, float* save_invstd, float* ab) {
  const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  bn_gather_partials(part, gy, C, c, lane, s, q);
  s = warp_sum_f64(s);
  q = warp_sum_f64(q);
  if (lane != 0) return;
  const double n = (double)rows;
  const double mean_d = s / n;
  double var_d = q / n - mean_d * mean_d;
  if (var_d < 0.0) var_d = 0.0;
  const float mean = (float)mean_d, var = (float)var_d;
  const float invstd = rsqrtf(var + eps);
  save_mean[c] = mean;
  save_invstd[c] = invstd;
  if (running_mean != nullptr) {
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (rows > 1 ? (float)(var_d * n / (n - 1.0)) : var);
  }
  const float a = param_at(gamma, pdt, c) * invstd;
  ab[c] = a;
  ab[C + c] = param_at(beta, pdt, c) - mean * a;
}

// Inference-mode coefficients from the running statistics.
__global__ void bn_eval_coef_kernel(const float* running_mean, const float* running_var, const void* gamma, const void* beta,
                                    int pdt, float eps, int C, float* ab) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float a = param_at(gamma, pdt, c) * rsqrtf(running_var[c] + eps);
  ab[c] = a;
  ab[C + c] = param_at(beta, pdt, c) - running_mean[c] * a;
}

// ---------------------------------------------------------------- forward transform (+ residual, + ReLU)
template <typename T>
__global__ void __launch_bounds__(kBnThreads) bn_apply_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                              const float* __restrict__ ab, T* __restrict__ y, long rows, int C,
                                                              int tx, long rpb, int relu) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  const int ty = kBnThreads / tx;
  const int lc = threadIdx.x % tx, lr = threadIdx.x / tx;
  const int cv = blockIdx.x * tx + lc;
  if (cv >= C / VN) return;
  float a[VN], b[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) {
    a[i] = ab[cv * VN + i];
    b[i] = ab[C + cv * VN + i];
  }
  const long r0 = (long)blockIdx.y * rpb, r1 = min(rows, r0 + rpb);
  const long off = (long)cv * VN;
#pragma unroll 4
  for (long r = r0 + lr; r < r1; r += ty) {
    V v, o;
    v.raw = *reinterpret_cast<const uint4*>(x + r * C + off);
    float f[VN];
#pragma unroll
    for (int i = 0; i < VN; ++i) f[i] = fmaf(v.get(i), a[i], b[i]);
    if (res != nullptr) {
      V rv;
      rv.raw = *reinterpret_cast<const uint4*>(res + r * C + off);
#pragma unroll
      for (int i = 0; i < VN; ++i) f[i] += rv.get(i);
    }
#pragma unroll
    for (int i = 0; i < VN; ++i) o.set(i, relu ? fmaxf(f[i], 0.f) : f[i]);
    *reinterpret_cast<uint4*>(y + r * C + off) = o.raw;
  }
}

// ---------------------------------------------------------------- backward reductions
template <typename T>
__global__ void __launch_bounds__(kBnThreads) bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                   const T* __restrict__ x, const float* __restrict__ mean,
                                                                   float* __restrict__ part, long rows, int C, int tx, long rpb,
                                                                   int relu) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  __shared__ float sh[2 * VN * kBnThreads];
  const int ty = kBnThreads / tx;
  const int lc = threadIdx.x % tx, lr = threadIdx.x / tx;
  const int cv = blockIdx.x * tx + lc;
  const bool active = cv < C / VN;
  const long r0 = (long)blockIdx.y * rpb, r1 = min(rows, r0 + rpb);
  float acc[2 * VN];
#pragma unroll
  for (int i = 0; i < 2 * VN; ++i) acc[i] = 0.f;
  if (active) {
    float mu[VN];
#pragma unroll
    for (int i = 0; i < VN; ++i) mu[i] = mean[cv * VN + i];
    const long off = (long)cv * VN;
#pragma unroll 2
    for (long r = r0 + lr; r < r1; r += ty) {
      V g, xv;
      g.raw = *reinterpret_cast<const uint4*>(dy + r * C + off);
      xv.raw = *reinterpret_cast<const uint4*>(x + r * C + off);
      if (relu) {
        V yv;
        yv.raw = *reinterpret_cast<const uint4*>(y + r * C + off);
#pragma unroll
        for (int i = 0; i < VN; ++i) {
          const float dz = yv.get(i) > 0.f ? g.get(i) : 0.f;
          acc[i] += dz;
          acc[VN + i] += dz * (xv.get(i) - mu[i]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < VN; ++i) {
          const float dz = g.get(i);
          acc[i] += dz;
          acc[VN + i] += dz * (xv.get(i) - mu[i]);
        }
      }
    }
  }
  float* out = part + (long)blockIdx.y * 2 * C;
  const int cbase = blockIdx.x * tx;
  bn_block_reduce<2 * VN>(acc, tx, ty, sh, [&](int c, int k, float s) {
    const int ch = (cbase + c) * VN + (k % VN);
    if (ch < C) out[(k / VN) * C + ch] = s;
  });
}

// coef: [0,C) k1 = gamma*invstd, [C,2C) k2 = sum(dz)/M, [2C,3C) k3 = invstd^2 * sum(dz*(x-mean))/M
__global__ void __launch_bounds__(256) bn_bwd_finalize_kernel(const float* __restrict__ part, int gy, long rows, int C,
                                                              const void* gamma, int pdt, const float* __restrict__ invstd,
                                                              void* dgamma, void* dbeta, float* coef) {
  const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;      // one warp per channel
  if (c >= C) return;
  double d1 = 0.0, d2 = 0.0;
  bn_gather_partials(part, gy, C, c, lane, d1, d2);
  const float s1 = (float)warp_sum_f64(d1), s2 = (float)warp_sum_f64(d2);
  if (lane != 0) return;
  const float is = invstd[c];
  param_store(dbeta, pdt, c, s1);
  param_store(dgamma, pdt, c, s2 * is);
  const float inv_m = 1.f / (float)rows;
  coef[c] = param_at(gamma, pdt, c) * is;
  coef[C + c] = s1 * inv_m;
  coef[2 * C + c] = s2 * is * is * inv_m;
}

template <typename T>
__global__ void __launch_bounds__(kBnThreads) bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                  const T* __restrict__ x, const float* __restrict__ mean,
                                                                  const float* __restrict__ coef, T* __restrict__ dx,
                                                                  T* __restrict__ dres, long rows, int C, int tx, long rpb,
                                                                  int relu) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  const int ty = kBnThreads / tx;
  const int lc = threadIdx.x % tx, lr = threadIdx.x / tx;
  const int cv = blockIdx.x * tx + lc;
  if (cv >= C / VN) return;
  float mu[VN], k1[VN], k2[VN], k3[VN];
#pragma unroll
  for (int i = 0; i < VN; ++i) {
    mu[i] = mean[cv * VN + i];
    k1[i] = coef[cv * VN + i];
    k2[i] = coef[C + cv * VN + i];
    k3[i] = coef[2 * C + cv * VN + i];
  }
  const long r0 = (long)blockIdx.y * rpb, r1 = min(rows, r0 + rpb);
  const long off = (long)cv * VN;
#pragma unroll 2
  for (long r = r0 + lr; r < r1; r += ty) {
    V g, xv, o, z;
    g.raw = *reinterpret_cast<const uint4*>(dy + r * C + off);
    xv.raw = *reinterpret_cast<const uint4*>(x + r * C + off);
    if (relu) {
      V yv;
      yv.raw = *reinterpret_cast<const uint4*>(y + r * C + off);
#pragma unroll
      for (int i = 0; i < VN; ++i) z.set(i, yv.get(i) > 0.f ? g.get(i) : 0.f);
    } else {
      z.raw = g.raw;
    }
#pragma unroll
    for (int i = 0; i < VN; ++i) o.set(i, k1[i] * (z.get(i) - k2[i] - (xv.get(i) - mu[i]) * k3[i]));
    *reinterpret_cast<uint4*>(dx + r * C + off) = o.raw;
    if (dres != nullptr) *reinterpret_cast<uint4*>(dres + r * C + off) = z.raw;
  }
}

static int bn_check(long rows, int C, int dtype, const void* a, const void* b) {
  DFB_REQUIRE(rows > 0 && C > 0, "batch_norm: empty problem");
  DFB_REQUIRE(dtype == DT_BF16 || dtype == DT_F32, "batch_norm: dtype");
  DFB_REQUIRE(C % (dtype == DT_BF16 ? 8 : 4) == 0, "batch_norm: C must be a multiple of %d", dtype == DT_BF16 ? 8 : 4);
  DFB_REQUIRE(((uintptr_t)a % 16) == 0 && ((uintptr_t)b % 16) == 0, "batch_norm: 16-byte alignment");
  return 0;
}

}  // namespace dfb

using namespace dfb;

// Row blocks of the partial-sum scratch: callers allocate `2 * C * dfb_bn_partial_blocks(rows, C, dtype)` floats.
DFB_API int dfb_bn_partial_blocks(long rows, int C, int dtype) { return bn_geom(rows, C, dtype, 4).gy; }

// Training forward.  x, res (nullable), y: [rows, C] of `dtype`; gamma/beta of `param_dtype`; running stats fp32 or null;
// save_mean / save_invstd: [C] fp32; part: scratch (see above); ab: [2*C] fp32 scratch.
DFB_API int dfb_bn_fwd_train(const void* x, const void* res, const void* gamma, const void* beta, float* running_mean,
                             float* running_var, void* y, float* save_mean, float* save_invstd, float* part, float* ab,
                             long rows, int C, float eps, float momentum, int relu, int dtype, int param_dtype,
                             cudaStream_t stream) {
  if (bn_check(rows, C, dtype, x, y)) return 1;
  const BnGeom g = bn_geom(rows, C, dtype, 4);
  dim3 grid(g.gx, g.gy);
  if (dtype == DT_BF16)
    bn_stats_kernel<__nv_bfloat16><<<grid, kBnThreads, 0, stream>>>((const __nv_bfloat16*)x, part, rows, C, g.tx, g.rpb);
  else
    bn_stats_kernel<float><<<grid, kBnThreads, 0, stream>>>((const float*)x, part, rows, C, g.tx, g.rpb);
  bn_stats_finalize_kernel<<<ceil_div(C, 8), 256, 0, stream>>>(part, g.gy, rows, g.rpb, C, gamma, beta, param_dtype, eps, momentum,
                                                                 running_mean, running_var, save_mean, save_invstd, ab);
  const BnGeom a = bn_geom(rows, C, dtype, 8);
  dim3 agrid(a.gx, a.gy);
  if (dtype == DT_BF16)
    bn_apply_kernel<__nv_bfloat16><<<agrid, kBnThreads, 0, stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)res, ab,
                                                                    (__nv_bfloat16*)y, rows, C, a.tx, a.rpb, relu);
  else
    bn_apply_kernel<float><<<agrid, kBnThreads, 0, stream>>>((const float*)x, (const float*)res, ab, (float*)y, rows, C, a.tx,
                                                            a.rpb, relu);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Inference forward (running statistics).
DFB_API int dfb_bn_fwd_eval(const void* x, const void* res, const void* gamma, const void* beta, const float* running_mean,
                            const float* running_var, void* y, float* ab, long rows, int C, float eps, int relu, int dtype,
                            int param_dtype, cudaStream_t stream) {
  if (bn_check(rows, C, dtype, x, y)) return 1;
  bn_eval_coef_kernel<<<ceil_div(C, 128), 128, 0, stream>>>(running_mean, running_var, gamma, beta, param_dtype, eps, C, ab);
  const BnGeom a = bn_geom(rows, C, dtype, 8);
  dim3 agrid(a.gx, a.gy);
  if (dtype == DT_BF16)
    bn_apply_kernel<__nv_bfloat16><<<agrid, kBnThreads, 0, stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)res, ab,
                                                                    (__nv_bfloat16*)y, rows, C, a.tx, a.rpb, relu);
  else
    bn_apply_kernel<float><<<agrid, kBnThreads, 0, stream>>>((const float*)x, (const float*)res, ab, (float*)y, rows, C, a.tx,
                                                            a.rpb, relu);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Training backward.  y is only read when relu != 0.  dres (nullable) receives dz = dy * (y > 0), the gradient of the
// residual input.  dgamma / dbeta: [C] of `param_dtype`.  part: as in the forward; coef: [3*C] fp32 scratch.
DFB_API int dfb_bn_bwd(const void* dy, const void* y, const void* x, const void* gamma, const float* save_mean,
                       const float* save_invstd, void* dx, void* dres, void* dgamma, void* dbeta, float* part, float* coef,
                       long rows, int C, int relu, int dtype, int param_dtype, cudaStream_t stream) {
  if (bn_check(rows, C, dtype, dy, dx)) return 1;
  DFB_REQUIRE(!relu || y != nullptr, "batch_norm backward: the ReLU mask needs y");
  const BnGeom g = bn_geom(rows, C, dtype, 4);
  dim3 grid(g.gx, g.gy);
  if (dtype == DT_BF16)
    bn_bwd_reduce_kernel<__nv_bfloat16><<<grid, kBnThreads, 0, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y,
                                                                        (const __nv_bfloat16*)x, save_mean, part, rows, C, g.tx,
                                                                        g.rpb, relu);
  else
    bn_bwd_reduce_kernel<float><<<grid, kBnThreads, 0, stream>>>((const float*)dy, (const float*)y, (const float*)x, save_mean,
                                                                part, rows, C, g.tx, g.rpb, relu);
  bn_bwd_finalize_kernel<<<ceil_div(C, 8), 256, 0, stream>>>(part, g.gy, rows, C, gamma, param_dtype, save_invstd, dgamma, dbeta,
                                                               coef);
  const BnGeom a = bn_geom(rows, C, dtype, 8);
  dim3 agrid(a.gx, a.gy);
  if (dtype == DT_BF16)
    bn_bwd_apply_kernel<__nv_bfloat16><<<agrid, kBnThreads, 0, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)y,
                                                                        (const __nv_bfloat16*)x, save_mean, coef,
                                                                        (__nv_bfloat16*)dx, (__nv_bfloat16*)dres, rows, C, a.tx,
                                                                        a.rpb, relu);
  else
    bn_bwd_apply_kernel<float><<<agrid, kBnThreads, 0, stream>>>((const float*)dy, (const float*)y, (const float*)x, save_mean,
                                                                coef, (float*)dx, (float*)dres, rows, C, a.tx, a.rpb, relu);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
