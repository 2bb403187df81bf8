// Fused residual-add + LayerNorm (forward and backward) and a 128-byte-row permutation kernel for attention layouts.
//
// Not in the reference (its models have neither normalisation nor attention); they belong to the GPT-2 / ResNet
// workloads named in BASELINE.json.  They exist because the step profile (profiles/step_profile_r1.md) showed the
// PyTorch LayerNorm backward (GammaBetaBackward) and the cat/copy kernels around attention costing more than a
// third of the GEMM time.  One warp per row, 16-byte loads, fp32 statistics.
//   forward : s = x (+ r);  y = (s - mean) * rstd * gamma + beta;  mean/rstd saved
//   backward: dx = (ds) + rstd * (g - mean_row(g) - xhat * mean_row(g * xhat)),  g = dy * gamma
//             dgamma = sum_rows dy * xhat,  dbeta = sum_rows dy   (per-block partials, then a reduce kernel)
#include "common.cuh"
#include "device_utils.cuh"
#include "ptx_sm100.cuh"

namespace dfb {

constexpr int kLnMaxVec = 8;      // up to 8 x 16 B per lane: D <= 2048 for bf16, 1024 for fp32

template <typename T, int NV>
__global__ void __launch_bounds__(128) add_ln_fwd_kernel(const T* __restrict__ x, const T* __restrict__ r, const T* __restrict__ gamma,
                                                         const T* __restrict__ beta, T* __restrict__ s_out, T* __restrict__ y,
                                                         float* __restrict__ mean, float* __restrict__ rstd, int rows, int D,
                                                         float eps) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const T* xr = x + (long)row * D;
  float v[NV][VN];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * VN;
    if (c < D) {
      V a;
      a.raw = *reinterpret_cast<const uint4*>(xr + c);
      if (r != nullptr) {
        V b;
        b.raw = *reinterpret_cast<const uint4*>(r + (long)row * D + c);
#pragma unroll
        for (int j = 0; j < VN; ++j) a.set(j, a.get(j) + b.get(j));     // rounded to T: the residual stream is stored in T
        *reinterpret_cast<uint4*>(s_out + (long)row * D + c) = a.raw;
      }
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        v[i][j] = a.get(j);
        sum += v[i][j];
      }
    }
  }
  const float mu = warp_sum(sum) / (float)D;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * VN;
    if (c < D) {
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        const float d = v[i][j] - mu;
        sq += d * d;
      }
    }
  }
  const float rs = rsqrtf(warp_sum(sq) / (float)D + eps);
  if (lane == 0) {
    mean[row] = mu;
    rstd[row] = rs;
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * VN;
    if (c < D) {
      V g, b, o;
      g.raw = *reinterpret_cast<const uint4*>(gamma + c);
      b.raw = *reinterpret_cast<const uint4*>(beta + c);
#pragma unroll
      for (int j = 0; j < VN; ++j) o.set(j, (v[i][j] - mu) * rs * g.get(j) + b.get(j));
      *reinterpret_cast<uint4*>(y + (long)row * D + c) = o.raw;
    }
  }
}

// Each block walks rows blockIdx.x*4 + warp, + gridDim.x*4, ... ; per-lane dgamma/dbeta partials live in registers and
// are combined across the 4 warps through shared memory, then written to part[blockIdx.x][2][D].
// All three row operands (dy, s, ds) are fetched up front as packed 16-byte vectors (9 loads in flight per lane at
// D = 768) and unpacked twice, which keeps the register count low enough for 4 blocks per SM.
template <typename T, int NV, bool kSum>
__global__ void __launch_bounds__(128, kSum ? 3 : 4) add_ln_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ ds, const T* __restrict__ s,
                                                            const T* __restrict__ gamma, const float* __restrict__ mean,
                                                            const float* __restrict__ rstd, T* __restrict__ dx,
                                                            float* __restrict__ part, int rows, int D) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  // kSum: a third partial, the column sums of dx.  dx is also the gradient of the residual branch `r` (s = x + r), i.e. of
  // the Dense layer that produced r, so its column sums ARE that layer's bias gradient: no separate pass over dx (K4).
  constexpr int kW = kSum ? 3 : 2;
  extern __shared__ float sh[];              // [4 warps][kW][D]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float dg[NV][VN], db[NV][VN], dxs[kSum ? NV : 1][VN];
  V gam[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * VN;
    if (c < D) gam[i].raw = *reinterpret_cast<const uint4*>(gamma + c);
    else gam[i].raw = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      dg[i][j] = 0.f;
      db[i][j] = 0.f;
      if (kSum) dxs[i][j] = 0.f;
    }
  }
  for (int row = blockIdx.x * 4 + warp; row < rows; row += gridDim.x * 4) {
    V a[NV], b[NV], r[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * VN;
      if (c < D) {
        a[i].raw = *reinterpret_cast<const uint4*>(s + (long)row * D + c);
        b[i].raw = *reinterpret_cast<const uint4*>(dy + (long)row * D + c);
        if (ds != nullptr) r[i].raw = *reinterpret_cast<const uint4*>(ds + (long)row * D + c);
      }
    }
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * VN;
      if (c < D) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          const float xhat = (a[i].get(j) - mu) * rs, d = b[i].get(j);
          dg[i][j] += d * xhat;
          db[i][j] += d;
          const float g = d * gam[i].get(j);
          s1 += g;
          s2 += g * xhat;
        }
      }
    }
    s1 = warp_sum(s1) / (float)D;
    s2 = warp_sum(s2) / (float)D;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * VN;
      if (c < D) {
        V o;
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          const float xhat = (a[i].get(j) - mu) * rs;
          float val = rs * (b[i].get(j) * gam[i].get(j) - s1 - xhat * s2);
          if (ds != nullptr) val += r[i].get(j);
          o.set(j, val);
          if (kSum) dxs[i][j] += o.get(j);             // the value as stored (rounded to T): equals a column sum over dx
        }
        *reinterpret_cast<uint4*>(dx + (long)row * D + c) = o.raw;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * VN;
    if (c < D) {
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        sh[(warp * kW + 0) * D + c + j] = dg[i][j];
        sh[(warp * kW + 1) * D + c + j] = db[i][j];
        if (kSum) sh[(warp * kW + 2) * D + c + j] = dxs[i][j];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < kW * D; c += 128) {
    const int which = c / D, col = c - which * D;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) t += sh[(w * kW + which) * D + col];
    part[((long)blockIdx.x * kW + which) * D + col] = t;
  }
}


// The same backward with the row operands staged through shared memory by the bulk-copy engine (r2: the register version
// above ran at 53 % of the copy roofline, 59 us per launch at 32768 x 768: a warp loads, reduces, stores in turn, and its
// 72 accumulator registers leave no room to prefetch the next row).  Every warp owns a ring of kStages rows x {s, dy, ds}
// and one mbarrier per stage; lane 0 refills a stage as soon as the warp has taken its row into registers, so each SM keeps
// blocks x 4 warps x kStages x 3 rows (166 KB at D = 768 bf16) in flight whatever the warps are doing.
template <typename T, int NV, bool kSum, int kStages>
__global__ void __launch_bounds__(128, kSum ? 3 : 4) add_ln_bwd_pipe_kernel(const T* __restrict__ dy, const T* __restrict__ ds,
                                                                         const T* __restrict__ s, const T* __restrict__ gamma,
                                                                         const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                         T* __restrict__ dx, float* __restrict__ part, int rows, int D) {
  using V = Vec16<T>;
  constexpr int VN = V::kN;
  constexpr int kW = kSum ? 3 : 2;
  extern __shared__ __align__(128) uint8_t sh_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t row_bytes = (uint32_t)D * sizeof(T);
  const int n_op = ds != nullptr ? 3 : 2;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sh_raw);                       // [4 warps][kStages]
  uint8_t* ring = sh_raw + 128 + (size_t)warp * kStages * 3 * row_bytes;      // this warp's stages: s | dy | ds
  uint64_t* my_bar = bars + warp * kStages;
  if (lane == 0) {
    for (int i = 0; i < kStages; ++i) mbar_init(&my_bar[i], 1);
    fence_mbar_init();
  }
  __syncwarp();
  const int row0 = blockIdx.x * 4 + warp, stride = gridDim.x * 4;
  const int n_mine = row0 < rows ? (rows - row0 + stride - 1) / stride : 0;
  auto issue = [&](int k) {                                                   // lane 0 only
    const long row = row0 + (long)k * stride;
    uint8_t* st = ring + (size_t)(k % kStages) * 3 * row_bytes;
    uint64_t* b = &my_bar[k % kStages];
    mbar_expect_tx(b, n_op * row_bytes);
    bulk_load_1d(st, s + row * D, row_bytes, b);
    bulk_load_1d(st + row_bytes, dy + row * D, row_bytes, b);
    if (n_op == 3) bulk_load_1d(st + 2 * row_bytes, ds + row * D, row_bytes, b);
  };
  if (lane == 0)
    for (int k = 0; k < kStages && k < n_mine; ++k) issue(k);
  float dg[NV][VN], db[NV][VN], dxs[kSum ? NV : 1][VN];
  V gam[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * VN;
    if (c < D) gam[i].raw = *reinterpret_cast<const uint4*>(gamma + c);
    else gam[i].raw = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int j = 0; j < VN; ++j) {
      dg[i][j] = 0.f;
      db[i][j] = 0.f;
      if (kSum) dxs[i][j] = 0.f;
    }
  }
  for (int k = 0; k < n_mine; ++k) {
    const long row = row0 + (long)k * stride;
    const uint8_t* st = ring + (size_t)(k % kStages) * 3 * row_bytes;
    const float mu = __ldg(mean + row), rs = __ldg(rstd + row);
    mbar_wait(&my_bar[k % kStages], (k / kStages) & 1);
    V a[NV], b[NV], r[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * VN;
      if (c < D) {
        a[i].raw = *reinterpret_cast<const uint4*>(st + c * sizeof(T));
        b[i].raw = *reinterpret_cast<const uint4*>(st + row_bytes + c * sizeof(T));
        if (n_op == 3) r[i].raw = *reinterpret_cast<const uint4*>(st + 2 * row_bytes + c * sizeof(T));
      }
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * VN;
      if (c < D) {
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          const float xhat = (a[i].get(j) - mu) * rs, d = b[i].get(j);
          dg[i][j] += d * xhat;
          db[i][j] += d;
          const float g = d * gam[i].get(j);
          s1 += g;
          s2 += g * xhat;
        }
      }
    }
    s1 = warp_sum(s1) / (float)D;             // every lane's shared-memory reads have returned: their values went into the sums
    s2 = warp_sum(s2) / (float)D;
    if (lane == 0 && k + kStages < n_mine) issue(k + kStages);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = (i * 32 + lane) * VN;
      if (c < D) {
        V o;
#pragma unroll
        for (int j = 0; j < VN; ++j) {
          const float xhat = (a[i].get(j) - mu) * rs;
          float val = rs * (b[i].get(j) * gam[i].get(j) - s1 - xhat * s2);
          if (n_op == 3) val += r[i].get(j);
          o.set(j, val);
          if (kSum) dxs[i][j] += o.get(j);
        }
        *reinterpret_cast<uint4*>(dx + row * D + c) = o.raw;
      }
    }
  }
  __syncthreads();                             // every warp is done with its ring: the partials reuse that memory
  float* sh = reinterpret_cast<float*>(sh_raw + 128);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (i * 32 + lane) * VN;
    if (c < D) {
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        sh[(warp * kW + 0) * D + c + j] = dg[i][j];
        sh[(warp * kW + 1) * D + c + j] = db[i][j];
        if (kSum) sh[(warp * kW + 2) * D + c + j] = dxs[i][j];
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < kW * D; c += 128) {
    const int which = c / D, col = c - which * D;
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) t += sh[(w * kW + which) * D + col];
    part[((long)blockIdx.x * kW + which) * D + col] = t;
  }
}

// out[which][col] = sum_b part[b][which][col]: 32 columns x 32 partial-groups per block, fixed order (deterministic).
template <typename T>
__global__ void __launch_bounds__(1024) ln_param_grad_reduce_kernel(const float* __restrict__ part, T* __restrict__ dgamma,
                                                                    T* __restrict__ dbeta, float* __restrict__ dx_colsum, int nblocks,
                                                                    int D) {
  const int kW = dx_colsum != nullptr ? 3 : 2;          // partial kinds per block: dgamma, dbeta [, column sums of dx]
  __shared__ float sh[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float t = 0.f;
  if (c < kW * D)
    for (int b = ty; b < nblocks; b += 32) t += part[(long)b * kW * D + c];
  sh[ty][tx] = t;
  __syncthreads();
  if (ty == 0 && c < kW * D) {
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) tot += sh[k][tx];
    const int which = c / D, col = c - which * D;
    if (which == 2) dx_colsum[col] = tot;
    else (which ? dbeta : dgamma)[col] = from_f32<T>(tot);
  }
}

// Rows of `row_bytes` (multiple of 16) indexed by (i0,i1,i2,i3): src + i.s bytes -> dst + i.d bytes.  The grid covers
// (i1, i0); the threads of a block sweep (i2, i3, 16-byte vector) with a single division per vector.  Both sides see whole
// 128-byte rows for the attention head layouts ([B,T,3,H,D] <-> [3,B,H,T,D], [B,H,T,D] <-> [B,T,H,D]).
__global__ void __launch_bounds__(256) permute_rows_kernel(const char* __restrict__ src, char* __restrict__ dst, int n0, int n1, int n2,
                                                           int n3, long s0, long s1, long s2, long s3, long d0, long d1, long d2,
                                                           long d3, int row_bytes) {
  const int vec_per_row = row_bytes >> 4;
  const int inner = n3 * vec_per_row;
  const int total = n2 * inner;
  for (int i1 = blockIdx.x; i1 < n1; i1 += gridDim.x) {
    const int i0 = blockIdx.y;
    const char* sp = src + i0 * s0 + i1 * s1;
    char* dp = dst + i0 * d0 + i1 * d1;
    for (int i = threadIdx.x; i < total; i += 256) {
      const int i2 = i / inner, rem = i - i2 * inner;
      const int i3 = rem / vec_per_row, v = rem - i3 * vec_per_row;
      const uint4 val = *reinterpret_cast<const uint4*>(sp + i2 * s2 + i3 * s3 + (long)v * 16);
      *reinterpret_cast<uint4*>(dp + i2 * d2 + i3 * d3 + (long)v * 16) = val;
    }
  }
}

template <typename T, int NV>
static int launch_ln_fwd(const void* x, const void* r, const void* gamma, const void* beta, void* s_out, void* y, float* mean,
                         float* rstd, int rows, int D, float eps, cudaStream_t st) {
  add_ln_fwd_kernel<T, NV><<<ceil_div(rows, 4), 128, 0, st>>>((const T*)x, (const T*)r, (const T*)gamma, (const T*)beta, (T*)s_out,
                                                           (T*)y, mean, rstd, rows, D, eps);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
static int g_ln_bwd_pipe = 0;          // dfb_layernorm_bwd_staged(1): rows staged through shared memory by bulk copies (r2 A/B: no faster, 51.1 vs 50.2 us)

template <typename T, int NV>
static int launch_ln_bwd(const void* dy, const void* ds, const void* s, const void* gamma, const float* mean, const float* rstd,
                         void* dx, float* part, int nblocks, int rows, int D, bool want_sum, cudaStream_t st) {
  // staged variant: ring of 4 (column-sum variant, 3 blocks per SM) / 3 (4 blocks per SM) rows per warp must fit beside its peers
  const int stages = want_sum ? 4 : 3, per_sm = want_sum ? 3 : 4;
  const long ring = 128 + 4L * stages * 3 * D * (long)sizeof(T), red = 128 + 4L * (want_sum ? 3 : 2) * D * (long)sizeof(float);
  const long pipe_smem = ring > red ? ring : red;
  if (g_ln_bwd_pipe && pipe_smem * per_sm + 1024 * per_sm <= 227 * 1024) {
    auto pk = want_sum ? add_ln_bwd_pipe_kernel<T, NV, true, 4> : add_ln_bwd_pipe_kernel<T, NV, false, 3>;
    DFB_CUDA_CHECK(cudaFuncSetAttribute(pk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pipe_smem));
    pk<<<nblocks, 128, pipe_smem, st>>>((const T*)dy, (const T*)ds, (const T*)s, (const T*)gamma, mean, rstd, (T*)dx, part, rows, D);
    DFB_CUDA_CHECK(cudaGetLastError());
    return 0;
  }
  auto kern = want_sum ? add_ln_bwd_kernel<T, NV, true> : add_ln_bwd_kernel<T, NV, false>;
  const int smem = 4 * (want_sum ? 3 : 2) * D * (int)sizeof(float);
  if (smem > 48 * 1024) DFB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<nblocks, 128, smem, st>>>((const T*)dy, (const T*)ds, (const T*)s, (const T*)gamma, mean, rstd, (T*)dx, part, rows, D);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace dfb

using namespace dfb;

static int ln_nv(int D, int vn) { return ceil_div(D, 32 * vn); }

// s_out = x + r (skipped when r is null: then s_out is unused), y = LayerNorm(s) * gamma + beta; mean/rstd fp32 [rows].
// D must be a multiple of 8 (bf16) / 4 (fp32) and at most 2048 / 1024.
DFB_API int dfb_add_layernorm_fwd(const void* x, const void* r, const void* gamma, const void* beta, void* s_out, void* y,
                                  float* mean, float* rstd, int rows, int D, float eps, int dtype, cudaStream_t stream) {
  DFB_REQUIRE(rows > 0 && D > 0, "layernorm: empty");
  const int vn = dtype == DT_BF16 ? 8 : 4;
  DFB_REQUIRE(dtype == DT_BF16 || dtype == DT_F32, "layernorm: dtype");
  DFB_REQUIRE(D % vn == 0 && ln_nv(D, vn) <= kLnMaxVec, "layernorm: D=%d unsupported (multiple of %d, <= %d)", D, vn, 32 * vn * kLnMaxVec);
  const int nv = ln_nv(D, vn);
#define DFB_LN_FWD(NV)                                                                                                     \
  if (nv <= NV)                                                                                                            \
    return dtype == DT_BF16 ? launch_ln_fwd<__nv_bfloat16, NV>(x, r, gamma, beta, s_out, y, mean, rstd, rows, D, eps, stream) \
                            : launch_ln_fwd<float, NV>(x, r, gamma, beta, s_out, y, mean, rstd, rows, D, eps, stream);
  DFB_LN_FWD(1) DFB_LN_FWD(2) DFB_LN_FWD(3) DFB_LN_FWD(4) DFB_LN_FWD(8)
#undef DFB_LN_FWD
  return 2;
}

DFB_API int dfb_layernorm_bwd_staged(int on) {
  g_ln_bwd_pipe = on;
  return 0;
}

DFB_API int dfb_layernorm_bwd_blocks(int rows) {
  int b = num_sms() * 4;
  int need = ceil_div(rows, 4);
  return b < need ? b : need;
}

// dx = ds + LayerNorm-backward(dy); dgamma/dbeta in the parameter dtype.  `dx_colsum` (nullable, fp32 [D]): column sums of
// dx, i.e. the bias gradient of the Dense layer whose output was the residual branch.  `part` is fp32 scratch of
// dfb_layernorm_bwd_blocks(rows) * (dx_colsum ? 3 : 2) * D floats supplied by the caller.
DFB_API int dfb_add_layernorm_bwd(const void* dy, const void* ds, const void* s, const void* gamma, const float* mean,
                                  const float* rstd, void* dx, void* dgamma, void* dbeta, float* dx_colsum, float* part, int rows,
                                  int D, int dtype, cudaStream_t stream) {
  const bool want_sum = dx_colsum != nullptr;
  DFB_REQUIRE(rows > 0 && D > 0, "layernorm_bwd: empty");
  const int vn = dtype == DT_BF16 ? 8 : 4;
  DFB_REQUIRE(dtype == DT_BF16 || dtype == DT_F32, "layernorm_bwd: dtype");
  DFB_REQUIRE(D % vn == 0 && ln_nv(D, vn) <= kLnMaxVec, "layernorm_bwd: D=%d unsupported", D);
  const int nv = ln_nv(D, vn);
  int nblocks = dfb_layernorm_bwd_blocks(rows);
  if (want_sum && nblocks > num_sms() * 3) nblocks = num_sms() * 3;      // the column-sum variant keeps 3 blocks per SM resident: one wave
  int rc = 2;
#define DFB_LN_BWD(NV)                                                                                                          \
  if (rc == 2 && nv <= NV)                                                                                                      \
    rc = dtype == DT_BF16 ? launch_ln_bwd<__nv_bfloat16, NV>(dy, ds, s, gamma, mean, rstd, dx, part, nblocks, rows, D, want_sum, stream) \
                          : launch_ln_bwd<float, NV>(dy, ds, s, gamma, mean, rstd, dx, part, nblocks, rows, D, want_sum, stream);
  DFB_LN_BWD(1) DFB_LN_BWD(2) DFB_LN_BWD(3) DFB_LN_BWD(4) DFB_LN_BWD(8)
#undef DFB_LN_BWD
  if (rc) return rc;
  if (dtype == DT_BF16)
    ln_param_grad_reduce_kernel<__nv_bfloat16><<<ceil_div((want_sum ? 3 : 2) * D, 32), 1024, 0, stream>>>(
        part, (__nv_bfloat16*)dgamma, (__nv_bfloat16*)dbeta, dx_colsum, nblocks, D);
  else
    ln_param_grad_reduce_kernel<float><<<ceil_div((want_sum ? 3 : 2) * D, 32), 1024, 0, stream>>>(part, (float*)dgamma, (float*)dbeta,
                                                                                                 dx_colsum, nblocks, D);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_permute_rows(const void* src, void* dst, int n0, int n1, int n2, int n3, long s0, long s1, long s2, long s3,
                             long d0, long d1, long d2, long d3, int row_bytes, cudaStream_t stream) {
  DFB_REQUIRE(row_bytes > 0 && (row_bytes % 16) == 0, "permute_rows: row_bytes must be a multiple of 16");
  DFB_REQUIRE(((s0 | s1 | s2 | s3 | d0 | d1 | d2 | d3) % 16) == 0, "permute_rows: strides must be multiples of 16 bytes");
  const long total = (long)n0 * n1 * n2 * n3 * (row_bytes / 16);
  DFB_REQUIRE(total > 0 && n0 <= 65535, "permute_rows: empty or n0 too large");
  dim3 blocks(n1 < 65535 ? n1 : 65535, n0);
  permute_rows_kernel<<<blocks, 256, 0, stream>>>((const char*)src, (char*)dst, n0, n1, n2, n3, s0, s1, s2, s3, d0, d1, d2,
                                                       d3, row_bytes);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
