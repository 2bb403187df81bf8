// Implicit-GEMM Conv2D on CUDA cores (NHWC activations, HWIO filters): forward, data gradient, filter gradient.
// General fallback for the shapes the tcgen05 path (conv_sm100.cu) does not take: fp32 models (MNIST CNN of
// /root/reference/experiment/mnist/model.json runs fp32), C_in < 16 (first layers), odd channel counts.
// The im2col matrix is never materialised: the A-operand loader computes the gather address per element.
//   FWD   : Y[p, co]        = sum_{r,s,ci} X[n, oh*sh-pt+r, ow*sw-pl+s, ci] * W[r,s,ci,co]       p = (n,oh,ow)
//   DGRAD : dX[q, ci]       = sum_{r,s,co} dY[n, (ih+pt-r)/sh, (iw+pl-s)/sw, co] * W[r,s,ci,co]  q = (n,ih,iw)
//   WGRAD : dW[(r,s,ci),co] = sum_p X[n, oh*sh-pt+r, ow*sw-pl+s, ci] * dY[p, co]                 (split over p, atomics)
#include "common.cuh"
#include "device_utils.cuh"

namespace dfb {

struct ConvShape {
  int N, H, W, C;        // input
  int OH, OW, K;         // output spatial, output channels
  int kh, kw, sh, sw, pt, pl;
};

enum ConvMode : int { CONV_FWD = 0, CONV_DGRAD = 1, CONV_WGRAD = 2 };

template <typename T, int MODE>
__global__ void __launch_bounds__(256)
conv_simt_kernel(const T* __restrict__ a_src, const T* __restrict__ b_src, void* __restrict__ out, const void* __restrict__ bias,
                 int bias_dtype, ConvShape s, int act, int ksplit, int out_is_f32) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  int GM, GN, GK;
  if (MODE == CONV_FWD) { GM = s.N * s.OH * s.OW; GN = s.K; GK = s.kh * s.kw * s.C; }
  else if (MODE == CONV_DGRAD) { GM = s.N * s.H * s.W; GN = s.C; GK = s.kh * s.kw * s.K; }
  else { GM = s.kh * s.kw * s.C; GN = s.K; GK = s.N * s.OH * s.OW; }
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int kchunk = ceil_div(ceil_div(GK, BK), ksplit) * BK;
  const int kbeg = blockIdx.z * kchunk, kend = min(GK, kbeg + kchunk);

  auto load_a = [&](int m, int k) -> float {
    if (m >= GM || k >= kend) return 0.f;
    if (MODE == CONV_FWD) {
      int ow = m % s.OW, t = m / s.OW, oh = t % s.OH, n = t / s.OH;
      int ci = k % s.C, rs = k / s.C, sx = rs % s.kw, r = rs / s.kw;
      int ih = oh * s.sh - s.pt + r, iw = ow * s.sw - s.pl + sx;
      if (ih < 0 || ih >= s.H || iw < 0 || iw >= s.W) return 0.f;
      return to_f32<T>(a_src[(((long)n * s.H + ih) * s.W + iw) * s.C + ci]);
    } else if (MODE == CONV_DGRAD) {
      int iw = m % s.W, t = m / s.W, ih = t % s.H, n = t / s.H;
      int co = k % s.K, rs = k / s.K, sx = rs % s.kw, r = rs / s.kw;
      int th = ih + s.pt - r, tw = iw + s.pl - sx;
      if (th < 0 || tw < 0 || (th % s.sh) != 0 || (tw % s.sw) != 0) return 0.f;
      int oh = th / s.sh, ow = tw / s.sw;
      if (oh >= s.OH || ow >= s.OW) return 0.f;
      return to_f32<T>(a_src[(((long)n * s.OH + oh) * s.OW + ow) * s.K + co]);
    } else {
      int ci = m % s.C, rs = m / s.C, sx = rs % s.kw, r = rs / s.kw;
      int ow = k % s.OW, t = k / s.OW, oh = t % s.OH, n = t / s.OH;
      int ih = oh * s.sh - s.pt + r, iw = ow * s.sw - s.pl + sx;
      if (ih < 0 || ih >= s.H || iw < 0 || iw >= s.W) return 0.f;
      return to_f32<T>(a_src[(((long)n * s.H + ih) * s.W + iw) * s.C + ci]);
    }
  };
  auto load_b = [&](int n, int k) -> float {
    if (n >= GN || k >= kend) return 0.f;
    if (MODE == CONV_FWD) return to_f32<T>(b_src[(long)k * s.K + n]);                 // W[(r,s,ci), co]
    if (MODE == CONV_DGRAD) {                                                        // W[r,s,ci=n,co]
      int co = k % s.K, rs = k / s.K;
      return to_f32<T>(b_src[((long)rs * s.C + n) * s.K + co]);
    }
    return to_f32<T>(b_src[(long)k * s.K + n]);                                       // dY[p, co]
  };

  float acc[4][4] = {};
  for (int k0 = kbeg; k0 < kend; k0 += BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int idx = tid + i * 256;
      // A: k fastest for FWD/DGRAD (channel-contiguous gathers); m fastest for WGRAD (ci contiguous along m)
      int kk = (MODE == CONV_WGRAD) ? (idx >> 6) : (idx & 15), mm = (MODE == CONV_WGRAD) ? (idx & 63) : (idx >> 4);
      As[kk][mm] = load_a(m0 + mm, k0 + kk);
      // B: n fastest for FWD/WGRAD (co contiguous); k fastest for DGRAD (co contiguous along k)
      int kb = (MODE == CONV_DGRAD) ? (idx & 15) : (idx >> 6), nn = (MODE == CONV_DGRAD) ? (idx >> 4) : (idx & 63);
      Bs[kb][nn] = load_b(n0 + nn, k0 + kb);
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
    if (gm >= GM) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int gn = n0 + tx * 4 + j;
      if (gn >= GN) continue;
      float v = acc[i][j];
      long o = (long)gm * GN + gn;
      if (ksplit > 1) {
        atomicAdd(reinterpret_cast<float*>(out) + o, v);
      } else {
        if (bias != nullptr)
          v += bias_dtype == DT_F32 ? reinterpret_cast<const float*>(bias)[gn]
                                    : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias)[gn]);
        if (act == ACT_RELU) v = fmaxf(v, 0.f);
        if (out_is_f32) reinterpret_cast<float*>(out)[o] = v;
        else reinterpret_cast<T*>(out)[o] = from_f32<T>(v);
      }
    }
  }
}

template <typename T>
static int launch_conv(int mode, const void* a, const void* b, void* out, const void* bias, int bias_dtype, const ConvShape& s,
                       int act, int ksplit, int out_is_f32, cudaStream_t stream) {
  int GM, GN;
  if (mode == CONV_FWD) { GM = s.N * s.OH * s.OW; GN = s.K; }
  else if (mode == CONV_DGRAD) { GM = s.N * s.H * s.W; GN = s.C; }
  else { GM = s.kh * s.kw * s.C; GN = s.K; }
  dim3 grid(ceil_div(GN, 64), ceil_div(GM, 64), ksplit);
  if (mode == CONV_FWD)
    conv_simt_kernel<T, CONV_FWD><<<grid, 256, 0, stream>>>((const T*)a, (const T*)b, out, bias, bias_dtype, s, act, ksplit, out_is_f32);
  else if (mode == CONV_DGRAD)
    conv_simt_kernel<T, CONV_DGRAD><<<grid, 256, 0, stream>>>((const T*)a, (const T*)b, out, bias, bias_dtype, s, act, ksplit, out_is_f32);
  else
    conv_simt_kernel<T, CONV_WGRAD><<<grid, 256, 0, stream>>>((const T*)a, (const T*)b, out, bias, bias_dtype, s, act, ksplit, out_is_f32);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace dfb

using namespace dfb;

// mode 0 fwd:   a = X [N,H,W,C],     b = W [kh,kw,C,K], out = Y  [N,OH,OW,K] (+bias, relu)
// mode 1 dgrad: a = dY [N,OH,OW,K],  b = W,             out = dX [N,H,W,C]
// mode 2 wgrad: a = X,               b = dY,            out = dW [kh,kw,C,K]; with ksplit > 1 `out` must be fp32
//               and is zero-filled here, partial sums are combined with atomicAdd.
DFB_API int dfb_conv2d_simt(int mode, const void* a, const void* b, void* out, const void* bias, int bias_dtype, int N, int H,
                            int W, int C, int OH, int OW, int K, int kh, int kw, int sh, int sw, int pt, int pl, int dtype,
                            int out_is_f32, int act, int ksplit, cudaStream_t stream) {
  DFB_REQUIRE(mode >= 0 && mode <= 2, "conv2d_simt: mode");
  DFB_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && OH > 0 && OW > 0 && K > 0, "conv2d_simt: empty problem");
  ConvShape s{N, H, W, C, OH, OW, K, kh, kw, sh, sw, pt, pl};
  if (ksplit < 1) ksplit = 1;
  if (ksplit > 1) {
    DFB_REQUIRE(out_is_f32 && bias == nullptr && act == ACT_NONE, "conv2d_simt: split needs fp32 output, no epilogue");
    size_t n = mode == CONV_WGRAD ? (size_t)kh * kw * C * K : (mode == CONV_FWD ? (size_t)N * OH * OW * K : (size_t)N * H * W * C);
    DFB_CUDA_CHECK(cudaMemsetAsync(out, 0, n * sizeof(float), stream));
  }
  if (dtype == DT_F32) return launch_conv<float>(mode, a, b, out, bias, bias_dtype, s, act, ksplit, 1, stream);
  if (dtype == DT_BF16) return launch_conv<__nv_bfloat16>(mode, a, b, out, bias, bias_dtype, s, act, ksplit, out_is_f32, stream);
  set_error("conv2d_simt: unsupported dtype %d", dtype);
  return 2;
}
