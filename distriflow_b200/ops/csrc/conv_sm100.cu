// Implicit-GEMM Conv2D on tcgen05 (NHWC activations, HWIO filters, bf16, fp32 accumulate in TMEM).
//
// Same persistent warp-specialised skeleton as gemm_sm100.cu (warp 0 TMA producer, warp 1 MMA issuer, warps 2..5
// epilogue, smem ring + double-buffered TMEM accumulator).  What makes it a convolution is the producer: there is no
// im2col buffer — for every filter tap (r, s) and 64-channel slice, ONE 4-D TMA box over the NHWC tensor fetches the
// shifted slab of activations of a whole pixel tile (tn images x th rows x tw columns <= 128 pixels).  Out-of-bounds
// coordinates are zero-filled by the TMA unit, which is exactly the convolution padding; the traversal stride of the
// tensor map (`elementStrides`) is the convolution stride.
//   FWD    Y[p, co]         = sum_{tap, ci} X[p (+) tap, ci] * W[tap, ci, co]     A K-major,  B MN-major (W rows = tap*C+ci)
//   DGRAD  dX[q, ci]        = sum_{tap, co} dY[q (-) tap, co] * W[tap, ci, co]    A K-major,  B K-major; strided convs: one
//          launch per parity class of the input pixel, written to a strided sub-grid of dX
//   WGRAD  dW[tap, ci, co] += sum_{p} X[p (+) tap, ci] * dY[p, co]                A, B MN-major, K = pixels, split over
//          pixel tiles and combined with TMA reduce-add (cp.reduce.async.bulk.tensor .add.f32) into a zeroed fp32 dW.
// Replaces the tf.js conv2d forward/backward of the Conv2D layers in /root/reference/experiment/mnist/model.json for
// the channel counts of the BASELINE.json ResNet configs (C % 64 == 0); other shapes run conv_simt.cu.
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace dfb {

constexpr int kCvThreads = 192;
constexpr int kCvEpi = 128;
constexpr int kCvBlockK = 64;
constexpr int kCvUmmaK = 16;
constexpr int kCvATile = 128 * 128;          // 128 rows x 128 B
constexpr int kCvOut = 128 * 128;
constexpr int kCvAtom = 64 * 128;

enum CvMode : int { CV_FWD = 0, CV_DGRAD = 1, CV_WGRAD = 2 };

template <int BLOCK_N>
struct CvCfg {
  static constexpr int kBTile = BLOCK_N * 128;
  static constexpr int kStage = kCvATile + kBTile;
  static constexpr int kStages = BLOCK_N == 256 ? 4 : 6;
  static constexpr int kTmemCols = 2 * BLOCK_N < 32 ? 32 : 2 * BLOCK_N;
  static constexpr int kSmem = kStages * kStage + 2 * kCvOut + 256 + 1024;
};

struct CvParams {
  int mode;
  int taps, kw;            // kh*kw, kw
  int Ck;                  // contraction channels per tap (FWD: C_in, DGRAD: C_out); WGRAD: unused
  int Cm, Cn;              // WGRAD: M = C_in (Cm), N = C_out (Cn);  FWD/DGRAD: Cn = output channels
  int tn, th, tw;          // pixel tile
  int tiles_h, tiles_n;    // pixel tiles along H and N (tw spans the full width)
  int stride_h, stride_w;  // FWD/WGRAD: convolution stride (start = o*stride + off + tap); DGRAD: 1
  int off_h, off_w;        // FWD/WGRAD: -pad_top/-pad_left;  DGRAD: +pad
  int tap_sign;            // +1 FWD/WGRAD, -1 DGRAD
  int tap_dh[9], tap_dw[9], tap_w[9];   // DGRAD: per contributing tap, the dY offset of output pixel (a, b) and the filter tap index
  int a_bytes;             // bytes one A box delivers (FWD/DGRAD: rows*128; WGRAD: per 64-channel atom)
  int b_bytes;             // WGRAD: bytes per B atom box
  int splits, kb_per_split;// WGRAD: split of the pixel-tile loop
  const void* bias;
  int bias_dtype;
  int act;
};

__device__ __forceinline__ void tma_reduce_add_2d(const void* tmap, const void* smem_src, int32_t c0, int32_t c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];\n" ::"l"(
                   reinterpret_cast<uint64_t>(tmap)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}

template <int MODE, int BLOCK_N>
__global__ void __launch_bounds__(kCvThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmD, const CvParams p) {
  using Cfg = CvCfg<BLOCK_N>;
  constexpr int kStages = Cfg::kStages;
  constexpr bool kF32Out = (MODE == CV_WGRAD);
  constexpr int kChunkN = kF32Out ? 32 : 64;
  constexpr int kChunks = BLOCK_N / kChunkN;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kCvATile;
  uint8_t* smem_out = smem + kStages * Cfg::kStage;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_out + 2 * kCvOut);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp_idx = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // tile space
  int m_tiles, n_blocks, kb_total;
  if (MODE == CV_WGRAD) {
    m_tiles = p.taps * ceil_div(p.Cm, 128);                 // (tap, ci block)
    n_blocks = ceil_div(p.Cn, BLOCK_N);
    kb_total = p.tiles_h * p.tiles_n;                       // pixel tiles
  } else {
    m_tiles = p.tiles_h * p.tiles_n;
    n_blocks = ceil_div(p.Cn, BLOCK_N);
    kb_total = p.taps * (p.Ck / kCvBlockK);
  }
  const int tiles_mn = m_tiles * n_blocks;
  const int num_tiles = tiles_mn * p.splits;

  if (MODE == CV_WGRAD) {       // pixel tiles may be shorter than 64 rows: the untouched tail rows must contribute zero
    uint4 z = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < kStages * Cfg::kStage / 16; i += kCvThreads) reinterpret_cast<uint4*>(smem)[i] = z;
    fence_proxy_async_smem();
  }
  if (warp_idx == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmD);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);
    }
    fence_mbar_init();
  }
  if (warp_idx == 1) tmem_alloc<Cfg::kTmemCols>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  auto decode = [&](int tile, int& mt, int& nb, int& split) {
    split = tile / tiles_mn;
    int t = tile - split * tiles_mn;
    nb = t / m_tiles;             // n-block outermost: CTAs running together share the same weight (B) tiles
    mt = t - nb * m_tiles;
  };

  if (warp_idx == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int mt, nb, split;
        decode(tile, mt, nb, split);
        const int n0 = nb * BLOCK_N;
        if (MODE == CV_WGRAD) {
          const int cblocks = ceil_div(p.Cm, 128);
          const int tap = mt / cblocks, ci0 = (mt - tap * cblocks) * 128;
          const int r = tap / p.kw, s = tap - r * p.kw;
          const int kb0 = split * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, kb_total);
          for (int kb = kb0; kb < kb1; ++kb) {
            const int img0 = (kb / p.tiles_h) * p.tn, oh0 = (kb % p.tiles_h) * p.th;
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_expect_tx(&full_bar[stage], 2 * p.a_bytes + (BLOCK_N / 64) * p.b_bytes);
            uint8_t* sa = smem_a + stage * kCvATile;
            uint8_t* sb = smem_b + stage * Cfg::kBTile;
            const int ih0 = oh0 * p.stride_h + p.off_h + r, iw0 = p.off_w + s;
#pragma unroll
            for (int i = 0; i < 2; ++i) tma_load_4d(sa + i * kCvAtom, &tmA, &full_bar[stage], ci0 + 64 * i, iw0, ih0, img0);
#pragma unroll
            for (int i = 0; i < BLOCK_N / 64; ++i) tma_load_4d(sb + i * kCvAtom, &tmB, &full_bar[stage], n0 + 64 * i, 0, oh0, img0);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        } else {
          const int img0 = (mt / p.tiles_h) * p.tn, h0 = (mt % p.tiles_h) * p.th;
          const int cblocks = p.Ck / kCvBlockK;
          for (int kb = 0; kb < kb_total; ++kb) {
            const int tap = kb / cblocks, c0 = (kb - tap * cblocks) * kCvBlockK;
            const int r = tap / p.kw, s = tap - r * p.kw;
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_expect_tx(&full_bar[stage], p.a_bytes + Cfg::kBTile);
            uint8_t* sa = smem_a + stage * kCvATile;
            uint8_t* sb = smem_b + stage * Cfg::kBTile;
            if (MODE == CV_DGRAD) tma_load_4d(sa, &tmA, &full_bar[stage], c0, p.tap_dw[tap], h0 + p.tap_dh[tap], img0);
            else tma_load_4d(sa, &tmA, &full_bar[stage], c0, p.off_w + s, h0 * p.stride_h + p.off_h + r, img0);
            if (MODE == CV_FWD) {       // W rows = tap*C_in + ci (K), columns = co (N): MN-major, 64x64 atoms
#pragma unroll
              for (int i = 0; i < BLOCK_N / 64; ++i)
                tma_load_2d(sb + i * kCvAtom, &tmB, &full_bar[stage], n0 + 64 * i, tap * p.Ck + c0);
            } else {                    // DGRAD: rows = tap*C_in + ci (N), columns = co (K): K-major
              tma_load_2d(sb, &tmB, &full_bar[stage], c0, p.tap_w[tap] * p.Cn + n0);
            }
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp_idx == 1) {
    if (elect_one()) {
      const uint32_t a_mn = MODE == CV_WGRAD ? 1u : 0u, b_mn = MODE == CV_DGRAD ? 0u : 1u;
      const uint32_t idesc = make_idesc_bf16(128, BLOCK_N, a_mn, b_mn);
      const uint32_t a_lbo = a_mn ? kCvAtom : 0, b_lbo = b_mn ? kCvAtom : 0;
      const uint32_t a_kstep = a_mn ? (kCvUmmaK * 128) >> 4 : (kCvUmmaK * 2) >> 4;
      const uint32_t b_kstep = b_mn ? (kCvUmmaK * 128) >> 4 : (kCvUmmaK * 2) >> 4;
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int mt, nb, split;
        decode(tile, mt, nb, split);
        int kb0 = 0, kb1 = kb_total;
        if (MODE == CV_WGRAD) { kb0 = split * p.kb_per_split; kb1 = min(kb0 + p.kb_per_split, kb_total); }
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t a_desc = make_smem_desc_sw128(smem_u32(smem_a + stage * kCvATile), a_lbo, 1024);
          const uint64_t b_desc = make_smem_desc_sw128(smem_u32(smem_b + stage * Cfg::kBTile), b_lbo, 1024);
#pragma unroll
          for (int k = 0; k < kCvBlockK / kCvUmmaK; ++k)
            umma_f16(tmem_d, a_desc + (uint64_t)(k * a_kstep), b_desc + (uint64_t)(k * b_kstep), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    const int quarter = warp_idx & 3, row = quarter * 32 + lane;
    const bool issuer = (threadIdx.x == 64);
    int acc = 0;
    uint32_t acc_phase = 0, out_iter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int mt, nb, split;
      decode(tile, mt, nb, split);
      const int n0 = nb * BLOCK_N;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + acc * BLOCK_N;
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c) {
        float v[kChunkN];
        {
          uint32_t r0[32];
          tmem_ld_32x32(taddr + c * kChunkN, r0);
          if constexpr (kChunkN == 64) {
            uint32_t r1[32];
            tmem_ld_32x32(taddr + c * kChunkN + 32, r1);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) v[32 + j] = __uint_as_float(r1[j]);
          } else {
            tmem_ld_wait();
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r0[j]);
        }
        if (c == kChunks - 1) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        uint8_t* stage_out = smem_out + (out_iter & 1) * kCvOut;
        if (issuer) tma_store_wait_read<1>();
        named_bar_sync(1, kCvEpi);
        if (MODE == CV_FWD && p.bias != nullptr) {
          const int nbase = n0 + c * kChunkN;
#pragma unroll
          for (int j = 0; j < kChunkN; ++j) {
            float b = 0.f;
            if (nbase + j < p.Cn)
              b = p.bias_dtype == DT_F32 ? __ldg(reinterpret_cast<const float*>(p.bias) + nbase + j)
                                         : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[nbase + j]);
            v[j] += b;
          }
        }
        if (MODE == CV_FWD && p.act == ACT_RELU) {
#pragma unroll
          for (int j = 0; j < kChunkN; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        uint8_t* row_ptr = stage_out + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint4 q;
          if constexpr (!kF32Out) {
            __nv_bfloat162 h0 = __floats2bfloat162_rn(v[8 * j + 0], v[8 * j + 1]);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(v[8 * j + 2], v[8 * j + 3]);
            __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * j + 4], v[8 * j + 5]);
            __nv_bfloat162 h3 = __floats2bfloat162_rn(v[8 * j + 6], v[8 * j + 7]);
            q.x = *reinterpret_cast<uint32_t*>(&h0);
            q.y = *reinterpret_cast<uint32_t*>(&h1);
            q.z = *reinterpret_cast<uint32_t*>(&h2);
            q.w = *reinterpret_cast<uint32_t*>(&h3);
          } else {
            q.x = __float_as_uint(v[4 * j + 0]);
            q.y = __float_as_uint(v[4 * j + 1]);
            q.z = __float_as_uint(v[4 * j + 2]);
            q.w = __float_as_uint(v[4 * j + 3]);
          }
          *reinterpret_cast<uint4*>(row_ptr + ((j ^ (row & 7)) << 4)) = q;
        }
        fence_proxy_async_smem();
        named_bar_sync(1, kCvEpi);
        if (issuer) {
          if (MODE == CV_WGRAD) {
            const int cblocks = ceil_div(p.Cm, 128);
            const int tap = mt / cblocks, ci0 = (mt - tap * cblocks) * 128;
            // rows tap*Cm + ci0 .. +127; rows past this tap's channels carry exact zeros (A was zero-filled there)
            tma_reduce_add_2d(&tmD, stage_out, n0 + c * kChunkN, tap * p.Cm + ci0);
          } else {
            const int img0 = (mt / p.tiles_h) * p.tn, h0 = (mt % p.tiles_h) * p.th;
            tma_store_4d(&tmD, stage_out, n0 + c * kChunkN, 0, h0, img0);
          }
          tma_store_commit();
        }
        ++out_iter;
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (issuer) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp_idx == 1) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

template <int MODE, int BLOCK_N>
static int launch_conv_tc(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& d, const CvParams& p, int num_tiles,
                          cudaStream_t stream) {
  using Cfg = CvCfg<BLOCK_N>;
  auto kern = conv_tc_kernel<MODE, BLOCK_N>;
  static bool configured = false;
  if (!configured) {
    DFB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem));
    configured = true;
  }
  int grid = num_tiles < num_sms() ? num_tiles : num_sms();
  kern<<<grid, kCvThreads, Cfg::kSmem, stream>>>(a, b, d, p);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// pixel tile: full output width, as many rows (and, when a whole image fits, images) as stay within `cap` pixels
static void pick_pixel_tile(int cap, int N, int OH, int OW, int& tn, int& th, int& tw) {
  tw = OW;
  th = cap / tw;
  if (th > OH) th = OH;
  if (th < 1) th = 1;
  tn = 1;
  if (th == OH) {
    tn = cap / (tw * th);
    if (tn > N) tn = N;
    if (tn < 1) tn = 1;
  }
}

static int nhwc_map(CUtensorMap* tm, const void* base, int N, int H, int W, int C, int bc, int bw, int bh, int bn, int sw, int sh) {
  uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  uint64_t str[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
  uint32_t box[4] = {(uint32_t)bc, (uint32_t)((bw - 1) * sw + 1), (uint32_t)((bh - 1) * sh + 1), (uint32_t)bn};
  uint32_t es[4] = {1, (uint32_t)sw, (uint32_t)sh, 1};
  return encode_tmap(tm, base, 0, 4, dims, str, box, es, 1);
}

// [N, H', W', C] view with explicit pixel strides (bytes) — the parity sub-grids of a strided data gradient
static int nhwc_map_strided(CUtensorMap* tm, const void* base, int N, int H, int W, int C, long sw_bytes, long sh_bytes, long sn_bytes,
                            int bc, int bw, int bh, int bn) {
  uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  uint64_t str[3] = {(uint64_t)sw_bytes, (uint64_t)sh_bytes, (uint64_t)sn_bytes};
  uint32_t box[4] = {(uint32_t)bc, (uint32_t)bw, (uint32_t)bh, (uint32_t)bn};
  return encode_tmap(tm, base, 0, 4, dims, str, box, nullptr, 1);
}

}  // namespace dfb

using namespace dfb;

// 1 if the tcgen05 path takes this problem (otherwise callers use dfb_conv2d_simt).
DFB_API int dfb_conv2d_tc_supported(int mode, int N, int H, int W, int C, int OH, int OW, int K, int kh, int kw, int sh, int sw) {
  if (C % 64 != 0 || K % 64 != 0) return 0;
  if (mode == CV_DGRAD && (kh * kw > 9 || sh > 2 || sw > 2)) return 0;
  if (mode == CV_DGRAD) { if (ceil_div(W, sw) > 128) return 0; } else if (OW > (mode == CV_WGRAD ? 64 : 128)) return 0;
  if ((OW - 1) * sw + 1 > 256 || (OH - 1) * sh + 1 > 100000) return 0;
  return 1;
}

DFB_API size_t dfb_conv2d_tc_workspace_bytes(int mode, int N, int H, int W, int C, int OH, int OW, int K, int kh, int kw) {
  (void)mode; (void)N; (void)H; (void)W; (void)C; (void)OH; (void)OW; (void)K; (void)kh; (void)kw;
  return 0;      // WGRAD partials are combined with TMA reduce-add straight into the fp32 output: no workspace
}

// mode 0: a = X [N,H,W,C] bf16, b = W [kh,kw,C,K] bf16, out = Y [N,OH,OW,K] bf16 (+bias, relu)
// mode 1: a = dY [N,OH,OW,K],   b = W,                  out = dX [N,H,W,C] bf16      (stride 1 only)
// mode 2: a = X,                b = dY,                 out = dW [kh,kw,C,K] fp32, zero-filled here
DFB_API int dfb_conv2d_tc(int mode, const void* a, const void* b, void* out, const void* bias, int bias_dtype, int N, int H,
                          int W, int C, int OH, int OW, int K, int kh, int kw, int sh, int sw, int pt, int pl, int dtype,
                          int out_is_f32, int act, int reserved, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  (void)reserved; (void)workspace; (void)workspace_bytes;
  DFB_REQUIRE(dtype == DT_BF16, "conv2d_tc: bf16 only");
  DFB_REQUIRE(dfb_conv2d_tc_supported(mode, N, H, W, C, OH, OW, K, kh, kw, sh, sw), "conv2d_tc: unsupported shape");
  DFB_REQUIRE((mode == CV_WGRAD) == (out_is_f32 != 0), "conv2d_tc: wgrad writes fp32, fwd/dgrad write bf16");
  CvParams p;
  memset(&p, 0, sizeof(p));
  p.mode = mode; p.taps = kh * kw; p.kw = kw; p.bias = bias; p.bias_dtype = bias_dtype; p.act = act; p.splits = 1;
  CUtensorMap tmA, tmB, tmD;
  if (mode == CV_FWD) {
    pick_pixel_tile(128, N, OH, OW, p.tn, p.th, p.tw);
    p.tiles_h = ceil_div(OH, p.th); p.tiles_n = ceil_div(N, p.tn);
    p.Ck = C; p.Cn = K; p.stride_h = sh; p.stride_w = sw; p.off_h = -pt; p.off_w = -pl; p.tap_sign = 1;
    p.a_bytes = p.tn * p.th * p.tw * 128;
    if (nhwc_map(&tmA, a, N, H, W, C, 64, p.tw, p.th, p.tn, sw, sh)) return 4;
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)kh * kw * C};
    uint64_t str[1] = {(uint64_t)K * 2};
    uint32_t box[2] = {64, 64};
    if (encode_tmap(&tmB, b, 0, 2, dims, str, box, nullptr, 1)) return 4;
    if (nhwc_map(&tmD, out, N, OH, OW, K, 64, p.tw, p.th, p.tn, 1, 1)) return 4;
    const int bn = K % 256 == 0 && p.tiles_h * p.tiles_n * (K / 256) >= num_sms() ? 256 : (K % 128 == 0 ? 128 : 64);
    const int tiles = p.tiles_h * p.tiles_n * ceil_div(K, bn);
    if (bn == 256) return launch_conv_tc<CV_FWD, 256>(tmA, tmB, tmD, p, tiles, stream);
    if (bn == 128) return launch_conv_tc<CV_FWD, 128>(tmA, tmB, tmD, p, tiles, stream);
    return launch_conv_tc<CV_FWD, 64>(tmA, tmB, tmD, p, tiles, stream);
  }
  if (mode == CV_DGRAD) {
    // dX[n, ih, iw, :] = sum over taps (r, s) with (ih + pt - r) % sh == 0 and (iw + pl - s) % sw == 0 of
    // dY[n, (ih+pt-r)/sh, (iw+pl-s)/sw, :] * W[r, s]^T.  Pixels of one parity class (ih % sh, iw % sw) share the same
    // tap subset, so each class is a stride-1 gather over dY written to a strided sub-grid of dX (strides in the map).
    p.Ck = K; p.Cn = C; p.stride_h = 1; p.stride_w = 1; p.tap_sign = -1;
    const int bn = C % 256 == 0 && ceil_div(N * H * W, 128 * sh * sw) * (C / 256) >= num_sms() ? 256 : (C % 128 == 0 ? 128 : 64);
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)kh * kw * C};
    uint64_t str[1] = {(uint64_t)K * 2};
    uint32_t box[2] = {64, (uint32_t)bn};
    if (encode_tmap(&tmB, b, 0, 2, dims, str, box, nullptr, 1)) return 4;
    if (sh > 1 || sw > 1) DFB_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)N * H * W * C * 2, stream));   // classes without taps stay zero
    for (int ph = 0; ph < sh; ++ph) {
      for (int pw = 0; pw < sw; ++pw) {
        const int Hc = (H - ph + sh - 1) / sh, Wc = (W - pw + sw - 1) / sw;
        if (Hc <= 0 || Wc <= 0) continue;
        int nt = 0;
        for (int r = 0; r < kh; ++r) {
          if ((ph + pt - r) % sh != 0) continue;
          for (int s2 = 0; s2 < kw; ++s2) {
            if ((pw + pl - s2) % sw != 0) continue;
            p.tap_dh[nt] = (ph + pt - r) / sh;       // exact: the numerator is a multiple of sh (C++ division of an exact multiple)
            p.tap_dw[nt] = (pw + pl - s2) / sw;
            p.tap_w[nt] = r * kw + s2;
            ++nt;
          }
        }
        if (nt == 0) continue;
        p.taps = nt;
        pick_pixel_tile(128, N, Hc, Wc, p.tn, p.th, p.tw);
        p.tiles_h = ceil_div(Hc, p.th); p.tiles_n = ceil_div(N, p.tn);
        p.a_bytes = p.tn * p.th * p.tw * 128;
        if (nhwc_map(&tmA, a, N, OH, OW, K, 64, p.tw, p.th, p.tn, 1, 1)) return 4;
        const char* dbase = reinterpret_cast<const char*>(out) + ((size_t)ph * W + pw) * C * 2;
        if (nhwc_map_strided(&tmD, dbase, N, Hc, Wc, C, (long)sw * C * 2, (long)sh * W * C * 2, (long)H * W * C * 2, 64, p.tw, p.th,
                             p.tn))
          return 4;
        const int tiles = p.tiles_h * p.tiles_n * ceil_div(C, bn);
        int rc = bn == 256 ? launch_conv_tc<CV_DGRAD, 256>(tmA, tmB, tmD, p, tiles, stream)
                 : bn == 128 ? launch_conv_tc<CV_DGRAD, 128>(tmA, tmB, tmD, p, tiles, stream)
                             : launch_conv_tc<CV_DGRAD, 64>(tmA, tmB, tmD, p, tiles, stream);
        if (rc) return rc;
      }
    }
    return 0;
  }
  // WGRAD
  pick_pixel_tile(64, N, OH, OW, p.tn, p.th, p.tw);
  p.tiles_h = ceil_div(OH, p.th); p.tiles_n = ceil_div(N, p.tn);
  p.Cm = C; p.Cn = K; p.stride_h = sh; p.stride_w = sw; p.off_h = -pt; p.off_w = -pl; p.tap_sign = 1;
  p.a_bytes = p.tn * p.th * p.tw * 128;
  p.b_bytes = p.a_bytes;
  if (nhwc_map(&tmA, a, N, H, W, C, 64, p.tw, p.th, p.tn, sw, sh)) return 4;
  if (nhwc_map(&tmB, b, N, OH, OW, K, 64, p.tw, p.th, p.tn, 1, 1)) return 4;
  {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)kh * kw * C};
    uint64_t str[1] = {(uint64_t)K * 4};
    uint32_t box[2] = {32, 128};
    if (encode_tmap(&tmD, out, 1, 2, dims, str, box, nullptr, 1)) return 4;
  }
  DFB_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)kh * kw * C * K * sizeof(float), stream));
  const int bn = K % 128 == 0 ? 128 : 64;
  const int out_tiles = p.taps * ceil_div(C, 128) * ceil_div(K, bn);
  const int kb_total = p.tiles_h * p.tiles_n;
  int splits = ceil_div(num_sms(), out_tiles);
  if (splits > kb_total) splits = kb_total;
  if (splits < 1) splits = 1;
  p.kb_per_split = ceil_div(kb_total, splits);
  p.splits = ceil_div(kb_total, p.kb_per_split);
  const int tiles = out_tiles * p.splits;
  if (bn == 128) return launch_conv_tc<CV_WGRAD, 128>(tmA, tmB, tmD, p, tiles, stream);
  return launch_conv_tc<CV_WGRAD, 64>(tmA, tmB, tmD, p, tiles, stream);
}
