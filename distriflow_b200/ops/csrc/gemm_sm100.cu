// Dense GEMM for sm_100a: persistent, warp-specialised, TMA -> shared memory -> tcgen05.mma with the
// accumulator in TMEM, epilogue TMEM -> registers -> swizzled shared memory -> TMA store.
//
//   D[M,N] = act(A[M,K] * B[N,K]^T + bias[N])          bf16 inputs, fp32 accumulate, bf16 or fp32 output
//
// Either operand may be K-major (contraction index contiguous) or MN-major (row index contiguous), which
// covers the three products of a Dense layer whose kernel is stored `[in, out]` as in tf.js:
//   forward  Y  = X  * W      : A = X  (K-major),  B = W  (MN-major)
//   dgrad    dX = dY * W^T    : A = dY (K-major),  B = W  (K-major)
//   wgrad    dW = X^T * dY    : A = X  (MN-major), B = dY (MN-major), long K -> split-K
// Reference call sites this replaces: `predictOnBatch` / `tf.variableGrads` on the Dense layers of
// /root/reference/src/common/models.ts:137-142.
//
// Roles (192 threads): warp 0 = TMA producer (one elected lane), warp 1 = TMEM allocator + MMA issuer
// (one elected lane), warps 2..5 = epilogue (warp w reads TMEM lanes 32*(w%4)..+31).
// Pipelines: smem ring full/empty (TMA <-> MMA), TMEM double buffer full/empty (MMA <-> epilogue).
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace dfb {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                 // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int kUmmaK = 16;
constexpr int kThreads = 192;
constexpr int kEpiThreads = 128;
constexpr int kATileBytes = kBlockM * kBlockK * 2;      // 16 KiB
constexpr int kStageOutBytes = 128 * 128;               // 128 rows x 128 B staging for the TMA store
constexpr int kAtomBytes = kBlockK * 128;               // one 64(MN) x 64(K) MN-major box = 8 KiB

template <int BLOCK_N>
struct GemmCfg {
  static constexpr int kBTileBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kATileBytes + kBTileBytes;
  static constexpr int kStages = (BLOCK_N == 256) ? 4 : 6;
  static constexpr int kTmemCols = 2 * BLOCK_N;         // double-buffered fp32 accumulator
  // <= 233472 - 2*1024 (two CTA reserves) so that a small comm/optimizer CTA can share the SM with a GEMM CTA
  static constexpr int kSmemBytes = kStages * kStageBytes + 2 * kStageOutBytes + 256 + 1024;
};

struct GemmParams {
  const void* bias;     // nullptr, fp32[N] or bf16[N]
  int bias_dtype;
  int M, N, K;
  int a_mn, b_mn;       // 1 = MN-major operand
  int act;
  int splits;           // split-K factor (1 = none)
  int kb_per_split;     // k-blocks per split
  int split_rows;       // row offset between splits in the fp32 workspace (multiple of 128)
  // fused all-gather of the B operand (kGather kernels only): B = W[K,N] MN-major, sharded over ranks along K
  const long long* pieces;    // [num_pieces][4] = {src (peer) address, dst (local) address, bytes, k-block}
  const int* piece_chunk0;    // [num_pieces + 1] prefix sums of 512-byte chunks
  const unsigned* expected;   // [k-blocks] remote bytes each k-block waits for
  unsigned* ready;            // [k-blocks] bytes landed so far (zeroed before the launch)
  int num_pieces;
  int rot_kb;                 // k-loop starts at this k-block (first locally owned one) and wraps around
  // auxiliary bf16 matrix [M, N] (pitch aux_ld) touched by the epilogue of bf16-output kernels:
  //   aux_mode 1: aux = accumulator + bias BEFORE the activation (the pre-activation a GELU backward needs)
  //   aux_mode 2: output = accumulator * gelu'(aux)              (GELU backward fused into the next layer's dgrad)
  void* aux;
  long aux_ld;
  int aux_mode;
  // 2-CTA aux kernel only: every epilogue warp also writes the column sums of its 32-row slab of the (rounded) output to
  // colsum_part[M/32 slabs][N]; a small reduce kernel adds the slabs in a fixed order.  This is the bias gradient of the
  // layer whose pre-activation gradient the GEMM produces (K4: bias-grad out of the epilogue, no pass over the output).
  float* colsum_part;
};

__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * k0 * (1.0f + 3.0f * k1 * x * x);
}

// Epilogue hook on one 64-column chunk of one accumulator row (bf16-output kernels only).
__device__ __forceinline__ void epilogue_aux(float (&v)[64], const GemmParams& a, int gm, int nb, int M, int N) {
  if (a.aux_mode == 0 || gm >= M) return;
  __nv_bfloat16* rowp = reinterpret_cast<__nv_bfloat16*>(a.aux) + (long)gm * a.aux_ld + nb;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (nb + 8 * j + 8 > N) break;
    if (a.aux_mode == 1) {
      __nv_bfloat162 h0 = __floats2bfloat162_rn(v[8 * j + 0], v[8 * j + 1]);
      __nv_bfloat162 h1 = __floats2bfloat162_rn(v[8 * j + 2], v[8 * j + 3]);
      __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * j + 4], v[8 * j + 5]);
      __nv_bfloat162 h3 = __floats2bfloat162_rn(v[8 * j + 6], v[8 * j + 7]);
      uint4 q = {*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1), *reinterpret_cast<uint32_t*>(&h2),
                 *reinterpret_cast<uint32_t*>(&h3)};
      *reinterpret_cast<uint4*>(rowp + 8 * j) = q;
    } else {
      const uint4 q = *reinterpret_cast<const uint4*>(rowp + 8 * j);
      const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&q);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[8 * j + i] *= gelu_tanh_grad(__bfloat162float(h[i]));
    }
  }
}

// bias for one chunk of columns [nb, nb + CH): every lane reads the same addresses (L1 broadcast), 16 bytes at a time
template <int CH>
__device__ __forceinline__ void add_bias_chunk(float (&v)[CH], const void* bias, int bias_dtype, int nb, int N) {
  if (bias_dtype == DT_F32) {
    const float* bp = reinterpret_cast<const float*>(bias) + nb;
    if (nb + CH <= N && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
      for (int j = 0; j < CH / 4; ++j) {
        const float4 b4 = __ldg(reinterpret_cast<const float4*>(bp) + j);
        v[4 * j] += b4.x; v[4 * j + 1] += b4.y; v[4 * j + 2] += b4.z; v[4 * j + 3] += b4.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < CH; ++j) v[j] += (nb + j < N) ? __ldg(bp + j) : 0.f;
    }
  } else {
    const __nv_bfloat16* bp = reinterpret_cast<const __nv_bfloat16*>(bias) + nb;
    if (nb + CH <= N && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
      for (int j = 0; j < CH / 8; ++j) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(bp) + j);
        const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&q);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[8 * j + i] += __bfloat162float(h[i]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < CH; ++j) v[j] += (nb + j < N) ? __bfloat162float(bp[j]) : 0.f;
    }
  }
}

__device__ __forceinline__ float gelu_tanh(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * x * (1.0f + t);
}

template <int BLOCK_N, typename OutT, bool kGather = false>
__global__ void __launch_bounds__(kThreads + (kGather ? 64 : 0), 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmD, const GemmParams p) {
  using Cfg = GemmCfg<BLOCK_N>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kChunkN = 128 / (int)sizeof(OutT);       // columns per 128-byte staging row
  constexpr int kChunks = BLOCK_N / kChunkN;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kATileBytes;
  uint8_t* smem_out = smem + kStages * Cfg::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_out + 2 * kStageOutBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp_idx = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_blocks = ceil_div(p.M, kBlockM);
  const int n_blocks = ceil_div(p.N, BLOCK_N);
  const int kb_total = ceil_div(p.K, kBlockK);
  const int tiles_mn = m_blocks * n_blocks;
  const int num_tiles = tiles_mn * p.splits;

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
      mbar_init(&tmem_empty[i], 4);     // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp_idx == 1) tmem_alloc<Cfg::kTmemCols>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // tile -> (m_blk, n_blk, split): groups of 8 m-blocks sweep n, so concurrently resident CTAs share A and B tiles in L2
  auto decode_tile = [&](int tile, int& m_blk, int& n_blk, int& split) {
    split = tile / tiles_mn;
    int t = tile - split * tiles_mn;
    const int group = 8 * n_blocks;
    int g = t / group;
    int first_m = g * 8;
    int gsz = min(8, m_blocks - first_m);
    int r = t - g * group;
    m_blk = first_m + r % gsz;
    n_blk = r / gsz;
  };

  if (warp_idx == 0) {
    // ===================================================================== TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk, split;
        decode_tile(tile, m_blk, n_blk, split);
        const int m0 = m_blk * kBlockM, n0 = n_blk * BLOCK_N;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, kb_total);
        for (int it = 0; it < kb1 - kb0; ++it) {
          int kb = kb0 + it;
          if constexpr (kGather) {
            kb = kb0 + (it + p.rot_kb) % (kb1 - kb0);
            const unsigned need = p.expected[kb];
            if (need) {                           // wait until the gather warps (of any CTA) have landed this k-block of W
              unsigned have;
              do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(have) : "l"(p.ready + kb) : "memory");
              } while (have < need);
              fence_proxy_async_all();            // generic-proxy stores -> visible to the TMA (async proxy) read below
            }
          }
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          uint8_t* sa = smem_a + stage * kATileBytes;
          uint8_t* sb = smem_b + stage * Cfg::kBTileBytes;
          const int k0 = kb * kBlockK;
          if (p.a_mn) {
#pragma unroll
            for (int i = 0; i < kBlockM / 64; ++i) tma_load_2d(sa + i * kAtomBytes, &tmA, &full_bar[stage], m0 + 64 * i, k0);
          } else {
            tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
          }
          if (p.b_mn) {
#pragma unroll
            for (int i = 0; i < BLOCK_N / 64; ++i) tma_load_2d(sb + i * kAtomBytes, &tmB, &full_bar[stage], n0 + 64 * i, k0);
          } else {
            tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp_idx == 1) {
    // ===================================================================== MMA issuer
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, (uint32_t)p.a_mn, (uint32_t)p.b_mn);
      // K-major SW128: 8-row groups 1024 B apart, one 128 B swizzle row per matrix row, k-step = 32 B.
      // MN-major SW128: 64-element MN atoms kAtomBytes apart (LBO), 8-k-row groups 1024 B apart (SBO), k-step = 16 rows.
      const uint32_t a_lbo = p.a_mn ? kAtomBytes : 0, b_lbo = p.b_mn ? kAtomBytes : 0;
      const uint32_t a_kstep = p.a_mn ? (kUmmaK * 128) >> 4 : (kUmmaK * 2) >> 4;
      const uint32_t b_kstep = p.b_mn ? (kUmmaK * 128) >> 4 : (kUmmaK * 2) >> 4;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk, split;
        decode_tile(tile, m_blk, n_blk, split);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, kb_total);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        for (int it = 0; it < kb1 - kb0; ++it) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t a_desc = make_smem_desc_sw128(smem_u32(smem_a + stage * kATileBytes), a_lbo, 1024);
          const uint64_t b_desc = make_smem_desc_sw128(smem_u32(smem_b + stage * Cfg::kBTileBytes), b_lbo, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            umma_f16(tmem_d, a_desc + (uint64_t)(k * a_kstep), b_desc + (uint64_t)(k * b_kstep), idesc,
                     (it > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);            // frees the smem slot once these MMAs have read it
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[acc]);                // accumulator complete -> epilogue
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (kGather && warp_idx >= 6) {
    // ===================================================================== gather warps: pull remote shards of W into the local buffer
    // Work is claimed dynamically (atomic ticket) so that progress never depends on a CTA that is not resident yet.
    const int total = p.piece_chunk0[p.num_pieces];
    unsigned* ticket = p.ready + ceil_div(p.K, kBlockK);
    while (true) {
      int c = 0;
      if (lane == 0) c = (int)atomicAdd(ticket, 1u);
      c = __shfl_sync(0xffffffff, c, 0);
      if (c >= total) break;
      int piece = 0;
      while (c >= p.piece_chunk0[piece + 1]) ++piece;
      const long long* pc = p.pieces + 4 * piece;
      const long base = (long)(c - p.piece_chunk0[piece]) * 512;
      const long off = base + lane * 16;
      const long bytes = pc[2];
      if (off < bytes) {
        uint4 v;
        asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(pc[0] + off) : "memory");
        *reinterpret_cast<uint4*>(pc[1] + off) = v;
      }
      __threadfence();
      __syncwarp();
      if (lane == 0) {
        const long left = bytes - base;
        const unsigned n = (unsigned)(left < 512 ? left : 512);
        asm volatile("red.release.gpu.global.add.u32 [%0], %1;\n" ::"l"(p.ready + (int)pc[3]), "r"(n) : "memory");
      }
    }
  } else {
    // ===================================================================== epilogue (warps 2..5)
    const int quarter = warp_idx & 3;                 // TMEM lane quarter this warp may read
    const int row = quarter * 32 + lane;              // row inside the 128-row tile
    const int epi_tid = threadIdx.x - 64;
    const bool issuer = (epi_tid == 0);
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t out_iter = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk, split;
      decode_tile(tile, m_blk, n_blk, split);
      const int m0 = m_blk * kBlockM, n0 = n_blk * BLOCK_N;
      const int out_row0 = m0 + split * p.split_rows;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + acc * BLOCK_N;
      // column chunks that lie entirely beyond N are never read, converted or stored (thin outputs: ResNet-20's 16-channel
      // layers use one chunk of a 128-wide tile, and their tiles are epilogue-bound: 3 k-blocks of MMA per tile)
      const int live_chunks = min(kChunks, ceil_div(p.N - n0, kChunkN));
#pragma unroll 1
      for (int c = 0; c < live_chunks; ++c) {
        float v[kChunkN];
        {
          uint32_t r[32];
          tmem_ld_32x32(taddr + c * kChunkN, r);
          if constexpr (kChunkN == 64) {
            uint32_t r2[32];
            tmem_ld_32x32(taddr + c * kChunkN + 32, r2);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) v[32 + j] = __uint_as_float(r2[j]);
          } else {
            tmem_ld_wait();
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        }
        if (c == live_chunks - 1) {                    // whole accumulator is in registers: release TMEM early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        uint8_t* stage_out = smem_out + (out_iter & 1) * kStageOutBytes;
        if (issuer) tma_store_wait_read<1>();          // the store that last used this buffer has drained
        named_bar_sync(1, kEpiThreads);
        if (p.bias != nullptr) add_bias_chunk<kChunkN>(v, p.bias, p.bias_dtype, n0 + c * kChunkN, p.N);
        if constexpr (kChunkN == 64) epilogue_aux(v, p, m0 + row, n0 + c * kChunkN, p.M, p.N);
        if (p.act == ACT_RELU) {
#pragma unroll
          for (int j = 0; j < kChunkN; ++j) v[j] = fmaxf(v[j], 0.f);
        } else if (p.act == ACT_GELU) {
#pragma unroll
          for (int j = 0; j < kChunkN; ++j) v[j] = gelu_tanh(v[j]);
        }
        uint8_t* row_ptr = stage_out + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {                  // eight 16-byte pieces per row, XOR-swizzled like SWIZZLE_128B
          uint4 q;
          if constexpr (sizeof(OutT) == 2) {
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
        named_bar_sync(1, kEpiThreads);
        if (issuer) {
          tma_store_2d(&tmD, stage_out, n0 + c * kChunkN, out_row0);
          tma_store_commit();
        }
        ++out_iter;
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
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

// ======================================================================================================
// 2-CTA variant: a cluster of two CTAs (one SM pair) computes a 256 x 256 tile with tcgen05.mma.cta_group::2.
// Each CTA stages its own 128 rows of A and its own 128-column half of B (32 KiB per k-block instead of 48 KiB),
// the leader CTA's elected thread issues the MMAs for both, and each CTA's TMEM receives its 128 accumulator rows.
// Per-SM shared-memory fill drops from 96 to 64 bytes per MMA cycle, which is what limited the 1-CTA kernel
// (profiles/ncu_kernels_r1.csv: tensor pipe 68-71 % active on the K = 768 products).
//   full[stage]   lives in the leader; both CTAs' TMA loads complete on it (peer-bit-masked barrier address)
//   empty[stage]  one per CTA; the leader's tcgen05.commit multicasts the arrival to both
//   tmem_full     one per CTA, multicast commit;  tmem_empty lives in the leader, 8 arrivals (4 epilogue warps x 2 CTAs)
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;
constexpr int k2Stages = 6;
constexpr int k2StageBytes = kATileBytes + 128 * kBlockK * 2;         // 16 KiB A + 16 KiB B half
constexpr int k2SmemBytes = k2Stages * k2StageBytes + 2 * kStageOutBytes + 256 + 1024;
constexpr int k2TmemCols = 512;
constexpr int k2EpiGroups = 2;                           // two groups of 4 epilogue warps, each drains half of the 256 columns
constexpr int k2Threads = 64 + k2EpiGroups * kEpiThreads;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const void* tmap, uint64_t* bar, int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_f16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {       // arrives on `bar` in BOTH CTAs of the pair
  const uint16_t mask = 3;
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t target_cta) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(smem_u32(bar)), "r"(target_cta));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(remote) : "memory");
}

// tcgen05.fence::before_thread_sync orders the warp's TMEM loads before this arrival; no generic-memory release is needed
// (the .release form costs a MEMBAR per tile and epilogue warp: 20 % of the epilogue's stall samples in ncu).
__device__ __forceinline__ void mbar_arrive_cluster_relaxed(uint64_t* bar, uint32_t target_cta) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(remote) : "r"(smem_u32(bar)), "r"(target_cta));
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(remote) : "memory");
}

// kAux: the epilogue also moves a second [M, N] bf16 tensor through TMA (aux_mode 1: stores the pre-activation next to the
// activated output; aux_mode 2: loads it and multiplies the accumulator by gelu'(aux)).  That variant trades one pipeline
// stage for a second 4 KiB staging slab per epilogue warp.
template <typename OutT, bool kAux>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(k2Threads, 1)
gemm2_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmX, const GemmParams p) {
  constexpr int k2Stages = kAux ? dfb::k2Stages - 1 : dfb::k2Stages;
  constexpr int kSlabBytes = 32 * 128;
  constexpr int kOutBytes = (kAux ? 2 : 1) * 8 * kSlabBytes;
  constexpr int BLOCK_N = 256;
  constexpr int kChunkN = 128 / (int)sizeof(OutT);
  constexpr int kChunks = BLOCK_N / kChunkN;
  constexpr int kBHalf = 128 * kBlockK * 2;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + k2Stages * kATileBytes;
  uint8_t* smem_out = smem + k2Stages * k2StageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem_out + kOutBytes);
  uint64_t* empty_bar = full_bar + k2Stages;
  uint64_t* tmem_full = empty_bar + k2Stages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* aux_bar = tmem_empty + 2;                       // one per epilogue warp (aux_mode 2 loads)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(aux_bar + 8);

  const int warp_idx = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const bool leader = (cta == 0);

  const int m_blocks = ceil_div(p.M, 256);
  const int n_blocks = ceil_div(p.N, BLOCK_N);
  const int kb_total = ceil_div(p.K, kBlockK);
  const int tiles_mn = m_blocks * n_blocks;
  const int num_tiles = tiles_mn * p.splits;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

  if (warp_idx == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmD);
    for (int i = 0; i < k2Stages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 2 * 4 * k2EpiGroups);       // one arrival per epilogue warp of both CTAs
    }
    for (int i = 0; i < 8; ++i) mbar_init(&aux_bar[i], 1);
    fence_mbar_init();
  }
  if (warp_idx == 1) {
    uint32_t ncols = k2TmemCols;
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_ptr)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  auto decode_tile = [&](int tile, int& m_blk, int& n_blk, int& split) {
    split = tile / tiles_mn;
    int t = tile - split * tiles_mn;
    const int group = 4 * n_blocks;               // groups of 4 x 256 rows sweep n: concurrently running pairs share tiles in L2
    int g = t / group;
    int first_m = g * 4;
    int gsz = min(4, m_blocks - first_m);
    int r = t - g * group;
    m_blk = first_m + r % gsz;
    n_blk = r / gsz;
  };

  if (warp_idx == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int m_blk, n_blk, split;
        decode_tile(tile, m_blk, n_blk, split);
        const int m0 = m_blk * 256 + (int)cta * 128, n0 = n_blk * BLOCK_N + (int)cta * 128;
        const int kb0 = split * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, kb_total);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * k2StageBytes);      // bytes of both CTAs land on the leader's barrier
          uint8_t* sa = smem_a + stage * kATileBytes;
          uint8_t* sb = smem_b + stage * kBHalf;
          const int k0 = kb * kBlockK;
          if (p.a_mn) {
#pragma unroll
            for (int i = 0; i < 2; ++i) tma_load_2d_2sm(sa + i * kAtomBytes, &tmA, &full_bar[stage], m0 + 64 * i, k0);
          } else {
            tma_load_2d_2sm(sa, &tmA, &full_bar[stage], k0, m0);
          }
          if (p.b_mn) {
#pragma unroll
            for (int i = 0; i < 2; ++i) tma_load_2d_2sm(sb + i * kAtomBytes, &tmB, &full_bar[stage], n0 + 64 * i, k0);
          } else {
            tma_load_2d_2sm(sb, &tmB, &full_bar[stage], k0, n0);
          }
          if (++stage == k2Stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp_idx == 1) {
    if (leader && elect_one()) {
      const uint32_t idesc = make_idesc_bf16(256, BLOCK_N, (uint32_t)p.a_mn, (uint32_t)p.b_mn);
      const uint32_t a_lbo = p.a_mn ? kAtomBytes : 0, b_lbo = p.b_mn ? kAtomBytes : 0;
      const uint32_t a_kstep = p.a_mn ? (kUmmaK * 128) >> 4 : (kUmmaK * 2) >> 4;
      const uint32_t b_kstep = p.b_mn ? (kUmmaK * 128) >> 4 : (kUmmaK * 2) >> 4;
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int m_blk, n_blk, split;
        decode_tile(tile, m_blk, n_blk, split);
        const int kb0 = split * p.kb_per_split, kb1 = min(kb0 + p.kb_per_split, kb_total);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
        for (int it = 0; it < kb1 - kb0; ++it) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t a_desc = make_smem_desc_sw128(smem_u32(smem_a + stage * kATileBytes), a_lbo, 1024);
          const uint64_t b_desc = make_smem_desc_sw128(smem_u32(smem_b + stage * kBHalf), b_lbo, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            umma_f16_2sm(tmem_d, a_desc + (uint64_t)(k * a_kstep), b_desc + (uint64_t)(k * b_kstep), idesc, (it > 0 || k > 0) ? 1u : 0u);
          umma_commit_2sm(&empty_bar[stage]);
          if (++stage == k2Stages) { stage = 0; phase ^= 1; }
        }
        umma_commit_2sm(&tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // warps 2-5 = group 0 (column chunks of the first half), warps 6-9 = group 1 (second half); a warp may only touch
    // TMEM lanes 32 * (warp % 4) .. + 31, so both groups cover all 128 rows.  Every warp stages and TMA-stores its own
    // 32-row slab: no barrier between epilogue warps, only __syncwarp.
    const int quarter = warp_idx & 3, row = quarter * 32 + lane;
    const int group = (warp_idx - 2) >> 2;
    constexpr int kGroupChunks = kChunks / k2EpiGroups;
    uint8_t* const slab = smem_out + (warp_idx - 2) * ((kAux ? 2 : 1) * kSlabBytes);   // output staging
    uint8_t* const slab_x = slab + kSlabBytes;                                          // aux staging (kAux only)
    uint64_t* const my_aux_bar = &aux_bar[warp_idx - 2];
    uint32_t aux_phase = 0;
    const bool aux_store = kAux && p.aux_mode == 1, aux_load = kAux && p.aux_mode == 2;
    auto issue_aux_load = [&](int tile, int c) {           // lane 0: fetch this warp's 32 x 64 slab of the aux tensor
      int m_blk, n_blk, split;
      decode_tile(tile, m_blk, n_blk, split);
      mbar_expect_tx(my_aux_bar, kSlabBytes);
      tma_load_2d(slab_x, &tmX, my_aux_bar, n_blk * BLOCK_N + c * kChunkN, m_blk * 256 + (int)cta * 128 + quarter * 32);
    };
    if (aux_load && pair < num_tiles && lane == 0) issue_aux_load(pair, group * kGroupChunks);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      int m_blk, n_blk, split;
      decode_tile(tile, m_blk, n_blk, split);
      const int m0 = m_blk * 256 + (int)cta * 128, n0 = n_blk * BLOCK_N;
      const int out_row0 = m0 + split * p.split_rows + quarter * 32;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t(quarter * 32) << 16) + acc * BLOCK_N;
#pragma unroll 1
      for (int c = group * kGroupChunks; c < (group + 1) * kGroupChunks; ++c) {
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
        if (c == (group + 1) * kGroupChunks - 1) {     // this warp's part of the accumulator is drained: tell the leader's MMA thread
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster_relaxed(&tmem_empty[acc], 0);
        }
        if (p.bias != nullptr) add_bias_chunk<kChunkN>(v, p.bias, p.bias_dtype, n0 + c * kChunkN, p.N);
        if constexpr (kAux && kChunkN == 64) {
          if (aux_load) {                              // v *= gelu'(aux): the slab was fetched while the accumulator was computed
            mbar_wait(my_aux_bar, aux_phase);
            aux_phase ^= 1;
            const uint8_t* xr = slab_x + lane * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint4 q = *reinterpret_cast<const uint4*>(xr + ((j ^ (lane & 7)) << 4));
              const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&q);
#pragma unroll
              for (int i = 0; i < 8; ++i) v[8 * j + i] *= gelu_tanh_grad(__bfloat162float(h[i]));
            }
            fence_proxy_async_smem();
            __syncwarp();                              // everybody has read the slab: the next fetch may overwrite it
            if (lane == 0) {
              if (c + 1 < (group + 1) * kGroupChunks) issue_aux_load(tile, c + 1);
              else if (tile + num_pairs < num_tiles) issue_aux_load(tile + num_pairs, group * kGroupChunks);
            }
          } else if (aux_store) {                      // pre-activation -> aux, through its own slab and TMA store
            if (lane == 0) tma_store_wait_read<1>();
            __syncwarp();
            uint8_t* xr = slab_x + lane * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              __nv_bfloat162 h0 = __floats2bfloat162_rn(v[8 * j + 0], v[8 * j + 1]);
              __nv_bfloat162 h1 = __floats2bfloat162_rn(v[8 * j + 2], v[8 * j + 3]);
              __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * j + 4], v[8 * j + 5]);
              __nv_bfloat162 h3 = __floats2bfloat162_rn(v[8 * j + 6], v[8 * j + 7]);
              uint4 q = {*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1), *reinterpret_cast<uint32_t*>(&h2),
                         *reinterpret_cast<uint32_t*>(&h3)};
              *reinterpret_cast<uint4*>(xr + ((j ^ (lane & 7)) << 4)) = q;
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmX, slab_x, n0 + c * kChunkN, m0 + quarter * 32);
              tma_store_commit();
            }
          }
        }
        if (p.act == ACT_RELU) {
#pragma unroll
          for (int j = 0; j < kChunkN; ++j) v[j] = fmaxf(v[j], 0.f);
        } else if (p.act == ACT_GELU) {
#pragma unroll
          for (int j = 0; j < kChunkN; ++j) v[j] = gelu_tanh(v[j]);
        }
        if (lane == 0) {                               // the previous store from the output slab has left it
          if (aux_store) tma_store_wait_read<1>(); else tma_store_wait_read<0>();
        }
        __syncwarp();
        uint8_t* row_ptr = slab + lane * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint4 q;
          if constexpr (sizeof(OutT) == 2) {
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
          *reinterpret_cast<uint4*>(row_ptr + ((j ^ (lane & 7)) << 4)) = q;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmD, slab, n0 + c * kChunkN, out_row0);
          tma_store_commit();
        }
        if constexpr (kAux && kChunkN == 64) {
          if (p.colsum_part != nullptr) {              // lane l: columns 2l, 2l+1 of the slab just staged (rows >= M hold zeros)
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int r = 0; r < 32; ++r) {
              const uint32_t w = *reinterpret_cast<const uint32_t*>(slab + r * 128 + ((((lane >> 2) ^ (r & 7)) << 4) | ((lane & 3) << 2)));
              s0 += __uint_as_float(w << 16);
              s1 += __uint_as_float(w & 0xffff0000u);
            }
            const int col = n0 + c * kChunkN + 2 * lane;
            if (col < p.N)
              *reinterpret_cast<float2*>(p.colsum_part + (long)(m_blk * 8 + (int)cta * 4 + quarter) * p.N + col) = make_float2(s0, s1);
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) tma_store_wait<0>();
  }

  tc_fence_before();
  cluster_sync_all();                 // nobody leaves (or frees TMEM) while the partner may still read its smem / signal its barriers
  if (warp_idx == 1) {
    tc_fence_after();
    uint32_t ncols = k2TmemCols;
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(ncols) : "memory");
  }
}

template <typename OutT, bool kAux = false>
static int launch_gemm2(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmD, const GemmParams& p, int max_ctas,
                        cudaStream_t stream, const CUtensorMap* tmX = nullptr) {
  auto kern = gemm2_tcgen05_kernel<OutT, kAux>;
  static bool configured = false;
  if (!configured) {
    DFB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, k2SmemBytes));
    configured = true;
  }
  const int tiles = ceil_div(p.M, 256) * ceil_div(p.N, 256) * p.splits;
  int pairs = num_sms() / 2;
  if (max_ctas > 0 && pairs > max_ctas / 2) pairs = max_ctas / 2;
  if (pairs > tiles) pairs = tiles;
  if (pairs < 1) pairs = 1;
  kern<<<2 * pairs, k2Threads, k2SmemBytes, stream>>>(tmA, tmB, tmD, tmX ? *tmX : tmD, p);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// out[m, n] = act(sum_s ws[s*split_rows + m, n] + bias[n]); ws fp32 with row pitch ldw.
// out[col] = sum over the 32-row slabs of part[slab][col], fixed order: 32 columns x 32 slab groups per block.
__global__ void __launch_bounds__(1024) colsum_part_reduce_kernel(const float* __restrict__ part, float* __restrict__ out, int slabs, int N) {
  __shared__ float sh[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float t = 0.f;
  if (c < N)
    for (int b = ty; b < slabs; b += 32) t += part[(long)b * N + c];
  sh[ty][tx] = t;
  __syncthreads();
  if (ty == 0 && c < N) {
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) tot += sh[k][tx];
    out[c] = tot;
  }
}

template <typename OutT>
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, OutT* __restrict__ out, const void* __restrict__ bias,
                                     int bias_dtype, int M, int N, int ldw, int ldd, int splits, int split_rows, int act) {
  const int n4 = N >> 2;
  const long total = (long)M * n4;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int m = (int)(i / n4), n = (int)(i - (long)m * n4) << 2;
    float4 acc = *reinterpret_cast<const float4*>(ws + (size_t)m * ldw + n);
    for (int s = 1; s < splits; ++s) {
      float4 t = *reinterpret_cast<const float4*>(ws + ((size_t)s * split_rows + m) * ldw + n);
      acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
    }
    float v[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (bias != nullptr)
        v[j] += bias_dtype == DT_F32 ? reinterpret_cast<const float*>(bias)[n + j]
                                     : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias)[n + j]);
      if (act == ACT_RELU) v[j] = fmaxf(v[j], 0.f);
      else if (act == ACT_GELU) v[j] = gelu_tanh(v[j]);
    }
    if constexpr (sizeof(OutT) == 2) {
      __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
      uint2 q = {*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1)};
      *reinterpret_cast<uint2*>(out + (size_t)m * ldd + n) = q;
    } else {
      *reinterpret_cast<float4*>(out + (size_t)m * ldd + n) = make_float4(v[0], v[1], v[2], v[3]);
    }
  }
}

template <int BLOCK_N, typename OutT, bool kGather = false>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmD, const GemmParams& p,
                       int max_ctas, cudaStream_t stream) {
  using Cfg = GemmCfg<BLOCK_N>;
  auto kern = gemm_tcgen05_kernel<BLOCK_N, OutT, kGather>;
  static bool configured = false;      // per instantiation; attribute is per device function (all devices of one kind)
  if (!configured) {
    DFB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured = true;
  }
  const int tiles = ceil_div(p.M, kBlockM) * ceil_div(p.N, BLOCK_N) * p.splits;
  int grid = tiles < num_sms() ? tiles : num_sms();
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  kern<<<grid, kThreads + (kGather ? 64 : 0), Cfg::kSmemBytes, stream>>>(tmA, tmB, tmD, p);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

static int make_operand_map(CUtensorMap* tm, const void* ptr, int rows, int K, int ld, int mn_major, int block_rows) {
  if (mn_major) {       // memory [K][rows]: inner = rows
    uint64_t dims[2] = {(uint64_t)rows, (uint64_t)K};
    uint64_t str[1] = {(uint64_t)ld * 2};
    uint32_t box[2] = {64, (uint32_t)kBlockK};
    return encode_tmap(tm, ptr, 0, 2, dims, str, box, nullptr, 1);
  }
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  uint64_t str[1] = {(uint64_t)ld * 2};
  uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)block_rows};
  return encode_tmap(tm, ptr, 0, 2, dims, str, box, nullptr, 1);
}

}  // namespace dfb

using namespace dfb;

// Bytes of fp32 scratch dfb_gemm_bf16 needs for a given problem and split factor (0 when splits <= 1).
DFB_API size_t dfb_gemm_workspace_bytes(int M, int N, int splits) {
  if (splits <= 1) return 0;
  size_t mpad = (size_t)ceil_div(M, 256) * 256;          // 256: the 2-CTA kernel works on 256-row tiles
  return (size_t)splits * mpad * (size_t)N * sizeof(float);
}

// Heuristic split-K factor: fill the SMs when the MN tile count is small and K is long.
DFB_API int dfb_gemm_pick_splits(int M, int N, int K, int block_n, int sms) {
  int tiles = ceil_div(M, kBlockM) * ceil_div(N, block_n);
  int kb = ceil_div(K, kBlockK);
  if (tiles * 2 > sms || kb < 16) return 1;
  int s = sms / tiles;
  if (s > kb / 8) s = kb / 8;
  if (s > 16) s = 16;
  return s < 1 ? 1 : s;
}

// See include/distriflow_b200.h.  `workspace` (fp32 scratch) is required iff splits > 1.
static int gemm_impl(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K, int lda, int ldb,
                     int ldd, int a_mn_major, int b_mn_major, int out_dtype, int act, int block_n, int splits, void* workspace,
                     size_t workspace_bytes, int max_ctas, void* aux, long aux_ld, int aux_mode, cudaStream_t stream,
                     float* colsum_part = nullptr) {
  DFB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %d x %d x %d", M, N, K);
  DFB_REQUIRE(colsum_part == nullptr || (block_n == 512 && aux_mode == 2 && bias == nullptr && out_dtype == DT_BF16),
              "gemm: epilogue column sums exist in the 2-CTA gelu'-epilogue kernel only (block_n 512, aux_mode 2, no bias)");
  DFB_REQUIRE(aux_mode == 0 || (aux != nullptr && out_dtype == DT_BF16 && (aux_ld % 8) == 0 && (N % 8) == 0 &&
                                ((uintptr_t)aux % 16) == 0),
              "gemm: the aux epilogue needs bf16 output, N and aux pitch multiples of 8, 16-byte aligned aux");
  DFB_REQUIRE(block_n == 128 || block_n == 256 || block_n == 512, "gemm: block_n must be 128, 256 or 512 (= 2-CTA 256x256)");
  const bool two_cta = (block_n == 512);
  if (two_cta) block_n = 256;
  DFB_REQUIRE(out_dtype == DT_BF16 || out_dtype == DT_F32, "gemm: out dtype");
  DFB_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "gemm: lda/ldb must be multiples of 8 elements (TMA 16-byte pitch)");
  DFB_REQUIRE((ldd * (out_dtype == DT_BF16 ? 2 : 4)) % 16 == 0, "gemm: output pitch must be a multiple of 16 bytes");
  DFB_REQUIRE(((uintptr_t)A % 16) == 0 && ((uintptr_t)B % 16) == 0 && ((uintptr_t)D % 16) == 0, "gemm: 16-byte alignment");
  const int kb_total = ceil_div(K, kBlockK);
  if (splits < 1) splits = 1;
  if (splits > kb_total) splits = kb_total;
  int kb_per = ceil_div(kb_total, splits);
  splits = ceil_div(kb_total, kb_per);

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.bias = bias; p.bias_dtype = bias_dtype; p.M = M; p.N = N; p.K = K; p.a_mn = a_mn_major ? 1 : 0; p.b_mn = b_mn_major ? 1 : 0;
  p.act = act; p.splits = splits; p.kb_per_split = kb_per; p.split_rows = 0;
  p.aux = aux; p.aux_ld = aux_ld; p.aux_mode = aux_mode; p.colsum_part = colsum_part;
  DFB_REQUIRE(aux_mode == 0 || splits == 1, "gemm: the aux epilogue cannot be combined with split-K");

  CUtensorMap tmA, tmB, tmD;
  if (make_operand_map(&tmA, A, M, K, lda, p.a_mn, kBlockM)) return 4;
  if (make_operand_map(&tmB, B, N, K, ldb, p.b_mn, two_cta ? 128 : block_n)) return 4;

  if (splits > 1) {
    DFB_REQUIRE((N % 4) == 0, "gemm split-K: N must be a multiple of 4");
    const size_t need = dfb_gemm_workspace_bytes(M, N, splits);
    DFB_REQUIRE(workspace != nullptr && workspace_bytes >= need, "gemm split-K: workspace of %zu bytes required, got %zu",
                need, workspace_bytes);
    const int mpad = ceil_div(M, 256) * 256;
    p.split_rows = mpad;
    GemmParams pk = p;
    pk.bias = nullptr; pk.act = ACT_NONE;
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)splits * mpad};
    uint64_t str[1] = {(uint64_t)N * 4};
    uint32_t box[2] = {32, two_cta ? 32u : 128u};          // the 2-CTA kernel stores one 32-row slab per epilogue warp
    if (encode_tmap(&tmD, workspace, 1, 2, dims, str, box, nullptr, 1)) return 4;
    int rc = two_cta ? launch_gemm2<float>(tmA, tmB, tmD, pk, max_ctas, stream)
             : block_n == 256 ? launch_gemm<256, float>(tmA, tmB, tmD, pk, max_ctas, stream)
                              : launch_gemm<128, float>(tmA, tmB, tmD, pk, max_ctas, stream);
    if (rc) return rc;
    const long total = (long)M * (N / 4);
    int blocks = (int)((total + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    if (out_dtype == DT_BF16)
      splitk_reduce_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>((const float*)workspace, (__nv_bfloat16*)D, bias,
                                                                      bias_dtype, M, N, N, ldd, splits, mpad, act);
    else
      splitk_reduce_kernel<float><<<blocks, 256, 0, stream>>>((const float*)workspace, (float*)D, bias, bias_dtype, M, N, N,
                                                              ldd, splits, mpad, act);
    DFB_CUDA_CHECK(cudaGetLastError());
    return 0;
  }

  {
    const int esz = out_dtype == DT_BF16 ? 2 : 4;
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)ldd * esz};
    uint32_t box[2] = {(uint32_t)(128 / esz), two_cta ? 32u : 128u};
    if (encode_tmap(&tmD, D, out_dtype == DT_BF16 ? 0 : 1, 2, dims, str, box, nullptr, 1)) return 4;
  }
  if (two_cta && aux_mode != 0) {                          // the aux tensor travels through TMA like the output
    CUtensorMap tmX;
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)aux_ld * 2};
    uint32_t box[2] = {64, 32};
    if (encode_tmap(&tmX, aux, 0, 2, dims, str, box, nullptr, 1)) return 4;
    return launch_gemm2<__nv_bfloat16, true>(tmA, tmB, tmD, p, max_ctas, stream, &tmX);
  }
  if (two_cta)
    return out_dtype == DT_BF16 ? launch_gemm2<__nv_bfloat16>(tmA, tmB, tmD, p, max_ctas, stream)
                                : launch_gemm2<float>(tmA, tmB, tmD, p, max_ctas, stream);
  if (out_dtype == DT_BF16)
    return block_n == 256 ? launch_gemm<256, __nv_bfloat16>(tmA, tmB, tmD, p, max_ctas, stream)
                          : launch_gemm<128, __nv_bfloat16>(tmA, tmB, tmD, p, max_ctas, stream);
  return block_n == 256 ? launch_gemm<256, float>(tmA, tmB, tmD, p, max_ctas, stream)
                        : launch_gemm<128, float>(tmA, tmB, tmD, p, max_ctas, stream);
}

// See include/distriflow_b200.h.  `workspace` (fp32 scratch) is required iff splits > 1.
DFB_API int dfb_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K,
                          int lda, int ldb, int ldd, int a_mn_major, int b_mn_major, int out_dtype, int act, int block_n,
                          int splits, void* workspace, size_t workspace_bytes, int max_ctas, cudaStream_t stream) {
  return gemm_impl(A, B, D, bias, bias_dtype, M, N, K, lda, ldb, ldd, a_mn_major, b_mn_major, out_dtype, act, block_n, splits,
                   workspace, workspace_bytes, max_ctas, nullptr, 0, 0, stream);
}

// Same GEMM with an auxiliary bf16 matrix in the epilogue: aux_mode 1 stores the pre-activation (accumulator + bias) to
// `aux` while D receives act(...); aux_mode 2 multiplies the accumulator by gelu'(aux) (GELU backward fused into the dgrad
// of the following layer).  No split-K.
DFB_API int dfb_gemm_bf16_aux(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N, int K,
                              int lda, int ldb, int ldd, int a_mn_major, int b_mn_major, int act, int block_n, void* aux,
                              long aux_ld, int aux_mode, int max_ctas, cudaStream_t stream) {
  return gemm_impl(A, B, D, bias, bias_dtype, M, N, K, lda, ldb, ldd, a_mn_major, b_mn_major, DT_BF16, act, block_n, 1, nullptr,
                   0, max_ctas, aux, aux_ld, aux_mode, stream);
}

// dgrad GEMM with the GELU backward in its epilogue (aux_mode 2) that ALSO produces the column sums of its output, i.e. the
// bias gradient of the layer in front of the GELU: D = (A B^T) * gelu'(aux), colsum[n] = sum_m D[m, n] over the stored
// (bf16-rounded) values.  `part`: fp32 scratch of dfb_gemm_colsum_part_bytes(M, N) bytes.  2-CTA kernel only.
DFB_API size_t dfb_gemm_colsum_part_bytes(int M, int N) { return (size_t)ceil_div(M, 256) * 8 * (size_t)N * sizeof(float); }

DFB_API int dfb_gemm_bf16_dgelu_colsum(const void* A, const void* B, void* D, int M, int N, int K, int lda, int ldb, int ldd,
                                       int a_mn_major, int b_mn_major, void* aux, long aux_ld, float* colsum, float* part,
                                       size_t part_bytes, int max_ctas, cudaStream_t stream) {
  DFB_REQUIRE(colsum != nullptr && part != nullptr && part_bytes >= dfb_gemm_colsum_part_bytes(M, N),
              "gemm dgelu+colsum: scratch of %zu bytes required, got %zu", dfb_gemm_colsum_part_bytes(M, N), part_bytes);
  int rc = gemm_impl(A, B, D, nullptr, 0, M, N, K, lda, ldb, ldd, a_mn_major, b_mn_major, DT_BF16, ACT_NONE, 512, 1, nullptr, 0,
                     max_ctas, aux, aux_ld, 2, stream, part);
  if (rc) return rc;
  colsum_part_reduce_kernel<<<ceil_div(N, 32), 1024, 0, stream>>>(part, colsum, ceil_div(M, 256) * 8, N);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// All-gather fused into the GEMM that first consumes the gathered weights (BASELINE north star, SURVEY K8+K1):
//   D[M,N] = act(A[M,K] * W[K,N] + bias), W row-major (MN-major B) and sharded over the ranks in contiguous byte ranges.
// `pieces` (device, int64 [num_pieces][4] = {peer source address, local destination address, bytes, k-block}) lists
// the remote ranges of W, cut at k-block (64-row) boundaries; `piece_chunk0` (device, int32 [num_pieces+1]) are prefix sums
// of their 512-byte chunk counts; `expected` (device, uint32 [k-blocks]) the remote bytes per k-block; `ready` (device,
// uint32 [k-blocks + 1]) is zeroed here.  Two extra warps per CTA copy the remote shards into the local W buffer over
// NVLink while the k-loop starts on the locally owned rows (`rot_kb`); the TMA producer waits on `ready[kb]` per k-block.
// The owners' slices must be final (the closing barrier of the bucket kernel guarantees it).
DFB_API int dfb_ag_gemm_bf16(const void* A, const void* W_local, void* D, const void* bias, int bias_dtype, int M, int N, int K,
                             int lda, int ldb, int ldd, int out_dtype, int act, int block_n, const long long* pieces,
                             const int* piece_chunk0, int num_pieces, const unsigned* expected, unsigned* ready, int rot_kb,
                             cudaStream_t stream) {
  DFB_REQUIRE(M > 0 && N > 0 && K > 0, "ag_gemm: empty problem");
  DFB_REQUIRE(block_n == 128 || block_n == 256, "ag_gemm: block_n must be 128 or 256");
  DFB_REQUIRE(out_dtype == DT_BF16, "ag_gemm: bf16 output only");
  DFB_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0 && (ldd % 8) == 0, "ag_gemm: pitches must be multiples of 8 elements");
  const int kb_total = ceil_div(K, kBlockK);
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.bias = bias; p.bias_dtype = bias_dtype; p.M = M; p.N = N; p.K = K; p.a_mn = 0; p.b_mn = 1; p.act = act; p.splits = 1;
  p.kb_per_split = kb_total; p.pieces = pieces; p.piece_chunk0 = piece_chunk0; p.expected = expected; p.ready = ready;
  p.num_pieces = num_pieces; p.rot_kb = ((rot_kb % kb_total) + kb_total) % kb_total;
  CUtensorMap tmA, tmB, tmD;
  if (make_operand_map(&tmA, A, M, K, lda, 0, kBlockM)) return 4;
  if (make_operand_map(&tmB, W_local, N, K, ldb, 1, block_n)) return 4;
  uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
  uint64_t str[1] = {(uint64_t)ldd * 2};
  uint32_t box[2] = {64, 128};
  if (encode_tmap(&tmD, D, 0, 2, dims, str, box, nullptr, 1)) return 4;
  DFB_CUDA_CHECK(cudaMemsetAsync(ready, 0, (size_t)(kb_total + 1) * sizeof(unsigned), stream));
  return block_n == 256 ? launch_gemm<256, __nv_bfloat16, true>(tmA, tmB, tmD, p, 0, stream)
                        : launch_gemm<128, __nv_bfloat16, true>(tmA, tmB, tmD, p, 0, stream);
}

// ---------------------------------------------------------------------------------------------------------
// Split-K GEMM for callers that must not allocate.
//
// Who this is for: code that runs where a device allocation is not an option — a step loop that has promised
// never to block the device (cudaMalloc and cudaFree synchronise it), a plug-in living inside a host program
// whose memory pool already owns the whole GPU and is not reachable from C (for instance a PyTorch process
// whose caching allocator holds the memory while the caller is a plain C/C++ extension without libtorch), or a
// job sized so close to the 180 GB limit that a late allocation could simply fail in the middle of training.
// Such a caller cannot hand dfb_gemm_bf16 the fp32 workspace that its split-K path needs, so this entry point
// brings the workspace with it: a fixed array in the library's own device data segment.
//
// Limit: kNoAllocScratchBytes (32 MiB) — dfb_gemm_workspace_bytes(M, N, splits) must not exceed it, otherwise
// the call fails with a message that states both numbers.  The array exists once per GPU that loads this
// library and stays until the process exits.
//
// Sharing rules: every call on a device uses the same array.  Nothing in it has to start at zero — each split
// writes its whole [ceil256(M) x N] partial tile grid (rows padded to 256, as in dfb_gemm_workspace_bytes) with plain stores before the reduce kernel reads it, so
// whatever the previous call left behind is overwritten, never accumulated into.  Two calls whose kernels can
// overlap on the GPU (different streams without an ordering between them) race on the array and corrupt each
// other; keep calls on one stream or order them with events.
//
// Set-up: in host code the name g_noalloc_scratch is only a host-side handle for the device array, not an
// address a kernel can use.  The first call on each device (with that device current) asks the runtime for the
// real device address with cudaGetSymbolAddress and remembers it; later calls reuse it.  Make that first call
// before starting any stream capture.
constexpr size_t kNoAllocScratchBytes = size_t(32) << 20;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchBytes / sizeof(float)];
// device address of the array, per device ordinal (0 = not resolved yet).  Filled without a lock: two host threads making
// the first call on one device both resolve the same address and store the same value (a benign duplicate store).
static float* g_noalloc_scratch_addr[64];

DFB_API int dfb_gemm_bf16_splitk_noalloc(const void* A, const void* B, void* D, const void* bias, int bias_dtype, int M, int N,
                                         int K, int lda, int ldb, int ldd, int a_mn_major, int b_mn_major, int out_dtype,
                                         int act, int block_n, int splits, int max_ctas, cudaStream_t stream) {
  DFB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm (no-alloc): empty problem %d x %d x %d", M, N, K);
  DFB_REQUIRE(splits >= 1, "gemm (no-alloc): splits must be >= 1");
  const size_t need = dfb_gemm_workspace_bytes(M, N, splits);
  DFB_REQUIRE(need <= kNoAllocScratchBytes,
              "gemm (no-alloc): M=%d N=%d splits=%d needs %zu bytes of split-K scratch, the built-in array holds %zu; "
              "lower `splits` or use dfb_gemm_bf16 with a caller-provided workspace",
              M, N, splits, need, kNoAllocScratchBytes);
  int dev = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  DFB_REQUIRE(dev >= 0 && dev < 64, "gemm (no-alloc): device ordinal %d out of range", dev);
  if (g_noalloc_scratch_addr[dev] == nullptr) {       // once per device, with that device current
    float* p = nullptr;
    DFB_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    DFB_REQUIRE(p != nullptr, "gemm (no-alloc): cudaGetSymbolAddress returned a null address");
    g_noalloc_scratch_addr[dev] = p;
  }
  return dfb_gemm_bf16(A, B, D, bias, bias_dtype, M, N, K, lda, ldb, ldd, a_mn_major, b_mn_major, out_dtype, act, block_n,
                       splits, g_noalloc_scratch_addr[dev], kNoAllocScratchBytes, max_ctas, stream);
}
