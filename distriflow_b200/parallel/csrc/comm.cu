// Communication runtime + fused collective kernels over NVLink 5 / NVSwitch peer memory.
//
// What the reference does with websocket messages becomes loads and stores on
// peer-mapped device memory issued from inside kernels:
//   * upload + mean + `update` + broadcast of new weights, synchronous mode
//     (/root/reference/src/server/federated_server.ts:71-116)          -> sync_rs_opt_ag_kernel:
//     every rank owns 1/R of each bucket, PULLS that slice of everybody's gradients with P2P loads, reduces in
//     fp32 in a fixed order, scales by 1/R, applies SGD / momentum / Adam to its fp32 master shard in place and
//     PUSHES the refreshed working weights to every rank with P2P stores (reduce-scatter + optimizer +
//     all-gather in ONE launch; no NCCL on the path).
//   * upload + immediate `update`, asynchronous mode
//     (/root/reference/src/server/asynchronousSGD_server.ts:65-108)    -> async_push / async_apply / async_pull:
//     workers add gradients into the owner shard's accumulation buffer with P2P reductions guarded by a
//     device-side staleness counter, the owner applies them in place, workers pull weights with P2P loads.
//   * `dataset.next()` / `completeBatch()` (/root/reference/src/server/dataset.ts:43-67) -> device work queue.
// Bootstrap (exchange of cudaIpc handles) is done by the Python side through torch.distributed.
#include "common.cuh"
#include "device_utils.cuh"

namespace dfb {

constexpr int kMaxRanks = 8;
static int g_close_per_bucket = 1;      // see dfb_sync_close_per_bucket

struct PeerPtrs {
  void* p[kMaxRanks];
};

// ---------------------------------------------------------------- system-scope signalling
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {       // peer data: never trust a cached line
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_v4(void* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void red_add_f32_sys(float* p, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;\n" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add_v4_f32_sys(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Signal pad layout (uint32 words) inside every rank's symmetric heap:
//   [0, 8)    arrive-in  flags, one per source rank      [8, 16)   arrive-out flags
//   [16]      local block counter                        [32, ...) async / work-queue words (see below)
//   [17]      barrier time-out record: 1 + the rank that never arrived (0 = healthy)      [18] generation it was waited for
constexpr int kSigIn = 0, kSigOut = 8, kSigBlocks = 16, kSigDead = 17, kSigDeadEpoch = 18;

// Barrier waits are bounded: a peer that died (or never launched its kernel) must not leave every other GPU spinning inside a
// kernel for ever, where not even the host watchdog can do more than print.  After `g_barrier_timeout_cycles` SM cycles the
// waiter records WHO it was waiting for in its own signal pad and carries on; the step's result is garbage, but the kernel
// ends, the stream drains and the host (`dfb_barrier_failures`, the engine's watchdog) can name the dead rank and stop.
__device__ unsigned long long g_barrier_timeout_cycles = 40000000000ull;       // ~20 s at 2 GHz; dfb_set_barrier_timeout_us

// Every block waits until all ranks have published `epoch` in this rank's `which` flags; block 0 publishes ours.
__device__ __forceinline__ void rank_barrier(const PeerPtrs& sig, int rank, int R, int which, uint32_t epoch, bool do_signal) {
  if (do_signal && threadIdx.x < R) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(sig.p[threadIdx.x]) + which + rank, epoch);
  }
  if (threadIdx.x < R) {
    uint32_t* pad = reinterpret_cast<uint32_t*>(sig.p[rank]);
    const uint32_t* mine = pad + which + threadIdx.x;
    const long long t0 = clock64();
    const unsigned long long limit = g_barrier_timeout_cycles;
    uint32_t polls = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if ((++polls & 1023u) == 0 && (unsigned long long)(clock64() - t0) > limit) {
        pad[kSigDeadEpoch] = epoch;
        st_release_sys(pad + kSigDead, 1u + threadIdx.x);        // rank threadIdx.x never published this generation
        break;
      }
    }
  }
  __syncthreads();
}

// ---------------------------------------------------------------- optimizer on 4/8 fp32 lanes
enum Opt : int { OPT_SGD = 0, OPT_MOMENTUM = 1, OPT_ADAM = 2, OPT_NONE = 3 };

struct OptParams {
  int kind;
  float lr, momentum, beta1, beta2, eps, weight_decay;
  float bias_c1, bias_c2;       // 1 - beta^t (host computed)
};

__device__ __forceinline__ float opt_step(float w, float g, float* m, float* v, const OptParams& o) {
  if (o.weight_decay != 0.f) g += o.weight_decay * w;
  if (o.kind == OPT_SGD) return w - o.lr * g;                    // /root/reference/src/common/models.ts:128-135
  if (o.kind == OPT_MOMENTUM) {
    float mm = o.momentum * (*m) + g;
    *m = mm;
    return w - o.lr * mm;
  }
  if (o.kind == OPT_ADAM) {
    float mm = o.beta1 * (*m) + (1.f - o.beta1) * g;
    float vv = o.beta2 * (*v) + (1.f - o.beta2) * g * g;
    *m = mm;
    *v = vv;
    return w - o.lr * (mm / o.bias_c1) / (sqrtf(vv / o.bias_c2) + o.eps);
  }
  return w;
}

// ---------------------------------------------------------------- sync: reduce-scatter + 1/R + optimizer + all-gather
struct SyncArgs {
  PeerPtrs grads;        // base of this bucket's gradient buffer on every rank (TG)
  PeerPtrs params;       // base of this bucket's working-weight buffer on every rank (TP)
  PeerPtrs sig;          // signal pads
  float* master;         // fp32 master weights of the OWNED slice (local)
  float* m;              // optimizer state of the owned slice (local; may be null)
  float* v;
  void* reduced_out;     // optional: owned slice of the mean gradient, TG (for tests / OPT_NONE)
  long n;                // elements in the bucket
  long shard;            // elements per rank slice (multiple of 8)
  int rank, R;
  uint32_t epoch;        // barrier generation, strictly increasing per launch; 0 = no cross-rank barriers (emulation)
  int push;              // 1 = push refreshed weights to every rank, 0 = only write the local copy
  long nopush_lo, nopush_hi;   // bucket-relative element range that is NOT pushed (its all-gather is fused into the consumer GEMM)
  // CUDA-graph replay support: values that change from step to step are read from device memory instead of being baked in
  const uint32_t* epoch_base;  // if set (and epoch != 0): barrier generation = *epoch_base + epoch
  const int* step_ptr;         // if set: optimizer step count for Adam's bias correction
  const float* lr_ptr;         // if set: learning rate (schedules keep working under a captured graph)
  OptParams opt;
};

template <typename TG, typename TP, int R_T>
__global__ void __launch_bounds__(256) sync_rs_opt_ag_kernel(const SyncArgs a) {   // 256 x <=128 regs: shares an SM with a GEMM CTA
  const int R = R_T > 0 ? R_T : a.R;
  uint32_t epoch = a.epoch;
  if (epoch && a.epoch_base) epoch += *a.epoch_base;
  OptParams opt = a.opt;
  if (a.lr_ptr != nullptr) opt.lr = *a.lr_ptr;
  if (a.step_ptr != nullptr && opt.kind == OPT_ADAM) {
    const float t = (float)max(*a.step_ptr, 1);
    opt.bias_c1 = 1.f - powf(opt.beta1, t);
    opt.bias_c2 = 1.f - powf(opt.beta2, t);
  }
  if (epoch) rank_barrier(a.sig, a.rank, R, kSigIn, epoch, blockIdx.x == 0);   // everybody's gradients are final
  using VG = Vec16<TG>;
  constexpr int VN = 8;                       // elements per vector (16 B of bf16, or 2 x 16 B of fp32)
  constexpr int GV = VN / VG::kN;             // 16-byte loads per peer per vector
  // vectors per thread-iteration: keep ~8 peer loads of 16 B in flight per thread whatever the world size, so that the
  // NVLink latency (~2 us) is covered at R = 2 as well as at R = 8
  constexpr int U = (R_T == 2 && GV == 1) ? 4 : ((R_T == 4 && GV == 1) || (R_T == 2)) ? 2 : 1;
  const long lo = (long)a.rank * a.shard;
  const long hi = min(a.n, lo + a.shard);
  const long nvec = hi > lo ? (hi - lo + VN - 1) / VN : 0;
  const float inv_r = 1.f / (float)R;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i0 = blockIdx.x * (long)blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
    VG g[U][kMaxRanks][GV];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = i0 + u * stride;
      if (i < nvec) {
        const long e = lo + i * VN;
#pragma unroll
        for (int q = 0; q < kMaxRanks; ++q) {
          if (q < R) {
            const TG* src = reinterpret_cast<const TG*>(a.grads.p[q]) + e;
#pragma unroll
            for (int h = 0; h < GV; ++h) g[u][q][h].raw = ld_peer_v4(src + h * VG::kN);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = i0 + u * stride;
      if (i >= nvec) continue;
      const long e = lo + i * VN;               // first element (bucket-relative)
      float acc[VN];
#pragma unroll
      for (int j = 0; j < VN; ++j) acc[j] = 0.f;
#pragma unroll
      for (int q = 0; q < kMaxRanks; ++q) {     // fixed order -> bitwise reproducible, identical on every rank
        if (q < R) {
#pragma unroll
          for (int h = 0; h < GV; ++h)
#pragma unroll
            for (int j = 0; j < VG::kN; ++j) acc[h * VG::kN + j] += g[u][q][h].get(j);
        }
      }
      const long s = i * VN;                    // shard-relative
      float w[VN];
      {
        const float4 w0 = *reinterpret_cast<const float4*>(a.master + s);
        const float4 w1 = *reinterpret_cast<const float4*>(a.master + s + 4);
        w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w; w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
      }
      float mm[VN], vv[VN];
      if (opt.kind == OPT_MOMENTUM || opt.kind == OPT_ADAM) {
        const float4 m0 = *reinterpret_cast<const float4*>(a.m + s), m1 = *reinterpret_cast<const float4*>(a.m + s + 4);
        mm[0] = m0.x; mm[1] = m0.y; mm[2] = m0.z; mm[3] = m0.w; mm[4] = m1.x; mm[5] = m1.y; mm[6] = m1.z; mm[7] = m1.w;
      }
      if (opt.kind == OPT_ADAM) {
        const float4 v0 = *reinterpret_cast<const float4*>(a.v + s), v1 = *reinterpret_cast<const float4*>(a.v + s + 4);
        vv[0] = v0.x; vv[1] = v0.y; vv[2] = v0.z; vv[3] = v0.w; vv[4] = v1.x; vv[5] = v1.y; vv[6] = v1.z; vv[7] = v1.w;
      }
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        acc[j] *= inv_r;
        w[j] = opt_step(w[j], acc[j], &mm[j], &vv[j], opt);
      }
      if (opt.kind != OPT_NONE) {
        *reinterpret_cast<float4*>(a.master + s) = make_float4(w[0], w[1], w[2], w[3]);
        *reinterpret_cast<float4*>(a.master + s + 4) = make_float4(w[4], w[5], w[6], w[7]);
        if (opt.kind == OPT_MOMENTUM || opt.kind == OPT_ADAM) {
          *reinterpret_cast<float4*>(a.m + s) = make_float4(mm[0], mm[1], mm[2], mm[3]);
          *reinterpret_cast<float4*>(a.m + s + 4) = make_float4(mm[4], mm[5], mm[6], mm[7]);
        }
        if (opt.kind == OPT_ADAM) {
          *reinterpret_cast<float4*>(a.v + s) = make_float4(vv[0], vv[1], vv[2], vv[3]);
          *reinterpret_cast<float4*>(a.v + s + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
        }
      }
      if (a.reduced_out != nullptr) {
        VG o[GV];
#pragma unroll
        for (int h = 0; h < GV; ++h) {
#pragma unroll
          for (int j = 0; j < VG::kN; ++j) o[h].set(j, acc[h * VG::kN + j]);
          reinterpret_cast<uint4*>(reinterpret_cast<TG*>(a.reduced_out) + s)[h] = o[h].raw;
        }
      }
      if (opt.kind != OPT_NONE) {
        using VP = Vec16<TP>;
        constexpr int PV = VN / VP::kN;
        VP o[PV];
#pragma unroll
        for (int h = 0; h < PV; ++h)
#pragma unroll
          for (int j = 0; j < VP::kN; ++j) o[h].set(j, w[h * VP::kN + j]);
        const bool pushed = (a.push & 1) && !(e >= a.nopush_lo && e < a.nopush_hi);
#pragma unroll
        for (int q = 0; q < kMaxRanks; ++q) {
          if (q < R && (pushed || q == a.rank)) {
            TP* dst = reinterpret_cast<TP*>(a.params.p[q]) + e;
#pragma unroll
            for (int h = 0; h < PV; ++h) st_peer_v4(dst + h * VP::kN, o[h].raw);
          }
        }
      }
    }
  }
  if (epoch && !(a.push & 2)) {
    // all blocks done -> the last one publishes "my slice is everywhere" and waits for everybody else's
    // (push bit 1 set: the step closes with ONE barrier after its last bucket instead, see step_close_kernel)
    __shared__ bool last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t* cnt = reinterpret_cast<uint32_t*>(a.sig.p[a.rank]) + kSigBlocks;
      last = (atomicAdd(cnt, 1u) == gridDim.x - 1);
      if (last) *cnt = 0;
    }
    __syncthreads();
    if (last) rank_barrier(a.sig, a.rank, R, kSigOut, epoch, true);
  }
}

// ---------------------------------------------------------------- sync over NVLS: the switch reduces and broadcasts
// Same contract as sync_rs_opt_ag_kernel for bf16 gradients and bf16 working weights, with the two transfers handed to the
// NVSwitch through a multicast mapping of the symmetric heap: ONE `multimem.ld_reduce` per 16 bytes returns the sum of
// all R ranks' gradients (added inside the switch with fp32 accumulation, delivered rounded to bf16), and ONE `multimem.st`
// writes the refreshed weights to all R ranks.  Per GPU the link then carries S/R in and S/R out per bucket instead of
// (R-1)/R x S each way.  Barriers stay on the unicast signal pads.
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_v4(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};\n" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

struct NvlsArgs {
  SyncArgs s;                      // grads unused; params = unicast bases (local / no-push stores)
  const __nv_bfloat16* mc_grads;   // multicast address of this bucket's gradient buffer
  __nv_bfloat16* mc_params;        // multicast address of this bucket's working-weight buffer
};

__global__ void __launch_bounds__(256) sync_nvls_kernel(const NvlsArgs na) {
  const SyncArgs& a = na.s;
  const int R = a.R;
  uint32_t epoch = a.epoch;
  if (epoch && a.epoch_base) epoch += *a.epoch_base;
  OptParams opt = a.opt;
  if (a.lr_ptr != nullptr) opt.lr = *a.lr_ptr;
  if (a.step_ptr != nullptr && opt.kind == OPT_ADAM) {
    const float t = (float)max(*a.step_ptr, 1);
    opt.bias_c1 = 1.f - powf(opt.beta1, t);
    opt.bias_c2 = 1.f - powf(opt.beta2, t);
  }
  rank_barrier(a.sig, a.rank, R, kSigIn, epoch, blockIdx.x == 0);
  using V = Vec16<__nv_bfloat16>;
  constexpr int VN = 8, U = 4;
  const long lo = (long)a.rank * a.shard;
  const long hi = min(a.n, lo + a.shard);
  const long nvec = hi > lo ? (hi - lo + VN - 1) / VN : 0;
  const float inv_r = 1.f / (float)R;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i0 = blockIdx.x * (long)blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
    V g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = i0 + u * stride;
      if (i < nvec) g[u].raw = multimem_ld_reduce_bf16x8(na.mc_grads + lo + i * VN);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long i = i0 + u * stride;
      if (i >= nvec) continue;
      const long e = lo + i * VN, s = i * VN;
      float w[VN], mm[VN], vv[VN];
      {
        const float4 w0 = *reinterpret_cast<const float4*>(a.master + s), w1 = *reinterpret_cast<const float4*>(a.master + s + 4);
        w[0] = w0.x; w[1] = w0.y; w[2] = w0.z; w[3] = w0.w; w[4] = w1.x; w[5] = w1.y; w[6] = w1.z; w[7] = w1.w;
      }
      if (opt.kind == OPT_MOMENTUM || opt.kind == OPT_ADAM) {
        const float4 m0 = *reinterpret_cast<const float4*>(a.m + s), m1 = *reinterpret_cast<const float4*>(a.m + s + 4);
        mm[0] = m0.x; mm[1] = m0.y; mm[2] = m0.z; mm[3] = m0.w; mm[4] = m1.x; mm[5] = m1.y; mm[6] = m1.z; mm[7] = m1.w;
      }
      if (opt.kind == OPT_ADAM) {
        const float4 v0 = *reinterpret_cast<const float4*>(a.v + s), v1 = *reinterpret_cast<const float4*>(a.v + s + 4);
        vv[0] = v0.x; vv[1] = v0.y; vv[2] = v0.z; vv[3] = v0.w; vv[4] = v1.x; vv[5] = v1.y; vv[6] = v1.z; vv[7] = v1.w;
      }
      V o;
#pragma unroll
      for (int j = 0; j < VN; ++j) {
        w[j] = opt_step(w[j], g[u].get(j) * inv_r, &mm[j], &vv[j], opt);
        o.set(j, w[j]);
      }
      *reinterpret_cast<float4*>(a.master + s) = make_float4(w[0], w[1], w[2], w[3]);
      *reinterpret_cast<float4*>(a.master + s + 4) = make_float4(w[4], w[5], w[6], w[7]);
      if (opt.kind == OPT_MOMENTUM || opt.kind == OPT_ADAM) {
        *reinterpret_cast<float4*>(a.m + s) = make_float4(mm[0], mm[1], mm[2], mm[3]);
        *reinterpret_cast<float4*>(a.m + s + 4) = make_float4(mm[4], mm[5], mm[6], mm[7]);
      }
      if (opt.kind == OPT_ADAM) {
        *reinterpret_cast<float4*>(a.v + s) = make_float4(vv[0], vv[1], vv[2], vv[3]);
        *reinterpret_cast<float4*>(a.v + s + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
      }
      if (e >= a.nopush_lo && e < a.nopush_hi)       // gathered by the consumer GEMM: only the owner's copy is written
        *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(a.params.p[a.rank]) + e) = o.raw;
      else
        multimem_st_v4(na.mc_params + e, o.raw);
    }
  }
  if (a.push & 2) return;          // closing barrier once per step (step_close_kernel)
  __shared__ bool last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* cnt = reinterpret_cast<uint32_t*>(a.sig.p[a.rank]) + kSigBlocks;
    last = (atomicAdd(cnt, 1u) == gridDim.x - 1);
    if (last) *cnt = 0;
  }
  __syncthreads();
  if (last) rank_barrier(a.sig, a.rank, R, kSigOut, epoch, true);
}

// ---------------------------------------------------------------- standalone all-gather (push own shard to everybody)
struct GatherArgs {
  PeerPtrs bufs;
  PeerPtrs sig;
  long n_bytes;          // bytes in the buffer
  long shard_bytes;      // bytes per rank slice (multiple of 16)
  int rank, R;
  uint32_t epoch;
};

__global__ void __launch_bounds__(512) allgather_push_kernel(const GatherArgs a) {
  const long lo = (long)a.rank * a.shard_bytes, hi = min(a.n_bytes, lo + a.shard_bytes);
  const long nvec = hi > lo ? (hi - lo) / 16 : 0;
  const char* src = reinterpret_cast<const char*>(a.bufs.p[a.rank]) + lo;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < nvec; i += (long)gridDim.x * blockDim.x) {
    uint4 v = *reinterpret_cast<const uint4*>(src + i * 16);
    for (int q = 0; q < a.R; ++q)
      if (q != a.rank) st_peer_v4(reinterpret_cast<char*>(a.bufs.p[q]) + lo + i * 16, v);
  }
  if (a.epoch) {
    __shared__ bool last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t* cnt = reinterpret_cast<uint32_t*>(a.sig.p[a.rank]) + kSigBlocks;
      last = (atomicAdd(cnt, 1u) == gridDim.x - 1);
      if (last) *cnt = 0;
    }
    __syncthreads();
    if (last) rank_barrier(a.sig, a.rank, a.R, kSigOut, a.epoch, true);
  }
}

// One closing barrier per STEP: launched behind the last bucket kernel of the step on the same stream.  When it completes on
// rank r, every rank has finished ALL its bucket kernels of the step (stream order + the release below), i.e. every peer's
// refreshed slices have landed in r's weight buffers and nobody reads r's gradient buffers any more.  Per-bucket opening
// barriers stay (a rank's slice may only be pulled once every rank's gradients of that bucket are final); the per-bucket
// closing barriers they replace were 6 of the 14 cross-rank barriers of a GPT-2 step at 8 GPUs.
__global__ void step_close_kernel(PeerPtrs sig, int rank, int R, uint32_t epoch, const uint32_t* epoch_base) {
  if (epoch_base) epoch += *epoch_base;
  __threadfence_system();
  rank_barrier(sig, rank, R, kSigOut, epoch, true);
}

__global__ void set_barrier_timeout_kernel(unsigned long long cycles) { g_barrier_timeout_cycles = cycles; }

// Cross-rank barrier on its own (bootstrap, tests, end of async runs).
__global__ void barrier_kernel(PeerPtrs sig, int rank, int R, uint32_t epoch) { rank_barrier(sig, rank, R, kSigIn, epoch, true); }

// ---------------------------------------------------------------- async: push / apply / pull
// Per-rank async control block (uint32 words at kAsyncBase in the signal pad of the OWNER rank):
//   [0] version   : number of gradients applied to this shard so far (the device-side staleness counter)
//   [1] active    : index (0/1) of the accumulation buffer currently open for pushes
//   [2],[3]       : in-flight pusher count per buffer
//   [4],[5]       : arrived (complete) pushes per buffer
//   [6] applied   : total gradients applied           [7] dropped_here : pushes refused by the staleness guard
constexpr int kAsyncBase = 32;
enum { A_VERSION = 0, A_ACTIVE = 1, A_INFLIGHT = 2, A_ARRIVED = 4, A_APPLIED = 6, A_DROPPED = 7, A_MAXSEEN = 8 };

struct AsyncPushArgs {
  const void* grad;      // this worker's full gradient, TG, n elements (local)
  PeerPtrs acc0, acc1;   // owner accumulation buffers (fp32, shard elements each) on every rank
  PeerPtrs sig;
  const uint32_t* seen_version;   // [R] shard versions this worker's weights were pulled at (local device memory)
  uint32_t* push_state;  // [R] scratch: chosen buffer per owner, or 0xffffffff if refused (local device memory)
  long n, shard;
  int rank, R;
  int max_staleness;
  float scale;           // multiplies the gradient (1.0, or 1/(1+staleness) is applied per owner when policy = scale)
  int policy_scale;
  const uint32_t* ticket;  // optional work-queue ticket of this step: 0xffffffff (dataset exhausted) turns the push into a no-op
};

// Phase 1 (one thread): read every owner's version, apply the staleness guard ONCE for the whole gradient (all owners or
// none, so the shards keep counting the same set of gradients), then register as in-flight pusher on every owner.
// push_state[0..R) = chosen buffer per owner or 0xffffffff, [8..8+R) = staleness, [15] = 1 if the gradient was accepted.
__global__ void async_push_begin_kernel(const AsyncPushArgs a) {
  if (threadIdx.x != 0) return;
  a.push_state[15] = 0;
  if (a.ticket != nullptr && a.ticket[0] == 0xffffffffu) {      // no batch was dispatched for this step: nothing to push
    for (int o = 0; o < a.R; ++o) a.push_state[o] = 0xffffffffu;
    return;
  }
  int worst = 0;
  for (int o = 0; o < a.R; ++o) {
    const uint32_t* ctl = reinterpret_cast<const uint32_t*>(a.sig.p[o]) + kAsyncBase;
    const int st = (int)(ld_acquire_sys(ctl + A_VERSION) - a.seen_version[o]);
    a.push_state[kMaxRanks + o] = (uint32_t)(st < 0 ? 0 : st);
    worst = max(worst, st);
  }
  if (worst > a.max_staleness) {
    for (int o = 0; o < a.R; ++o) {
      atomicAdd_system(reinterpret_cast<uint32_t*>(a.sig.p[o]) + kAsyncBase + A_DROPPED, 1u);
      a.push_state[o] = 0xffffffffu;
    }
    return;
  }
  for (int o = 0; o < a.R; ++o) {
    uint32_t* ctl = reinterpret_cast<uint32_t*>(a.sig.p[o]) + kAsyncBase;
    uint32_t b;
    while (true) {                                        // register on the buffer that is open *now*
      b = ld_acquire_sys(ctl + A_ACTIVE) & 1u;
      atomicAdd_system(ctl + A_INFLIGHT + b, 1u);
      __threadfence_system();
      if ((ld_acquire_sys(ctl + A_ACTIVE) & 1u) == b) break;
      atomicSub_system(ctl + A_INFLIGHT + b, 1u);         // the owner flipped in between: retry on the new one
    }
    a.push_state[o] = b;
    atomicMax_system(ctl + A_MAXSEEN, (uint32_t)worst);
  }
  a.push_state[15] = 1;
}

// Phase 2: add every slice of the gradient into its owner's open accumulation buffer with P2P reductions.
template <typename TG>
__global__ void __launch_bounds__(512) async_push_data_kernel(const AsyncPushArgs a) {
  using VG = Vec16<TG>;
  const long nv = (a.n + VG::kN - 1) / VG::kN;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
    const long e = i * VG::kN;
    const int o = (int)(e / a.shard);
    const uint32_t b = a.push_state[o];
    if (b == 0xffffffffu) continue;
    float sc = a.scale;
    if (a.policy_scale) sc *= 1.f / (1.f + (float)a.push_state[kMaxRanks + o]);
    VG g;
    g.raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const TG*>(a.grad) + e);
    float* dst = reinterpret_cast<float*>(b ? a.acc1.p[o] : a.acc0.p[o]) + (e - (long)o * a.shard);
    if constexpr (VG::kN == 8) {
      red_add_v4_f32_sys(dst, g.get(0) * sc, g.get(1) * sc, g.get(2) * sc, g.get(3) * sc);
      red_add_v4_f32_sys(dst + 4, g.get(4) * sc, g.get(5) * sc, g.get(6) * sc, g.get(7) * sc);
    } else {
      red_add_v4_f32_sys(dst, g.get(0) * sc, g.get(1) * sc, g.get(2) * sc, g.get(3) * sc);
    }
  }
}

// Phase 3 (one block): publish completion — arrived++ then inflight-- on every owner we pushed to.
__global__ void async_push_end_kernel(const AsyncPushArgs a) {
  const int o = threadIdx.x;
  if (o >= a.R) return;
  const uint32_t b = a.push_state[o];
  if (b == 0xffffffffu) return;
  uint32_t* ctl = reinterpret_cast<uint32_t*>(a.sig.p[o]) + kAsyncBase;
  __threadfence_system();
  atomicAdd_system(ctl + A_ARRIVED + b, 1u);
  __threadfence_system();
  atomicSub_system(ctl + A_INFLIGHT + b, 1u);
}

struct AsyncApplyArgs {
  float* acc0;           // local accumulation buffers (fp32, shard elements)
  float* acc1;
  uint32_t* ctl;         // local control block (signal pad + kAsyncBase)
  float* master;         // fp32 master shard
  float* m;
  float* v;
  void* param_shard;     // this rank's slice inside its own working-weight buffer (TP)
  uint32_t* scratch;     // [2] device scratch: {buffer to drain, count}
  long shard_elems;
  const int* step_ptr;   // nullable: optimizer step count in device memory (Adam bias correction under CUDA-graph replay)
  int version_per_apply; // 1: the shard version counts optimizer UPDATES (+1 per apply round), 0: gradients applied (+count)
  OptParams opt;
};

// Phase 1 (one block, one thread): flip the open buffer, wait until nobody is still adding into the closed one.
__global__ void async_apply_begin_kernel(const AsyncApplyArgs a) {
  if (threadIdx.x != 0) return;
  const uint32_t old = ld_acquire_sys(a.ctl + A_ACTIVE) & 1u;
  if (ld_acquire_sys(a.ctl + A_ARRIVED + old) == 0 && ld_acquire_sys(a.ctl + A_INFLIGHT + old) == 0) {
    a.scratch[0] = old;
    a.scratch[1] = 0;                                   // nothing to apply
    return;
  }
  st_release_sys(a.ctl + A_ACTIVE, old ^ 1u);
  __threadfence_system();
  while (ld_acquire_sys(a.ctl + A_INFLIGHT + old) != 0) {
  }                                                     // pushers are running kernels that never wait: bounded
  a.scratch[0] = old;
  a.scratch[1] = ld_acquire_sys(a.ctl + A_ARRIVED + old);
}

// Phase 2: apply the accumulated gradient sum in place, refresh the working-weight slice, clear the buffer.
template <typename TP>
__global__ void __launch_bounds__(512) async_apply_data_kernel(const AsyncApplyArgs a) {
  const uint32_t cnt = a.scratch[1];
  if (cnt == 0) return;
  float* acc = a.scratch[0] ? a.acc1 : a.acc0;
  OptParams opt = a.opt;
  if (a.step_ptr != nullptr && opt.kind == OPT_ADAM) {
    const float t = (float)max(*a.step_ptr, 1);
    opt.bias_c1 = 1.f - powf(opt.beta1, t);
    opt.bias_c2 = 1.f - powf(opt.beta2, t);
  }
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < a.shard_elems; i += (long)gridDim.x * blockDim.x) {
    float g = acc[i];
    acc[i] = 0.f;
    float mm = a.m ? a.m[i] : 0.f, vv = a.v ? a.v[i] : 0.f;
    float w = opt_step(a.master[i], g, &mm, &vv, opt);
    a.master[i] = w;
    if (a.m && opt.kind != OPT_SGD) a.m[i] = mm;
    if (a.v && opt.kind == OPT_ADAM) a.v[i] = vv;
    reinterpret_cast<TP*>(a.param_shard)[i] = from_f32<TP>(w);
  }
}

// Phase 3 (one thread): version += count, arrived = 0.
__global__ void async_apply_end_kernel(const AsyncApplyArgs a) {
  if (threadIdx.x != 0) return;
  const uint32_t cnt = a.scratch[1];
  if (cnt == 0) return;
  const uint32_t b = a.scratch[0];
  atomicExch_system(a.ctl + A_ARRIVED + b, 0u);
  atomicAdd_system(a.ctl + A_APPLIED, cnt);
  __threadfence_system();
  // publish last: a reader that sees the version sees the weights.  All gradients that met in the buffer were applied as ONE
  // optimizer step on their sum, i.e. one model update: `version_per_apply` counts it as one version, so `maximumStaleness`
  // bounds the number of UPDATES a gradient may lag behind (with +count, 8 lock-step workers are 7 "versions" stale each).
  atomicAdd_system(a.ctl + A_VERSION, a.version_per_apply ? 1u : cnt);
}

struct AsyncPullArgs {
  PeerPtrs params;       // working-weight buffers on every rank (each rank's OWN slice is authoritative)
  PeerPtrs sig;
  void* dst;             // local working-weight buffer (full model)
  uint32_t* seen_version;  // [R] out
  long n_bytes, shard_bytes;
  int rank, R;
};

__global__ void __launch_bounds__(512) async_pull_kernel(const AsyncPullArgs a) {
  if (blockIdx.x == 0 && threadIdx.x < a.R)             // version first: staleness is then never under-estimated
    a.seen_version[threadIdx.x] = ld_acquire_sys(reinterpret_cast<const uint32_t*>(a.sig.p[threadIdx.x]) + kAsyncBase + A_VERSION);
  __syncthreads();
  const long nv = a.n_bytes / 16;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < nv; i += (long)gridDim.x * blockDim.x) {
    const long off = i * 16;
    const int o = (int)(off / a.shard_bytes);
    if (o == a.rank) continue;                          // own slice is already in place
    uint4 v = ld_peer_v4(reinterpret_cast<const char*>(a.params.p[o]) + off);
    *reinterpret_cast<uint4*>(reinterpret_cast<char*>(a.dst) + off) = v;
  }
}

// ---------------------------------------------------------------- device work queue (dataset dispatcher)
// Words at kWqBase in the signal pad of the queue's home rank:
//   [0] ticket (next fresh batch of this epoch)   [1] completed count   [2] epoch   [3] batches   [4] epochs
//   [5] done flag   [6] reissue cursor            [16 ...] completion bitmap (1 bit per batch)
constexpr int kWqBase = 64;
enum { W_TICKET = 0, W_COMPLETED = 1, W_EPOCH = 2, W_BATCHES = 3, W_EPOCHS = 4, W_DONE = 5, W_CURSOR = 6, W_LOCK = 7, W_BITMAP = 16 };

__global__ void wq_init_kernel(uint32_t* wq, uint32_t batches, uint32_t epochs, int bitmap_words) {
  for (int i = threadIdx.x; i < W_BITMAP + bitmap_words; i += blockDim.x) wq[i] = 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    wq[W_BATCHES] = batches;
    wq[W_EPOCHS] = epochs;
  }
}

// The epoch word carries a "rolling" marker (top bit) while the completer that finished an epoch resets the queue, and all
// completions are serialised by a spin lock (W_LOCK): check-epoch / set-bit / count / roll is one critical section, so a
// duplicate completion of a re-issued batch can never land in the freshly cleared bitmap of the next epoch (which would
// leave that batch "done" before it ran and the epoch counter one short for ever).  Completions are rare (one per batch) and
// the section is a few dozen instructions of one thread; holders are always running kernels, so the lock cannot be held by
// a kernel that is waiting to be scheduled.
constexpr uint32_t kWqRolling = 0x80000000u;
constexpr uint32_t kWqSpinLimit = 1u << 22;            // bounded waits: a wedged queue reports "exhausted" instead of hanging the GPU

// One thread takes a ticket.  out[0] = batch index or 0xffffffff when the dataset is exhausted, out[1] = epoch.
// FCFS in index order; when the fresh tickets of the epoch are gone, still-incomplete batches are re-issued
// (round-robin scan of the bitmap) exactly like DistributedDataset.next() (dataset.ts:57-60).
__global__ void wq_take_kernel(uint32_t* wq, uint32_t* out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  for (uint32_t spins = 0; spins < kWqSpinLimit; ++spins) {
    if (ld_acquire_sys(wq + W_DONE)) { out[0] = 0xffffffffu; out[1] = ld_acquire_sys(wq + W_EPOCH) & ~kWqRolling; return; }
    const uint32_t batches = wq[W_BATCHES];
    const uint32_t epoch = ld_acquire_sys(wq + W_EPOCH);
    if (epoch & kWqRolling) { __nanosleep(200); continue; }   // the queue is being reset for the next epoch
    const uint32_t t = atomicAdd_system(wq + W_TICKET, 1u);
    if (t < batches) {
      if (ld_acquire_sys(wq + W_EPOCH) == epoch) { out[0] = t; out[1] = epoch; return; }
      continue;                                       // epoch rolled over under us: take again (the lost ticket is re-issued)
    }
    // fresh ones ran out: re-issue something still incomplete
    const uint32_t start = atomicAdd_system(wq + W_CURSOR, 1u);
    for (uint32_t k = 0; k < batches; ++k) {
      const uint32_t b = (start + k) % batches;
      const uint32_t word = ld_acquire_sys(wq + W_BITMAP + (b >> 5));
      if (!((word >> (b & 31)) & 1u)) {
        if (ld_acquire_sys(wq + W_EPOCH) != epoch) break;     // includes the rolling marker: the zero we read is the reset
        out[0] = b; out[1] = epoch; return;
      }
    }
    __nanosleep(500);                                 // everything complete (or epoch rolled): look again
  }
  out[0] = 0xffffffffu;                               // gave up: see kWqSpinLimit
  out[1] = ld_acquire_sys(wq + W_EPOCH) & ~kWqRolling;
}

// Mark `batch` of `epoch` complete (idempotent).  The completion that finishes the epoch advances it.
__global__ void wq_complete_kernel(uint32_t* wq, const uint32_t* ticket, const uint32_t* accepted) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (accepted != nullptr && *accepted == 0) return;    // the gradient of this batch was refused: leave it incomplete -> re-issued
  const uint32_t b = ticket[0], epoch = ticket[1];
  if (b == 0xffffffffu) return;
  uint32_t spins = 0;
  while (atomicCAS_system(wq + W_LOCK, 0u, 1u) != 0u) {
    if (++spins > kWqSpinLimit) return;
    __nanosleep(100);
  }
  __threadfence_system();
  if (ld_acquire_sys(wq + W_EPOCH) == epoch) {          // else: stale completion from a previous epoch, ignored
    const uint32_t bit = 1u << (b & 31);
    const uint32_t old = atomicOr_system(wq + W_BITMAP + (b >> 5), bit);
    if (!(old & bit)) {                                 // else: duplicate
      const uint32_t batches = wq[W_BATCHES];
      const uint32_t done = atomicAdd_system(wq + W_COMPLETED, 1u) + 1;
      if (done == batches) {                            // last one of the epoch: roll over
        const int words = (batches + 31) / 32;
        if (epoch + 1 >= wq[W_EPOCHS]) {
          st_release_sys(wq + W_DONE, 1u);
        } else {
          st_release_sys(wq + W_EPOCH, epoch | kWqRolling);     // publish "rolling" BEFORE anything is cleared
          __threadfence_system();
          for (int i = 0; i < words; ++i) atomicExch_system(wq + W_BITMAP + i, 0u);
          atomicExch_system(wq + W_COMPLETED, 0u);
          atomicExch_system(wq + W_CURSOR, 0u);
          atomicExch_system(wq + W_TICKET, 0u);
          __threadfence_system();
          st_release_sys(wq + W_EPOCH, epoch + 1);
        }
      }
    }
  }
  __threadfence_system();
  atomicExch_system(wq + W_LOCK, 0u);
}

// Gather the rows of batch `ticket[0]` from a device-resident dataset: dst[r, :] = src[(b*bs + r) % rows, :].
template <typename U>
__global__ void __launch_bounds__(256) gather_batch_kernel(const char* __restrict__ src, char* __restrict__ dst,
                                                           const uint32_t* __restrict__ ticket, long row_bytes, int batch_size,
                                                           long rows) {
  const uint32_t b = ticket[0];
  if (b == 0xffffffffu) return;
  const long vec_per_row = row_bytes / (long)sizeof(U);
  const long total = vec_per_row * batch_size;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += (long)gridDim.x * 256) {
    const long r = i / vec_per_row, c = i - r * vec_per_row;
    const long sr = ((long)b * batch_size + r) % rows;
    reinterpret_cast<U*>(dst + r * row_bytes)[c] = reinterpret_cast<const U*>(src + sr * row_bytes)[c];
  }
}

}  // namespace dfb

using namespace dfb;

// ============================================================================================ C ABI
DFB_API int dfb_symm_alloc(size_t bytes, void** base_out) {
  void* p = nullptr;
  DFB_CUDA_CHECK(cudaMalloc(&p, bytes));
  DFB_CUDA_CHECK(cudaMemset(p, 0, bytes));
  DFB_CUDA_CHECK(cudaDeviceSynchronize());
  *base_out = p;
  return 0;
}
DFB_API int dfb_symm_free(void* base) {
  DFB_CUDA_CHECK(cudaFree(base));
  return 0;
}
DFB_API int dfb_symm_get_handle(void* base, void* handle_out64) {
  cudaIpcMemHandle_t h;
  DFB_CUDA_CHECK(cudaIpcGetMemHandle(&h, base));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t size");
  memcpy(handle_out64, &h, 64);
  return 0;
}
DFB_API int dfb_symm_open_handle(const void* handle64, void** peer_ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DFB_CUDA_CHECK(cudaIpcOpenMemHandle(peer_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
DFB_API int dfb_symm_close_handle(void* peer_ptr) {
  DFB_CUDA_CHECK(cudaIpcCloseMemHandle(peer_ptr));
  return 0;
}
DFB_API int dfb_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  DFB_CUDA_CHECK(e);
  return 0;
}

static void fill_peers(PeerPtrs& d, void* const* src, int R) {
  for (int i = 0; i < kMaxRanks; ++i) d.p[i] = i < R ? src[i] : nullptr;
}

// Fused reduce-scatter + 1/R + optimizer + all-gather for one bucket.  grads/params/sig: arrays of R base pointers
// (peer-mapped).  master/m/v: local fp32 slices.  grad_dtype/param_dtype: DT_*.  epoch = 0 disables the cross-rank
// barriers (single-GPU emulation: launch once per rank, in any order, gradients must already be final).
static int sync_impl(void* const* grads, void* const* params, void* const* sig, float* master, float* m, float* v,
                     void* reduced_out, long n, long shard, int rank, int R, unsigned epoch, int push, int grad_dtype,
                     int param_dtype, int opt_kind, float lr, float momentum, float beta1, float beta2, float eps,
                     float weight_decay, int step, int ctas, long nopush_lo, long nopush_hi, const unsigned* epoch_base,
                     const int* step_ptr, const float* lr_ptr, cudaStream_t stream) {
  DFB_REQUIRE(R >= 1 && R <= kMaxRanks && rank >= 0 && rank < R, "sync: bad rank %d / world %d", rank, R);
  DFB_REQUIRE(shard > 0 && (shard % 8) == 0, "sync: shard must be a positive multiple of 8 elements");
  DFB_REQUIRE(shard * R >= n, "sync: shard * R must cover the bucket");
  DFB_REQUIRE(opt_kind >= 0 && opt_kind <= 3, "sync: optimizer kind");
  SyncArgs a;
  fill_peers(a.grads, grads, R);
  fill_peers(a.params, params, R);
  if (sig) fill_peers(a.sig, sig, R); else memset(&a.sig, 0, sizeof(a.sig));
  DFB_REQUIRE(epoch == 0 || sig != nullptr, "sync: barriers need signal pads");
  a.master = master; a.m = m; a.v = v; a.reduced_out = reduced_out; a.n = n; a.shard = shard; a.rank = rank; a.R = R;
  a.epoch = epoch; a.push = (push & 1) | (g_close_per_bucket ? 0 : 2); a.nopush_lo = nopush_lo; a.nopush_hi = nopush_hi; a.epoch_base = epoch_base; a.step_ptr = step_ptr;
  a.lr_ptr = lr_ptr;
  a.opt.kind = opt_kind; a.opt.lr = lr; a.opt.momentum = momentum; a.opt.beta1 = beta1; a.opt.beta2 = beta2; a.opt.eps = eps;
  a.opt.weight_decay = weight_decay;
  a.opt.bias_c1 = 1.f - powf(beta1, (float)(step < 1 ? 1 : step));
  a.opt.bias_c2 = 1.f - powf(beta2, (float)(step < 1 ? 1 : step));
  DFB_REQUIRE(opt_kind == OPT_SGD || opt_kind == OPT_NONE || m != nullptr, "sync: optimizer state missing");
  DFB_REQUIRE(opt_kind != OPT_ADAM || v != nullptr, "sync: Adam second moment missing");
  const long lo = (long)rank * shard, hi = lo + shard < n ? lo + shard : n;
  long nvec = hi > lo ? (hi - lo + 7) / 8 : 0;
  int grid = ctas > 0 ? ctas : 64;
  long need = (nvec + 255) / 256;
  if (need < 1) need = 1;
  if (grid > need) grid = (int)need;
#define DFB_SYNC_LAUNCH(TG, TP)                                                     \
  do {                                                                               \
    if (R == 8) sync_rs_opt_ag_kernel<TG, TP, 8><<<grid, 256, 0, stream>>>(a);       \
    else if (R == 4) sync_rs_opt_ag_kernel<TG, TP, 4><<<grid, 256, 0, stream>>>(a);  \
    else if (R == 2) sync_rs_opt_ag_kernel<TG, TP, 2><<<grid, 256, 0, stream>>>(a);  \
    else sync_rs_opt_ag_kernel<TG, TP, 0><<<grid, 256, 0, stream>>>(a);              \
  } while (0)
  if (grad_dtype == DT_BF16 && param_dtype == DT_BF16) DFB_SYNC_LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (grad_dtype == DT_F32 && param_dtype == DT_F32) DFB_SYNC_LAUNCH(float, float);
  else if (grad_dtype == DT_F32 && param_dtype == DT_BF16) DFB_SYNC_LAUNCH(float, __nv_bfloat16);
  else if (grad_dtype == DT_BF16 && param_dtype == DT_F32) DFB_SYNC_LAUNCH(__nv_bfloat16, float);
  else {
    set_error("sync: unsupported dtypes grad=%d param=%d", grad_dtype, param_dtype);
    return 2;
  }
#undef DFB_SYNC_LAUNCH
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// 1 (default): every bucket kernel ends with its own cross-rank barrier.  0: bucket kernels skip it and the caller closes the
// step once with dfb_step_close (process-wide switch, set by the engine at construction).
DFB_API int dfb_sync_close_per_bucket(int on) {
  dfb::g_close_per_bucket = on ? 1 : 0;
  return 0;
}
DFB_API int dfb_step_close(void* const* sig, int rank, int R, unsigned epoch, const unsigned* epoch_base, cudaStream_t stream) {
  DFB_REQUIRE(R >= 1 && R <= kMaxRanks && epoch != 0, "step_close: bad arguments");
  PeerPtrs s;
  fill_peers(s, sig, R);
  step_close_kernel<<<1, 32, 0, stream>>>(s, rank, R, epoch, epoch_base);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_sync_rs_opt_ag(void* const* grads, void* const* params, void* const* sig, float* master, float* m, float* v,
                               void* reduced_out, long n, long shard, int rank, int R, unsigned epoch, int push, int grad_dtype,
                               int param_dtype, int opt_kind, float lr, float momentum, float beta1, float beta2, float eps,
                               float weight_decay, int step, int ctas, long nopush_lo, long nopush_hi, cudaStream_t stream) {
  return sync_impl(grads, params, sig, master, m, v, reduced_out, n, shard, rank, R, epoch, push, grad_dtype, param_dtype, opt_kind,
                   lr, momentum, beta1, beta2, eps, weight_decay, step, ctas, nopush_lo, nopush_hi, nullptr, nullptr, nullptr, stream);
}

// Same kernel for CUDA-graph replay: the barrier generation is *epoch_base + epoch, Adam's step count is *step_ptr and the
// learning rate is *lr_ptr (nullable: then `lr`), all read on the device at run time, so one captured launch stays valid for
// every later step and learning-rate schedules keep working.
DFB_API int dfb_sync_rs_opt_ag_replayable(void* const* grads, void* const* params, void* const* sig, float* master, float* m,
                                          float* v, long n, long shard, int rank, int R, unsigned epoch, int grad_dtype,
                                          int param_dtype, int opt_kind, float lr, float momentum, float beta1, float beta2,
                                          float eps, float weight_decay, int ctas, long nopush_lo, long nopush_hi,
                                          const unsigned* epoch_base, const int* step_ptr, const float* lr_ptr,
                                          cudaStream_t stream) {
  return sync_impl(grads, params, sig, master, m, v, nullptr, n, shard, rank, R, epoch, 1, grad_dtype, param_dtype, opt_kind, lr,
                   momentum, beta1, beta2, eps, weight_decay, 1, ctas, nopush_lo, nopush_hi, epoch_base, step_ptr, lr_ptr, stream);
}

// NVLS variant (bf16 gradients and weights; `mc_grads` / `mc_params`: multicast addresses of the bucket's buffers, from
// dfb_mc_bind_map; `params`: unicast bases for the ranges that are not pushed).  epoch_base / step_ptr may be null
// (eager launches) or device words (CUDA-graph replay), as in dfb_sync_rs_opt_ag_replayable.
DFB_API int dfb_sync_nvls(const void* mc_grads, void* mc_params, void* const* params, void* const* sig, float* master, float* m,
                          float* v, long n, long shard, int rank, int R, unsigned epoch, int opt_kind, float lr, float momentum,
                          float beta1, float beta2, float eps, float weight_decay, int step, int ctas, long nopush_lo,
                          long nopush_hi, const unsigned* epoch_base, const int* step_ptr, const float* lr_ptr,
                          cudaStream_t stream) {
  DFB_REQUIRE(R >= 2 && R <= kMaxRanks && rank >= 0 && rank < R, "nvls: bad rank %d / world %d", rank, R);
  DFB_REQUIRE(shard > 0 && (shard % 8) == 0 && shard * R >= n, "nvls: shard must be a multiple of 8 elements and cover the bucket");
  DFB_REQUIRE(mc_grads != nullptr && mc_params != nullptr && sig != nullptr && epoch != 0, "nvls: multicast addresses, signal pads and a barrier generation are required");
  DFB_REQUIRE(opt_kind >= 0 && opt_kind <= 2, "nvls: optimizer kind");
  DFB_REQUIRE(opt_kind == OPT_SGD || m != nullptr, "nvls: optimizer state missing");
  DFB_REQUIRE(opt_kind != OPT_ADAM || v != nullptr, "nvls: Adam second moment missing");
  NvlsArgs na;
  SyncArgs& a = na.s;
  memset(&a.grads, 0, sizeof(a.grads));
  fill_peers(a.params, params, R);
  fill_peers(a.sig, sig, R);
  a.master = master; a.m = m; a.v = v; a.reduced_out = nullptr; a.n = n; a.shard = shard; a.rank = rank; a.R = R;
  a.epoch = epoch; a.push = g_close_per_bucket ? 1 : 3; a.nopush_lo = nopush_lo; a.nopush_hi = nopush_hi; a.epoch_base = epoch_base; a.step_ptr = step_ptr;
  a.lr_ptr = lr_ptr;
  a.opt.kind = opt_kind; a.opt.lr = lr; a.opt.momentum = momentum; a.opt.beta1 = beta1; a.opt.beta2 = beta2; a.opt.eps = eps;
  a.opt.weight_decay = weight_decay;
  a.opt.bias_c1 = 1.f - powf(beta1, (float)(step < 1 ? 1 : step));
  a.opt.bias_c2 = 1.f - powf(beta2, (float)(step < 1 ? 1 : step));
  na.mc_grads = reinterpret_cast<const __nv_bfloat16*>(mc_grads);
  na.mc_params = reinterpret_cast<__nv_bfloat16*>(mc_params);
  const long lo = (long)rank * shard, hi = lo + shard < n ? lo + shard : n;
  long need = hi > lo ? ((hi - lo + 7) / 8 + 255) / 256 : 1;
  int grid = ctas > 0 ? ctas : 64;
  if (grid > need) grid = (int)need;
  sync_nvls_kernel<<<grid, 256, 0, stream>>>(na);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_allgather_push(void* const* bufs, void* const* sig, long n_bytes, long shard_bytes, int rank, int R,
                               unsigned epoch, int ctas, cudaStream_t stream) {
  DFB_REQUIRE(R >= 1 && R <= kMaxRanks && (shard_bytes % 16) == 0, "allgather: bad arguments");
  GatherArgs a;
  fill_peers(a.bufs, bufs, R);
  if (sig) fill_peers(a.sig, sig, R); else memset(&a.sig, 0, sizeof(a.sig));
  a.n_bytes = n_bytes; a.shard_bytes = shard_bytes; a.rank = rank; a.R = R; a.epoch = epoch;
  int grid = ctas > 0 ? ctas : 32;
  allgather_push_kernel<<<grid, 512, 0, stream>>>(a);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Bound of every in-kernel cross-rank wait on the current device, in microseconds of SM time at the device's nominal clock.
DFB_API int dfb_set_barrier_timeout_us(long long us, cudaStream_t stream) {
  int dev = 0, khz = 0;
  DFB_CUDA_CHECK(cudaGetDevice(&dev));
  DFB_CUDA_CHECK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  if (khz <= 0) khz = 1900000;
  const unsigned long long cycles = (unsigned long long)us * (unsigned long long)khz / 1000ull;
  DFB_REQUIRE(us > 0 && cycles > 0, "set_barrier_timeout: non-positive time-out");
  set_barrier_timeout_kernel<<<1, 1, 0, stream>>>(cycles);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
// out2[0] = 0 (healthy) or 1 + the rank a barrier of this device gave up on, out2[1] = the generation; clears the record.
DFB_API int dfb_barrier_failures(unsigned* pad, unsigned* host_out2, cudaStream_t stream) {
  DFB_CUDA_CHECK(cudaMemcpyAsync(host_out2, pad + kSigDead, 2 * sizeof(unsigned), cudaMemcpyDeviceToHost, stream));
  DFB_CUDA_CHECK(cudaStreamSynchronize(stream));
  DFB_CUDA_CHECK(cudaMemsetAsync(pad + kSigDead, 0, 2 * sizeof(unsigned), stream));
  return 0;
}

DFB_API int dfb_rank_barrier(void* const* sig, int rank, int R, unsigned epoch, cudaStream_t stream) {
  PeerPtrs s;
  fill_peers(s, sig, R);
  barrier_kernel<<<1, 32, 0, stream>>>(s, rank, R, epoch);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_async_push(const void* grad, void* const* acc0, void* const* acc1, void* const* sig,
                           const unsigned* seen_version, unsigned* push_state, long n, long shard, int rank, int R,
                           int max_staleness, float scale, int policy_scale, int grad_dtype, int ctas, const unsigned* ticket,
                           cudaStream_t stream) {
  DFB_REQUIRE(R >= 1 && R <= kMaxRanks && (shard % 8) == 0, "async_push: bad arguments");
  AsyncPushArgs a;
  a.grad = grad;
  fill_peers(a.acc0, acc0, R);
  fill_peers(a.acc1, acc1, R);
  fill_peers(a.sig, sig, R);
  a.seen_version = seen_version; a.push_state = push_state; a.n = n; a.shard = shard; a.rank = rank; a.R = R;
  a.max_staleness = max_staleness; a.scale = scale; a.policy_scale = policy_scale; a.ticket = ticket;
  async_push_begin_kernel<<<1, 32, 0, stream>>>(a);
  int grid = ctas > 0 ? ctas : 32;
  if (grad_dtype == DT_BF16) async_push_data_kernel<__nv_bfloat16><<<grid, 512, 0, stream>>>(a);
  else if (grad_dtype == DT_F32) async_push_data_kernel<float><<<grid, 512, 0, stream>>>(a);
  else {
    set_error("async_push: dtype");
    return 2;
  }
  async_push_end_kernel<<<1, 32, 0, stream>>>(a);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_async_apply(float* acc0, float* acc1, unsigned* ctl_pad, float* master, float* m, float* v, void* param_shard,
                            unsigned* scratch, long shard_elems, int param_dtype, int opt_kind, float lr, float momentum,
                            float beta1, float beta2, float eps, float weight_decay, int step, int ctas, const int* step_ptr,
                            int version_per_apply, cudaStream_t stream) {
  AsyncApplyArgs a;
  a.step_ptr = step_ptr; a.version_per_apply = version_per_apply;
  a.acc0 = acc0; a.acc1 = acc1; a.ctl = ctl_pad + kAsyncBase; a.master = master; a.m = m; a.v = v; a.param_shard = param_shard;
  a.scratch = scratch; a.shard_elems = shard_elems;
  a.opt.kind = opt_kind; a.opt.lr = lr; a.opt.momentum = momentum; a.opt.beta1 = beta1; a.opt.beta2 = beta2; a.opt.eps = eps;
  a.opt.weight_decay = weight_decay;
  a.opt.bias_c1 = 1.f - powf(beta1, (float)(step < 1 ? 1 : step));
  a.opt.bias_c2 = 1.f - powf(beta2, (float)(step < 1 ? 1 : step));
  async_apply_begin_kernel<<<1, 32, 0, stream>>>(a);
  int grid = ctas > 0 ? ctas : 32;
  if (param_dtype == DT_BF16) async_apply_data_kernel<__nv_bfloat16><<<grid, 512, 0, stream>>>(a);
  else if (param_dtype == DT_F32) async_apply_data_kernel<float><<<grid, 512, 0, stream>>>(a);
  else {
    set_error("async_apply: dtype");
    return 2;
  }
  async_apply_end_kernel<<<1, 32, 0, stream>>>(a);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

DFB_API int dfb_async_pull(void* const* params, void* const* sig, void* dst, unsigned* seen_version, long n_bytes,
                           long shard_bytes, int rank, int R, int ctas, cudaStream_t stream) {
  DFB_REQUIRE(R >= 1 && R <= kMaxRanks && (shard_bytes % 16) == 0 && (n_bytes % 16) == 0, "async_pull: bad arguments");
  AsyncPullArgs a;
  fill_peers(a.params, params, R);
  fill_peers(a.sig, sig, R);
  a.dst = dst; a.seen_version = seen_version; a.n_bytes = n_bytes; a.shard_bytes = shard_bytes; a.rank = rank; a.R = R;
  int grid = ctas > 0 ? ctas : 32;
  async_pull_kernel<<<grid, 512, 0, stream>>>(a);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Read the async counters of a control block into host-visible words: out[0..8] = version, active, inflight0/1,
// arrived0/1, applied, dropped, max staleness seen.  (cudaMemcpyAsync from the pad; the caller synchronises.)
DFB_API int dfb_async_read_ctl(const unsigned* ctl_pad, unsigned* host_out9, cudaStream_t stream) {
  DFB_CUDA_CHECK(cudaMemcpyAsync(host_out9, ctl_pad + kAsyncBase, 9 * sizeof(unsigned), cudaMemcpyDeviceToHost, stream));
  return 0;
}

// `wq` is the queue's control block in the home rank's heap (any rank's mapping of it): either the slot inside the signal
// pad (`pad + dfb_wq_pad_offset_words()`, room for `dfb_wq_pad_capacity()` batches) or, for larger datasets, a block of
// `dfb_wq_words(batches)` 32-bit words allocated from the heap.
DFB_API int dfb_wq_init(unsigned* wq, unsigned batches, unsigned epochs, cudaStream_t stream) {
  DFB_REQUIRE(batches >= 1 && epochs < kWqRolling, "wq_init: bad arguments");
  const int words = (int)((batches + 31) / 32);
  wq_init_kernel<<<1, 256, 0, stream>>>(wq, batches, epochs, words);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
DFB_API int dfb_wq_take(unsigned* wq, unsigned* ticket_out2, cudaStream_t stream) {
  wq_take_kernel<<<1, 32, 0, stream>>>(wq, ticket_out2);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
DFB_API int dfb_wq_complete(unsigned* wq, const unsigned* ticket2, const unsigned* accepted, cudaStream_t stream) {
  wq_complete_kernel<<<1, 32, 0, stream>>>(wq, ticket2, accepted);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
DFB_API int dfb_wq_words(unsigned batches) { return W_BITMAP + (int)((batches + 31) / 32); }
DFB_API int dfb_wq_pad_offset_words(void) { return kWqBase; }
DFB_API int dfb_wq_pad_capacity(void) { return (4096 / 4 - kWqBase - W_BITMAP) * 32; }
DFB_API int dfb_gather_batch(const void* src, void* dst, const unsigned* ticket, long row_bytes, int batch_size, long rows,
                             cudaStream_t stream) {
  DFB_REQUIRE((row_bytes % 4) == 0, "gather_batch: row size must be a multiple of 4 bytes");
  const bool wide = (row_bytes % 16) == 0 && ((uintptr_t)src % 16) == 0 && ((uintptr_t)dst % 16) == 0;
  long total = row_bytes / (wide ? 16 : 4) * batch_size;
  long blocks = (total + 255) / 256;
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  if (blocks < 1) blocks = 1;
  if (wide) gather_batch_kernel<uint4><<<(int)blocks, 256, 0, stream>>>((const char*)src, (char*)dst, ticket, row_bytes, batch_size, rows);
  else gather_batch_kernel<uint32_t><<<(int)blocks, 256, 0, stream>>>((const char*)src, (char*)dst, ticket, row_bytes, batch_size, rows);
  DFB_CUDA_CHECK(cudaGetLastError());
  return 0;
}
