// Fused gradient all-reduce / broadcast / barrier over NVLink 5 + NVSwitch, written for sm_100a.
//
// These kernels replace the NCCL ring all-reduce that the reference reaches through
// smdistributed.dataparallel (notebooks/code/cifar10-distributed-smddp-gpu.py:148,167; job log
// notebooks/2_pytorch_dist_smddp_gpu.ipynb:350-400 "24 coll channels ... via P2P/IPC/read") and the
// separate "divide by world size" pass of the DDP Reducer.  There is no NCCL call here: every rank
// maps every peer's arena (csrc/symm.cu) and the kernels issue the NVLink loads/stores themselves.
//
//   one-shot  (W1)  every rank reads all peers' copies, reduces in registers (fp32, fixed rank order
//                   -> bitwise identical on all ranks), scales, writes its own copy.  In place, with a
//                   mid barrier between "all reads done" and "first write".
//   two-shot  (W2)  rank r reduces shard r from all peers (reduce-scatter by pull) and PUSHES the scaled
//                   result into every peer's arena (all-gather by posted stores): 2 barriers total.
//   NVLS      (W3)  rank r issues multimem.ld_reduce on the multicast address of shard r (the switch
//                   sums all N copies), scales, multimem.st broadcasts the result to all N arenas.
//   broadcast (W4)  pull from the root's arena.       barrier (W5)  flag exchange on the signal pads.
//
//   fused     (W3+W13)  reduce-scatter -> SGD on the owned shard -> all-gather of the WEIGHTS in one kernel
//                   (allreduce_sgd_shard_kernel): the separate optimizer pass and the gradient broadcast disappear.
//
// Synchronisation: per-CTA flag barrier on the signal pads (first MiB of every arena).  Block b of rank
// a writes epoch e into slot [e&1][b][a] of every peer and spins on its own slots until all peers wrote e.
// The polls are relaxed system-scope loads (a fence per poll cost microseconds); the poll that finally sees the
// epoch is followed by one ld.acquire.sys of the same flag, which is what orders the CTA's later reads of peer data
// after the peer's flag store (the CTA barrier behind it extends that to every thread).  On the release side, where a
// CTA has pushed data into peer memory, all threads meet at a CTA barrier and only the flag-writing threads issue a
// system-scope fence before their flag store (cumulative over the CTA barrier) instead of one fence per thread.
// Epochs live in device memory, so the kernels are CUDA-graph replayable.  A spin that exceeds timeout_ns raises the
// host-mapped error flag and falls through instead of hanging (SURVEY §5.3 failure detection); the host polls that flag
// after every training step (runtime/engine.py) and before a checkpoint is written (workloads/cifar10_train.py).
#include "common.cuh"
#include <stdlib.h>

namespace {

constexpr int kThreads = 512;
// layout of the signal MiB:  [0, 2 * B200_MAX_COMM_BLOCKS * B200_MAX_RANKS * 4) flags (uint32) | [64 KiB, 64 KiB + blocks*4) epochs (local only)
constexpr size_t kEpochOff = 64 * 1024;

__device__ __forceinline__ uint32_t* sig_slot(char* base, uint32_t phase, int block, int src) {
  return reinterpret_cast<uint32_t*>(base) + ((phase * B200_MAX_COMM_BLOCKS + block) * B200_MAX_RANKS + src);
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Per-CTA barrier across all ranks.  All threads of the CTA must call it.  `epoch` is the CTA's barrier
// counter, loaded once per kernel (barrier_begin) and stored back once (barrier_end).
//
// RELEASE = true : this CTA wrote peer memory (pushed / multicast stores) that the peers consume after the
//                  barrier -> every thread fences at system scope before the flag goes out.
// RELEASE = false: the barrier only orders "everyone arrived" (start of a collective: the data was written
//                  by earlier kernels, complete at the kernel boundary) or "everyone finished reading".  Flags
//                  are relaxed system-scope stores/loads; system-scope MEMBARs cost microseconds each and are
//                  what dominated small-message latency.
__device__ __forceinline__ uint32_t barrier_begin(const b200_comm& c) {
  return reinterpret_cast<uint32_t*>(c.peer[c.rank] + kEpochOff)[blockIdx.x];
}
__device__ __forceinline__ void barrier_end(const b200_comm& c, uint32_t epoch) {
  if (threadIdx.x == 0) reinterpret_cast<uint32_t*>(c.peer[c.rank] + kEpochOff)[blockIdx.x] = epoch;
}
template <bool RELEASE>
__device__ __forceinline__ void cta_barrier_all_ranks(const b200_comm& c, uint32_t& epoch) {
  __syncthreads();
  const uint32_t e = ++epoch;
  if (threadIdx.x < c.world) {
    const int r = threadIdx.x;
    // release: every thread's peer stores happen-before this point through the CTA barrier above; one system-scope
    // fence by each flag writer (cumulative) publishes them before the flag
    if (RELEASE) __threadfence_system();
    st_relaxed_sys(sig_slot(c.peer[r], e & 1u, blockIdx.x, c.rank), e);
    const uint32_t* mine = sig_slot(c.peer[c.rank], e & 1u, blockIdx.x, r);
    if (ld_relaxed_sys(mine) != e) {
      const unsigned long long t0 = globaltimer_ns();
      unsigned spins = 0;
      while (ld_relaxed_sys(mine) != e) {
        if ((++spins & 0x3ff) == 0 && globaltimer_ns() - t0 > c.timeout_ns) {
          if (c.err_flag) {
            c.err_flag[0] = 1;
            c.err_flag[1] = r;            // which peer never arrived
            c.err_flag[2] = (int)blockIdx.x;
          }
          break;
        }
      }
    }
    (void)ld_acquire_sys(mine);     // acquire: reads of this peer's data issued after the barrier see what it published
  }
  __syncthreads();
}

// ---- 16-byte vector traits -----------------------------------------------------------------------
template <int DT>
struct Vec;
template <>
struct Vec<B200_DTYPE_F32> {
  static constexpr int kElems = 4;
  __device__ static void accumulate(float* acc, const int4& raw) {
    const float* f = reinterpret_cast<const float*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] += f[i];
  }
  __device__ static int4 pack(const float* acc, float scale) {
    int4 o;
    float* f = reinterpret_cast<float*>(&o);
#pragma unroll
    for (int i = 0; i < 4; ++i) f[i] = acc[i] * scale;
    return o;
  }
  __device__ static int4 mc_ld_reduce(const void* mc) {
    int4 o;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(o.x), "=r"(o.y), "=r"(o.z), "=r"(o.w)
                 : "l"(mc)
                 : "memory");
    return o;
  }
};
template <>
struct Vec<B200_DTYPE_BF16> {
  static constexpr int kElems = 8;
  __device__ static void accumulate(float* acc, const int4& raw) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      acc[2 * i] += f.x;
      acc[2 * i + 1] += f.y;
    }
  }
  __device__ static int4 pack(const float* acc, float scale) {
    int4 o;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(acc[2 * i] * scale, acc[2 * i + 1] * scale);
    return o;
  }
  __device__ static int4 mc_ld_reduce(const void* mc) {
    int4 o;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(o.x), "=r"(o.y), "=r"(o.z), "=r"(o.w)
                 : "l"(mc)
                 : "memory");
    return o;
  }
};

__device__ __forceinline__ int4 ld_peer(const void* p) {  // peer data is written by other GPUs: no .nc
  int4 v;
  asm volatile("ld.relaxed.sys.global.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(void* p, const int4& v) {
  asm volatile("st.relaxed.sys.global.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_st(void* mc, const int4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// ---- W1 one-shot ----------------------------------------------------------------------------------
// `out` is a local buffer (may be the local arena itself: in place).  Reads of a round complete on all
// ranks (mid barrier) before anyone overwrites, so in-place use is race free.
template <int DT, int WORLD, int UNROLL>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_oneshot_kernel(b200_comm c, size_t offset, char* out, size_t chunks, float scale, int need_mid_barrier) {
  using V = Vec<DT>;
  uint32_t epoch = barrier_begin(c);
  cta_barrier_all_ranks<false>(c, epoch);  // every peer entered: its gradients for this bucket are final
  const size_t per_round = (size_t)gridDim.x * kThreads * UNROLL;
  for (size_t base = (size_t)blockIdx.x * kThreads * UNROLL; base < chunks; base += per_round) {
    float acc[UNROLL][V::kElems];
    int4 raw[WORLD][UNROLL];
#pragma unroll
    for (int r = 0; r < WORLD; ++r) {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
        raw[r][u] = make_int4(0, 0, 0, 0);
        if (idx < chunks) raw[r][u] = ld_peer(c.peer[r] + offset + idx * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
      for (int i = 0; i < V::kElems; ++i) acc[u][i] = 0.f;
#pragma unroll
      for (int r = 0; r < WORLD; ++r) V::accumulate(acc[u], raw[r][u]);  // fixed order 0..N-1
    }
    if (need_mid_barrier) cta_barrier_all_ranks<false>(c, epoch);   // all ranks hold this round in registers
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
      if (idx < chunks) *reinterpret_cast<int4*>(out + idx * 16) = V::pack(acc[u], scale);
    }
  }
  if (!need_mid_barrier) cta_barrier_all_ranks<false>(c, epoch);  // out-of-place: protect the arena from early reuse
  barrier_end(c, epoch);
}

// ---- W2 two-shot (pull-reduce + push-broadcast) and W3 NVLS ---------------------------------------
template <int DT, int WORLD, int UNROLL, bool NVLS>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_shard_kernel(b200_comm c, size_t offset, size_t chunks, float scale) {
  using V = Vec<DT>;
  // equal 16-byte-chunk shards, independent of tensor boundaries ("balanced fusion buffer")
  const size_t per = (chunks + WORLD - 1) / WORLD;
  const size_t lo = (size_t)c.rank * per;
  const size_t hi = lo + per < chunks ? lo + per : chunks;
  uint32_t epoch = barrier_begin(c);
  cta_barrier_all_ranks<false>(c, epoch);
  const size_t per_round = (size_t)gridDim.x * kThreads * UNROLL;
  for (size_t base = lo + (size_t)blockIdx.x * kThreads * UNROLL; base < hi; base += per_round) {
    if constexpr (NVLS) {
      int4 red[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
        if (idx < hi) red[u] = V::mc_ld_reduce(c.mc + offset + idx * 16);
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
        if (idx < hi) {
          float acc[V::kElems];
#pragma unroll
          for (int i = 0; i < V::kElems; ++i) acc[i] = 0.f;
          V::accumulate(acc, red[u]);
          mc_st(c.mc + offset + idx * 16, V::pack(acc, scale));
        }
      }
    } else {
      int4 raw[WORLD][UNROLL];
#pragma unroll
      for (int r = 0; r < WORLD; ++r) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
          raw[r][u] = make_int4(0, 0, 0, 0);
          if (idx < hi) raw[r][u] = ld_peer(c.peer[r] + offset + idx * 16);
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
        if (idx < hi) {
          float acc[V::kElems];
#pragma unroll
          for (int i = 0; i < V::kElems; ++i) acc[i] = 0.f;
#pragma unroll
          for (int r = 0; r < WORLD; ++r) V::accumulate(acc, raw[r][u]);
          const int4 o = V::pack(acc, scale);
#pragma unroll
          for (int i = 0; i < WORLD; ++i) {
            const int p = (c.rank + i) % WORLD;  // stagger destinations across ranks
            st_peer(c.peer[p] + offset + idx * 16, o);
          }
        }
      }
    }
  }
  cta_barrier_all_ranks<true>(c, epoch);  // all pushed / multicast stores have landed everywhere
  barrier_end(c, epoch);
}

// ---- fused: reduce-scatter -> SGD on the owned shard -> all-gather of the WEIGHTS, one kernel -------------------
// The step the reference takes as `loss.backward()` (NCCL all-reduce of every bucket) followed by `optimizer.step()`
// (notebooks/code/cifar10-distributed-smddp-gpu.py:167-168), and the idea of SMDDP's "balanced fusion buffer" (a bucket
// cut into N equal shards regardless of tensor boundaries, each reduced by a different owner:
// static/images/training/training24.png), as ONE kernel per bucket over peer memory:
//   1. rank r reduces shard r of the bucket's gradients: multimem.ld_reduce on the multicast address (the switch adds
//      the N copies) or, without NVLS, 16-byte loads from every peer's arena summed in fixed rank order;
//   2. applies SGD with momentum to shard r of the fp32 master weights (momentum state exists only for the owned shard);
//   3. writes the NEW WEIGHTS (and, in bf16 mode, their bf16 shadow) into every rank's parameter arena: multimem.st, or
//      posted stores to each peer.  Every rank therefore holds bit-identical parameters by construction;
//   4. after the exit barrier (all reads of everyone's gradients are done) zeroes its local copy of the bucket's
//      gradients for the next backward.
// Compared with all-reduce + a separate optimizer pass this moves the same bytes over NVLink (weights instead of averaged
// gradients on the way back) and drops the optimizer's read of p, g, m and write of p, m, shadow for (N-1)/N of the model.
// fp32 gradients / masters only (the arena layout of DistributedDataParallel).
__device__ __forceinline__ void mc_st_b64(void* mc, const uint2& v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(mc), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void st_peer_b64(void* p, const uint2& v) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

struct SgdArgs {
  size_t param_off;     // byte offset (in the arena) of the bucket's fp32 master weights
  size_t shadow_off;    // byte offset of the bucket's bf16 shadow weights, 0 = none
  float* momentum;      // LOCAL fp32 momentum buffer of the bucket (only the owned shard is ever touched), may be null
  float lr, momentum_coef, one_minus_damp, wd;
  int nesterov;
};

template <int WORLD, int UNROLL, bool NVLS>
__global__ void __launch_bounds__(kThreads, 1)
allreduce_sgd_shard_kernel(b200_comm c, size_t offset, size_t chunks, float scale, SgdArgs a) {
  using V = Vec<B200_DTYPE_F32>;
  const size_t per = (chunks + WORLD - 1) / WORLD;
  const size_t lo = (size_t)c.rank * per;
  const size_t hi = lo + per < chunks ? lo + per : chunks;
  char* const local = c.peer[c.rank];
  uint32_t epoch = barrier_begin(c);
  cta_barrier_all_ranks<false>(c, epoch);          // every rank's gradients for this bucket are final
  const size_t per_round = (size_t)gridDim.x * kThreads * UNROLL;
  for (size_t base = lo + (size_t)blockIdx.x * kThreads * UNROLL; base < hi; base += per_round) {
    float g[UNROLL][4];
    if constexpr (NVLS) {
      int4 red[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
        if (idx < hi) red[u] = V::mc_ld_reduce(c.mc + offset + idx * 16);
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const float* f = reinterpret_cast<const float*>(&red[u]);
#pragma unroll
        for (int i = 0; i < 4; ++i) g[u][i] = f[i];
      }
    } else {
      int4 raw[WORLD][UNROLL];
#pragma unroll
      for (int r = 0; r < WORLD; ++r) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
          raw[r][u] = make_int4(0, 0, 0, 0);
          if (idx < hi) raw[r][u] = ld_peer(c.peer[r] + offset + idx * 16);
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
        for (int i = 0; i < 4; ++i) g[u][i] = 0.f;
#pragma unroll
        for (int r = 0; r < WORLD; ++r) V::accumulate(g[u], raw[r][u]);   // fixed order 0..N-1
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
      if (idx >= hi) continue;
      const float4 pv = *reinterpret_cast<const float4*>(local + a.param_off + idx * 16);
      float p[4] = {pv.x, pv.y, pv.z, pv.w};
      float m[4] = {0.f, 0.f, 0.f, 0.f};
      if (a.momentum_coef != 0.f) {
        const float4 mv = reinterpret_cast<const float4*>(a.momentum)[idx];
        m[0] = mv.x; m[1] = mv.y; m[2] = mv.z; m[3] = mv.w;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float d = fmaf(g[u][k], scale, a.wd * p[k]);
        if (a.momentum_coef != 0.f) {
          m[k] = fmaf(a.momentum_coef, m[k], a.one_minus_damp * d);
          d = a.nesterov ? fmaf(a.momentum_coef, m[k], d) : m[k];
        }
        p[k] = fmaf(-a.lr, d, p[k]);
      }
      if (a.momentum_coef != 0.f) reinterpret_cast<float4*>(a.momentum)[idx] = make_float4(m[0], m[1], m[2], m[3]);
      int4 o;
      float* of = reinterpret_cast<float*>(&o);
      of[0] = p[0]; of[1] = p[1]; of[2] = p[2]; of[3] = p[3];
      uint2 sh;
      if (a.shadow_off != 0) {
        __nv_bfloat162 l2 = __floats2bfloat162_rn(p[0], p[1]), h2 = __floats2bfloat162_rn(p[2], p[3]);
        sh.x = *reinterpret_cast<uint32_t*>(&l2);
        sh.y = *reinterpret_cast<uint32_t*>(&h2);
      }
      if constexpr (NVLS) {
        mc_st(c.mc + a.param_off + idx * 16, o);
        if (a.shadow_off != 0) mc_st_b64(c.mc + a.shadow_off + idx * 8, sh);
      } else {
#pragma unroll
        for (int i = 0; i < WORLD; ++i) {
          const int q = (c.rank + i) % WORLD;       // stagger destinations across ranks
          st_peer(c.peer[q] + a.param_off + idx * 16, o);
          if (a.shadow_off != 0) st_peer_b64(c.peer[q] + a.shadow_off + idx * 8, sh);
        }
      }
    }
  }
  cta_barrier_all_ranks<true>(c, epoch);           // all new weights have landed everywhere; nobody reads gradients any more
  barrier_end(c, epoch);
  // Zero the local copy of the bucket's gradients for the next backward.  The barrier above only pairs CTA b with CTA b of
  // the other ranks, so this CTA may clear exactly the chunks that "CTA b" of each owner has read: the same index set as the
  // loop above, taken for every shard.
  const int4 z = make_int4(0, 0, 0, 0);
#pragma unroll 1
  for (int r = 0; r < WORLD; ++r) {
    const size_t rlo = (size_t)r * per;
    const size_t rhi = rlo + per < chunks ? rlo + per : chunks;
    for (size_t base = rlo + (size_t)blockIdx.x * kThreads * UNROLL; base < rhi; base += per_round) {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const size_t idx = base + (size_t)u * kThreads + threadIdx.x;
        if (idx < rhi) *reinterpret_cast<int4*>(local + offset + idx * 16) = z;
      }
    }
  }
}

// ---- W4 broadcast (pull) --------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1) broadcast_kernel(b200_comm c, size_t offset, size_t chunks, int root) {
  uint32_t epoch = barrier_begin(c);
  cta_barrier_all_ranks<false>(c, epoch);  // root's data is final, receivers' buffers are free
  if (c.rank != root) {
    const char* src = c.peer[root] + offset;
    char* dst = c.peer[c.rank] + offset;
    const size_t stride = (size_t)gridDim.x * kThreads;
    for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < chunks; i += stride)
      *reinterpret_cast<int4*>(dst + i * 16) = ld_peer(src + i * 16);
  }
  cta_barrier_all_ranks<false>(c, epoch);  // root may overwrite again
  barrier_end(c, epoch);
}

__global__ void __launch_bounds__(kThreads, 1) barrier_kernel(b200_comm c) {
  uint32_t epoch = barrier_begin(c);
  cta_barrier_all_ranks<true>(c, epoch);
  barrier_end(c, epoch);
}

template <typename T>
__global__ void scale_kernel(T* p, size_t n, float scale);
template <>
__global__ void scale_kernel<float>(float* p, size_t n, float scale) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] *= scale;
}
template <>
__global__ void scale_kernel<__nv_bfloat16>(__nv_bfloat16* p, size_t n, float scale) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = __float2bfloat16(__bfloat162float(p[i]) * scale);
}

int clamp_blocks(size_t work_chunks, int unroll, int max_blocks) {
  size_t need = (work_chunks + (size_t)kThreads * unroll - 1) / ((size_t)kThreads * unroll);
  if (max_blocks <= 0 || max_blocks > B200_MAX_COMM_BLOCKS) max_blocks = B200_MAX_COMM_BLOCKS;
  if (need < 1) need = 1;
  if (need > (size_t)max_blocks) need = max_blocks;
  return (int)need;
}

template <int DT, int WORLD>
int launch_world(const b200_comm& c, size_t offset, size_t chunks, float scale, int algo, int max_blocks,
                 cudaStream_t s) {
  if (algo == B200_ALGO_ONESHOT) {
    constexpr int U = (WORLD <= 4) ? 4 : 2;
    int blocks = clamp_blocks(chunks, U, max_blocks);
    allreduce_oneshot_kernel<DT, WORLD, U><<<blocks, kThreads, 0, s>>>(c, offset, c.peer[c.rank] + offset, chunks, scale, 1);
  } else if (algo == B200_ALGO_NVLS) {
    // 16-byte multimem.ld_reduce in flight per thread (the switch round trip is ~3 us): 8 by default, B200DDP_NVLS_UNROLL=4|8|16
    static const int unroll = []() { const char* e = getenv("B200DDP_NVLS_UNROLL"); const int v = e ? atoi(e) : 8; return v; }();
    if (unroll >= 16) {
      int blocks = clamp_blocks((chunks + WORLD - 1) / WORLD, 16, max_blocks);
      allreduce_shard_kernel<DT, WORLD, 16, true><<<blocks, kThreads, 0, s>>>(c, offset, chunks, scale);
    } else if (unroll <= 4) {
      int blocks = clamp_blocks((chunks + WORLD - 1) / WORLD, 4, max_blocks);
      allreduce_shard_kernel<DT, WORLD, 4, true><<<blocks, kThreads, 0, s>>>(c, offset, chunks, scale);
    } else {
      int blocks = clamp_blocks((chunks + WORLD - 1) / WORLD, 8, max_blocks);
      allreduce_shard_kernel<DT, WORLD, 8, true><<<blocks, kThreads, 0, s>>>(c, offset, chunks, scale);
    }
  } else {
    constexpr int U = (WORLD <= 4) ? 4 : 2;
    int blocks = clamp_blocks((chunks + WORLD - 1) / WORLD, U, max_blocks);
    allreduce_shard_kernel<DT, WORLD, U, false><<<blocks, kThreads, 0, s>>>(c, offset, chunks, scale);
  }
  B200_CHECK_LAUNCH();
  return 0;
}

template <int DT>
int launch_dtype(const b200_comm& c, size_t offset, size_t chunks, float scale, int algo, int max_blocks,
                 cudaStream_t s) {
  switch (c.world) {
    case 2: return launch_world<DT, 2>(c, offset, chunks, scale, algo, max_blocks, s);
    case 3: return launch_world<DT, 3>(c, offset, chunks, scale, algo, max_blocks, s);
    case 4: return launch_world<DT, 4>(c, offset, chunks, scale, algo, max_blocks, s);
    case 5: return launch_world<DT, 5>(c, offset, chunks, scale, algo, max_blocks, s);
    case 6: return launch_world<DT, 6>(c, offset, chunks, scale, algo, max_blocks, s);
    case 7: return launch_world<DT, 7>(c, offset, chunks, scale, algo, max_blocks, s);
    case 8: return launch_world<DT, 8>(c, offset, chunks, scale, algo, max_blocks, s);
    default: b200_set_error("world size %d not in 2..8", c.world); return B200_ERR_ARG;
  }
}

template <int DT, int WORLD>
int launch_oneshot_out(const b200_comm& c, size_t offset, char* out, size_t chunks, float scale, int max_blocks,
                       cudaStream_t s) {
  constexpr int U = (WORLD <= 4) ? 4 : 2;
  int blocks = clamp_blocks(chunks, U, max_blocks);
  const int in_place = (out == c.peer[c.rank] + offset) ? 1 : 0;
  allreduce_oneshot_kernel<DT, WORLD, U><<<blocks, kThreads, 0, s>>>(c, offset, out, chunks, scale, in_place);
  B200_CHECK_LAUNCH();
  return 0;
}

}  // namespace

static size_t elem_size(int dtype) { return dtype == B200_DTYPE_BF16 ? 2 : 4; }

// THE algorithm selector (the reducer, b200_allreduce's AUTO and the Python layer all come here): one-shot up to
// oneshot_max_bytes (0 = default max(64 KiB, 2 MiB / (world - 1)): one-shot pulls (N-1) x S per GPU, so its cut-over shrinks
// with N), then NVLS when a multicast mapping is bound, the world is larger than 2 (with two ranks the switch reduction buys
// nothing) and the message has at least nvls_min_bytes (0 = no lower limit), else two-shot.
extern "C" int b200_pick_algo(int world, int has_multicast, size_t bytes, size_t oneshot_max_bytes, size_t nvls_min_bytes) {
  if (oneshot_max_bytes == 0) {
    const size_t scaled = world > 1 ? ((size_t)2 << 20) / (size_t)(world - 1) : ((size_t)2 << 20);
    oneshot_max_bytes = scaled > 64 * 1024 ? scaled : 64 * 1024;
  }
  if (bytes <= oneshot_max_bytes) return B200_ALGO_ONESHOT;
  if (has_multicast && world > 2 && bytes >= nvls_min_bytes) return B200_ALGO_NVLS;
  return B200_ALGO_TWOSHOT;
}

extern "C" int b200_allreduce(const b200_comm* comm, size_t offset, size_t count, int dtype, float scale, int algo,
                              int max_blocks, void* stream) {
  B200_REQUIRE(comm != nullptr, "null communicator");
  B200_REQUIRE(dtype == B200_DTYPE_F32 || dtype == B200_DTYPE_BF16, "dtype must be f32 or bf16");
  B200_REQUIRE(offset % 16 == 0, "offset must be 16-byte aligned");
  B200_REQUIRE((count * elem_size(dtype)) % 16 == 0, "byte count must be a multiple of 16 (pad the bucket)");
  B200_REQUIRE(offset >= B200_SIGNAL_BYTES, "offset overlaps the signal region");
  cudaStream_t s = (cudaStream_t)stream;
  if (count == 0) return 0;
  const b200_comm& c = *comm;
  if (c.world == 1) {
    if (scale != 1.0f) {
      unsigned g = b200_grid_for(count, 256, 8);
      if (dtype == B200_DTYPE_F32)
        scale_kernel<float><<<g, 256, 0, s>>>((float*)(c.peer[0] + offset), count, scale);
      else
        scale_kernel<__nv_bfloat16><<<g, 256, 0, s>>>((__nv_bfloat16*)(c.peer[0] + offset), count, scale);
      B200_CHECK_LAUNCH();
    }
    return 0;
  }
  const size_t bytes = count * elem_size(dtype);
  const size_t chunks = bytes / 16;
  if (algo == B200_ALGO_AUTO) algo = b200_pick_algo(c.world, c.mc != nullptr, bytes, 0, 0);
  if (algo == B200_ALGO_NVLS && c.mc == nullptr) {
    b200_set_error("NVLS requested but no multicast mapping is bound");
    return B200_ERR_UNSUPPORTED;
  }
  return dtype == B200_DTYPE_F32 ? launch_dtype<B200_DTYPE_F32>(c, offset, chunks, scale, algo, max_blocks, s)
                                 : launch_dtype<B200_DTYPE_BF16>(c, offset, chunks, scale, algo, max_blocks, s);
}

// Fused bucket step: gradient reduce-scatter + SGD(momentum) on the owned shard + all-gather of the new weights (and bf16
// shadow) + zeroing of the local gradients, one kernel.  grad_offset / param_offset / shadow_offset are byte offsets of the
// bucket inside the symmetric arena (shadow_offset 0 = no shadow); `momentum` is this rank's LOCAL fp32 buffer for the
// bucket (count elements; only the owned shard is used) or NULL when momentum == 0.  algo: B200_ALGO_NVLS or _TWOSHOT
// (AUTO: NVLS when a multicast mapping is bound and world > 2).
extern "C" int b200_allreduce_sgd(const b200_comm* comm, size_t grad_offset, size_t param_offset, size_t shadow_offset,
                                  float* momentum_buf, size_t count, float scale, float lr, float momentum, float dampening,
                                  float weight_decay, int nesterov, int algo, int max_blocks, void* stream) {
  B200_REQUIRE(comm != nullptr, "null communicator");
  B200_REQUIRE(comm->world >= 2 && comm->world <= B200_MAX_RANKS, "the fused bucket step needs 2..8 ranks");
  B200_REQUIRE(grad_offset % 16 == 0 && param_offset % 16 == 0 && shadow_offset % 8 == 0, "offsets must be 16-byte (shadow: 8-byte) aligned");
  B200_REQUIRE(count % 4 == 0, "element count must be a multiple of 4");
  B200_REQUIRE(grad_offset >= B200_SIGNAL_BYTES && param_offset >= B200_SIGNAL_BYTES, "offset overlaps the signal region");
  B200_REQUIRE(momentum == 0.f || momentum_buf != nullptr, "momentum buffer required");
  if (count == 0) return 0;
  const b200_comm& c = *comm;
  if (algo == B200_ALGO_AUTO) algo = (c.mc != nullptr && c.world > 2) ? B200_ALGO_NVLS : B200_ALGO_TWOSHOT;
  B200_REQUIRE(algo == B200_ALGO_NVLS || algo == B200_ALGO_TWOSHOT, "fused bucket step: algo must be NVLS or two-shot");
  if (algo == B200_ALGO_NVLS && c.mc == nullptr) {
    b200_set_error("NVLS requested but no multicast mapping is bound");
    return B200_ERR_UNSUPPORTED;
  }
  SgdArgs a;
  a.param_off = param_offset; a.shadow_off = shadow_offset; a.momentum = momentum_buf;
  a.lr = lr; a.momentum_coef = momentum; a.one_minus_damp = 1.f - dampening; a.wd = weight_decay; a.nesterov = nesterov;
  const size_t chunks = count / 4;
  cudaStream_t s = (cudaStream_t)stream;
#define B200_FS(W)                                                                                               \
  case W: {                                                                                                      \
    if (algo == B200_ALGO_NVLS) {                                                                                \
      int blocks = clamp_blocks((chunks + W - 1) / W, 4, max_blocks);                                            \
      allreduce_sgd_shard_kernel<W, 4, true><<<blocks, kThreads, 0, s>>>(c, grad_offset, chunks, scale, a);      \
    } else {                                                                                                     \
      constexpr int U = (W <= 4) ? 2 : 1;                                                                        \
      int blocks = clamp_blocks((chunks + W - 1) / W, U, max_blocks);                                            \
      allreduce_sgd_shard_kernel<W, U, false><<<blocks, kThreads, 0, s>>>(c, grad_offset, chunks, scale, a);     \
    }                                                                                                            \
    break;                                                                                                       \
  }
  switch (c.world) {
    B200_FS(2) B200_FS(3) B200_FS(4) B200_FS(5) B200_FS(6) B200_FS(7) B200_FS(8)
    default: b200_set_error("world size %d not in 2..8", c.world); return B200_ERR_ARG;
  }
#undef B200_FS
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_allreduce_oneshot_out(const b200_comm* comm, size_t offset, void* out, size_t count, int dtype,
                                          float scale, int max_blocks, void* stream) {
  B200_REQUIRE(comm != nullptr && out != nullptr, "null argument");
  B200_REQUIRE(offset % 16 == 0 && ((uintptr_t)out) % 16 == 0, "16-byte alignment required");
  B200_REQUIRE((count * elem_size(dtype)) % 16 == 0, "byte count must be a multiple of 16");
  const b200_comm& c = *comm;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t chunks = count * elem_size(dtype) / 16;
  if (c.world == 1) {
    B200_CHECK_CUDA(cudaMemcpyAsync(out, c.peer[0] + offset, chunks * 16, cudaMemcpyDeviceToDevice, s));
    return 0;
  }
#define B200_OS(DT)                                                                                         \
  switch (c.world) {                                                                                        \
    case 2: return launch_oneshot_out<DT, 2>(c, offset, (char*)out, chunks, scale, max_blocks, s);          \
    case 3: return launch_oneshot_out<DT, 3>(c, offset, (char*)out, chunks, scale, max_blocks, s);          \
    case 4: return launch_oneshot_out<DT, 4>(c, offset, (char*)out, chunks, scale, max_blocks, s);          \
    case 5: return launch_oneshot_out<DT, 5>(c, offset, (char*)out, chunks, scale, max_blocks, s);          \
    case 6: return launch_oneshot_out<DT, 6>(c, offset, (char*)out, chunks, scale, max_blocks, s);          \
    case 7: return launch_oneshot_out<DT, 7>(c, offset, (char*)out, chunks, scale, max_blocks, s);          \
    case 8: return launch_oneshot_out<DT, 8>(c, offset, (char*)out, chunks, scale, max_blocks, s);          \
    default: break;                                                                                         \
  }
  if (dtype == B200_DTYPE_F32) {
    B200_OS(B200_DTYPE_F32)
  } else {
    B200_OS(B200_DTYPE_BF16)
  }
#undef B200_OS
  b200_set_error("world size %d not in 1..8", c.world);
  return B200_ERR_ARG;
}

extern "C" int b200_broadcast(const b200_comm* comm, size_t offset, size_t bytes, int root, int max_blocks,
                              void* stream) {
  B200_REQUIRE(comm != nullptr, "null communicator");
  B200_REQUIRE(offset % 16 == 0 && bytes % 16 == 0, "offset/bytes must be multiples of 16");
  B200_REQUIRE(root >= 0 && root < comm->world, "bad root");
  if (comm->world == 1 || bytes == 0) return 0;
  int blocks = clamp_blocks(bytes / 16, 4, max_blocks);
  broadcast_kernel<<<blocks, kThreads, 0, (cudaStream_t)stream>>>(*comm, offset, bytes / 16, root);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_barrier(const b200_comm* comm, void* stream) {
  B200_REQUIRE(comm != nullptr, "null communicator");
  if (comm->world == 1) return 0;
  barrier_kernel<<<1, kThreads, 0, (cudaStream_t)stream>>>(*comm);
  B200_CHECK_LAUNCH();
  return 0;
}
