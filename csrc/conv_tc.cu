// Implicit-GEMM convolution / linear layers on the 5th-gen tensor cores (tcgen05 + TMEM), sm_100a.
// SURVEY W6 (conv fwd / dgrad / wgrad) and W9 (linear == 1x1 convolution on a [M,1,1,K] tensor).
//
// Replaces the cuDNN convolutions and cuBLAS GEMMs the reference reaches through nn.Conv2d / nn.Linear
// (torchvision resnet18 under notebooks/code/cifar10-distributed-smddp-gpu.py:32,165-167; MNTD models
// notebooks/code/model_lib/cifar10_cnn_model.py:11-18, mnist_cnn_model.py:11-15).
//
// One kernel template, three modes.  Activations are NHWC bf16, weights are [Cout][R][S][Cin] bf16 (the
// channels-last physical layout of the flat parameter arena), accumulation is fp32 in TMEM.
//
//   FWD    D[m=(n,ho,wo)][co]   = sum_{r,s,ci} X[n, ho*st-p+r, wo*st-p+s, ci] * W[co][r][s][ci]
//          A: gathered K-major,  B: weight rows K-major.            D -> bf16 (+bias, ReLU)
//   DGRAD  D[m=(n,h,w)][ci]     = sum_{r,s,co} dY[n, (h+p-r)/st, (w+p-s)/st, co] * W[co][r][s][ci]
//          A: gathered K-major,  B: weight rows *MN-major* (no transposed weight copy).  D -> bf16
//   WGRAD  D[m=co][(r,s,ci)]    = sum_{p=(n,ho,wo)} dY[p][co] * X[n, ho*st-p+r, wo*st-p+s, ci]
//          A and B both MN-major (the reduction runs over positions).  D -> fp32 red.add into the flat
//          gradient arena (split-K over positions across CTAs; the arena is zeroed by the optimizer).
//
// CTA = 5 warps.  Warps 0-3 gather operand tiles with 16-byte cp.async (zero-fill implements padding,
// stride holes and tile tails) into 128B-swizzled shared memory, a 4-stage ring guarded by mbarriers;
// warp 4 allocates TMEM and one elected lane issues tcgen05.mma (M=128, N=64/128, K=16 per instruction,
// 4 per stage), tcgen05.commit releases ring slots; warps 0-3 then read the accumulator back with
// tcgen05.ld (32x32b, one row per thread) and run the epilogue.
#include "common.cuh"

namespace {

constexpr int MODE_FWD = 0, MODE_DGRAD = 1, MODE_WGRAD = 2;
constexpr int BM = 128;
constexpr int BK = 64;                 // bf16 elements per k-block = one 128-byte swizzle row
constexpr int STAGES = 4;
constexpr int LAG = 2;                 // cp.async groups kept in flight behind the newest one
constexpr int PRODUCERS = 128;
constexpr int THREADS = PRODUCERS + 32;

struct FastDiv {
  uint32_t d, magic, shift;
};
__host__ FastDiv make_fastdiv(uint32_t d) {
  FastDiv f;
  f.d = d;
  uint32_t s = 0;
  while ((1ull << s) < d) ++s;
  f.shift = s;
  f.magic = (uint32_t)((((1ull << 32) * ((1ull << s) - d)) / d) + 1);
  return f;
}
__device__ __forceinline__ uint32_t fdiv(uint32_t x, const FastDiv& f) { return (__umulhi(x, f.magic) + x) >> f.shift; }

struct ConvParams {
  const __nv_bfloat16* x;    // image-side tensor (FWD/WGRAD: input, DGRAD: unused)
  const __nv_bfloat16* w;    // [Cout][R][S][Cin]
  const __nv_bfloat16* dy;   // output-side tensor (DGRAD/WGRAD)
  void* out;                 // FWD: y bf16 [M][Cout]; DGRAD: dx bf16 [M][Cin]; WGRAD: dw fp32 [Cout][R*S*Cin]
  const float* bias;         // FWD only, may be null
  int N, H, W, C;            // input geometry (C = Cin)
  int Kc, R, S, stride, pad; // Kc = Cout
  int Ho, Wo;
  int relu;
  uint32_t gemm_m, gemm_n, gemm_k;   // per mode
  uint32_t kblocks, kb_per_split;
  FastDiv div_c, div_s, div_kc, div_wo, div_ho, div_w, div_h, div_rs;
};

// ------------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done = 0;
  uint32_t spins = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
    if (done) break;
    if (++spins > (1u << 26)) __trap();   // a broken pipeline must fault, never hang the GPU
  }
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;   // src-size 0 => 16 bytes of zeros (padding / tails)
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor, SWIZZLE_128B, version 1 (Blackwell).  Addresses/offsets in 16-byte units.
//   K-major : rows of 128 B, 8-row groups SBO=1024 B apart (LBO unused -> 1)
//   MN-major: K rows of 128 B (64 MN elements), 8-row groups SBO=1024 B apart, 64-element MN groups LBO apart
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;   // version
  d |= (uint64_t)2 << 61;   // SWIZZLE_128B
  return d;
}
// Instruction descriptor for kind::f16: bf16 x bf16 -> fp32
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t swz(uint32_t row, uint32_t chunk) { return row * 128u + ((chunk ^ (row & 7u)) << 4); }

// ------------------------------------------------------------------------------------------------
template <int MODE, int BN>
struct Cfg {
  static constexpr bool A_MN = (MODE == MODE_WGRAD);
  static constexpr bool B_MN = (MODE != MODE_FWD);
  static constexpr int A_BYTES = BM * BK * 2;          // 16 KiB either way
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

template <int MODE, int BN>
__global__ void __launch_bounds__(THREADS, 1) conv_tc_kernel(const ConvParams p) {
  using C = Cfg<MODE, BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const uint32_t kb_begin = blockIdx.z * p.kb_per_split;
  const uint32_t kb_end = min(p.kblocks, kb_begin + p.kb_per_split);
  const uint32_t nkb = kb_end > kb_begin ? kb_end - kb_begin : 0;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], PRODUCERS);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (nkb > 0) {
    if (warp < 4) {
      // =============================== PRODUCERS (cp.async gather) ===============================
      const uint32_t smem_base = smem_u32(smem);
      // ---- per-thread A row info
      // K-major A (FWD/DGRAD): thread owns chunk column jA = tid&7 of rows (tid>>3) + 16*i, i<8
      // MN-major A (WGRAD): 16 chunks per K row: column cA = tid&15, rows (tid>>4) + 8*i, i<8
      int a_n[8], a_h[8], a_w[8];   // FWD: (n, ho*st-pad, wo*st-pad); DGRAD: (n, h+pad, w+pad); -1 => row out of range
      if (MODE != MODE_WGRAD) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t m = m0 + (tid >> 3) + 16 * i;
          if (m < p.gemm_m) {
            if (MODE == MODE_FWD) {
              const uint32_t t = fdiv(m, p.div_wo), wo = m - t * p.Wo;
              const uint32_t n = fdiv(t, p.div_ho), ho = t - n * p.Ho;
              a_n[i] = (int)n;
              a_h[i] = (int)ho * p.stride - p.pad;
              a_w[i] = (int)wo * p.stride - p.pad;
            } else {
              const uint32_t t = fdiv(m, p.div_w), w = m - t * p.W;
              const uint32_t n = fdiv(t, p.div_h), h = t - n * p.H;
              a_n[i] = (int)n;
              a_h[i] = (int)h + p.pad;
              a_w[i] = (int)w + p.pad;
            }
          } else {
            a_n[i] = -1; a_h[i] = 0; a_w[i] = 0;
          }
        }
      }
      // ---- per-thread B info for WGRAD: chunk column -> (r, s, ci) is fixed for the whole tile
      int wb_r = 0, wb_s = 0, wb_c = 0;
      bool wb_ok = false;
      if (MODE == MODE_WGRAD) {
        constexpr int CPR = BN / 8;
        const uint32_t nn = n0 + (tid % CPR) * 8;
        wb_ok = nn < p.gemm_n;
        const uint32_t tap = fdiv(nn, p.div_c);
        wb_c = (int)(nn - tap * p.C);
        wb_r = (int)fdiv(tap, p.div_s);
        wb_s = (int)(tap - wb_r * p.S);
      }

      for (uint32_t it = 0; it < nkb; ++it) {
        const uint32_t kb = kb_begin + it;
        const uint32_t stage = it % STAGES, ph = (it / STAGES) & 1;
        mbar_wait(&empty_bar[stage], ph ^ 1);
        const uint32_t sa = smem_base + stage * C::STAGE_BYTES, sb = sa + C::A_BYTES;

        // ------------------------------- A operand -------------------------------
        if (MODE == MODE_FWD || MODE == MODE_DGRAD) {
          const uint32_t j = tid & 7;
          const uint32_t k = kb * BK + j * 8;
          const bool k_ok = k < p.gemm_k;
          // k -> (tap, channel)
          const uint32_t cdim = (MODE == MODE_FWD) ? (uint32_t)p.C : (uint32_t)p.Kc;
          const uint32_t tap = fdiv(k, MODE == MODE_FWD ? p.div_c : p.div_kc);
          const int ch = (int)(k - tap * cdim);
          const int r = (int)fdiv(tap, p.div_s), s = (int)tap - r * p.S;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t row = (tid >> 3) + 16 * i;
            bool ok = k_ok && a_n[i] >= 0;
            const __nv_bfloat16* src = (MODE == MODE_FWD) ? p.x : p.dy;
            if (MODE == MODE_FWD) {
              const int h = a_h[i] + r, w = a_w[i] + s;
              ok = ok && h >= 0 && h < p.H && w >= 0 && w < p.W;
              if (ok) src = p.x + ((((size_t)a_n[i] * p.H + h) * p.W + w) * p.C + ch);
            } else {
              const int hn = a_h[i] - r, wn = a_w[i] - s;   // = ho*stride, wo*stride
              int ho = hn, wo = wn;
              bool div_ok = hn >= 0 && wn >= 0;
              if (p.stride != 1) {
                ho = hn / p.stride; wo = wn / p.stride;
                div_ok = div_ok && (ho * p.stride == hn) && (wo * p.stride == wn);
              }
              ok = ok && div_ok && ho < p.Ho && wo < p.Wo;
              if (ok) src = p.dy + ((((size_t)a_n[i] * p.Ho + ho) * p.Wo + wo) * p.Kc + ch);
            }
            cp_async16(sa + swz(row, j), src, ok);
          }
        } else {  // WGRAD A^T: rows are positions, 16 chunks (128 co) per row
          const uint32_t c = tid & 15;
          const uint32_t co = m0 + c * 8;
          const bool co_ok = co < p.gemm_m;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t row = (tid >> 4) + 8 * i;
            const uint32_t pos = kb * BK + row;
            const bool ok = co_ok && pos < p.gemm_k;
            const __nv_bfloat16* src = ok ? p.dy + ((size_t)pos * p.Kc + co) : p.dy;
            cp_async16(sa + (c >> 3) * (BK * 128) + swz(row, c & 7), src, ok);
          }
        }

        // ------------------------------- B operand -------------------------------
        if (MODE == MODE_FWD) {          // weight rows, K-major: BN rows x 8 chunks
          const uint32_t j = tid & 7;
          const uint32_t k = kb * BK + j * 8;
          const bool k_ok = k < p.gemm_k;
#pragma unroll
          for (int i = 0; i < BN / 16; ++i) {
            const uint32_t row = (tid >> 3) + 16 * i;
            const uint32_t nn = n0 + row;
            const bool ok = k_ok && nn < p.gemm_n;
            const __nv_bfloat16* src = ok ? p.w + ((size_t)nn * p.gemm_k + k) : p.w;
            cp_async16(sb + swz(row, j), src, ok);
          }
        } else if (MODE == MODE_DGRAD) {  // B^T rows k=(r,s,co) -> weight row (co,r,s); BN/8 chunks of ci
          constexpr int CPR = BN / 8;
          constexpr int RPP = PRODUCERS / CPR;   // rows per pass
          const uint32_t c = tid % CPR;
          const uint32_t ci = n0 + c * 8;
          const bool ci_ok = ci < p.gemm_n;
#pragma unroll
          for (int i = 0; i < BK / RPP; ++i) {
            const uint32_t row = (tid / CPR) + RPP * i;
            const uint32_t k = kb * BK + row;
            const bool ok = ci_ok && k < p.gemm_k;
            const uint32_t tap = fdiv(k, p.div_kc);
            const uint32_t co = k - tap * p.Kc;
            const __nv_bfloat16* src = ok ? p.w + (((size_t)co * (p.R * p.S) + tap) * p.C + ci) : p.w;
            cp_async16(sb + (c >> 3) * (BK * 128) + swz(row, c & 7), src, ok);
          }
        } else {  // WGRAD B^T: rows are positions, chunk column fixed (r,s,ci)
          constexpr int CPR = BN / 8;
          constexpr int RPP = PRODUCERS / CPR;
          const uint32_t c = tid % CPR;
#pragma unroll
          for (int i = 0; i < BK / RPP; ++i) {
            const uint32_t row = (tid / CPR) + RPP * i;
            const uint32_t pos = kb * BK + row;
            bool ok = wb_ok && pos < p.gemm_k;
            const __nv_bfloat16* src = p.x;
            if (ok) {
              const uint32_t t = fdiv(pos, p.div_wo), wo = pos - t * p.Wo;
              const uint32_t n = fdiv(t, p.div_ho), ho = t - n * p.Ho;
              const int h = (int)ho * p.stride - p.pad + wb_r, w = (int)wo * p.stride - p.pad + wb_s;
              ok = h >= 0 && h < p.H && w >= 0 && w < p.W;
              if (ok) src = p.x + ((((size_t)n * p.H + h) * p.W + w) * p.C + wb_c);
            }
            cp_async16(sb + (c >> 3) * (BK * 128) + swz(row, c & 7), src, ok);
          }
        }
        cp_async_commit();
        if (it >= LAG) {
          cp_async_wait<LAG>();          // group (it - LAG) has landed
          fence_proxy_async();           // generic-proxy writes -> visible to the async proxy (tcgen05)
          mbar_arrive(&full_bar[(it - LAG) % STAGES]);
        }
      }
      // drain the last min(LAG, nkb) groups
      cp_async_wait<0>();
      fence_proxy_async();
      {
        const uint32_t first = nkb > LAG ? nkb - LAG : 0;
        for (uint32_t it = first; it < nkb; ++it) mbar_arrive(&full_bar[it % STAGES]);
      }

      // =============================== EPILOGUE (TMEM -> registers -> global) ===============================
      mbar_wait(tmem_full, 0);
      tc_fence_after();
      const uint32_t row = warp * 32 + lane;          // accumulator row == TMEM lane
      const uint32_t m = m0 + row;
#pragma unroll 1
      for (int cb = 0; cb < BN / 32; ++cb) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + cb * 32, v);
        const uint32_t nn0 = n0 + cb * 32;
        if (m < p.gemm_m) {
          if (MODE == MODE_WGRAD) {
            float* dst = reinterpret_cast<float*>(p.out) + (size_t)m * p.gemm_n + nn0;
            if (nn0 + 32 <= p.gemm_n) {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 4 * q), "f"(__uint_as_float(v[4 * q])),
                             "f"(__uint_as_float(v[4 * q + 1])), "f"(__uint_as_float(v[4 * q + 2])),
                             "f"(__uint_as_float(v[4 * q + 3]))
                             : "memory");
            } else {
              for (int q = 0; q < 32; ++q)
                if (nn0 + q < p.gemm_n) atomicAdd(dst + q, __uint_as_float(v[q]));
            }
          } else {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)m * p.gemm_n + nn0;
            float f[32];
#pragma unroll
            for (int q = 0; q < 32; ++q) {
              float a = __uint_as_float(v[q]);
              if (MODE == MODE_FWD) {
                if (p.bias != nullptr && nn0 + q < p.gemm_n) a += __ldg(p.bias + nn0 + q);
                if (p.relu) a = fmaxf(a, 0.f);
              }
              f[q] = a;
            }
            if (nn0 + 32 <= p.gemm_n) {
#pragma unroll
              for (int q = 0; q < 4; ++q) *reinterpret_cast<bf16x8*>(dst + 8 * q) = float_to_bf16x8(f + 8 * q);
            } else {
              for (int q = 0; q < 32; ++q)
                if (nn0 + q < p.gemm_n) dst[q] = __float2bfloat16(f[q]);
            }
          }
        }
      }
    } else if (lane == 0) {
      // =============================== MMA ISSUER (one thread) ===============================
      constexpr uint32_t idesc = make_idesc(BM, BN, C::A_MN ? 1 : 0, C::B_MN ? 1 : 0);
      const uint32_t smem_base = smem_u32(smem);
      for (uint32_t it = 0; it < nkb; ++it) {
        const uint32_t stage = it % STAGES, ph = (it / STAGES) & 1;
        mbar_wait(&full_bar[stage], ph);
        tc_fence_after();
        const uint32_t sa = smem_base + stage * C::STAGE_BYTES, sb = sa + C::A_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          // K-major: +32 B per K=16 inside the 128-B swizzle row.  MN-major: +2 row groups (2048 B).
          const uint64_t da = C::A_MN ? make_desc(sa + k * 2048, BK * 128, 1024) : make_desc(sa + k * 32, 16, 1024);
          const uint64_t db = C::B_MN ? make_desc(sb + k * 2048, BK * 128, 1024) : make_desc(sb + k * 32, 16, 1024);
          umma_bf16(tmem_base, da, db, idesc, (it | (uint32_t)k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);            // ring slot free once these MMAs retire
        if (it + 1 == nkb) umma_commit(tmem_full);  // accumulator complete
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    __syncwarp();
    tmem_dealloc(tmem_base, BN);
  }
}

// ------------------------------------------------------------------------------------------------
template <int MODE, int BN>
int launch(const ConvParams& p, dim3 grid, cudaStream_t s) {
  using C = Cfg<MODE, BN>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<MODE, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set = true;
  }
  conv_tc_kernel<MODE, BN><<<grid, THREADS, C::SMEM_BYTES, s>>>(p);
  B200_CHECK_LAUNCH();
  return 0;
}

int fill_common(ConvParams& p, const b200_conv_geom* g) {
  B200_REQUIRE(g != nullptr, "null geometry");
  B200_REQUIRE(g->C % 8 == 0 && g->K % 8 == 0, "channel counts must be multiples of 8 (16-byte gathers)");
  B200_REQUIRE(g->stride >= 1 && g->R >= 1 && g->S >= 1 && g->N >= 1, "bad geometry");
  p.N = g->N; p.H = g->H; p.W = g->W; p.C = g->C;
  p.Kc = g->K; p.R = g->R; p.S = g->S; p.stride = g->stride; p.pad = g->pad;
  p.Ho = (g->H + 2 * g->pad - g->R) / g->stride + 1;
  p.Wo = (g->W + 2 * g->pad - g->S) / g->stride + 1;
  B200_REQUIRE(p.Ho >= 1 && p.Wo >= 1, "empty output");
  B200_REQUIRE((size_t)g->N * g->H * g->W < (1ull << 31) && (size_t)g->N * p.Ho * p.Wo < (1ull << 31), "tensor too large for 32-bit tile indices");
  p.div_c = make_fastdiv(p.C); p.div_s = make_fastdiv(p.S); p.div_kc = make_fastdiv(p.Kc);
  p.div_wo = make_fastdiv(p.Wo); p.div_ho = make_fastdiv(p.Ho);
  p.div_w = make_fastdiv(p.W); p.div_h = make_fastdiv(p.H); p.div_rs = make_fastdiv(p.R * p.S);
  p.relu = 0; p.bias = nullptr; p.x = nullptr; p.w = nullptr; p.dy = nullptr; p.out = nullptr;
  return 0;
}

}  // namespace

extern "C" int b200_conv_fwd(const void* x, const void* w, const float* bias, void* y, const b200_conv_geom* g,
                             int relu, void* stream) {
  ConvParams p;
  int rc = fill_common(p, g);
  if (rc) return rc;
  B200_REQUIRE(x && w && y, "null tensor");
  p.x = (const __nv_bfloat16*)x; p.w = (const __nv_bfloat16*)w; p.out = y; p.bias = bias; p.relu = relu;
  p.gemm_m = (uint32_t)p.N * p.Ho * p.Wo; p.gemm_n = p.Kc; p.gemm_k = (uint32_t)p.R * p.S * p.C;
  p.kblocks = (p.gemm_k + BK - 1) / BK; p.kb_per_split = p.kblocks;
  const bool narrow = p.gemm_n <= 64;
  dim3 grid((p.gemm_m + BM - 1) / BM, (p.gemm_n + (narrow ? 64 : 128) - 1) / (narrow ? 64 : 128), 1);
  return narrow ? launch<MODE_FWD, 64>(p, grid, (cudaStream_t)stream) : launch<MODE_FWD, 128>(p, grid, (cudaStream_t)stream);
}

extern "C" int b200_conv_dgrad(const void* dy, const void* w, void* dx, const b200_conv_geom* g, void* stream) {
  ConvParams p;
  int rc = fill_common(p, g);
  if (rc) return rc;
  B200_REQUIRE(dy && w && dx, "null tensor");
  p.dy = (const __nv_bfloat16*)dy; p.w = (const __nv_bfloat16*)w; p.out = dx;
  p.gemm_m = (uint32_t)p.N * p.H * p.W; p.gemm_n = p.C; p.gemm_k = (uint32_t)p.R * p.S * p.Kc;
  p.kblocks = (p.gemm_k + BK - 1) / BK; p.kb_per_split = p.kblocks;
  const bool narrow = p.gemm_n <= 64;
  dim3 grid((p.gemm_m + BM - 1) / BM, (p.gemm_n + (narrow ? 64 : 128) - 1) / (narrow ? 64 : 128), 1);
  return narrow ? launch<MODE_DGRAD, 64>(p, grid, (cudaStream_t)stream) : launch<MODE_DGRAD, 128>(p, grid, (cudaStream_t)stream);
}

extern "C" int b200_conv_wgrad(const void* x, const void* dy, float* dw, const b200_conv_geom* g, int max_splits,
                               void* stream) {
  ConvParams p;
  int rc = fill_common(p, g);
  if (rc) return rc;
  B200_REQUIRE(x && dy && dw, "null tensor");
  p.x = (const __nv_bfloat16*)x; p.dy = (const __nv_bfloat16*)dy; p.out = dw;
  p.gemm_m = p.Kc; p.gemm_n = (uint32_t)p.R * p.S * p.C; p.gemm_k = (uint32_t)p.N * p.Ho * p.Wo;
  p.kblocks = (p.gemm_k + BK - 1) / BK;
  const bool narrow = p.gemm_n <= 64;
  const uint32_t bn = narrow ? 64 : 128;
  const uint32_t tiles = ((p.gemm_m + BM - 1) / BM) * ((p.gemm_n + bn - 1) / bn);
  // split the position reduction so the grid covers the machine ~2x; >= 4 k-blocks per CTA
  uint32_t splits = (2u * (uint32_t)b200_num_sms() + tiles - 1) / tiles;
  if (max_splits > 0 && splits > (uint32_t)max_splits) splits = (uint32_t)max_splits;
  const uint32_t max_by_k = (p.kblocks + 3) / 4;
  if (splits > max_by_k) splits = max_by_k;
  if (splits < 1) splits = 1;
  p.kb_per_split = (p.kblocks + splits - 1) / splits;
  splits = (p.kblocks + p.kb_per_split - 1) / p.kb_per_split;
  dim3 grid((p.gemm_m + BM - 1) / BM, (p.gemm_n + bn - 1) / bn, splits);
  return narrow ? launch<MODE_WGRAD, 64>(p, grid, (cudaStream_t)stream) : launch<MODE_WGRAD, 128>(p, grid, (cudaStream_t)stream);
}
