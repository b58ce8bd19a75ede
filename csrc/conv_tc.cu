// Implicit-GEMM convolution / linear layers on the 5th-gen tensor cores (tcgen05 + TMEM), sm_100a.
// SURVEY W6 (conv fwd / dgrad / wgrad) and W9 (linear == 1x1 convolution on a [M,1,1,K] tensor).
//
// Replaces the cuDNN convolutions and cuBLAS GEMMs the reference reaches through nn.Conv2d / nn.Linear
// (torchvision resnet18 under notebooks/code/cifar10-distributed-smddp-gpu.py:32,165-167; MNTD models
// notebooks/code/model_lib/cifar10_cnn_model.py:11-18, mnist_cnn_model.py:11-15).
//
// One kernel template, three modes, two operand types.  Activations are NHWC, weights are [Cout][R][S][Cin] (the
// channels-last physical layout of the flat parameter arena), accumulation is fp32 in TMEM.  T = __nv_bfloat16 runs
// tcgen05.mma kind::f16 (K = 16 per instruction); T = float keeps activations, weights and gradients in fp32 storage and
// runs kind::tf32 (K = 8 per instruction: the tensor core reads the top 19 bits of every fp32 operand) -- the precision
// of the reference's own run, which trains fp32 with cuDNN's TF32 convolutions and no autocast
// (notebooks/code/cifar10-distributed-smddp-gpu.py:145,155-168).  A 128-byte swizzle row holds 64 bf16 or 32 fp32, so the
// shared-memory tiles, the ring and the descriptors' byte arithmetic are the same for both; what changes is elements per
// 16-byte chunk (8 / 4), k-blocks per filter tap, and the MN-major (transposed) operand layout, which for 32-bit elements
// must be the "128B swizzle with 32-byte atomicity" form (4-row atoms, 32-byte units XORed with the row index).
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
// stride holes and tile tails) into 128B-swizzled shared memory, a 6-stage ring guarded by mbarriers;
// warp 4 allocates TMEM and one elected lane issues tcgen05.mma (M=128, N=64/128, K=16 per instruction,
// 4 per stage), tcgen05.commit releases ring slots; warps 0-3 then read the accumulator back with
// tcgen05.ld (32x32b, one row per thread) and run the epilogue.
//
// Small-problem machinery (the reference's resnet18 runs at 8x8 .. 1x1 spatial):
//   * filter taps that can never touch a real pixel (e.g. 8 of the 9 taps of a 3x3 conv on a 1x1 map)
//     are dropped from the K loop (FWD/DGRAD) or from the grid (WGRAD) — decided on the host;
//   * split-K inside a thread-block cluster when the tile grid cannot fill 148 SMs: the 2/4/8 CTAs of a
//     cluster (hardware co-scheduled) each accumulate a K slice in their own TMEM, stage the fp32 partial
//     in shared memory, and after a cluster barrier every CTA reduces its BN/ZC column slice over all peers
//     through distributed shared memory (ld.shared::cluster) and runs the epilogue for it: no atomics, no
//     global scratch, deterministic.  WGRAD additionally spreads clusters over K and joins them with red.add.
#include <cuda.h>
#include <stdlib.h>

#include <algorithm>

#include "tc_ptx.cuh"
#include "act8.cuh"

using namespace b200tc;

namespace {

constexpr int MODE_FWD = 0, MODE_DGRAD = 1, MODE_WGRAD = 2;
constexpr int BM = 128;
// per operand type: bytes per element, elements per 16-byte chunk, K elements per k-block (= one 128-byte swizzle row),
// K per tcgen05.mma, and the A/B format code of the instruction descriptor (1 = BF16, 2 = TF32)
template <typename T> struct El;
template <> struct El<__nv_bfloat16> { static constexpr int ES = 2, EPC = 8, BK = 64, UK = 16; static constexpr uint32_t FMT = 1; };
template <> struct El<float> { static constexpr int ES = 4, EPC = 4, BK = 32, UK = 8; static constexpr uint32_t FMT = 2; };
// Operand ring depth ST (template parameter): 6 stages / 1 CTA per SM for the latency-bound small problems, 3 stages /
// 2 CTAs per SM for grids of many tiles (one tile's load latency and epilogue hide behind the other CTA's MMAs).
constexpr int lag_for(int st) { return st >= 6 ? 4 : 2; }   // cp.async groups kept in flight behind the newest one
constexpr int PRODUCERS = 128;
constexpr int THREADS = PRODUCERS + 32;
long long* g_prof = nullptr;           // b200_conv_set_profile_buffer
bool g_split_k = true;                 // b200_conv_set_split_k(0) pins FWD/DGRAD to one CTA per tile (reproducibility checks)

struct ConvParams {
  const void* x;             // image-side tensor (FWD/WGRAD: input, DGRAD: unused); element type T of the kernel
  const void* w;             // [Cout][R][S][Cin]
  const void* dy;            // output-side tensor (DGRAD/WGRAD)
  void* out;                 // FWD: y T [M][Cout]; DGRAD: dx T [M][Cin]; WGRAD: dw fp32 [Cout][R*S*Cin]
  const float* bias;         // FWD only, may be null
  const void* residual;      // DGRAD only, may be null: added to the output (the other gradient of a forked tensor)
  float* bn_sums;            // FWD only, may be null: per-channel sum [gemm_n] and sum of squares [gemm_n] of the output
  int N, H, W, C;            // input geometry (C = Cin)
  int Kc, R, S, stride, pad; // Kc = Cout
  int Ho, Wo;
  int relu;
  uint32_t gemm_m, gemm_n, gemm_k;   // per mode
  uint32_t nvk;              // number of (virtual) k-blocks that are executed
  uint32_t kb_per_split;     // virtual k-blocks per grid.z slice
  uint32_t cpt;              // k-blocks per filter tap (0: no tap skipping, virtual == real)
  uint32_t a_tma, b_tma;     // operand staged by TMA (plain 2-D matrix) instead of the cp.async gather
  long long* prof;           // optional: CTA (0,0,0) drops clock64() stamps of its phases here (tuning aid)
  unsigned long long tapmask;  // bit t set: tap t can touch a real pixel somewhere in the tensor
  uint8_t tap_list[64];      // the set bits of tapmask, ascending
  FastDiv div_c, div_s, div_kc, div_wo, div_ho, div_w, div_h, div_cpt;
};

template <typename T, int MODE, int BN, int ST = 6>
struct Cfg {
  static constexpr bool A_MN = (MODE == MODE_WGRAD);
  static constexpr bool B_MN = (MODE != MODE_FWD);
  static constexpr int A_BYTES = BM * 128;             // BM rows (K-major) or BK k-rows x BM/(128/ES) groups (MN-major) of 128 B
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int SMEM_BYTES = ST * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + 2048 /*row info*/;
};

__device__ __forceinline__ uint32_t virt2kb(const ConvParams& p, uint32_t v) {
  if (p.cpt == 0) return v;
  const uint32_t q = fdiv(v, p.div_cpt);
  return (uint32_t)p.tap_list[q] * p.cpt + (v - q * p.cpt);
}

// BatchNorm statistics in the conv epilogue: every thread holds one output row (NCOLS fp32 accumulators); a butterfly of
// NCOLS-1 shuffles leaves lane l with the warp's column-l total (instead of 5 shuffles per column), then one red.add
// per column per warp into sums[channel] / sums[gemm_n + channel].  Rows beyond gemm_m are exact zeros (zero-filled
// operands, no bias), so they need no mask.
template <int NCOLS>
__device__ __forceinline__ void bn_stats_accumulate(const ConvParams& p, uint32_t nn0, const float* f, int lane) {
  static_assert(NCOLS == 8 || NCOLS == 16 || NCOLS == 32 || NCOLS == 64 || NCOLS == 128, "column block");
  constexpr int W = NCOLS < 32 ? NCOLS : 32;
#pragma unroll
  for (int c0 = 0; c0 < NCOLS; c0 += 32) {
    float s[W], q[W];
#pragma unroll
    for (int i = 0; i < W; ++i) { s[i] = f[c0 + i]; q[i] = s[i] * s[i]; }
#pragma unroll
    for (int st = W / 2; st >= 1; st >>= 1) {
      const bool up = (lane & st) != 0;
#pragma unroll
      for (int i = 0; i < st; ++i) {
        const float ks = up ? s[i + st] : s[i], ss = up ? s[i] : s[i + st];
        const float kq = up ? q[i + st] : q[i], sq = up ? q[i] : q[i + st];
        s[i] = ks + __shfl_xor_sync(0xffffffffu, ss, st);
        q[i] = kq + __shfl_xor_sync(0xffffffffu, sq, st);
      }
    }
#pragma unroll
    for (int st = W; st < 32; st <<= 1) {                     // narrow blocks: the lanes above W hold the other partials
      s[0] += __shfl_xor_sync(0xffffffffu, s[0], st);
      q[0] += __shfl_xor_sync(0xffffffffu, q[0], st);
    }
    const uint32_t ch = nn0 + c0 + (lane & (W - 1));
    if (lane < W && ch < p.gemm_n) {
      atomicAdd(p.bn_sums + ch, s[0]);
      atomicAdd(p.bn_sums + p.gemm_n + ch, q[0]);
    }
  }
}

// Write NCOLS accumulator columns [nn0, nn0+NCOLS) of GEMM row m, 8 columns (one 16-byte bf16 vector) at a
// time; gemm_n is a multiple of 8 in every mode.  FWD: +bias, ReLU, bf16.  DGRAD: (+residual) bf16.  WGRAD: fp32 red.add
// into the gradient arena.  Kept small on purpose: a tile runs this code once, so its size is fetch latency.
template <typename T, int MODE, int NCOLS>
__device__ __forceinline__ void epilogue_store(const ConvParams& p, uint32_t m, uint32_t nn0, float* f) {
  if (m >= p.gemm_m) return;
  Act8<T> resv[NCOLS / 8];
  if (MODE == MODE_DGRAD && p.residual != nullptr) {        // all of this row's residual vectors in flight at once
#pragma unroll
    for (int q = 0; q < NCOLS / 8; ++q)
      if (nn0 + 8 * q < p.gemm_n) resv[q] = ld_act8(reinterpret_cast<const T*>(p.residual) + (size_t)m * p.gemm_n + nn0 + 8 * q);
  }
#pragma unroll
  for (int q = 0; q < NCOLS / 8; ++q) {
    const uint32_t nn = nn0 + 8 * q;
    if (nn >= p.gemm_n) break;
    float* v = f + 8 * q;
    if (MODE == MODE_WGRAD) {
      float* dst = reinterpret_cast<float*>(p.out) + (size_t)m * p.gemm_n + nn;
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 4), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
    } else {
      if (MODE == MODE_FWD) {
        if (p.bias != nullptr) {
          const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + nn));
          const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + nn) + 1);
          v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
          v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
        }
        if (p.relu) {
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
        }
      }
      if (MODE == MODE_DGRAD && p.residual != nullptr) {
        float r[8];
        act8_to_float(resv[q], r);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] += r[k];
      }
      st_act8(reinterpret_cast<T*>(p.out) + (size_t)m * p.gemm_n + nn, v);
    }
  }
}

template <typename T, int MODE, int BN, int ZC, int ST>
__global__ void __launch_bounds__(THREADS, (ST <= 3 ? 2 : 1))
conv_tc_kernel(const __grid_constant__ ConvParams p, const __grid_constant__ CUtensorMap tmA,
               const __grid_constant__ CUtensorMap tmB) {
  using C = Cfg<T, MODE, BN, ST>;
  constexpr int STAGES = ST, LAG = lag_for(ST);
  constexpr int BK = El<T>::BK, EPC = El<T>::EPC;
  const T* const px = reinterpret_cast<const T*>(p.x);
  const T* const pw = reinterpret_cast<const T*>(p.w);
  const T* const pdy = reinterpret_cast<const T*>(p.dy);
  pdl_trigger();                  // the next kernel of the stream may be scheduled now; it waits for us in its own pdl_wait()
  const uint32_t m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  if (MODE == MODE_WGRAD && p.cpt != 0) {
    // N tile = columns (r,s,ci): drop tiles whose taps never see a real pixel (their dW is exactly 0)
    const uint32_t t_lo = fdiv(n0, p.div_c);
    const uint32_t n_hi = min(n0 + (uint32_t)BN, p.gemm_n) - 1;
    const uint32_t t_hi = fdiv(n_hi, p.div_c);
    bool any = false;
    for (uint32_t t = t_lo; t <= t_hi; ++t) any = any || ((p.tapmask >> t) & 1ull);
    if (!any) return;
  }
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  const T** s_rowbase = reinterpret_cast<const T**>(smem + STAGES * C::STAGE_BYTES + 256);
  unsigned long long* s_rowmask = reinterpret_cast<unsigned long long*>(smem + STAGES * C::STAGE_BYTES + 256 + 1024);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool prof = p.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
#define STAMP(i, t) do { if (prof && tid == (t)) p.prof[i] = clock64(); } while (0)
  STAMP(0, 0);
  const uint32_t v_begin = blockIdx.z * p.kb_per_split;
  const uint32_t v_end = min(p.nvk, v_begin + p.kb_per_split);
  const uint32_t nkb = v_end > v_begin ? v_end - v_begin : 0;

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      // arrivals per stage: the 128 gather threads (if any operand still uses cp.async) + the TMA-issuing thread
      mbar_init(&full_bar[s], ((p.a_tma & p.b_tma) ? 0u : (uint32_t)PRODUCERS) + ((p.a_tma | p.b_tma) ? 1u : 0u));
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    if (p.a_tma) tma_prefetch_desc(&tmA);
    if (p.b_tma) tma_prefetch_desc(&tmB);
  }
  if (warp == 4) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                     // barriers, TMEM and descriptors are set up; from here on we read the previous kernel's output
  STAMP(1, 0);

  if (warp < 4) {
    if (nkb > 0) {
      // =============================== PRODUCERS (cp.async gather) ===============================
      const uint32_t smem_base = smem_u32(smem);
      // K-major A (FWD/DGRAD): thread owns chunk column tid&7 of rows (tid>>3) + 16*i, i<8
      // MN-major A (WGRAD): BM/EPC chunks per K row (16 bf16 / 32 fp32): column tid % CPRA, rows tid / CPRA + RPPA*i, i<8
      // Per row: a base pointer and a bit mask of the filter taps that land on a real pixel, computed once per
      // tile by ONE thread per row (through shared memory).  In the K loop an A chunk is then
      // `rowbase[i] + tap_offset + channel` guarded by one bit test.
      //   FWD   base = &x[n, ho*st-pad, wo*st-pad, 0],   tap (r,s) adds  (r*W + s)*C
      //   DGRAD base = &dy[n, (h+pad)>>sh, (w+pad)>>sh, 0], tap (r,s) subtracts ((r>>sh)*Wo + (s>>sh))*Kc, sh = stride==2;
      //         a tap is live when (h+pad-r), (w+pad-s) are multiples of the stride and inside the output map.
      const T* rowbase[8];
      unsigned long long rowmask[8];
      if (MODE != MODE_WGRAD && !p.a_tma) {
        const int sh = (p.stride == 2) ? 1 : 0;
        {
          const uint32_t m = m0 + tid;                      // tid < 128: this thread's row
          const T* base = (MODE == MODE_FWD) ? px : pdy;
          unsigned long long mask = 0ull;
          if (m < p.gemm_m) {
            int h0, w0, hlim, wlim;
            if (MODE == MODE_FWD) {
              const uint32_t t = fdiv(m, p.div_wo), wo = m - t * p.Wo;
              const uint32_t n = fdiv(t, p.div_ho), ho = t - n * p.Ho;
              h0 = (int)ho * p.stride - p.pad; w0 = (int)wo * p.stride - p.pad; hlim = p.H; wlim = p.W;
              base = px + (((long long)n * p.H + h0) * p.W + w0) * p.C;
            } else {
              const uint32_t t = fdiv(m, p.div_w), w = m - t * p.W;
              const uint32_t n = fdiv(t, p.div_h), h = t - n * p.H;
              h0 = (int)h + p.pad; w0 = (int)w + p.pad; hlim = p.Ho; wlim = p.Wo;
              base = pdy + (((long long)n * p.Ho + (h0 >> sh)) * p.Wo + (w0 >> sh)) * p.Kc;
            }
            unsigned long long sbits = 0ull;
#pragma unroll 1
            for (int s2 = 0; s2 < p.S; ++s2) {
              const int wn = (MODE == MODE_FWD) ? w0 + s2 : w0 - s2;
              const bool ok = (MODE == MODE_FWD) ? ((unsigned)wn < (unsigned)wlim) : (wn >= 0 && (wn & sh) == 0 && (wn >> sh) < wlim);
              if (ok) sbits |= 1ull << s2;
            }
#pragma unroll 1
            for (int r = 0; r < p.R; ++r) {
              const int hn = (MODE == MODE_FWD) ? h0 + r : h0 - r;
              const bool ok = (MODE == MODE_FWD) ? ((unsigned)hn < (unsigned)hlim) : (hn >= 0 && (hn & sh) == 0 && (hn >> sh) < hlim);
              if (ok) mask |= sbits << (r * p.S);
            }
          }
          s_rowbase[tid] = base;
          s_rowmask[tid] = mask;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          rowbase[i] = s_rowbase[(tid >> 3) + 16 * i];
          rowmask[i] = s_rowmask[(tid >> 3) + 16 * i];
        }
      }
      // WGRAD B: chunk column -> (r, s, ci) is fixed for the whole tile
      int wb_r = 0, wb_s = 0, wb_c = 0;
      bool wb_ok = false;
      if (MODE == MODE_WGRAD) {
        constexpr int CPR = BN / EPC;
        const uint32_t nn = n0 + (tid % CPR) * EPC;
        wb_ok = nn < p.gemm_n;
        const uint32_t tap = fdiv(nn, p.div_c);
        wb_c = (int)(nn - tap * p.C);
        wb_r = (int)fdiv(tap, p.div_s);
        wb_s = (int)(tap - wb_r * p.S);
      }

      for (uint32_t it = 0; it < nkb; ++it) {
        const uint32_t kb = (MODE == MODE_WGRAD) ? (v_begin + it) : virt2kb(p, v_begin + it);
        const uint32_t stage = it % STAGES, ph = (it / STAGES) & 1;
        mbar_wait(&empty_bar[stage], ph ^ 1);
        const uint32_t sa = smem_base + stage * C::STAGE_BYTES, sb = sa + C::A_BYTES;

        // ------------------------------- A operand -------------------------------
        if (p.a_tma) {
          // plain 2-D operand: fetched by TMA below
        } else if (MODE == MODE_FWD || MODE == MODE_DGRAD) {
          const uint32_t j = tid & 7;
          const uint32_t k = kb * BK + j * EPC;
          const bool k_ok = k < p.gemm_k;
          const uint32_t cdim = (MODE == MODE_FWD) ? (uint32_t)p.C : (uint32_t)p.Kc;
          const uint32_t tap = fdiv(k, MODE == MODE_FWD ? p.div_c : p.div_kc);
          const int ch = (int)(k - tap * cdim);
          const int r = (int)fdiv(tap, p.div_s), s2 = (int)tap - r * p.S;
          long long off;     // same for every row of the tile
          if (MODE == MODE_FWD) {
            off = ((long long)r * p.W + s2) * p.C + ch;
          } else {
            const int sh = (p.stride == 2) ? 1 : 0;
            off = -((long long)(r >> sh) * p.Wo + (s2 >> sh)) * p.Kc + ch;
          }
          const unsigned long long bit = k_ok ? (1ull << (tap & 63u)) : 0ull;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t row = (tid >> 3) + 16 * i;
            const bool ok = (rowmask[i] & bit) != 0ull;
            // the dummy source of a zero-filled chunk must be a mapped address: rowbase may point before the tensor
            cp_async16(sa + swz(row, j), ok ? (rowbase[i] + off) : (MODE == MODE_FWD ? px : pdy), ok);
          }
        } else {  // WGRAD A^T: rows are positions, BM/EPC chunks (128 co) per row
          constexpr int CPRA = BM / EPC, RPPA = PRODUCERS / CPRA;
          const uint32_t c = tid % CPRA;
          const uint32_t co = m0 + c * EPC;
          const bool co_ok = co < p.gemm_m;
#pragma unroll
          for (int i = 0; i < BK / RPPA; ++i) {
            const uint32_t row = (tid / CPRA) + RPPA * i;
            const uint32_t pos = kb * BK + row;
            const bool ok = co_ok && pos < p.gemm_k;
            const T* src = ok ? pdy + ((size_t)pos * p.Kc + co) : pdy;
            cp_async16(sa + (c >> 3) * (BK * 128) + swz_mn<T>(row, c & 7), src, ok);
          }
        }

        // ------------------------------- B operand -------------------------------
        if (p.b_tma) {
          // plain 2-D operand: fetched by TMA below
        } else if (MODE == MODE_FWD) {          // weight rows, K-major: BN rows x 8 chunks
          const uint32_t j = tid & 7;
          const uint32_t k = kb * BK + j * EPC;
          const bool k_ok = k < p.gemm_k;
#pragma unroll
          for (int i = 0; i < BN / 16; ++i) {
            const uint32_t row = (tid >> 3) + 16 * i;
            const uint32_t nn = n0 + row;
            const bool ok = k_ok && nn < p.gemm_n;
            const T* src = ok ? pw + ((size_t)nn * p.gemm_k + k) : pw;
            cp_async16(sb + swz(row, j), src, ok);
          }
        } else if (MODE == MODE_DGRAD) {  // B^T rows k=(r,s,co) -> weight row (co,r,s); BN/EPC chunks of ci
          constexpr int CPR = BN / EPC;
          constexpr int RPP = PRODUCERS / CPR;   // rows per pass
          const uint32_t c = tid % CPR;
          const uint32_t ci = n0 + c * EPC;
          const bool ci_ok = ci < p.gemm_n;
          if (p.cpt != 0 || (p.Kc % BK) == 0) {
            // Cout multiple of BK: the BK rows of this k-block share one tap and walk consecutive co
            const uint32_t k0 = kb * BK;
            const uint32_t tap = fdiv(k0, p.div_kc);
            const uint32_t co0 = k0 - tap * p.Kc;
            const size_t row_stride = (size_t)(p.R * p.S) * p.C;
            const T* base = pw + ((size_t)co0 * (p.R * p.S) + tap) * p.C + ci;
            const bool blk_ok = ci_ok && k0 < p.gemm_k;
#pragma unroll
            for (int i = 0; i < BK / RPP; ++i) {
              const uint32_t row = (tid / CPR) + RPP * i;
              cp_async16(sb + (c >> 3) * (BK * 128) + swz_mn<T>(row, c & 7), blk_ok ? base + row * row_stride : pw, blk_ok);
            }
          } else {
#pragma unroll 1
            for (int i = 0; i < BK / RPP; ++i) {
              const uint32_t row = (tid / CPR) + RPP * i;
              const uint32_t k = kb * BK + row;
              const bool ok = ci_ok && k < p.gemm_k;
              const uint32_t tap = fdiv(k, p.div_kc);
              const uint32_t co = k - tap * p.Kc;
              const T* src = ok ? pw + (((size_t)co * (p.R * p.S) + tap) * p.C + ci) : pw;
              cp_async16(sb + (c >> 3) * (BK * 128) + swz_mn<T>(row, c & 7), src, ok);
            }
          }
        } else {  // WGRAD B^T: rows are positions, chunk column fixed (r,s,ci)
          constexpr int CPR = BN / EPC;
          constexpr int RPP = PRODUCERS / CPR;
          const uint32_t c = tid % CPR;
#pragma unroll
          for (int i = 0; i < BK / RPP; ++i) {
            const uint32_t row = (tid / CPR) + RPP * i;
            const uint32_t pos = kb * BK + row;
            bool ok = wb_ok && pos < p.gemm_k;
            const T* src = px;
            if (ok) {
              const uint32_t t = fdiv(pos, p.div_wo), wo = pos - t * p.Wo;
              const uint32_t n = fdiv(t, p.div_ho), ho = t - n * p.Ho;
              const int h = (int)ho * p.stride - p.pad + wb_r, w = (int)wo * p.stride - p.pad + wb_s;
              ok = h >= 0 && h < p.H && w >= 0 && w < p.W;
              if (ok) src = px + ((((size_t)n * p.H + h) * p.W + w) * p.C + wb_c);
            }
            cp_async16(sb + (c >> 3) * (BK * 128) + swz_mn<T>(row, c & 7), src, ok);
          }
        }
        if ((p.a_tma | p.b_tma) && tid == 0) {
          // ---- TMA: one thread arms the stage barrier with the byte count and issues the tile loads; the
          // hardware swizzles (128B) on the way in and zero-fills rows/columns beyond the tensor
          const uint32_t bytes = (p.a_tma ? (uint32_t)C::A_BYTES : 0u) + (p.b_tma ? (uint32_t)C::B_BYTES : 0u);
          mbar_arrive_expect_tx(&full_bar[stage], bytes);
          if (p.a_tma) {
            if (MODE == MODE_WGRAD) {         // A^T tile: BK positions x 128 co as 128-byte-wide groups (2 bf16 / 4 fp32)
#pragma unroll
              for (int gidx = 0; gidx < BM / BK; ++gidx)
                tma_load_2d(sa + gidx * (BK * 128), &tmA, (int)m0 + gidx * BK, (int)(kb * BK), &full_bar[stage]);
            } else {                          // [M][K] matrix (1x1 conv, stride 1 / linear)
              tma_load_2d(sa, &tmA, (int)(kb * BK), (int)m0, &full_bar[stage]);
            }
          }
          if (p.b_tma) tma_load_2d(sb, &tmB, (int)(kb * BK), (int)n0, &full_bar[stage]);   // weight rows [Cout][K]
        }
        if (!(p.a_tma & p.b_tma)) {
          cp_async_commit();
          if (it >= LAG) {
            cp_async_wait<LAG>();          // group (it - LAG) has landed
            fence_proxy_async();           // generic-proxy writes -> visible to the async proxy (tcgen05)
            mbar_arrive(&full_bar[(it - LAG) % STAGES]);
          }
        }
      }
      if (!(p.a_tma & p.b_tma)) {
        // drain the last min(LAG, nkb) groups
        cp_async_wait<0>();
        fence_proxy_async();
        const uint32_t first = nkb > LAG ? nkb - LAG : 0;
        for (uint32_t it = first; it < nkb; ++it) mbar_arrive(&full_bar[it % STAGES]);
      }
      STAMP(2, 0);     // all loads issued and landed
      mbar_wait(tmem_full, 0);
      tc_fence_after();
      STAMP(3, 0);     // accumulator complete
    }

    // =============================== EPILOGUE (TMEM -> registers -> global) ===============================
    const uint32_t row = warp * 32 + lane;          // accumulator row == TMEM lane
    const uint32_t m = m0 + row;
    if (ZC == 1) {
      if (nkb > 0 || MODE != MODE_WGRAD) {
#pragma unroll 1
        for (int cb = 0; cb < BN / 32; ++cb) {
          float f[32];
          if (nkb > 0) {
            uint32_t v[32];
            tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + cb * 32, v);
#pragma unroll
            for (int q = 0; q < 32; ++q) f[q] = __uint_as_float(v[q]);
          } else {
#pragma unroll
            for (int q = 0; q < 32; ++q) f[q] = 0.f;
          }
          if (MODE == MODE_FWD && p.bn_sums != nullptr) bn_stats_accumulate<32>(p, n0 + cb * 32, f, lane);
          epilogue_store<T, MODE, 32>(p, m, n0 + cb * 32, f);
        }
      }
    } else {
      // ---- cluster split-K: the ZC CTAs of a cluster hold partial sums of the same tile.  Stage the
      // partial in shared memory (column-major: conflict-free, one row per thread) ...
      // (Measured alternative, rejected: row-major staging with 16-byte stores and 16-byte ld.shared::cluster reads, rows
      // BN + 4 floats apart -- a quarter of the load instructions, but every lane of a remote load then touches its own
      // 128-byte line: the join went from 2.4 to 5.2 us per kernel and the resnet18 step from 1.27 to 1.47 ms.  Remote
      // shared memory wants the lanes of one load on consecutive words, like global memory.)
      float* stage = reinterpret_cast<float*>(smem);        // the operand ring is idle now: all MMAs retired
#pragma unroll 1
      for (int cb = 0; cb < BN / 32; ++cb) {
        uint32_t v[32];
        if (nkb > 0) {
          tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + cb * 32, v);
        } else {
#pragma unroll
          for (int q = 0; q < 32; ++q) v[q] = 0u;
        }
#pragma unroll
        for (int q = 0; q < 32; ++q) stage[(cb * 32 + q) * BM + row] = __uint_as_float(v[q]);
      }
    }
  } else if (lane == 0 && nkb > 0) {
    // =============================== MMA ISSUER (one thread) ===============================
    constexpr uint32_t idesc = make_idesc(BM, BN, C::A_MN ? 1 : 0, C::B_MN ? 1 : 0, El<T>::FMT);
    const uint32_t smem_base = smem_u32(smem);
    for (uint32_t it = 0; it < nkb; ++it) {
      const uint32_t stage = it % STAGES, ph = (it / STAGES) & 1;
      mbar_wait(&full_bar[stage], ph);
      tc_fence_after();
      const uint32_t sa = smem_base + stage * C::STAGE_BYTES, sb = sa + C::A_BYTES;
#pragma unroll
      for (int k = 0; k < BK / El<T>::UK; ++k) {
        // K-major: +32 B per MMA (16 bf16 / 8 fp32) inside the 128-B swizzle row.  MN-major: the UK k-rows of one MMA are
        // 2048 B (bf16: two 8-row atoms) or 1024 B (fp32: two 4-row atoms) further on; 128-byte MN groups are BK*128 B apart.
        const uint64_t da = C::A_MN ? make_desc_mn<T>(sa + k * (El<T>::UK * 128), BK * 128) : make_desc(sa + k * 32, 16, 1024);
        const uint64_t db = C::B_MN ? make_desc_mn<T>(sb + k * (El<T>::UK * 128), BK * 128) : make_desc(sb + k * 32, 16, 1024);
        umma<T>(tmem_base, da, db, idesc, (it | (uint32_t)k) != 0 ? 1u : 0u);
      }
      umma_commit(&empty_bar[stage]);            // ring slot free once these MMAs retire
      if (it + 1 == nkb) umma_commit(tmem_full);  // accumulator complete
    }
  }
  STAMP(4, 0);         // epilogue phase 1 done (direct store, or partial staged)
  static_assert(ZC == 1 || (size_t)BM * BN * 4 <= (size_t)ST * C::STAGE_BYTES, "split-K staging must fit the operand ring");
  if (ZC > 1) {
    // ... then every CTA reduces its BN/ZC column slice over all peers through distributed shared memory
    // (ld.shared::cluster) and runs the epilogue for that slice.  No atomics, no global scratch.
    cluster_sync_all();
    STAMP(5, 0);
    if (warp < 4) {
      constexpr int SLICE = BN / ZC;
      const uint32_t rank = cluster_ctarank();
      const uint32_t row = warp * 32 + lane;
      const uint32_t m = m0 + row;
      const uint32_t stage_u32 = smem_u32(smem);
      float f[SLICE];
#pragma unroll
      for (int q = 0; q < SLICE; ++q) f[q] = 0.f;
#pragma unroll 2
      for (uint32_t z = 0; z < (uint32_t)ZC; ++z) {
        const uint32_t base = dsmem_addr(stage_u32 + ((rank * SLICE) * BM + row) * 4u, z);
#pragma unroll
        for (int q = 0; q < SLICE; ++q) f[q] += ld_dsmem_f32(base + (uint32_t)q * BM * 4u);
      }
      STAMP(8, 0);     // peers' partial sums read
      if (MODE == MODE_FWD && p.bn_sums != nullptr) bn_stats_accumulate<SLICE>(p, n0 + rank * SLICE, f, lane);
      STAMP(9, 0);
      if (MODE != MODE_WGRAD || p.nvk > 0) epilogue_store<T, MODE, SLICE>(p, m, n0 + rank * SLICE, f);
    }
    STAMP(6, 0);
    cluster_sync_relaxed();  // nobody leaves while a peer may still read its shared memory (no memory ordering needed)
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    __syncwarp();
    tmem_dealloc(tmem_base, BN);
  }
  STAMP(7, 0);
#undef STAMP
}

// ------------------------------------------------------------------------------------------------
// tiles per SM from which the 3-stage / 2-CTAs-per-SM variant is used (B200DDP_TWO_CTA_TILES; 0 = never)
static const uint32_t g_two_cta_tiles = []() { const char* e = getenv("B200DDP_TWO_CTA_TILES"); const int v = e ? atoi(e) : 2; return (uint32_t)(v < 1 ? 1000000 : v); }();

template <typename T, int MODE, int BN, int ZC, int ST = 6>
int launch_zc(const ConvParams& p, const CUtensorMap& tmA, const CUtensorMap& tmB, dim3 grid, cudaStream_t s) {
  using C = Cfg<T, MODE, BN, ST>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(conv_tc_kernel<T, MODE, BN, ZC, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = dim3(THREADS, 1, 1);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (ZC > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 1;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = ZC;
    ++na;
  }
  if (b200_pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_tc_kernel<T, MODE, BN, ZC, ST>, p, tmA, tmB));
  return 0;
}

template <typename T, int MODE, int BN>
int launch(const ConvParams& p, const CUtensorMap& tmA, const CUtensorMap& tmB, dim3 grid, uint32_t zc, cudaStream_t s) {
  switch (zc) {
    case 1:
      // many tiles per SM: two smaller CTAs per SM overlap each other's load latency and epilogue
      if ((size_t)grid.x * grid.y >= (size_t)g_two_cta_tiles * b200_num_sms()) return launch_zc<T, MODE, BN, 1, 3>(p, tmA, tmB, grid, s);
      return launch_zc<T, MODE, BN, 1>(p, tmA, tmB, grid, s);
    case 2: return launch_zc<T, MODE, BN, 2>(p, tmA, tmB, grid, s);
    case 4: return launch_zc<T, MODE, BN, 4>(p, tmA, tmB, grid, s);
    case 8: return launch_zc<T, MODE, BN, 8>(p, tmA, tmB, grid, s);
    default: b200_set_error("cluster split %u is not 1, 2, 4 or 8", zc); return B200_ERR_ARG;
  }
}

// ---- TMA descriptors (cuTensorMapEncodeTiled through the runtime's driver entry point) -------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
bool g_use_tma = true;
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* f = nullptr;
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &st) == cudaSuccess &&
        st == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
    else
      cudaGetLastError();
    const char* e = getenv("B200DDP_TMA");
    if (e && (e[0] == '0')) g_use_tma = false;
  }
  return g_use_tma ? fn : nullptr;
}
// row-major matrix [outer][inner] of T -> boxes of box_outer rows x 128 bytes (64 bf16 / 32 fp32), 128B swizzle.
// mn_major: the box is an MN-major operand tile (rows are k): 32-bit elements then need the 32-byte-atom swizzle.
template <typename T>
bool make_tmap_2d(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint32_t box_outer, bool mn_major = false) {
  EncodeTiledFn fn = encode_tiled_fn();
  constexpr uint64_t ES = sizeof(T);
  if (fn == nullptr || (reinterpret_cast<uintptr_t>(base) & 15) != 0 || (inner * ES) % 16 != 0) return false;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {inner * ES};
  cuuint32_t box[2] = {(cuuint32_t)(128 / ES), box_outer};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapSwizzle sw = (mn_major && ES == 4) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  return fn(m, ES == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims,
            strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int fill_common(ConvParams& p, const b200_conv_geom* g) {
  B200_REQUIRE(g != nullptr, "null geometry");
  B200_REQUIRE(g->C % 8 == 0 && g->K % 8 == 0, "channel counts must be multiples of 8 (16-byte gathers)");
  B200_REQUIRE(g->stride >= 1 && g->R >= 1 && g->S >= 1 && g->N >= 1, "bad geometry");
  memset(&p, 0, sizeof(p));
  p.prof = g_prof;
  p.N = g->N; p.H = g->H; p.W = g->W; p.C = g->C;
  p.Kc = g->K; p.R = g->R; p.S = g->S; p.stride = g->stride; p.pad = g->pad;
  p.Ho = (g->H + 2 * g->pad - g->R) / g->stride + 1;
  p.Wo = (g->W + 2 * g->pad - g->S) / g->stride + 1;
  B200_REQUIRE(p.Ho >= 1 && p.Wo >= 1, "empty output");
  B200_REQUIRE(g->R * g->S <= 64, "filters larger than 64 taps are not supported");
  B200_REQUIRE((size_t)g->N * g->H * g->W < (1ull << 31) && (size_t)g->N * p.Ho * p.Wo < (1ull << 31), "tensor too large for 32-bit tile indices");
  p.div_c = make_fastdiv(p.C); p.div_s = make_fastdiv(p.S); p.div_kc = make_fastdiv(p.Kc);
  p.div_wo = make_fastdiv(p.Wo); p.div_ho = make_fastdiv(p.Ho);
  p.div_w = make_fastdiv(p.W); p.div_h = make_fastdiv(p.H);
  p.div_cpt = make_fastdiv(1);
  // ---- filter taps that can touch a real pixel.  The relation "input row h pairs with output row ho
  // through tap r" (h = ho*stride - pad + r) is the same for fwd, dgrad and wgrad.
  unsigned long long mask = 0;
  if (p.R * p.S <= 64) {
    for (int r = 0; r < p.R; ++r) {
      bool r_ok = false;
      for (int ho = 0; ho < p.Ho && !r_ok; ++ho) { const int h = ho * p.stride - p.pad + r; r_ok = h >= 0 && h < p.H; }
      for (int s = 0; s < p.S; ++s) {
        bool s_ok = false;
        for (int wo = 0; wo < p.Wo && !s_ok; ++wo) { const int w = wo * p.stride - p.pad + s; s_ok = w >= 0 && w < p.W; }
        if (r_ok && s_ok) mask |= 1ull << (r * p.S + s);
      }
    }
  } else {
    mask = ~0ull;
  }
  p.tapmask = mask;
  int n = 0;
  for (int t = 0; t < p.R * p.S && t < 64; ++t)
    if ((mask >> t) & 1ull) p.tap_list[n++] = (uint8_t)t;
  return 0;
}

// K-loop plan for FWD/DGRAD: `cdim` = channels per tap along K, BK = K elements per k-block
void plan_k(ConvParams& p, uint32_t cdim, uint32_t BK) {
  const uint32_t kblocks = (p.gemm_k + BK - 1) / BK;
  int ntaps_valid = 0;
  for (int t = 0; t < 64; ++t) ntaps_valid += (int)((p.tapmask >> t) & 1ull);
  if (p.R * p.S <= 64 && cdim % BK == 0 && ntaps_valid < p.R * p.S && ntaps_valid > 0) {
    p.cpt = cdim / BK;
    p.div_cpt = make_fastdiv(p.cpt);
    p.nvk = (uint32_t)ntaps_valid * p.cpt;
  } else {
    p.cpt = 0;
    p.nvk = kblocks;
  }
  p.kb_per_split = p.nvk;
}


// K split inside a thread-block cluster (DSMEM reduction): largest power of two <= min(8, want)
// Clusters of 4 pack the 16..20-SM GPCs of a B200 almost fully (36 x 4 = 144 of 148 SMs); clusters of 8
// leave SMs stranded and measured ~2x slower on the resnet18 layer shapes.  B200DDP_CLUSTER_MAX overrides.
uint32_t cluster_cap() {
  static uint32_t cap = 0;
  if (cap == 0) {
    const char* e = getenv("B200DDP_CLUSTER_MAX");
    const int v = e ? atoi(e) : 4;
    cap = (v >= 8) ? 8u : (v >= 4) ? 4u : (v >= 2) ? 2u : 1u;
  }
  return cap;
}
// ... but with only a handful of tiles there is nothing to strand: up to B200DDP_CLUSTER8_TILES tiles (default 8) split 8 ways.
// Measured on the fp32 resnet18 step, one B200 (profiles/r2/conv_tile_tuning.md): batch 32, 0.700 -> 0.658 ms; batch 256 unchanged.
uint32_t cluster8_tiles() {
  static int thr = -1;
  if (thr < 0) {
    const char* e = getenv("B200DDP_CLUSTER8_TILES");
    thr = e ? atoi(e) : 8;
    if (thr < 0) thr = 0;
  }
  return (uint32_t)thr;
}
uint32_t cluster_split(uint32_t want, uint32_t tiles = 0xffffffffu) {
  const uint32_t cap = tiles <= cluster8_tiles() ? 8u : cluster_cap();
  uint32_t zc = 1;
  while (zc * 2 <= want && zc * 2 <= cap) zc *= 2;
  return zc;
}

// FWD/DGRAD with few 128-wide tiles: 64-wide tiles instead -- twice the CTAs streaming operands, half the columns per CTA to
// stage and to pull from the cluster peers in the split-K join.  `tiles128` = tile count with BN = 128; the threshold comes
// from B200DDP_NARROW_TILES (default 18; 0 = only when Cout <= 64).  Measured on the fp32 resnet18 step, one B200: batch 32,
// 0.766 -> 0.700 ms; batch 256, 1.273 -> 1.262 ms; thresholds of 40 and 150 change nothing further.
bool prefer_narrow(uint32_t tiles128) {
  static int thr = -1;
  if (thr < 0) {
    const char* e = getenv("B200DDP_NARROW_TILES");
    thr = e ? atoi(e) : 18;
    if (thr < 0) thr = 0;
  }
  return thr > 0 && tiles128 <= (uint32_t)thr;
}

// FWD/DGRAD: split K over a cluster when the tile grid alone cannot fill the machine
uint32_t plan_cluster(ConvParams& p, uint32_t tiles) {
  const uint32_t sms = (uint32_t)b200_num_sms();
  p.kb_per_split = p.nvk;
  if (!g_split_k || tiles * 2 > sms || p.nvk < 4) return 1;
  uint32_t want = sms / tiles;
  if (want > p.nvk / 2) want = p.nvk / 2;            // >= 2 k-blocks per CTA
  const uint32_t zc = cluster_split(want, tiles);
  p.kb_per_split = (p.nvk + zc - 1) / zc;
  return zc;
}

}  // namespace

extern "C" int b200_conv_set_profile_buffer(long long* dev_buf16) {   // 16 slots (8 phase boundaries + 2 inside the split-K join)
  g_prof = dev_buf16;
  return 0;
}

extern "C" int b200_conv_set_split_k(int enabled) {
  g_split_k = enabled != 0;
  return 0;
}

template <typename T>
static int conv_fwd_impl(const void* x, const void* w, const float* bias, void* y, const b200_conv_geom* g, int relu,
                         float* bn_sums, void* stream) {
  ConvParams p;
  int rc = fill_common(p, g);
  if (rc) return rc;
  B200_REQUIRE(x && w && y, "null tensor");
  B200_REQUIRE(bn_sums == nullptr || (bias == nullptr && !relu), "BN statistics are taken of the raw conv output: no bias, no ReLU");
  p.x = x; p.w = w; p.out = y; p.bias = bias; p.relu = relu; p.bn_sums = bn_sums;
  p.gemm_m = (uint32_t)p.N * p.Ho * p.Wo; p.gemm_n = p.Kc; p.gemm_k = (uint32_t)p.R * p.S * p.C;
  plan_k(p, p.C, El<T>::BK);
  const bool narrow = p.gemm_n <= 64 || prefer_narrow(((p.gemm_m + BM - 1) / BM) * ((p.gemm_n + 127) / 128));
  const uint32_t bn = narrow ? 64 : 128;
  dim3 grid((p.gemm_m + BM - 1) / BM, (p.gemm_n + bn - 1) / bn, 1);
  const uint32_t zc = plan_cluster(p, grid.x * grid.y);
  grid.z = zc;
  CUtensorMap tmA, tmB;
  memset(&tmA, 0, sizeof(tmA));
  memset(&tmB, 0, sizeof(tmB));
  p.b_tma = make_tmap_2d<T>(&tmB, w, p.gemm_k, p.gemm_n, bn) ? 1u : 0u;                       // weights [Cout][R*S*C]
  const bool plain = p.R == 1 && p.S == 1 && p.stride == 1 && p.pad == 0;
  p.a_tma = (plain && make_tmap_2d<T>(&tmA, x, p.gemm_k, p.gemm_m, BM)) ? 1u : 0u;             // x as [M][C]
  return narrow ? launch<T, MODE_FWD, 64>(p, tmA, tmB, grid, zc, (cudaStream_t)stream)
                : launch<T, MODE_FWD, 128>(p, tmA, tmB, grid, zc, (cudaStream_t)stream);
}

extern "C" int b200_conv_fwd(const void* x, const void* w, const float* bias, void* y, const b200_conv_geom* g,
                             int relu, void* stream) {
  return conv_fwd_impl<__nv_bfloat16>(x, w, bias, y, g, relu, nullptr, stream);
}
extern "C" int b200_conv_fwd_f32(const void* x, const void* w, const float* bias, void* y, const b200_conv_geom* g,
                                 int relu, void* stream) {
  return conv_fwd_impl<float>(x, w, bias, y, g, relu, nullptr, stream);
}

// Same convolution (no bias, no ReLU), and the epilogue also adds the per-channel sum and sum of squares of the fp32
// output into bn_sums[0..Kc) / bn_sums[Kc..2Kc): the statistics pass of a following training-mode BatchNorm for free.
// bn_sums must be zero on entry (b200_bn_finalize leaves it zero again).
extern "C" int b200_conv_fwd_stats(const void* x, const void* w, void* y, const b200_conv_geom* g, float* bn_sums,
                                   void* stream) {
  B200_REQUIRE(bn_sums != nullptr, "null bn_sums");
  return conv_fwd_impl<__nv_bfloat16>(x, w, nullptr, y, g, 0, bn_sums, stream);
}
extern "C" int b200_conv_fwd_stats_f32(const void* x, const void* w, void* y, const b200_conv_geom* g, float* bn_sums,
                                       void* stream) {
  B200_REQUIRE(bn_sums != nullptr, "null bn_sums");
  return conv_fwd_impl<float>(x, w, nullptr, y, g, 0, bn_sums, stream);
}

template <typename T>
static int conv_dgrad_impl(const void* dy, const void* w, void* dx, const b200_conv_geom* g, const void* residual,
                           void* stream) {
  ConvParams p;
  int rc = fill_common(p, g);
  if (rc) return rc;
  B200_REQUIRE(dy && w && dx, "null tensor");
  B200_REQUIRE(p.stride <= 2, "dgrad supports stride 1 and 2");
  p.dy = dy; p.w = w; p.out = dx; p.residual = residual;
  p.gemm_m = (uint32_t)p.N * p.H * p.W; p.gemm_n = p.C; p.gemm_k = (uint32_t)p.R * p.S * p.Kc;
  plan_k(p, p.Kc, El<T>::BK);
  const bool narrow = p.gemm_n <= 64 || prefer_narrow(((p.gemm_m + BM - 1) / BM) * ((p.gemm_n + 127) / 128));
  const uint32_t bn = narrow ? 64 : 128;
  dim3 grid((p.gemm_m + BM - 1) / BM, (p.gemm_n + bn - 1) / bn, 1);
  const uint32_t zc = plan_cluster(p, grid.x * grid.y);
  grid.z = zc;
  CUtensorMap tmA, tmB;
  memset(&tmA, 0, sizeof(tmA));
  memset(&tmB, 0, sizeof(tmB));
  const bool plain = p.R == 1 && p.S == 1 && p.stride == 1 && p.pad == 0;
  p.a_tma = (plain && make_tmap_2d<T>(&tmA, dy, p.gemm_k, p.gemm_m, BM)) ? 1u : 0u;            // dy as [M][Cout]
  return narrow ? launch<T, MODE_DGRAD, 64>(p, tmA, tmB, grid, zc, (cudaStream_t)stream)
                : launch<T, MODE_DGRAD, 128>(p, tmA, tmB, grid, zc, (cudaStream_t)stream);
}

extern "C" int b200_conv_dgrad(const void* dy, const void* w, void* dx, const b200_conv_geom* g, void* stream) {
  return conv_dgrad_impl<__nv_bfloat16>(dy, w, dx, g, nullptr, stream);
}
extern "C" int b200_conv_dgrad_f32(const void* dy, const void* w, void* dx, const b200_conv_geom* g, void* stream) {
  return conv_dgrad_impl<float>(dy, w, dx, g, nullptr, stream);
}

// dx = dgrad(dy, w) + residual (same dtype and shape as dx): when the conv's input also feeds another branch (a ResNet skip
// connection), that branch's gradient is added in the epilogue instead of by a separate elementwise kernel.
extern "C" int b200_conv_dgrad_add(const void* dy, const void* w, void* dx, const b200_conv_geom* g, const void* residual,
                                   void* stream) {
  B200_REQUIRE(residual != nullptr, "null residual");
  return conv_dgrad_impl<__nv_bfloat16>(dy, w, dx, g, residual, stream);
}
extern "C" int b200_conv_dgrad_add_f32(const void* dy, const void* w, void* dx, const b200_conv_geom* g, const void* residual,
                                       void* stream) {
  B200_REQUIRE(residual != nullptr, "null residual");
  return conv_dgrad_impl<float>(dy, w, dx, g, residual, stream);
}

template <typename T>
static int conv_wgrad_impl(const void* x, const void* dy, float* dw, const b200_conv_geom* g, int max_splits, void* stream) {
  constexpr uint32_t BK = El<T>::BK;
  ConvParams p;
  int rc = fill_common(p, g);
  if (rc) return rc;
  B200_REQUIRE(x && dy && dw, "null tensor");
  p.x = x; p.dy = dy; p.out = dw;
  p.gemm_m = p.Kc; p.gemm_n = (uint32_t)p.R * p.S * p.C; p.gemm_k = (uint32_t)p.N * p.Ho * p.Wo;
  const uint32_t kblocks = (p.gemm_k + BK - 1) / BK;
  p.nvk = kblocks;
  int ntaps_valid = 0;
  for (int t = 0; t < 64; ++t) ntaps_valid += (int)((p.tapmask >> t) & 1ull);
  p.cpt = (p.R * p.S <= 64 && ntaps_valid < p.R * p.S) ? 1u : 0u;   // non-zero enables the N-tile skip; K is never remapped
  const bool narrow = p.gemm_n <= 64;
  const uint32_t bn = narrow ? 64 : 128;
  uint32_t n_tiles = (p.gemm_n + bn - 1) / bn;
  uint32_t live_n_tiles = n_tiles;
  if (p.cpt) {   // only the live tiles count when sizing the split
    live_n_tiles = 0;
    for (uint32_t t = 0; t < n_tiles; ++t) {
      const uint32_t lo = (t * bn) / p.C, hi = (std::min((t + 1) * bn, p.gemm_n) - 1) / p.C;
      bool any = false;
      for (uint32_t q = lo; q <= hi && q < 64; ++q) any = any || ((p.tapmask >> q) & 1ull);
      live_n_tiles += any ? 1 : 0;
    }
    if (live_n_tiles == 0) return 0;
  }
  const uint32_t tiles = ((p.gemm_m + BM - 1) / BM) * live_n_tiles;
  // Split the position reduction over about half the SMs (B200DDP_WGRAD_SM_PCT, default 60), >= 2 k-blocks per CTA.
  // The kernel's cost is per-CTA latency, not math: clusters of 4 at 1 CTA/SM only fit ~32 at a time (GPCs of 18-20
  // SMs), so a grid of 148+ CTAs runs in two or three waves (23-31 us on the resnet18 shapes) where 74 CTAs take
  // one (~11 us) and leave SMs to the dgrad / BN kernels that run concurrently on the main stream.  The `zc` CTAs of a
  // cluster reduce through DSMEM so only one red.add per element leaves the cluster; `zo` clusters meet in the atomics.
  static const uint32_t sm_pct = []() { const char* e = getenv("B200DDP_WGRAD_SM_PCT"); const int v = e ? atoi(e) : 60; return (uint32_t)(v < 1 ? 1 : v); }();
  // max_splits < 0: the caller knows nothing runs next to this wgrad (the first layer's, at the tail of backward) and
  // gives the SM share to use as -percent
  const uint32_t pct = max_splits < 0 ? (uint32_t)(-max_splits) : sm_pct;
  uint32_t splits = ((uint32_t)b200_num_sms() * pct / 100u) / tiles;
  if (max_splits > 0 && splits > (uint32_t)max_splits) splits = (uint32_t)max_splits;
  const uint32_t max_by_k = (kblocks + 1) / 2;
  if (splits > max_by_k) splits = max_by_k;
  if (splits < 1) splits = 1;
  const uint32_t zc = cluster_split(splits);
  const uint32_t zo = (splits + zc - 1) / zc;
  p.kb_per_split = (kblocks + zo * zc - 1) / (zo * zc);
  p.div_cpt = make_fastdiv(1);
  dim3 grid((p.gemm_m + BM - 1) / BM, n_tiles, zo * zc);
  CUtensorMap tmA, tmB;
  memset(&tmA, 0, sizeof(tmA));
  memset(&tmB, 0, sizeof(tmB));
  p.a_tma = make_tmap_2d<T>(&tmA, dy, p.gemm_m, p.gemm_k, BK, /*mn_major=*/true) ? 1u : 0u;   // dy as [positions][Cout], BK x 128-byte boxes
  return narrow ? launch<T, MODE_WGRAD, 64>(p, tmA, tmB, grid, zc, (cudaStream_t)stream)
                : launch<T, MODE_WGRAD, 128>(p, tmA, tmB, grid, zc, (cudaStream_t)stream);
}

extern "C" int b200_conv_wgrad(const void* x, const void* dy, float* dw, const b200_conv_geom* g, int max_splits,
                               void* stream) {
  return conv_wgrad_impl<__nv_bfloat16>(x, dy, dw, g, max_splits, stream);
}
extern "C" int b200_conv_wgrad_f32(const void* x, const void* dy, float* dw, const b200_conv_geom* g, int max_splits,
                                   void* stream) {
  return conv_wgrad_impl<float>(x, dy, dw, g, max_splits, stream);
}
