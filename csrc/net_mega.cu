// Workshop `Net` megakernel (subsuming W7 conv_smallc and W10 linear_small for this model):
// forward + loss + backward of the LeNet-style CIFAR-10 CNN the workshop trains
// (notebooks/code/cifar10-distributed-native-cpu.py:22-39,146-156; cifar10-distributed-smddp-gpu.py:34-51) in TWO
// launches instead of ~45, all in fp32 on the CUDA cores — the model is 62,006 parameters and 1.3 MFLOP per
// image, so a training step is pure launch latency for any layer-per-kernel design.
//
//   net_fwd_bwd_kernel   persistent CTAs, one image at a time entirely in shared memory:
//                        conv5x5+bias+ReLU+pool -> conv5x5+bias+ReLU+pool -> fc400-120-84-10 -> softmax-CE ->
//                        backward through every layer.  Weight gradients of conv1, conv2, fc2, fc3 and all biases
//                        are accumulated in shared memory over the CTA's images and flushed once with red.add
//                        into the flat gradient arena; fc1's activations/gradients go to a small global scratch.
//   net_fc1_wgrad_kernel dW_fc1[120][400] += dh1^T p2 over the batch (the one gradient too large for smem).
//
// Weight layouts: conv weights are either torch's [co][ci][r][s] or the channels-last [co][r][s][ci] of the
// DDP flat arena (`krsc`); gradients are written in the same layout.  The arena is zeroed by the optimizer, so
// everything here accumulates.  Mean loss: last-CTA ticket (scratch counter, self-resetting), deterministic.
#include "common.cuh"

namespace {

#ifndef NET_T
#define NET_T 512
#endif
constexpr int T = NET_T;
constexpr int C1 = 6, C2 = 16, KS = 5;
constexpr int H0 = 32, P1 = 14, P2 = 5;      // 32 -conv5-> 28 -pool-> 14 -conv5-> 10 -pool-> 5
constexpr int F0 = 400, F1 = 120, F2 = 84, F3 = 10;

struct NetParams {
  const float* x;            // [B][3][32][32] fp32 NCHW (normalised)
  const int64_t* target;     // [B]
  const uint8_t* xu8;        // alternative input: [B][32][32][3] uint8 NHWC, augmented + normalised in the kernel
  const int32_t* offsets;    // [B][3] (dy, dx, flip) of RandomCrop(32, padding=4) + RandomHorizontalFlip, or NULL
  const float *mean, *inv_std;   // [3] each (device)
  const float *w1, *b1, *w2, *b2, *fw1, *fb1, *fw2, *fb2, *fw3, *fb3;
  float *gw1, *gb1, *gw2, *gb2, *gfw1, *gfb1, *gfw2, *gfb2, *gfw3, *gfb3;
  float* scratch;            // per image: dh1[120], p2[400]
  float* row_loss;           // [B]
  float* loss_mean;          // scalar
  float* logits;             // optional [B][10]
  unsigned* counter;         // ticket
  long long* prof;           // optional phase stamps (clock64) of CTA 0, first image
  int B, krsc, want_grad;
  float grad_scale;          // 1/B
};

__device__ __forceinline__ int widx(int krsc, int co, int ci, int r, int s, int cin) {
  return krsc ? ((co * KS + r) * KS + s) * cin + ci : ((co * cin + ci) * KS + r) * KS + s;
}

// ---- shared-memory layouts (chosen so that the 32 lanes of a warp hit different banks in the hot loops)
constexpr int IN_RS = 37;                    // input row stride: the 25 taps (r*37+q) of one channel fall in 25 banks
constexpr int IN_PS = H0 * IN_RS + 25;       // plane stride = 25 (mod 32): the next channel's taps take the banks left over
constexpr int DP = 18;                       // conv2's output gradient, zero-padded by 4 on every side (10 + 8)
constexpr int NPOS1 = P1 * P1, NPOS2 = P2 * P2;
constexpr int P1W_PS = P1 * IN_RS + 19;      // 537 = 25 (mod 32)
constexpr int W1G = 5;                       // conv1's weight gradient is summed in 5 groups of pooled positions
constexpr int G2S = 26;                      // g2 row stride in float2 (25 padded so that every row is 16-byte aligned)
// per-image slice of the global scratch: what the batched fc weight-gradient kernel needs
constexpr int SC_DH1 = 0, SC_P2 = 120, SC_DH2 = 520, SC_H1 = 604, SC_DLG = 724, SC_H2 = 736, SC_STRIDE = 832;

struct Smem {
  // weights re-laid out for vector loads: w1t[ci][r][q][8 co], w2t[ci][r][q][16 co], w2d[co][r][q][8 ci]
  alignas(16) float w1t[3 * 25 * 8];
  alignas(16) float w2t[C1 * 25 * C2];
  alignas(16) float w2d[C2 * 25 * 8];
  alignas(16) float fw2[F2 * F1];            // small fc weights, resident for all images of the CTA
  alignas(16) float fw3[F3 * F2];
  alignas(16) float da2p[C2 * DP * DP];      // dense dL/d(conv2 out), zero-padded; only 400 entries are ever non-zero
  alignas(16) float2 g1[C1 * NPOS1];         // {masked gradient at a pooled position, offset of its window in `in`}
  alignas(16) float2 g2[C2 * G2S];           // same for conv2 (offset into a p1 plane)
  alignas(16) float p1[C1 * NPOS1];
  alignas(16) float p2[F0];
  alignas(16) float part[NET_T / 32][F0];    // scratch: conv2 half sums / per-warp partials of W1^T dh1 / dgrad quarter sums
  alignas(16) uint8_t raw[2][3 * H0 * H0];   // uint8 input mode: the raw image, prefetched with cp.async
  float in[2][3 * IN_PS];                    // double buffer: the next image is prefetched with cp.async
  float p1w[C1 * P1W_PS];                    // second copy of p1 in the bank-conflict-free geometry of `in` (for dW2)
  float h1[F1], h2[F2], lg[F3], dlg[F3], dh2[F2], dh1[F1];
  float b1[8], b2[C2], fb1[F1], fb2[F2], fb3[F3];
  // gradient accumulators over this CTA's images (torch order); the fc weight gradients are done by the second kernel
  float gw1[W1G][C1 * 3 * 25], gb1[C1], gw2[C2 * C1 * 25], gb2[C2], gfb1[F1], gfb2[F2], gfb3[F3];
  float red[32];
  uint8_t i1[C1 * NPOS1], i2[F0];
  bool last;
};
static_assert(sizeof(Smem) <= 227 * 1024, "Smem exceeds the 227 KB a CTA can opt in to");
static_assert(sizeof(float) * (NET_T / 32) * F0 >= sizeof(float) * 4 * C1 * NPOS1, "part[] too small for the dgrad quarter sums");

// pooled maximum of a 2x2 window given as v[dy*2+dx]; ties go to the first element in scan order (torch's choice)
__device__ __forceinline__ float pool4(float v0, float v1, float v2, float v3, int& bi) {
  float best = v0; bi = 0;
  if (v1 > best) { best = v1; bi = 1; }
  if (v2 > best) { best = v2; bi = 2; }
  if (v3 > best) { best = v3; bi = 3; }
  return best;
}

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}

// image -> smem in the padded geometry (row stride 37, plane stride 1209), asynchronously
__device__ __forceinline__ void prefetch_image(float* dst, const float* __restrict__ src, int tid) {
  for (int e = tid; e < 3 * H0 * H0; e += T) cp_async4(dst + (e >> 10) * IN_PS + ((e >> 5) & 31) * IN_RS + (e & 31), src + e);
}

__device__ __forceinline__ void prefetch_raw(uint8_t* dst, const uint8_t* __restrict__ src, int tid) {
  if (tid < 3 * H0 * H0 / 16) cp_async16(dst + tid * 16, src + tid * 16);
}

__device__ __forceinline__ float dot4(float4 a, float4 b, float acc) {
  acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); acc = fmaf(a.z, b.z, acc); return fmaf(a.w, b.w, acc);
}
__device__ __forceinline__ void axpy4(float4& acc, float4 w, float d) {
  acc.x = fmaf(w.x, d, acc.x); acc.y = fmaf(w.y, d, acc.y); acc.z = fmaf(w.z, d, acc.z); acc.w = fmaf(w.w, d, acc.w);
}

#define NET_STAMP(k) do { if (p.prof && blockIdx.x == 0 && tid == 0 && (k < 2 || img == 0)) p.prof[k] = clock64(); } while (0)

__global__ void __launch_bounds__(T, 1) net_fwd_bwd_kernel(const NetParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = T / 32;
  constexpr int ROWV = F0 / 4;                                // fc1 row length in float4
  int img = 0;
  pdl_trigger();          // the fc weight-gradient kernel may be scheduled as SMs free up; it waits for this grid in pdl_wait()
  pdl_wait();
  NET_STAMP(0);
  if (blockIdx.x < p.B) {
    if (p.xu8) prefetch_raw(s.raw[0], p.xu8 + (size_t)blockIdx.x * 3 * H0 * H0, tid);
    else prefetch_image(s.in[0], p.x + (size_t)blockIdx.x * 3 * H0 * H0, tid);
  }

  // ---- weights -> smem (vector-friendly layouts), accumulators -> 0
  for (int i = tid; i < 3 * 25 * 8; i += T) {
    const int co = i & 7, t = i >> 3, q = t % 5, r = (t / 5) % 5, ci = t / 25;
    s.w1t[i] = co < C1 ? p.w1[widx(p.krsc, co, ci, r, q, 3)] : 0.f;
  }
  for (int i = tid; i < C1 * 25 * C2; i += T) {
    const int co = i & 15, t = i >> 4, q = t % 5, r = (t / 5) % 5, ci = t / 25;
    s.w2t[i] = p.w2[widx(p.krsc, co, ci, r, q, C1)];
  }
  for (int i = tid; i < C2 * 25 * 8; i += T) {
    const int ci = i & 7, t = i >> 3, q = t % 5, r = (t / 5) % 5, co = t / 25;
    s.w2d[i] = ci < C1 ? p.w2[widx(p.krsc, co, ci, r, q, C1)] : 0.f;
  }
  for (int i = tid; i < F2 * F1 / 4; i += T) reinterpret_cast<float4*>(s.fw2)[i] = __ldg(reinterpret_cast<const float4*>(p.fw2) + i);
  for (int i = tid; i < F3 * F2 / 4; i += T) reinterpret_cast<float4*>(s.fw3)[i] = __ldg(reinterpret_cast<const float4*>(p.fw3) + i);
  for (int i = tid; i < C2 * DP * DP / 4; i += T) reinterpret_cast<float4*>(s.da2p)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = tid; i < W1G * C1 * 3 * 25; i += T) (&s.gw1[0][0])[i] = 0.f;
  for (int i = tid; i < C2 * C1 * 25; i += T) s.gw2[i] = 0.f;
  if (tid < 8) s.b1[tid] = tid < C1 ? p.b1[tid] : 0.f;
  if (tid < C1) s.gb1[tid] = 0.f;
  if (tid < C2) { s.b2[tid] = p.b2[tid]; s.gb2[tid] = 0.f; }
  if (tid < F1) { s.fb1[tid] = p.fb1[tid]; s.gfb1[tid] = 0.f; }
  if (tid < F2) { s.fb2[tid] = p.fb2[tid]; s.gfb2[tid] = 0.f; }
  if (tid < F3) { s.fb3[tid] = p.fb3[tid]; s.gfb3[tid] = 0.f; }
  __syncthreads();
  NET_STAMP(1);

  for (img = blockIdx.x; img < p.B; img += gridDim.x) {
    // ---------------------------------------------------------------- forward
    const int buf = ((img - blockIdx.x) / gridDim.x) & 1;
    const float* in = p.xu8 ? s.in[0] : s.in[buf];
    const int tgt = (int)__ldg(p.target + img);
    cp_async_wait_all();
    __syncthreads();
    if (p.xu8) {
      // the reference's transform (RandomCrop(32, 4) + flip + ToTensor + Normalize), same arithmetic as b200_augment
      if (img + gridDim.x < p.B) prefetch_raw(s.raw[buf ^ 1], p.xu8 + (size_t)(img + gridDim.x) * 3 * H0 * H0, tid);
      int dy = 4, dx = 4, flip = 0;
      if (p.offsets) { dy = __ldg(p.offsets + img * 3); dx = __ldg(p.offsets + img * 3 + 1); flip = __ldg(p.offsets + img * 3 + 2); }
      const uint8_t* raw = s.raw[buf];
      for (int e = tid; e < 3 * H0 * H0; e += T) {
        const int c = e >> 10, y = (e >> 5) & 31, x = e & 31;
        const int sh = y + dy - 4, sw = (flip ? (H0 - 1 - x) : x) + dx - 4;
        const bool inside = sh >= 0 && sh < H0 && sw >= 0 && sw < H0;
        const float v = inside ? (float)raw[(sh * H0 + sw) * 3 + c] * (1.f / 255.f) : 0.f;
        s.in[0][c * IN_PS + y * IN_RS + x] = (v - __ldg(p.mean + c)) * __ldg(p.inv_std + c);
      }
      __syncthreads();
    } else if (img + gridDim.x < p.B) {
      prefetch_image(s.in[buf ^ 1], p.x + (size_t)(img + gridDim.x) * 3 * H0 * H0, tid);
    }
    NET_STAMP(2);
    if (tid < NPOS1) {                                        // conv1 + bias + ReLU + 2x2 maxpool: one pooled position,
      const int py = tid / P1, px = tid % P1;                 // all 6 channels, 24 accumulators in registers
      float acc[4][C1];
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int c = 0; c < C1; ++c) acc[k][c] = s.b1[c];
#pragma unroll 1
      for (int ci = 0; ci < 3; ++ci) {
        const float* a = in + ci * IN_PS + (2 * py) * IN_RS + 2 * px;
        float win[6][6];
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
          for (int j = 0; j < 6; ++j) win[i][j] = a[i * IN_RS + j];
#pragma unroll
        for (int r = 0; r < 5; ++r)
#pragma unroll
          for (int q = 0; q < 5; ++q) {
            const float* wp = s.w1t + ((ci * 5 + r) * 5 + q) * 8;
            const float4 wa = *reinterpret_cast<const float4*>(wp);
            const float2 wb = *reinterpret_cast<const float2*>(wp + 4);
            const float w[C1] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float v = win[(k >> 1) + r][(k & 1) + q];
#pragma unroll
              for (int c = 0; c < C1; ++c) acc[k][c] = fmaf(v, w[c], acc[k][c]);
            }
          }
      }
#pragma unroll
      for (int c = 0; c < C1; ++c) {
        int bi;
        const float best = pool4(acc[0][c], acc[1][c], acc[2][c], acc[3][c], bi);
        s.p1[c * NPOS1 + tid] = fmaxf(best, 0.f);
        s.p1w[c * P1W_PS + py * IN_RS + px] = fmaxf(best, 0.f);
        s.i1[c * NPOS1 + tid] = (uint8_t)bi;
      }
    }
    __syncthreads();
    NET_STAMP(3);
    {                                                         // conv2 + bias + ReLU + pool: warp = 4 output channels x half
      float acc[4][4];                                        // of the input channels, lane = pooled position
      const int c0 = (warp & 3) * 4, half = warp >> 2;
      const bool active = warp < 8 && lane < NPOS2;
      if (active) {
        const int py = lane / P2, px = lane % P2;
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[k][c] = 0.f;
#pragma unroll 1
        for (int ci = half * 3; ci < half * 3 + 3; ++ci) {
          const float* a = s.p1 + ci * NPOS1 + (2 * py) * P1 + 2 * px;
          float win[6][6];
#pragma unroll
          for (int i = 0; i < 6; ++i)
#pragma unroll
            for (int j = 0; j < 6; j += 2) {
              const float2 v = *reinterpret_cast<const float2*>(a + i * P1 + j);
              win[i][j] = v.x; win[i][j + 1] = v.y;
            }
#pragma unroll
          for (int r = 0; r < 5; ++r)
#pragma unroll
            for (int q = 0; q < 5; ++q) {
              const float4 w4 = *reinterpret_cast<const float4*>(s.w2t + ((ci * 5 + r) * 5 + q) * C2 + c0);
              const float w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float v = win[(k >> 1) + r][(k & 1) + q];
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[k][c] = fmaf(v, w[c], acc[k][c]);
              }
            }
        }
        if (half) {
          float* dst = &s.part[0][0] + (warp & 3) * 32 + lane;
#pragma unroll
          for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int c = 0; c < 4; ++c) dst[(k * 4 + c) * 128] = acc[k][c];
        }
      }
      __syncthreads();
      if (active && !half) {
        const float* src = &s.part[0][0] + (warp & 3) * 32 + lane;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float v[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = acc[k][c] + src[(k * 4 + c) * 128] + s.b2[c0 + c];
          int bi;
          const float best = pool4(v[0], v[1], v[2], v[3], bi);
          const int o = (c0 + c) * NPOS2 + lane;              // flatten order (c, h, w)
          s.p2[o] = fmaxf(best, 0.f);
          s.i2[o] = (uint8_t)bi;
        }
      }
    }
    __syncthreads();
    NET_STAMP(4);
    {                                                         // fc1 + ReLU: weights streamed from L2, two rows (8 x 16 B per
      const float4* p2v = reinterpret_cast<const float4*>(s.p2);   // lane) in flight
      const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
      float4 xv[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) xv[k] = (lane + 32 * k < ROWV) ? p2v[lane + 32 * k] : zero4;
      for (int o = warp; o < F1; o += 2 * NW) {
        const bool two = o + NW < F1;
        const float4* r0 = reinterpret_cast<const float4*>(p.fw1 + (size_t)o * F0) + lane;
        const float4* r1 = r0 + (two ? NW * ROWV : 0);
        float4 a[4], b[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const bool in_row = lane + 32 * k < ROWV;
          a[k] = in_row ? __ldg(r0 + 32 * k) : zero4;
          b[k] = in_row ? __ldg(r1 + 32 * k) : zero4;
        }
        float v0 = 0.f, v1 = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) { v0 = dot4(a[k], xv[k], v0); v1 = dot4(b[k], xv[k], v1); }
        v0 = warp_sum(v0); v1 = warp_sum(v1);
        if (lane == 0) {
          s.h1[o] = fmaxf(v0 + s.fb1[o], 0.f);
          if (two) s.h1[o + NW] = fmaxf(v1 + s.fb1[o + NW], 0.f);
        }
      }
    }
    __syncthreads();
    NET_STAMP(5);
    for (int o = warp; o < F2; o += 2 * NW) {                 // fc2 + ReLU, two rows per warp at a time
      const bool two = o + NW < F2;
      const float* w0 = s.fw2 + o * F1;
      const float* w1 = w0 + (two ? NW * F1 : 0);
      float v0 = 0.f, v1 = 0.f;
#pragma unroll
      for (int i = lane; i < F1; i += 32) { const float h = s.h1[i]; v0 = fmaf(w0[i], h, v0); v1 = fmaf(w1[i], h, v1); }
      v0 = warp_sum(v0); v1 = warp_sum(v1);
      if (lane == 0) {
        s.h2[o] = fmaxf(v0 + s.fb2[o], 0.f);
        if (two) s.h2[o + NW] = fmaxf(v1 + s.fb2[o + NW], 0.f);
      }
    }
    __syncthreads();
    if (warp < F3) {                                          // fc3, then softmax cross-entropy by warp 0
      float v = 0.f;
      for (int i = lane; i < F2; i += 32) v = fmaf(s.fw3[warp * F2 + i], s.h2[i], v);
      v = warp_sum(v);
      if (lane == 0) s.lg[warp] = v + s.fb3[warp];
    }
    __syncthreads();
    if (warp == 0) {
      const float v = lane < F3 ? s.lg[lane] : -INFINITY;
      const float mx = warp_max(v);
      const float e = lane < F3 ? __expf(v - mx) : 0.f;
      const float se = warp_sum(e);
      const int t = tgt;
      if (lane < F3) {
        s.dlg[lane] = (e / se - (lane == t ? 1.f : 0.f)) * p.grad_scale;
        if (p.logits) p.logits[(size_t)img * F3 + lane] = v;
      }
      if (lane == 0) p.row_loss[img] = (t >= 0 && t < F3) ? (mx + __logf(se) - s.lg[t]) : 0.f;
    }
    __syncthreads();
    NET_STAMP(6);
    if (!p.want_grad) continue;

    // ---------------------------------------------------------------- backward (activations)
    for (int j = tid; j < F2; j += T) {                       // dh2 = W3^T dlg, masked by ReLU(h2)
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < F3; ++k) acc = fmaf(s.fw3[k * F2 + j], s.dlg[k], acc);
      s.dh2[j] = s.h2[j] > 0.f ? acc : 0.f;
    }
    __syncthreads();
    for (int j = tid; j < F1; j += T) {                       // dh1 = W2^T dh2, masked
      float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 6
      for (int k = 0; k < F2; k += 2) {
        acc0 = fmaf(s.fw2[k * F1 + j], s.dh2[k], acc0);
        acc1 = fmaf(s.fw2[(k + 1) * F1 + j], s.dh2[k + 1], acc1);
      }
      s.dh1[j] = s.h1[j] > 0.f ? acc0 + acc1 : 0.f;
    }
    __syncthreads();
    NET_STAMP(7);
    {                                                         // dp2 = W1^T dh1: every warp streams whole rows of W1 (two rows,
      const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);   // 8 x 16 B per lane in flight); per-warp partials meet in smem
      float4 acc[4] = {zero4, zero4, zero4, zero4};
      for (int o = warp; o < F1; o += 2 * NW) {
        const bool two = o + NW < F1;
        const float d0 = s.dh1[o], d1 = two ? s.dh1[o + NW] : 0.f;
        const float4* r0 = reinterpret_cast<const float4*>(p.fw1 + (size_t)o * F0) + lane;
        const float4* r1 = r0 + (two ? NW * ROWV : 0);
        float4 a[4], b[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const bool in_row = lane + 32 * k < ROWV;
          a[k] = in_row ? __ldg(r0 + 32 * k) : zero4;
          b[k] = in_row ? __ldg(r1 + 32 * k) : zero4;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { axpy4(acc[k], a[k], d0); axpy4(acc[k], b[k], d1); }
      }
      float4* dst = reinterpret_cast<float4*>(s.part[warp]);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (lane + 32 * k < ROWV) dst[lane + 32 * k] = acc[k];
    }
    __syncthreads();
    NET_STAMP(8);
    int dpos = 0;                                             // where this thread's conv2 gradient sits in da2p
    if (tid < F0) {                                           // dp2 -> through pool2 + ReLU: one non-zero per 2x2 window
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) acc += s.part[w][tid];
      const int co = tid / NPOS2, pos = tid % NPOS2, sel = s.i2[tid];
      const int Y = 2 * (pos / P2) + (sel >> 1), X = 2 * (pos % P2) + (sel & 1);
      const float g = s.p2[tid] > 0.f ? acc : 0.f;
      s.g2[co * G2S + pos] = make_float2(g, __int_as_float(Y * IN_RS + X));
      dpos = co * DP * DP + (Y + 4) * DP + X + 4;
      s.da2p[dpos] = g;
    }
    __syncthreads();
    NET_STAMP(9);
    float* dp1q = &s.part[0][0];                              // [4 channel groups][6][196] partial sums of dL/dp1
    if (warp < 4) {
      // conv2 dgrad on 4 warps: lane = (row y, half row), 7 x 6 outputs in registers, 4 of conv2's output channels per warp
      if (lane < 2 * P1) {
        const int y = lane >> 1, x0 = (lane & 1) * 7;
        float acc[7][C1];
#pragma unroll
        for (int i = 0; i < 7; ++i)
#pragma unroll
          for (int c = 0; c < C1; ++c) acc[i][c] = 0.f;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          const int co = warp * 4 + c;
#pragma unroll 1
          for (int r = 0; r < 5; ++r) {
            const float* d = s.da2p + co * DP * DP + (y + 4 - r) * DP + x0;
            float row[11];
#pragma unroll
            for (int j = 0; j < 11; ++j) row[j] = d[j];
            const float* w = s.w2d + (co * 25 + r * 5) * 8;
#pragma unroll
            for (int q = 0; q < 5; ++q) {
              const float4 wa = *reinterpret_cast<const float4*>(w + q * 8);
              const float2 wb = *reinterpret_cast<const float2*>(w + q * 8 + 4);
#pragma unroll
              for (int i = 0; i < 7; ++i) {
                const float v = row[i + 4 - q];
                acc[i][0] = fmaf(v, wa.x, acc[i][0]); acc[i][1] = fmaf(v, wa.y, acc[i][1]); acc[i][2] = fmaf(v, wa.z, acc[i][2]);
                acc[i][3] = fmaf(v, wa.w, acc[i][3]); acc[i][4] = fmaf(v, wb.x, acc[i][4]); acc[i][5] = fmaf(v, wb.y, acc[i][5]);
              }
            }
          }
        }
#pragma unroll
        for (int i = 0; i < 7; ++i)
#pragma unroll
          for (int c = 0; c < C1; ++c) dp1q[(warp * C1 + c) * NPOS1 + y * P1 + x0 + i] = acc[i][c];
      }
    } else {
      // meanwhile the other 12 warps: conv2's weight gradient (25 non-zeros per output channel), bias gradients, and
      // this image's fc activations / gradients for the batched fc weight-gradient kernel
      const int t2 = tid - 128;
      constexpr int T2 = T - 128;
      for (int co = warp - 4; co < C2; co += NW - 4) {          // one output channel per warp, lane = (ci, r), 5 taps each:
        if (lane < C1 * 5) {                                  // 7 smem words per 5 FMAs
          const int ci = lane / 5, r = lane % 5;
          const float* a = s.p1w + ci * P1W_PS + r * IN_RS;
          const float2* g = s.g2 + co * G2S;
          float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 5
          for (int i = 0; i < NPOS2; ++i) {
            const float2 gg = g[i];
            const float* ap = a + __float_as_int(gg.y);
#pragma unroll
            for (int q = 0; q < 5; ++q) acc[q] = fmaf(gg.x, ap[q], acc[q]);
          }
          float* dst = s.gw2 + (co * C1 + ci) * 25 + r * 5;
#pragma unroll
          for (int q = 0; q < 5; ++q) dst[q] += acc[q];
        }
      }
      for (int c = warp - 4; c < C2; c += NW - 4) {
        float acc = lane < NPOS2 ? s.g2[c * G2S + lane].x : 0.f;
        acc = warp_sum(acc);
        if (lane == 0) s.gb2[c] += acc;
      }
      if (t2 < F1) s.gfb1[t2] += s.dh1[t2];
      if (t2 < F2) s.gfb2[t2] += s.dh2[t2];
      if (t2 < F3) s.gfb3[t2] += s.dlg[t2];
      float* sc = p.scratch + (size_t)img * SC_STRIDE;
      for (int i = t2; i < F0; i += T2) sc[SC_P2 + i] = s.p2[i];
      if (t2 < F1) { sc[SC_DH1 + t2] = s.dh1[t2]; sc[SC_H1 + t2] = s.h1[t2]; }
      if (t2 < F2) { sc[SC_DH2 + t2] = s.dh2[t2]; sc[SC_H2 + t2] = s.h2[t2]; }
      if (t2 < F3) sc[SC_DLG + t2] = s.dlg[t2];
    }
    __syncthreads();
    NET_STAMP(10);
    if (tid < F0) s.da2p[dpos] = 0.f;                         // back to all-zero for the next image
    if (tid < NPOS1) {                                        // through pool1 + ReLU -> sparse gradient of conv1's output
      const int py = tid / P1, px = tid % P1;
#pragma unroll
      for (int c = 0; c < C1; ++c) {
        const int o = c * NPOS1 + tid, sel = s.i1[o];
        const float sum = (dp1q[o] + dp1q[C1 * NPOS1 + o]) + (dp1q[2 * C1 * NPOS1 + o] + dp1q[3 * C1 * NPOS1 + o]);
        s.g1[o] = make_float2(s.p1[o] > 0.f ? sum : 0.f,
                              __int_as_float((2 * py + (sel >> 1)) * IN_RS + 2 * px + (sel & 1)));
      }
    }
    __syncthreads();
    NET_STAMP(11);

    // ---------------------------------------------------------------- conv1 weight gradient (accumulated in smem)
    if (warp < 15) {                                          // half-warp = (output channel, group of ~40 pooled positions),
      const int unit = warp * 2 + (lane >> 4), l = lane & 15; // lane = (ci, r) with 5 taps each: 7 smem words per 5 FMAs
      if (l < 15) {
        const int co = unit / W1G, grp = unit % W1G, ci = l / 5, r = l % 5;
        const float* a = in + ci * IN_PS + r * IN_RS;
        const float2* g = s.g1 + co * NPOS1;
        const int i0 = grp * 40, i1 = min(NPOS1, i0 + 40);
        float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
        for (int i = i0; i < i1; ++i) {
          const float2 gg = g[i];
          const float* ap = a + __float_as_int(gg.y);
#pragma unroll
          for (int q = 0; q < 5; ++q) acc[q] = fmaf(gg.x, ap[q], acc[q]);
        }
        float* dst = s.gw1[grp] + co * 75 + ci * 25 + r * 5;
#pragma unroll
        for (int q = 0; q < 5; ++q) dst[q] += acc[q];
      }
    } else {
      for (int c = 0; c < C1; ++c) {                          // conv1 bias: sum of the sparse gradient map
        float acc = 0.f;
        for (int i = lane; i < NPOS1; i += 32) acc += s.g1[c * NPOS1 + i].x;
        acc = warp_sum(acc);
        if (lane == 0) s.gb1[c] += acc;
      }
    }
    NET_STAMP(12);
    __syncthreads();
  }

  img = 0;
  NET_STAMP(13);
  // ---- flush this CTA's accumulated gradients into the arena
  if (p.want_grad && blockIdx.x < p.B) {
    for (int i = tid; i < C1 * 3 * 25; i += T) {
      const int s2 = i % 5, r = (i / 5) % 5, ci = (i / 25) % 3, co = i / 75;
      atomicAdd(p.gw1 + widx(p.krsc, co, ci, r, s2, 3), (s.gw1[0][i] + s.gw1[1][i]) + (s.gw1[2][i] + s.gw1[3][i]) + s.gw1[4][i]);
    }
    for (int i = tid; i < C2 * C1 * 25; i += T) {
      const int s2 = i % 5, r = (i / 5) % 5, ci = (i / 25) % C1, co = i / (25 * C1);
      atomicAdd(p.gw2 + widx(p.krsc, co, ci, r, s2, C1), s.gw2[i]);
    }
    if (tid < C1) atomicAdd(p.gb1 + tid, s.gb1[tid]);
    if (tid < C2) atomicAdd(p.gb2 + tid, s.gb2[tid]);
    if (tid < F1) atomicAdd(p.gfb1 + tid, s.gfb1[tid]);
    if (tid < F2) atomicAdd(p.gfb2 + tid, s.gfb2[tid]);
    if (tid < F3) atomicAdd(p.gfb3 + tid, s.gfb3[tid]);
  }
  NET_STAMP(14);
  // ---- deterministic mean loss by the last CTA
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned ticket = atomicAdd(p.counter, 1u);
    s.last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (s.last) {
    __threadfence();
    float acc = 0.f;
    for (int i = tid; i < p.B; i += T) acc += __ldcg(p.row_loss + i);
    acc = block_sum(acc, s.red);
    if (tid == 0) {
      *p.loss_mean = acc / (float)p.B;
      *p.counter = 0u;
    }
  }
  NET_STAMP(15);
}

// Batched weight gradients of the three fc layers: dW[o][i] += sum_img d[img][o] * a[img][i], read from the per-image
// scratch slices.  blockIdx.x: 15 blocks of 8 rows for fc1 (120x400), 11 for fc2 (84x120), 2 for fc3 (10x84);
// blockIdx.y: a chunk of images; 416 threads = one per column (fc1 has 400).
__global__ void __launch_bounds__(416) net_fc_wgrad_kernel(const float* __restrict__ scratch, float* __restrict__ g1,
                                                           float* __restrict__ g2, float* __restrict__ g3, int B, int imgs_per_cta) {
  __shared__ float dh[8][64];
  pdl_trigger();
  pdl_wait();
  int bx = blockIdx.x, d_off, a_off, rows, cols;
  float* out;
  if (bx < 15) { d_off = SC_DH1; a_off = SC_P2; rows = F1; cols = F0; out = g1; }
  else if (bx < 26) { bx -= 15; d_off = SC_DH2; a_off = SC_H1; rows = F2; cols = F1; out = g2; }
  else { bx -= 26; d_off = SC_DLG; a_off = SC_H2; rows = F3; cols = F2; out = g3; }
  const int o0 = bx * 8;
  const int i0 = blockIdx.y * imgs_per_cta, i1 = min(B, i0 + imgs_per_cta);
  const int i = threadIdx.x;
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  for (int base = i0; base < i1; base += 64) {
    const int n = min(64, i1 - base);
    __syncthreads();
    for (int t = threadIdx.x; t < 8 * n; t += blockDim.x) {
      const int k = t / n, j = t % n;
      dh[k][j] = (o0 + k < rows) ? scratch[(size_t)(base + j) * SC_STRIDE + d_off + o0 + k] : 0.f;
    }
    __syncthreads();
    if (i < cols) {
      const float* ap = scratch + (size_t)base * SC_STRIDE + a_off + i;
      int j = 0;
      for (; j + 8 <= n; j += 8) {                            // 8 independent (coalesced) row loads in flight
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldg(ap + (size_t)(j + u) * SC_STRIDE);
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] = fmaf(dh[k][j + u], v[u], acc[k]);
      }
      for (; j < n; ++j) {
        const float v = __ldg(ap + (size_t)j * SC_STRIDE);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(dh[k][j], v, acc[k]);
      }
    }
  }
  if (i < cols) {
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (o0 + k < rows) atomicAdd(out + (size_t)(o0 + k) * cols + i, acc[k]);
  }
}

}  // namespace

static long long* g_net_prof = nullptr;
// Debug: 16 clock64 stamps of CTA 0 (setup, each phase of its first image, flush) go to `buf` (device memory, or NULL to stop).
extern "C" void b200_net_set_profile_buffer(long long* buf) { g_net_prof = buf; }

extern "C" size_t b200_net_scratch_bytes(int batch) { return (size_t)batch * SC_STRIDE * sizeof(float) + 256; }

// params / grads: 10 pointers each in state_dict order (conv1.w, conv1.b, conv2.w, conv2.b, fc1.w, fc1.b, fc2.w, fc2.b,
// fc3.w, fc3.b); grads may be NULL (forward / evaluation only).  scratch: b200_net_scratch_bytes(batch), its last 256 bytes
// hold the ticket counter and must be zero before the first use.  The fc weights must be 16-byte aligned.
static int net_launch(NetParams& p, const float* const* params, float* const* grads, int batch, int krsc, float* row_loss,
                      float* loss_mean, float* logits, void* scratch, size_t scratch_bytes, void* stream) {
  B200_REQUIRE(p.target && params && row_loss && loss_mean && scratch && batch > 0, "bad argument");
  B200_REQUIRE(scratch_bytes >= b200_net_scratch_bytes(batch), "scratch too small (b200_net_scratch_bytes)");
  B200_REQUIRE((((uintptr_t)params[4] | (uintptr_t)params[6] | (uintptr_t)params[8]) & 15) == 0, "fc weights must be 16-byte aligned");
  p.B = batch; p.krsc = krsc; p.want_grad = grads != nullptr;
  p.w1 = params[0]; p.b1 = params[1]; p.w2 = params[2]; p.b2 = params[3]; p.fw1 = params[4]; p.fb1 = params[5];
  p.fw2 = params[6]; p.fb2 = params[7]; p.fw3 = params[8]; p.fb3 = params[9];
  if (grads) {
    p.gw1 = grads[0]; p.gb1 = grads[1]; p.gw2 = grads[2]; p.gb2 = grads[3]; p.gfw1 = grads[4]; p.gfb1 = grads[5];
    p.gfw2 = grads[6]; p.gfb2 = grads[7]; p.gfw3 = grads[8]; p.gfb3 = grads[9];
  }
  p.scratch = (float*)scratch;
  p.counter = (unsigned*)((char*)scratch + (size_t)batch * SC_STRIDE * sizeof(float));
  p.row_loss = row_loss; p.loss_mean = loss_mean; p.logits = logits;
  p.grad_scale = 1.f / (float)batch;
  p.prof = g_net_prof;
  static bool attr = false;
  if (!attr) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(net_fwd_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
    attr = true;
  }
  const int grid = batch < b200_num_sms() ? batch : b200_num_sms();
  cudaStream_t s = (cudaStream_t)stream;
  net_fwd_bwd_kernel<<<grid, T, sizeof(Smem), s>>>(p);
  B200_CHECK_LAUNCH();
  if (grads) {
    const int chunks = batch >= 128 ? 16 : (batch >= 32 ? 8 : 1);
    const int per = (batch + chunks - 1) / chunks;
    b200_launch(net_fc_wgrad_kernel, dim3(28, chunks), dim3(416), 0, s, (const float*)p.scratch, p.gfw1, p.gfw2, p.gfw3, batch, per);
    B200_CHECK_LAUNCH();
  }
  return 0;
}

extern "C" int b200_net_fwd_bwd(const float* x_nchw, const int64_t* targets, int batch, const float* const* params,
                                float* const* grads, int krsc, float* row_loss, float* loss_mean, float* logits,
                                void* scratch, size_t scratch_bytes, void* stream) {
  B200_REQUIRE(x_nchw && ((uintptr_t)x_nchw & 15) == 0, "input must be non-null and 16-byte aligned");
  NetParams p;
  memset(&p, 0, sizeof(p));
  p.x = x_nchw; p.target = targets;
  return net_launch(p, params, grads, batch, krsc, row_loss, loss_mean, logits, scratch, scratch_bytes, stream);
}

// Same, fed with the raw uint8 NHWC batch: crop/flip (offsets [B][3] = dy, dx, flip; NULL = centre, no flip) and
// normalisation happen while the image is staged into shared memory, so a training step needs no augmentation launch.
extern "C" int b200_net_fwd_bwd_u8(const uint8_t* img_nhwc, const int32_t* offsets, const float* mean, const float* inv_std,
                                   const int64_t* targets, int batch, const float* const* params, float* const* grads, int krsc,
                                   float* row_loss, float* loss_mean, float* logits, void* scratch, size_t scratch_bytes,
                                   void* stream) {
  B200_REQUIRE(img_nhwc && mean && inv_std && ((uintptr_t)img_nhwc & 15) == 0, "uint8 input must be non-null and 16-byte aligned");
  NetParams p;
  memset(&p, 0, sizeof(p));
  p.xu8 = img_nhwc; p.offsets = offsets; p.mean = mean; p.inv_std = inv_std; p.target = targets;
  return net_launch(p, params, grads, batch, krsc, row_loss, loss_mean, logits, scratch, scratch_bytes, stream);
}
