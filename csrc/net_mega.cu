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
constexpr int H0 = 32, H1 = 28, P1 = 14, H2 = 10, P2 = 5;
constexpr int F0 = 400, F1 = 120, F2 = 84, F3 = 10;

struct NetParams {
  const float* x;            // [B][3][32][32] fp32 NCHW (normalised)
  const int64_t* target;     // [B]
  const float *w1, *b1, *w2, *b2, *fw1, *fb1, *fw2, *fb2, *fw3, *fb3;
  float *gw1, *gb1, *gw2, *gb2, *gfw1, *gfb1, *gfw2, *gfb2, *gfw3, *gfb3;
  float* scratch;            // per image: dh1[120], p2[400]
  float* row_loss;           // [B]
  float* loss_mean;          // scalar
  float* logits;             // optional [B][10]
  unsigned* counter;         // ticket
  int B, krsc, want_grad;
  float grad_scale;          // 1/B
};

__device__ __forceinline__ int widx(int krsc, int co, int ci, int r, int s, int cin) {
  return krsc ? ((co * KS + r) * KS + s) * cin + ci : ((co * cin + ci) * KS + r) * KS + s;
}

struct Smem {
  // per-image activations (first: `in` is filled with 16-byte vector stores)
  float in[3 * H0 * H0], a1[C1 * H1 * H1], p1[C1 * P1 * P1], a2[C2 * H2 * H2], p2[F0];
  float h1[F1], h2[F2], lg[F3], dlg[F3], dh2[F2], dh1[F1], dp2[F0], da2[C2 * H2 * H2], dp1[C1 * P1 * P1];
  // weights kept in torch order regardless of the global layout
  float w1[C1 * 3 * 25], b1[C1], w2[C2 * C1 * 25], b2[C2];
  // gradient accumulators over this CTA's images
  float gw1[C1 * 3 * 25], gb1[C1], gw2[C2 * C1 * 25], gb2[C2], gfw2[F2 * F1], gfb2[F2], gfw3[F3 * F2], gfb3[F3], gfb1[F1];
  float fw2[F2 * F1], fw3[F3 * F2];          // small fc weights, resident for all images of the CTA
  float part[NET_T / 32][F0];                // per-warp partial sums of W1^T dh1
  float red[32];
  uint8_t i1[C1 * P1 * P1], i2[F0];
  bool last;
};

__device__ __forceinline__ float warp_dot_global(const float* __restrict__ wrow, const float* __restrict__ v, int n, int lane) {
  float acc = 0.f;
  for (int i = lane; i < n; i += 32) acc = fmaf(__ldg(wrow + i), v[i], acc);
  return warp_sum(acc);
}

__global__ void __launch_bounds__(T, 1) net_fwd_bwd_kernel(const NetParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = T / 32;

  // ---- weights -> smem (torch order), accumulators -> 0
  for (int i = tid; i < C1 * 3 * 25; i += T) {
    const int s2 = i % 5, r = (i / 5) % 5, ci = (i / 25) % 3, co = i / 75;
    s.w1[i] = p.w1[widx(p.krsc, co, ci, r, s2, 3)];
    s.gw1[i] = 0.f;
  }
  for (int i = tid; i < C2 * C1 * 25; i += T) {
    const int s2 = i % 5, r = (i / 5) % 5, ci = (i / 25) % C1, co = i / (25 * C1);
    s.w2[i] = p.w2[widx(p.krsc, co, ci, r, s2, C1)];
    s.gw2[i] = 0.f;
  }
  for (int i = tid; i < F2 * F1; i += T) { s.gfw2[i] = 0.f; s.fw2[i] = p.fw2[i]; }
  for (int i = tid; i < F3 * F2; i += T) { s.gfw3[i] = 0.f; s.fw3[i] = p.fw3[i]; }
  if (tid < C1) { s.b1[tid] = p.b1[tid]; s.gb1[tid] = 0.f; }
  if (tid < C2) { s.b2[tid] = p.b2[tid]; s.gb2[tid] = 0.f; }
  if (tid < F1) s.gfb1[tid] = 0.f;
  if (tid < F2) s.gfb2[tid] = 0.f;
  if (tid < F3) s.gfb3[tid] = 0.f;
  __syncthreads();

  for (int img = blockIdx.x; img < p.B; img += gridDim.x) {
    // ---------------------------------------------------------------- forward
    const float* xin = p.x + (size_t)img * 3 * H0 * H0;
    for (int i = tid; i < 3 * H0 * H0 / 4; i += T) reinterpret_cast<float4*>(s.in)[i] = __ldg(reinterpret_cast<const float4*>(xin) + i);
    __syncthreads();
    for (int o = tid; o < C1 * H1 * H1; o += T) {            // conv1 + bias + ReLU
      const int x = o % H1, y = (o / H1) % H1, co = o / (H1 * H1);
      float acc = s.b1[co];
      const float* w = s.w1 + co * 75;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int r = 0; r < 5; ++r)
#pragma unroll
          for (int q = 0; q < 5; ++q) acc = fmaf(s.in[(ci * H0 + y + r) * H0 + x + q], w[(ci * 5 + r) * 5 + q], acc);
      s.a1[o] = fmaxf(acc, 0.f);
    }
    __syncthreads();
    for (int o = tid; o < C1 * P1 * P1; o += T) {            // maxpool 2x2
      const int x = o % P1, y = (o / P1) % P1, c = o / (P1 * P1);
      const float* a = s.a1 + (c * H1 + 2 * y) * H1 + 2 * x;
      float best = a[0]; int bi = 0;
      if (a[1] > best) { best = a[1]; bi = 1; }
      if (a[H1] > best) { best = a[H1]; bi = 2; }
      if (a[H1 + 1] > best) { best = a[H1 + 1]; bi = 3; }
      s.p1[o] = best; s.i1[o] = (uint8_t)bi;
    }
    __syncthreads();
    for (int o = tid; o < C2 * H2 * H2; o += T) {            // conv2 + bias + ReLU
      const int x = o % H2, y = (o / H2) % H2, co = o / (H2 * H2);
      float acc = s.b2[co];
      const float* w = s.w2 + co * (C1 * 25);
#pragma unroll 2
      for (int ci = 0; ci < C1; ++ci)
#pragma unroll
        for (int r = 0; r < 5; ++r)
#pragma unroll
          for (int q = 0; q < 5; ++q) acc = fmaf(s.p1[(ci * P1 + y + r) * P1 + x + q], w[(ci * 5 + r) * 5 + q], acc);
      s.a2[o] = fmaxf(acc, 0.f);
    }
    __syncthreads();
    for (int o = tid; o < F0; o += T) {                       // maxpool 2x2 -> flatten (c,h,w)
      const int x = o % P2, y = (o / P2) % P2, c = o / (P2 * P2);
      const float* a = s.a2 + (c * H2 + 2 * y) * H2 + 2 * x;
      float best = a[0]; int bi = 0;
      if (a[1] > best) { best = a[1]; bi = 1; }
      if (a[H2] > best) { best = a[H2]; bi = 2; }
      if (a[H2 + 1] > best) { best = a[H2 + 1]; bi = 3; }
      s.p2[o] = best; s.i2[o] = (uint8_t)bi;
    }
    __syncthreads();
    for (int o = warp; o < F1; o += NW) {                     // fc1 + ReLU (weights streamed from L2)
      const float v = warp_dot_global(p.fw1 + (size_t)o * F0, s.p2, F0, lane);
      if (lane == 0) s.h1[o] = fmaxf(v + __ldg(p.fb1 + o), 0.f);
    }
    __syncthreads();
    for (int o = warp; o < F2; o += NW) {                     // fc2 + ReLU
      float v = 0.f;
      for (int i = lane; i < F1; i += 32) v = fmaf(s.fw2[o * F1 + i], s.h1[i], v);
      v = warp_sum(v);
      if (lane == 0) s.h2[o] = fmaxf(v + __ldg(p.fb2 + o), 0.f);
    }
    __syncthreads();
    for (int o = warp; o < F3; o += NW) {                     // fc3
      float v = 0.f;
      for (int i = lane; i < F2; i += 32) v = fmaf(s.fw3[o * F2 + i], s.h2[i], v);
      v = warp_sum(v);
      if (lane == 0) s.lg[o] = v + __ldg(p.fb3 + o);
    }
    __syncthreads();
    if (warp == 0) {                                          // softmax cross-entropy
      const float v = lane < F3 ? s.lg[lane] : -INFINITY;
      const float mx = warp_max(v);
      const float e = lane < F3 ? __expf(v - mx) : 0.f;
      const float se = warp_sum(e);
      const int t = (int)p.target[img];
      if (lane < F3) {
        s.dlg[lane] = (e / se - (lane == t ? 1.f : 0.f)) * p.grad_scale;
        if (p.logits) p.logits[(size_t)img * F3 + lane] = v;
      }
      if (lane == 0) p.row_loss[img] = (t >= 0 && t < F3) ? (mx + __logf(se) - s.lg[t]) : 0.f;
    }
    __syncthreads();
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
      float acc = 0.f;
#pragma unroll 4
      for (int k = 0; k < F2; ++k) acc = fmaf(s.fw2[k * F1 + j], s.dh2[k], acc);
      s.dh1[j] = s.h1[j] > 0.f ? acc : 0.f;
    }
    __syncthreads();
    {                                                         // dp2 = W1^T dh1: warps stream whole rows (coalesced,
      float acc[13];                                          // 13 independent loads in flight), partials meet in smem
#pragma unroll
      for (int k = 0; k < 13; ++k) acc[k] = 0.f;
      for (int o = warp; o < F1; o += NW) {
        const float d = s.dh1[o];
        const float* wr = p.fw1 + (size_t)o * F0 + lane;
#pragma unroll
        for (int k = 0; k < 13; ++k)
          if (lane + 32 * k < F0) acc[k] = fmaf(__ldg(wr + 32 * k), d, acc[k]);
      }
#pragma unroll
      for (int k = 0; k < 13; ++k)
        if (lane + 32 * k < F0) s.part[warp][lane + 32 * k] = acc[k];
    }
    __syncthreads();
    for (int i = tid; i < F0; i += T) {
      float acc = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) acc += s.part[w][i];
      s.dp2[i] = acc;
    }
    __syncthreads();
    for (int o = tid; o < C2 * H2 * H2; o += T) {             // unpool + ReLU mask -> da2
      const int x = o % H2, y = (o / H2) % H2, c = o / (H2 * H2);
      const int po = (c * P2 + (y >> 1)) * P2 + (x >> 1);
      const bool hit = s.i2[po] == (uint8_t)(((y & 1) << 1) | (x & 1));
      s.da2[o] = (hit && s.a2[o] > 0.f) ? s.dp2[po] : 0.f;
    }
    __syncthreads();
    for (int o = tid; o < C1 * P1 * P1; o += T) {             // conv2 dgrad -> dp1
      const int x = o % P1, y = (o / P1) % P1, ci = o / (P1 * P1);
      float acc = 0.f;
      for (int co = 0; co < C2; ++co) {
        const float* w = s.w2 + (co * C1 + ci) * 25;
        const float* d = s.da2 + co * H2 * H2;
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          const int yy = y - r;
          if (yy < 0 || yy >= H2) continue;
#pragma unroll
          for (int q = 0; q < 5; ++q) {
            const int xx = x - q;
            if (xx >= 0 && xx < H2) acc = fmaf(d[yy * H2 + xx], w[r * 5 + q], acc);
          }
        }
      }
      s.dp1[o] = acc;
    }
    __syncthreads();
    for (int o = tid; o < C1 * H1 * H1; o += T) {             // unpool + ReLU mask -> da1 (in place of a1)
      const int x = o % H1, y = (o / H1) % H1, c = o / (H1 * H1);
      const int po = (c * P1 + (y >> 1)) * P1 + (x >> 1);
      const bool hit = s.i1[po] == (uint8_t)(((y & 1) << 1) | (x & 1));
      s.a1[o] = (hit && s.a1[o] > 0.f) ? s.dp1[po] : 0.f;
    }
    __syncthreads();

    // ---------------------------------------------------------------- backward (weights, accumulated in smem)
    for (int o = tid; o < C1 * 3 * 25; o += T) {              // dW1[co][ci][r][q] += sum_{y,x} da1 * in
      const int q = o % 5, r = (o / 5) % 5, ci = (o / 25) % 3, co = o / 75;
      const float* d = s.a1 + co * H1 * H1;
      const float* a = s.in + (ci * H0 + r) * H0 + q;
      float acc = 0.f;
      for (int y = 0; y < H1; ++y)
#pragma unroll 4
        for (int x = 0; x < H1; ++x) acc = fmaf(d[y * H1 + x], a[y * H0 + x], acc);
      s.gw1[o] += acc;
    }
    for (int o = tid; o < C2 * C1 * 25; o += T) {             // dW2
      const int q = o % 5, r = (o / 5) % 5, ci = (o / 25) % C1, co = o / (25 * C1);
      const float* d = s.da2 + co * H2 * H2;
      const float* a = s.p1 + (ci * P1 + r) * P1 + q;
      float acc = 0.f;
#pragma unroll 2
      for (int y = 0; y < H2; ++y)
#pragma unroll
        for (int x = 0; x < H2; ++x) acc = fmaf(d[y * H2 + x], a[y * P1 + x], acc);
      s.gw2[o] += acc;
    }
    for (int o = tid; o < F2 * F1; o += T) s.gfw2[o] = fmaf(s.dh2[o / F1], s.h1[o % F1], s.gfw2[o]);
    for (int o = tid; o < F3 * F2; o += T) s.gfw3[o] = fmaf(s.dlg[o / F2], s.h2[o % F2], s.gfw3[o]);
    if (tid < F1) s.gfb1[tid] += s.dh1[tid];
    if (tid < F2) s.gfb2[tid] += s.dh2[tid];
    if (tid < F3) s.gfb3[tid] += s.dlg[tid];
    for (int c = warp; c < C1; c += NW) {                     // conv biases: sum of the gradient maps
      float acc = 0.f;
      for (int i = lane; i < H1 * H1; i += 32) acc += s.a1[c * H1 * H1 + i];
      acc = warp_sum(acc);
      if (lane == 0) s.gb1[c] += acc;
    }
    for (int c = warp; c < C2; c += NW) {
      float acc = 0.f;
      for (int i = lane; i < H2 * H2; i += 32) acc += s.da2[c * H2 * H2 + i];
      acc = warp_sum(acc);
      if (lane == 0) s.gb2[c] += acc;
    }
    float* sc = p.scratch + (size_t)img * (F1 + F0);          // fc1's outer product is done over the batch later
    for (int i = tid; i < F1; i += T) sc[i] = s.dh1[i];
    for (int i = tid; i < F0; i += T) sc[F1 + i] = s.p2[i];
    __syncthreads();
  }

  // ---- flush this CTA's accumulated gradients into the arena
  if (p.want_grad && blockIdx.x < p.B) {
    for (int i = tid; i < C1 * 3 * 25; i += T) {
      const int s2 = i % 5, r = (i / 5) % 5, ci = (i / 25) % 3, co = i / 75;
      atomicAdd(p.gw1 + widx(p.krsc, co, ci, r, s2, 3), s.gw1[i]);
    }
    for (int i = tid; i < C2 * C1 * 25; i += T) {
      const int s2 = i % 5, r = (i / 5) % 5, ci = (i / 25) % C1, co = i / (25 * C1);
      atomicAdd(p.gw2 + widx(p.krsc, co, ci, r, s2, C1), s.gw2[i]);
    }
    for (int i = tid; i < F2 * F1; i += T) atomicAdd(p.gfw2 + i, s.gfw2[i]);
    for (int i = tid; i < F3 * F2; i += T) atomicAdd(p.gfw3 + i, s.gfw3[i]);
    if (tid < C1) atomicAdd(p.gb1 + tid, s.gb1[tid]);
    if (tid < C2) atomicAdd(p.gb2 + tid, s.gb2[tid]);
    if (tid < F1) atomicAdd(p.gfb1 + tid, s.gfb1[tid]);
    if (tid < F2) atomicAdd(p.gfb2 + tid, s.gfb2[tid]);
    if (tid < F3) atomicAdd(p.gfb3 + tid, s.gfb3[tid]);
  }
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
}

// dW_fc1[o][i] += sum_img dh1[img][o] * p2[img][i];  grid (F1/8 row blocks, image chunks), 416 threads
__global__ void __launch_bounds__(416) net_fc1_wgrad_kernel(const float* __restrict__ scratch, float* __restrict__ gfw1, int B,
                                                            int imgs_per_cta) {
  __shared__ float dh[8][64];
  const int o0 = blockIdx.x * 8;
  const int i0 = blockIdx.y * imgs_per_cta, i1 = min(B, i0 + imgs_per_cta);
  const int i = threadIdx.x;
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  for (int base = i0; base < i1; base += 64) {
    const int n = min(64, i1 - base);
    __syncthreads();
    for (int t = threadIdx.x; t < 8 * n; t += blockDim.x) dh[t / n][t % n] = scratch[(size_t)(base + t % n) * (F1 + F0) + o0 + t / n];
    __syncthreads();
    if (i < F0) {
      for (int j = 0; j < n; ++j) {
        const float v = __ldg(scratch + (size_t)(base + j) * (F1 + F0) + F1 + i);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(dh[k][j], v, acc[k]);
      }
    }
  }
  if (i < F0) {
#pragma unroll
    for (int k = 0; k < 8; ++k) atomicAdd(gfw1 + (size_t)(o0 + k) * F0 + i, acc[k]);
  }
}

}  // namespace

extern "C" size_t b200_net_scratch_bytes(int batch) { return (size_t)batch * (F1 + F0) * sizeof(float) + 256; }

// params / grads: 10 pointers each in state_dict order (conv1.w, conv1.b, conv2.w, conv2.b, fc1.w, fc1.b, fc2.w, fc2.b,
// fc3.w, fc3.b); grads may be NULL (forward / evaluation only).  scratch: b200_net_scratch_bytes(batch), its last 256 bytes
// hold the ticket counter and must be zero before the first use.
extern "C" int b200_net_fwd_bwd(const float* x_nchw, const int64_t* targets, int batch, const float* const* params,
                                float* const* grads, int krsc, float* row_loss, float* loss_mean, float* logits,
                                void* scratch, size_t scratch_bytes, void* stream) {
  B200_REQUIRE(x_nchw && targets && params && row_loss && loss_mean && scratch && batch > 0, "bad argument");
  B200_REQUIRE(scratch_bytes >= b200_net_scratch_bytes(batch), "scratch too small (b200_net_scratch_bytes)");
  NetParams p;
  memset(&p, 0, sizeof(p));
  p.x = x_nchw; p.target = targets; p.B = batch; p.krsc = krsc; p.want_grad = grads != nullptr;
  p.w1 = params[0]; p.b1 = params[1]; p.w2 = params[2]; p.b2 = params[3]; p.fw1 = params[4]; p.fb1 = params[5];
  p.fw2 = params[6]; p.fb2 = params[7]; p.fw3 = params[8]; p.fb3 = params[9];
  if (grads) {
    p.gw1 = grads[0]; p.gb1 = grads[1]; p.gw2 = grads[2]; p.gb2 = grads[3]; p.gfw1 = grads[4]; p.gfb1 = grads[5];
    p.gfw2 = grads[6]; p.gfb2 = grads[7]; p.gfw3 = grads[8]; p.gfb3 = grads[9];
  }
  p.scratch = (float*)scratch;
  p.counter = (unsigned*)((char*)scratch + (size_t)batch * (F1 + F0) * sizeof(float));
  p.row_loss = row_loss; p.loss_mean = loss_mean; p.logits = logits;
  p.grad_scale = 1.f / (float)batch;
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
    const int chunks = batch >= 64 ? 8 : 1;
    const int per = (batch + chunks - 1) / chunks;
    net_fc1_wgrad_kernel<<<dim3(F1 / 8, chunks), 416, 0, s>>>(p.scratch, p.gfw1, batch, per);
    B200_CHECK_LAUNCH();
  }
  return 0;
}
