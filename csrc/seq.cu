// Sequence-model kernels for the trojan-detection (MNTD) model zoo, all fp32:
//
//   sgemm           strided fp32 GEMM on the CUDA cores for the small products around these models (gate
//                   pre-activations, LSTM / text-CNN weight gradients, the (k x 300) convolutions as GEMMs over
//                   overlapping embedding windows).  The models are a few hundred KFLOP per sample; tensor cores
//                   in bf16 would cost accuracy the meta-classifier's query tuning needs and win nothing here.
//   lstm_rec_fwd    the recurrence of one LSTM layer (model_lib/audio_rnn_model.py:11, nn.LSTM(40, 100, 2)):
//                   one CTA per sequence, every thread keeps one row of W_hh in registers for all T steps,
//                   h_t lives in shared memory; gates, cell states and h are saved for the backward pass.
//   lstm_rec_bwd    back-propagation through time of the same layer: produces the pre-activation gate
//                   gradients for every step (the weight gradients are then three GEMMs over [B*T]).
//   mel_frontend    reflect-padded framing + Hann window + 2048-point FFT + power + 40-band mel projection
//                   + dB scaling of audio_rnn_model.py:25-35, one CTA per frame, nothing but the features
//                   ever leaves shared memory.
//   gather_rows     embedding lookup (rtNLP_cnn_model.py:17-19).
//   textcnn_pool    bias + ReLU + max-over-time of the conv GEMM output and its backward
//                   (rtNLP_cnn_model.py:49-52).
#include "common.cuh"

namespace {

// ------------------------------------------------------------------------------------------------
// sgemm: C[m][n] = sum_k A(m,k) B(k,n) (+ bias[n]) (+ beta C) (ReLU) with A(m,k) = A[m*sam + k*sak],
// B(k,n) = B[k*sbk + n*sbn]: transposes and overlapping rows are just strides.
// ------------------------------------------------------------------------------------------------
constexpr int GBM = 64, GBN = 64, GBK = 16;

struct GemmArgs {
  const float *A, *B, *bias;
  float* C;
  int M, N, K, ldc, relu, splitk, kchunk;
  long long sam, sak, sbk, sbn;
  float beta;
};

__global__ void __launch_bounds__(256) sgemm_kernel(const GemmArgs g) {
  __shared__ __align__(16) float As[GBK][GBM + 4];
  __shared__ __align__(16) float Bs[GBK][GBN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * GBM, n0 = blockIdx.x * GBN;
  const int kbeg = blockIdx.z * g.kchunk, kend = min(g.K, kbeg + g.kchunk);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = kbeg; k0 < kend; k0 += GBK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 256;
      int m, k;
      if (g.sak == 1) { m = idx >> 4; k = idx & 15; } else { k = idx >> 6; m = idx & 63; }
      const int gm = m0 + m, gk = k0 + k;
      As[k][m] = (gm < g.M && gk < kend) ? __ldg(g.A + gm * g.sam + gk * g.sak) : 0.f;
      int n, kb;
      if (g.sbn == 1) { kb = idx >> 6; n = idx & 63; } else { n = idx >> 4; kb = idx & 15; }
      const int gn = n0 + n, gkb = k0 + kb;
      Bs[kb][n] = (gn < g.N && gkb < kend) ? __ldg(g.B + gkb * g.sbk + gn * g.sbn) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GBK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= g.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= g.N) continue;
      float v = acc[i][j];
      float* c = g.C + (size_t)m * g.ldc + n;
      if (g.splitk > 1) {                                     // C was zeroed (or holds beta*C) by the host side
        if (blockIdx.z == 0 && g.bias) v += g.bias[n];
        atomicAdd(c, v);
      } else {
        if (g.bias) v += g.bias[n];
        if (g.beta != 0.f) v += g.beta * *c;
        *c = g.relu ? fmaxf(v, 0.f) : v;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// LSTM recurrence (PyTorch gate order i, f, g, o; rows of W_hh are [i | f | g | o] x H)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

template <int H>
__global__ void __launch_bounds__(4 * H + 16, 1)
lstm_rec_fwd_kernel(const float* __restrict__ pre, const float* __restrict__ whh, float* __restrict__ hseq,
                    float* __restrict__ gates, float* __restrict__ cseq, int T) {
  static_assert(H % 4 == 0, "H must be a multiple of 4");
  __shared__ __align__(16) float h_s[H];
  __shared__ float g_s[4 * H];
  const int b = blockIdx.x, j = threadIdx.x;
  const bool row = j < 4 * H;
  float w[H];
#pragma unroll
  for (int k = 0; k < H; ++k) w[k] = row ? __ldg(whh + (size_t)j * H + k) : 0.f;
  if (j < H) h_s[j] = 0.f;
  float c = 0.f;
  const float* pb = pre + (size_t)b * T * 4 * H;
  float nxt = row ? __ldg(pb + j) : 0.f;
  __syncthreads();
  for (int t = 0; t < T; ++t) {
    float a = nxt;
    if (row && t + 1 < T) nxt = __ldg(pb + (size_t)(t + 1) * 4 * H + j);      // prefetch the next step's input projection
    float a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int k = 0; k < H; k += 4) {
      const float4 hv = *reinterpret_cast<const float4*>(h_s + k);
      a = fmaf(w[k], hv.x, a); a1 = fmaf(w[k + 1], hv.y, a1); a2 = fmaf(w[k + 2], hv.z, a2); a3 = fmaf(w[k + 3], hv.w, a3);
    }
    a = (a + a1) + (a2 + a3);
    if (row) {
      const float act = (j >= 2 * H && j < 3 * H) ? tanhf(a) : sigmoidf_(a);
      g_s[j] = act;
      gates[((size_t)b * T + t) * 4 * H + j] = act;
    }
    __syncthreads();
    if (j < H) {
      c = g_s[H + j] * c + g_s[j] * g_s[2 * H + j];
      const float h = g_s[3 * H + j] * tanhf(c);
      h_s[j] = h;
      hseq[((size_t)b * T + t) * H + j] = h;
      cseq[((size_t)b * T + t) * H + j] = c;
    }
    __syncthreads();
  }
}

// thread (q, k) = (threadIdx / H, threadIdx % H) keeps W_hh[q*H + jj][k], jj < H, in registers: a quarter of column k
template <int H>
__global__ void __launch_bounds__(4 * H + 16, 1)
lstm_rec_bwd_kernel(const float* __restrict__ dh_out, const float* __restrict__ gates, const float* __restrict__ cseq,
                    const float* __restrict__ whh, float* __restrict__ dgates, int T) {
  __shared__ __align__(16) float dg_s[4 * H];
  __shared__ float part[4][H];
  const int b = blockIdx.x, tid = threadIdx.x;
  const bool act = tid < 4 * H;
  const int q = act ? tid / H : 0, k = act ? tid % H : 0;
  float w[H];
#pragma unroll
  for (int jj = 0; jj < H; ++jj) w[jj] = act ? __ldg(whh + (size_t)(q * H + jj) * H + k) : 0.f;
  float dc_next = 0.f;
  if (act) part[q][k] = 0.f;
  __syncthreads();
  for (int t = T - 1; t >= 0; --t) {
    if (tid < H) {
      const size_t base = ((size_t)b * T + t);
      const float dh = dh_out[base * H + tid] + (part[0][tid] + part[1][tid]) + (part[2][tid] + part[3][tid]);
      const float* gt = gates + base * 4 * H;
      const float i = gt[tid], f = gt[H + tid], g = gt[2 * H + tid], o = gt[3 * H + tid];
      const float c = cseq[base * H + tid];
      const float cp = t > 0 ? cseq[(base - 1) * H + tid] : 0.f;
      const float tc = tanhf(c);
      const float dc = dc_next + dh * o * (1.f - tc * tc);
      const float dai = dc * g * i * (1.f - i), daf = dc * cp * f * (1.f - f), dag = dc * i * (1.f - g * g),
                  dao = dh * tc * o * (1.f - o);
      dc_next = dc * f;
      dg_s[tid] = dai; dg_s[H + tid] = daf; dg_s[2 * H + tid] = dag; dg_s[3 * H + tid] = dao;
      float* out = dgates + base * 4 * H;
      out[tid] = dai; out[H + tid] = daf; out[2 * H + tid] = dag; out[3 * H + tid] = dao;
    }
    __syncthreads();
    if (act) {                                                // dh_{t-1}[k] = sum_j W_hh[j][k] * dgate[j], a quarter per thread
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
      const float* d = dg_s + q * H;
#pragma unroll
      for (int jj = 0; jj < H; jj += 4) {
        const float4 dv = *reinterpret_cast<const float4*>(d + jj);
        a0 = fmaf(w[jj], dv.x, a0); a1 = fmaf(w[jj + 1], dv.y, a1); a2 = fmaf(w[jj + 2], dv.z, a2); a3 = fmaf(w[jj + 3], dv.w, a3);
      }
      part[q][k] = (a0 + a1) + (a2 + a3);
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// mel front end: one CTA per (clip, frame); torch.stft(n_fft=2048, hop=512, center=True, reflect) semantics
// ------------------------------------------------------------------------------------------------
constexpr int NFFT = 2048, LOGN = 11, NBIN = NFFT / 2 + 1, HOP = NFFT / 4;

__global__ void __launch_bounds__(256)
mel_frontend_kernel(const float* __restrict__ x, const float* __restrict__ window, const float* __restrict__ melfb,
                    float* __restrict__ out, int L, int frames, int n_mels) {
  __shared__ float2 a[NFFT];
  __shared__ float2 tw[NFFT / 2];
  __shared__ float pw[NBIN + 3];
  const int f = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  for (int k = tid; k < NFFT / 2; k += 256) {                 // exp(-2 pi i k / N)
    float s, c;
    sincospif(-(float)k / (float)(NFFT / 2), &s, &c);
    tw[k] = make_float2(c, s);
  }
  const float* xb = x + (size_t)b * L;
  for (int n = tid; n < NFFT; n += 256) {                     // frame with reflect padding, windowed, bit-reversed order
    int src = f * HOP - NFFT / 2 + n;
    if (src < 0) src = -src;
    if (src >= L) src = 2 * (L - 1) - src;
    a[__brev((unsigned)n) >> (32 - LOGN)] = make_float2(xb[src] * window[n], 0.f);
  }
  __syncthreads();
#pragma unroll 1
  for (int s = 1; s <= LOGN; ++s) {                           // radix-2 decimation in time
    const int half = 1 << (s - 1);
    for (int idx = tid; idx < NFFT / 2; idx += 256) {
      const int grp = idx >> (s - 1), pos = idx & (half - 1);
      const int i0 = (grp << s) + pos, i1 = i0 + half;
      const float2 w = tw[pos << (LOGN - s)], u = a[i0], v = a[i1];
      const float2 t = make_float2(w.x * v.x - w.y * v.y, w.x * v.y + w.y * v.x);
      a[i0] = make_float2(u.x + t.x, u.y + t.y);
      a[i1] = make_float2(u.x - t.x, u.y - t.y);
    }
    __syncthreads();
  }
  for (int k = tid; k < NBIN; k += 256) pw[k] = a[k].x * a[k].x + a[k].y * a[k].y;
  __syncthreads();
  const int lane = tid & 31, warp = tid >> 5;
  for (int m = warp; m < n_mels; m += 8) {                    // mel projection + dB + the model's (x + 50) / 50
    const float* fb = melfb + (size_t)m * NBIN;
    float acc = 0.f;
    for (int k = lane; k < NBIN; k += 32) acc = fmaf(__ldg(fb + k), pw[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) out[((size_t)b * frames + f) * n_mels + m] = (10.f * log10f(fmaxf(acc, 1e-10f)) + 50.f) / 50.f;
  }
}

// Backward of the mel front end w.r.t. the waveform (the meta-classifier tunes input queries through the shadow model,
// meta_classifier.py:13,21-24).  Per frame: recompute X = FFT(frame * window) and the mel powers, then
//   dmel[m] = dout[m] * (10 / ln 10) / 50 / mel[m]            (0 where the 1e-10 clamp is active)
//   dpw[k]  = sum_m fb[m][k] dmel[m],   G[k] = 2 dpw[k] X[k]  (k <= N/2, 0 above: only those bins reach the output)
//   d(frame*window)[n] = Re sum_k G[k] e^{+2 pi i k n / N}     (the same radix-2 network with conjugated twiddles)
// and scatter-add dframe[n] * window[n] through the reflect padding into dx (atomics: frames overlap 4x).
__device__ __forceinline__ void fft_inplace(float2* a, const float2* tw, bool conj, int tid) {
#pragma unroll 1
  for (int s = 1; s <= LOGN; ++s) {
    const int half = 1 << (s - 1);
    for (int idx = tid; idx < NFFT / 2; idx += 256) {
      const int grp = idx >> (s - 1), pos = idx & (half - 1);
      const int i0 = (grp << s) + pos, i1 = i0 + half;
      float2 w = tw[pos << (LOGN - s)];
      if (conj) w.y = -w.y;
      const float2 u = a[i0], v = a[i1];
      const float2 t = make_float2(w.x * v.x - w.y * v.y, w.x * v.y + w.y * v.x);
      a[i0] = make_float2(u.x + t.x, u.y + t.y);
      a[i1] = make_float2(u.x - t.x, u.y - t.y);
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
mel_frontend_bwd_kernel(const float* __restrict__ x, const float* __restrict__ window, const float* __restrict__ melfb,
                        const float* __restrict__ dout, float* __restrict__ dx, int L, int frames, int n_mels) {
  __shared__ float2 a[NFFT];
  __shared__ float2 g[NFFT];
  __shared__ float2 tw[NFFT / 2];
  __shared__ float pw[NBIN + 3];
  __shared__ float dmel[64];
  const int f = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int k = tid; k < NFFT / 2; k += 256) {
    float s, c;
    sincospif(-(float)k / (float)(NFFT / 2), &s, &c);
    tw[k] = make_float2(c, s);
  }
  const float* xb = x + (size_t)b * L;
  for (int n = tid; n < NFFT; n += 256) {
    int src = f * HOP - NFFT / 2 + n;
    if (src < 0) src = -src;
    if (src >= L) src = 2 * (L - 1) - src;
    a[__brev((unsigned)n) >> (32 - LOGN)] = make_float2(xb[src] * window[n], 0.f);
  }
  __syncthreads();
  fft_inplace(a, tw, false, tid);
  for (int k = tid; k < NBIN; k += 256) pw[k] = a[k].x * a[k].x + a[k].y * a[k].y;
  __syncthreads();
  for (int m = warp; m < n_mels; m += 8) {
    const float* fb = melfb + (size_t)m * NBIN;
    float acc = 0.f;
    for (int k = lane; k < NBIN; k += 32) acc = fmaf(__ldg(fb + k), pw[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      const float go = dout[((size_t)b * frames + f) * n_mels + m];
      dmel[m] = acc > 1e-10f ? go * (10.f / 2.302585093f / 50.f) / acc : 0.f;
    }
  }
  __syncthreads();
  for (int k = tid; k < NFFT; k += 256) {                     // G in bit-reversed order, ready for the in-place network
    float2 v = make_float2(0.f, 0.f);
    if (k < NBIN) {
      float d = 0.f;
      for (int m = 0; m < n_mels; ++m) d = fmaf(__ldg(melfb + (size_t)m * NBIN + k), dmel[m], d);
      v = make_float2(2.f * d * a[k].x, 2.f * d * a[k].y);
    }
    g[__brev((unsigned)k) >> (32 - LOGN)] = v;
  }
  __syncthreads();
  fft_inplace(g, tw, true, tid);
  float* dxb = dx + (size_t)b * L;
  for (int n = tid; n < NFFT; n += 256) {
    int src = f * HOP - NFFT / 2 + n;
    if (src < 0) src = -src;
    if (src >= L) src = 2 * (L - 1) - src;
    atomicAdd(dxb + src, g[n].x * window[n]);
  }
}

// ------------------------------------------------------------------------------------------------
// text CNN pieces
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ table, const int64_t* __restrict__ idx, float* __restrict__ out, size_t rows,
                   int dim, int64_t vocab) {
  const size_t total = rows * dim;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / dim;
    const int64_t id = idx[r];
    out[i] = (id >= 0 && id < vocab) ? __ldg(table + (size_t)id * dim + (i - r * dim)) : 0.f;
  }
}

// feat[b][col + f] = max_{t <= L-k} relu(C[b*L + t][f] + bias[f]); arg[b][f] = the first maximising t
__global__ void __launch_bounds__(128)
textcnn_pool_fwd_kernel(const float* __restrict__ C, const float* __restrict__ bias, float* __restrict__ feat,
                        int32_t* __restrict__ arg, int B, int L, int N, int k, int ldf, int col) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * N) return;
  const int b = i / N, f = i % N;
  const float bf = bias[f];
  float best = -INFINITY;
  int bi = 0;
  for (int t = 0; t + k <= L; ++t) {
    const float v = fmaxf(C[((size_t)b * L + t) * N + f] + bf, 0.f);
    if (v > best) { best = v; bi = t; }
  }
  feat[(size_t)b * ldf + col + f] = best;
  arg[i] = bi;
}

__global__ void __launch_bounds__(256)
textcnn_pool_bwd_kernel(const float* __restrict__ dfeat, const float* __restrict__ feat, const int32_t* __restrict__ arg,
                        float* __restrict__ dC, int B, int L, int N, int ldf, int col) {
  const size_t total = (size_t)B * L * N;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int f = (int)(i % N);
    const size_t bt = i / N;
    const int t = (int)(bt % L), b = (int)(bt / L);
    const size_t o = (size_t)b * ldf + col + f;
    dC[i] = (arg[b * N + f] == t && feat[o] > 0.f) ? dfeat[o] : 0.f;
  }
}

}  // namespace

// ================================================================================================
// C ABI
// ================================================================================================
extern "C" int b200_sgemm(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, long long sam,
                          long long sak, long long sbk, long long sbn, int ldc, float beta, int relu, void* stream) {
  B200_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0 && ldc >= N, "bad sgemm argument");
  GemmArgs g;
  g.A = A; g.B = B; g.C = C; g.bias = bias; g.M = M; g.N = N; g.K = K; g.ldc = ldc; g.relu = relu;
  g.sam = sam; g.sak = sak; g.sbk = sbk; g.sbn = sbn; g.beta = beta;
  const int tiles = ((M + GBM - 1) / GBM) * ((N + GBN - 1) / GBN);
  int splitk = 1;
  if (!relu && beta == 0.f && ldc == N && K >= 1024 && tiles < b200_num_sms() / 2) {      // few tiles, long K: spread K over CTAs
    splitk = b200_num_sms() / tiles;
    if (splitk > K / 256) splitk = K / 256;
    if (splitk < 1) splitk = 1;
  }
  g.kchunk = ((K + splitk - 1) / splitk + GBK - 1) / GBK * GBK;
  splitk = (K + g.kchunk - 1) / g.kchunk;
  g.splitk = splitk;
  cudaStream_t s = (cudaStream_t)stream;
  if (splitk > 1) B200_CHECK_CUDA(cudaMemsetAsync(C, 0, (size_t)M * N * sizeof(float), s));
  sgemm_kernel<<<dim3((N + GBN - 1) / GBN, (M + GBM - 1) / GBM, splitk), 256, 0, s>>>(g);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_lstm_rec_fwd(const float* pre, const float* w_hh, float* hseq, float* gates, float* cseq, int B, int T,
                                 int H, void* stream) {
  B200_REQUIRE(pre && w_hh && hseq && gates && cseq && B > 0 && T > 0, "bad lstm argument");
  B200_REQUIRE(H == 100, "lstm kernels are instantiated for hidden size 100 (audio_rnn_model.py:11)");
  lstm_rec_fwd_kernel<100><<<B, 416, 0, (cudaStream_t)stream>>>(pre, w_hh, hseq, gates, cseq, T);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_lstm_rec_bwd(const float* dh, const float* gates, const float* cseq, const float* w_hh, float* dgates,
                                 int B, int T, int H, void* stream) {
  B200_REQUIRE(dh && gates && cseq && w_hh && dgates && B > 0 && T > 0, "bad lstm argument");
  B200_REQUIRE(H == 100, "lstm kernels are instantiated for hidden size 100 (audio_rnn_model.py:11)");
  lstm_rec_bwd_kernel<100><<<B, 416, 0, (cudaStream_t)stream>>>(dh, gates, cseq, w_hh, dgates, T);
  B200_CHECK_LAUNCH();
  return 0;
}

// x [B][L] fp32 waveform -> out [B][1 + L/512][n_mels]; window [2048] (Hann), melfb [n_mels][1025]
extern "C" int b200_mel_frontend(const float* x, const float* window, const float* melfb, float* out, int B, int L,
                                 int n_mels, void* stream) {
  B200_REQUIRE(x && window && melfb && out && B > 0 && n_mels > 0, "bad mel_frontend argument");
  B200_REQUIRE(L > NFFT / 2, "clip shorter than the reflect padding (1024 samples)");
  const int frames = 1 + L / HOP;
  mel_frontend_kernel<<<dim3(frames, B), 256, 0, (cudaStream_t)stream>>>(x, window, melfb, out, L, frames, n_mels);
  B200_CHECK_LAUNCH();
  return 0;
}

// dx [B][L] (zeroed by the caller, accumulated into) = d(loss)/d(waveform) given dout [B][1 + L/512][n_mels]
extern "C" int b200_mel_frontend_bwd(const float* x, const float* window, const float* melfb, const float* dout, float* dx,
                                     int B, int L, int n_mels, void* stream) {
  B200_REQUIRE(x && window && melfb && dout && dx && B > 0 && n_mels > 0 && n_mels <= 64, "bad mel_frontend_bwd argument");
  B200_REQUIRE(L > NFFT / 2, "clip shorter than the reflect padding (1024 samples)");
  const int frames = 1 + L / HOP;
  mel_frontend_bwd_kernel<<<dim3(frames, B), 256, 0, (cudaStream_t)stream>>>(x, window, melfb, dout, dx, L, frames, n_mels);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_gather_rows(const float* table, const int64_t* idx, float* out, long long rows, int dim, long long vocab,
                                void* stream) {
  B200_REQUIRE(table && idx && out && rows > 0 && dim > 0, "bad gather argument");
  gather_rows_kernel<<<b200_grid_for((size_t)rows * dim, 256, 8), 256, 0, (cudaStream_t)stream>>>(table, idx, out, (size_t)rows,
                                                                                                  dim, vocab);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_textcnn_pool_fwd(const float* C, const float* bias, float* feat, int32_t* arg, int B, int L, int N, int k,
                                     int ldf, int col, void* stream) {
  B200_REQUIRE(C && bias && feat && arg && B > 0 && N > 0 && k > 0 && L >= k, "bad textcnn argument (sequence shorter than the kernel?)");
  textcnn_pool_fwd_kernel<<<(B * N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(C, bias, feat, arg, B, L, N, k, ldf, col);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_textcnn_pool_bwd(const float* dfeat, const float* feat, const int32_t* arg, float* dC, int B, int L, int N,
                                     int ldf, int col, void* stream) {
  B200_REQUIRE(dfeat && feat && arg && dC && B > 0 && N > 0 && L > 0, "bad textcnn argument");
  textcnn_pool_bwd_kernel<<<b200_grid_for((size_t)B * L * N, 256, 8), 256, 0, (cudaStream_t)stream>>>(dfeat, feat, arg, dC, B, L, N,
                                                                                                      ldf, col);
  B200_CHECK_LAUNCH();
  return 0;
}
