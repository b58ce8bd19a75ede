// Pooling, ReLU backward, GPU-side augmentation and trojan stamping.
//
//  maxpool / avgpool  NHWC bf16; replace ATen MaxPool2d / AdaptiveAvgPool2d (torchvision resnet18 stem
//                     + head; nn.MaxPool2d(2,2) in the workshop Net, …native-cpu.py:26,33-34).
//  augment            RandomCrop(32, padding=4) + RandomHorizontalFlip + ToTensor + Normalize of the
//                     reference's transform (…native-cpu.py:42-49) as ONE kernel on a uint8 batch.
//  trojan_stamp       batched alpha-blend of a trigger pattern (model_lib/cifar10_cnn_model.py:77-84,
//                     mnist_cnn_model.py:74-81, audio_rnn_model.py:77-83).
#include "common.cuh"
#include "act8.cuh"

namespace {

// one thread per (n, ho, wo, 8-channel vector)
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ amax,
                   int N, int H, int W, int C, int Ho, int Wo, int k, int stride, int pad) {
  pdl_trigger();
  pdl_wait();
  const int CV = C / 8;
  const size_t total = (size_t)N * Ho * Wo * CV;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    size_t t = i / CV;
    const int wo = (int)(t % Wo);
    t /= Wo;
    const int ho = (int)(t % Ho);
    const int n = (int)(t / Ho);
    float best[8];
    uint8_t bi[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) { best[q] = -INFINITY; bi[q] = 0; }
    for (int r = 0; r < k; ++r) {
      const int h = ho * stride - pad + r;
      if (h < 0 || h >= H) continue;
      for (int s = 0; s < k; ++s) {
        const int w = wo * stride - pad + s;
        if (w < 0 || w >= W) continue;
        float f[8];
        act8_to_float(ld_act8(x + (((size_t)n * H + h) * W + w) * C + cv * 8), f);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          if (f[q] > best[q]) { best[q] = f[q]; bi[q] = (uint8_t)(r * k + s); }
      }
    }
    st_act8(y + i * 8, best);
    if (amax) {
      uint2 packed;
      packed.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | (bi[3] << 24);
      packed.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | (bi[7] << 24);
      *reinterpret_cast<uint2*>(amax + i * 8) = packed;
    }
  }
}

// gather form: one thread per input (n,h,w,8ch); visits the <= ceil(k/stride)^2 windows that cover it
template <typename T>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ amax,
                   T* __restrict__ dx, int N, int H, int W, int C, int Ho, int Wo, int k, int stride,
                   int pad) {
  pdl_trigger();
  pdl_wait();
  const int CV = C / 8;
  const size_t total = (size_t)N * H * W * CV;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    size_t t = i / CV;
    const int w = (int)(t % W);
    t /= W;
    const int h = (int)(t % H);
    const int n = (int)(t / H);
    float acc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] = 0.f;
    const int ho_lo = max(0, (h + pad - k + stride) / stride), ho_hi = min(Ho - 1, (h + pad) / stride);
    const int wo_lo = max(0, (w + pad - k + stride) / stride), wo_hi = min(Wo - 1, (w + pad) / stride);
    for (int ho = ho_lo; ho <= ho_hi; ++ho) {
      const int r = h - (ho * stride - pad);
      if (r < 0 || r >= k) continue;
      for (int wo = wo_lo; wo <= wo_hi; ++wo) {
        const int s = w - (wo * stride - pad);
        if (s < 0 || s >= k) continue;
        const size_t o = ((((size_t)n * Ho + ho) * Wo + wo) * CV + cv) * 8;
        const uint2 packed = *reinterpret_cast<const uint2*>(amax + o);
        float g[8];
        act8_to_float(ld_act8(dy + o), g);
        const uint8_t me = (uint8_t)(r * k + s);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint8_t a = (uint8_t)(((q < 4 ? packed.x : packed.y) >> (8 * (q & 3))) & 0xff);
          if (a == me) acc[q] += g[q];
        }
      }
    }
    st_act8(dx + i * 8, acc);
  }
}

// x [N][HW][C] -> y [N][C] (mean over HW); one thread per (n, 8ch)
template <typename T>
__global__ void __launch_bounds__(256)
avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int HW, int C) {
  pdl_trigger();
  pdl_wait();
  const int CV = C / 8;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * CV) return;
  const int cv = i % CV, n = i / CV;
  float acc[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) acc[q] = 0.f;
  for (int p = 0; p < HW; ++p) {
    float f[8];
    act8_to_float(ld_act8(x + ((size_t)n * HW + p) * C + cv * 8), f);
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] += f[q];
  }
  const float inv = 1.f / (float)HW;
#pragma unroll
  for (int q = 0; q < 8; ++q) acc[q] *= inv;
  st_act8(y + (size_t)n * C + cv * 8, acc);
}

template <typename T>
__global__ void __launch_bounds__(256)
avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int N, int HW, int C) {
  pdl_trigger();
  pdl_wait();
  const int CV = C / 8;
  const size_t total = (size_t)N * HW * CV;
  const float inv = 1.f / (float)HW;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % CV);
    const int n = (int)(i / ((size_t)HW * CV));
    float f[8];
    act8_to_float(ld_act8(dy + (size_t)n * C + cv * 8), f);
#pragma unroll
    for (int q = 0; q < 8; ++q) f[q] *= inv;
    st_act8(dx + i * 8, f);
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
relu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y, T* __restrict__ dx, size_t n8) {
  pdl_trigger();
  pdl_wait();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    float g[8], v[8];
    act8_to_float(ld_act8(dy + i * 8), g);
    act8_to_float(ld_act8(y + i * 8), v);
#pragma unroll
    for (int q = 0; q < 8; ++q) g[q] = v[q] > 0.f ? g[q] : 0.f;
    st_act8(dx + i * 8, g);
  }
}

// layout 0: out NHWC bf16 with C_out >= C channels (extra channels zero) ; layout 1: out NCHW fp32 (C_out == C) ;
// layout 2: out NHWC fp32 with C_out >= C channels (the fp32 activation path)
__global__ void __launch_bounds__(256)
augment_kernel(const uint8_t* __restrict__ img, const int32_t* __restrict__ offsets, void* __restrict__ out, int N,
               int H, int W, int C, int C_out, int pad, const float* __restrict__ mean,
               const float* __restrict__ inv_std, int layout) {
  pdl_trigger();
  pdl_wait();
  const size_t total = (size_t)N * H * W;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int w = (int)(i % W);
    const int h = (int)((i / W) % H);
    const int n = (int)(i / ((size_t)W * H));
    int dy = pad, dx = pad, flip = 0;
    if (offsets != nullptr) {
      dy = offsets[n * 3 + 0];
      dx = offsets[n * 3 + 1];
      flip = offsets[n * 3 + 2];
    }
    const int wc = flip ? (W - 1 - w) : w;           // flip acts on the cropped image
    const int sh = h + dy - pad, sw = wc + dx - pad;  // source pixel in the unpadded image
    const bool inside = sh >= 0 && sh < H && sw >= 0 && sw < W;
    const uint8_t* src = img + (((size_t)n * H + sh) * W + sw) * C;
    if (layout == 2) {
      float* o = reinterpret_cast<float*>(out) + i * C_out;
      for (int c = 0; c < C_out; ++c) {
        float v = 0.f;
        if (c < C) v = ((inside ? (float)src[c] * (1.f / 255.f) : 0.f) - mean[c]) * inv_std[c];
        o[c] = v;
      }
    } else if (layout == 0) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out) + i * C_out;
      for (int c = 0; c < C_out; ++c) {
        float v = 0.f;
        if (c < C) v = ((inside ? (float)src[c] * (1.f / 255.f) : 0.f) - mean[c]) * inv_std[c];
        o[c] = __float2bfloat16(v);
      }
    } else {
      float* o = reinterpret_cast<float*>(out);
      for (int c = 0; c < C; ++c)
        o[(((size_t)n * C + c) * H + h) * W + w] = ((inside ? (float)src[c] * (1.f / 255.f) : 0.f) - mean[c]) * inv_std[c];
    }
  }
}

// x [N][C][H][W] fp32 ; pattern [PC][ph][pw] with PC in {1, C} (PC==1 stamps channel 0 only, as MNIST does)
__global__ void __launch_bounds__(256)
trojan_stamp_kernel(float* __restrict__ x, int N, int C, int H, int W, const float* __restrict__ pattern, int PC,
                    int ph, int pw, int loc_h, int loc_w, float alpha, const uint8_t* __restrict__ mask) {
  pdl_trigger();
  pdl_wait();
  const size_t total = (size_t)N * PC * ph * pw;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int s = (int)(i % pw);
    const int r = (int)((i / pw) % ph);
    const int c = (int)((i / ((size_t)pw * ph)) % PC);
    const int n = (int)(i / ((size_t)pw * ph * PC));
    if (mask != nullptr && mask[n] == 0) continue;
    const int h = loc_h + r, w = loc_w + s;
    if (h >= H || w >= W) continue;
    float* p = x + (((size_t)n * C + c) * H + h) * W + w;
    *p = alpha * pattern[((size_t)c * ph + r) * pw + s] + (1.f - alpha) * (*p);
  }
}


// dbias[c] += sum_rows dy[row][c]   (bias gradient of conv / linear layers; fp32 atomics into the arena)
template <typename T>
__global__ void __launch_bounds__(256)
colsum_accum_kernel(const T* __restrict__ dy, size_t rows, int C, int c_valid, float* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sm[32][64];
  const int v = threadIdx.x & 7, rl = threadIdx.x >> 3;
  const int c = blockIdx.y * 64 + v * 8;
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  if (c < C) {
    for (size_t r = (size_t)blockIdx.x * 32 + rl; r < rows; r += (size_t)gridDim.x * 32) {
      float f[8];
      act8_to_float(ld_act8(dy + r * C + c), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) sm[rl][v * 8 + k] = acc[k];
  __syncthreads();
  if (threadIdx.x < 64) {
    float t = 0.f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) t += sm[r][threadIdx.x];
    const int ch = blockIdx.y * 64 + threadIdx.x;
    if (ch < c_valid) atomicAdd(out + ch, t);
  }
}

}  // namespace

// Every activation-touching entry point exists for both activation dtypes: `name` (NHWC bf16) and `name_f32` (NHWC fp32).
#define B200_POOL_BOTH(name, params, args)                                                \
  extern "C" int name params { return name##_impl<__nv_bfloat16> args; }                  \
  extern "C" int name##_f32 params { return name##_impl<float> args; }

template <typename T>
static int b200_maxpool_fwd_impl(const void* x, void* y, uint8_t* argmax, int N, int H, int W, int C, int k, int stride, int pad,
                                 void* stream) {
  B200_REQUIRE(x && y, "null argument");
  B200_REQUIRE(C % 8 == 0 && k * k <= 255 && stride > 0, "C must be a multiple of 8");
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const size_t total = (size_t)N * Ho * Wo * (C / 8);
  b200_launch(maxpool_fwd_kernel<T>, dim3(b200_grid_for(total, 256, 8)), dim3(256), 0, (cudaStream_t)stream, (const T*)x, (T*)y,
              argmax, N, H, W, C, Ho, Wo, k, stride, pad);
  B200_CHECK_LAUNCH();
  return 0;
}
B200_POOL_BOTH(b200_maxpool_fwd,
               (const void* x, void* y, uint8_t* argmax, int N, int H, int W, int C, int k, int stride, int pad, void* stream),
               (x, y, argmax, N, H, W, C, k, stride, pad, stream))

template <typename T>
static int b200_maxpool_bwd_impl(const void* dy, const uint8_t* argmax, void* dx, int N, int H, int W, int C, int k, int stride,
                                 int pad, void* stream) {
  B200_REQUIRE(dy && argmax && dx, "null argument");
  B200_REQUIRE(C % 8 == 0, "C must be a multiple of 8");
  const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
  const size_t total = (size_t)N * H * W * (C / 8);
  b200_launch(maxpool_bwd_kernel<T>, dim3(b200_grid_for(total, 256, 8)), dim3(256), 0, (cudaStream_t)stream, (const T*)dy, argmax,
              (T*)dx, N, H, W, C, Ho, Wo, k, stride, pad);
  B200_CHECK_LAUNCH();
  return 0;
}
B200_POOL_BOTH(b200_maxpool_bwd,
               (const void* dy, const uint8_t* argmax, void* dx, int N, int H, int W, int C, int k, int stride, int pad, void* stream),
               (dy, argmax, dx, N, H, W, C, k, stride, pad, stream))

template <typename T>
static int b200_avgpool_fwd_impl(const void* x, void* y, int N, int HW, int C, void* stream) {
  B200_REQUIRE(x && y && C % 8 == 0, "bad argument");
  const int total = N * (C / 8);
  b200_launch(avgpool_fwd_kernel<T>, dim3((total + 255) / 256), dim3(256), 0, (cudaStream_t)stream, (const T*)x, (T*)y, N, HW, C);
  B200_CHECK_LAUNCH();
  return 0;
}
B200_POOL_BOTH(b200_avgpool_fwd, (const void* x, void* y, int N, int HW, int C, void* stream), (x, y, N, HW, C, stream))

template <typename T>
static int b200_avgpool_bwd_impl(const void* dy, void* dx, int N, int HW, int C, void* stream) {
  B200_REQUIRE(dy && dx && C % 8 == 0, "bad argument");
  const size_t total = (size_t)N * HW * (C / 8);
  b200_launch(avgpool_bwd_kernel<T>, dim3(b200_grid_for(total, 256, 8)), dim3(256), 0, (cudaStream_t)stream, (const T*)dy, (T*)dx, N,
              HW, C);
  B200_CHECK_LAUNCH();
  return 0;
}
B200_POOL_BOTH(b200_avgpool_bwd, (const void* dy, void* dx, int N, int HW, int C, void* stream), (dy, dx, N, HW, C, stream))

template <typename T>
static int b200_relu_bwd_impl(const void* dy, const void* y, void* dx, size_t n, void* stream) {
  B200_REQUIRE(dy && y && dx && n % 8 == 0, "n must be a multiple of 8");
  if (n == 0) return 0;
  b200_launch(relu_bwd_kernel<T>, dim3(b200_grid_for(n / 8, 256, 8)), dim3(256), 0, (cudaStream_t)stream, (const T*)dy, (const T*)y,
              (T*)dx, n / 8);
  B200_CHECK_LAUNCH();
  return 0;
}
B200_POOL_BOTH(b200_relu_bwd, (const void* dy, const void* y, void* dx, size_t n, void* stream), (dy, y, dx, n, stream))

extern "C" int b200_augment(const uint8_t* img_u8, const int32_t* offsets, void* out, int N, int H, int W, int C,
                            int C_out, int pad, const float* mean, const float* inv_std, int layout, void* stream) {
  B200_REQUIRE(img_u8 && out && mean && inv_std, "null argument");
  B200_REQUIRE(layout == 0 || layout == 2 || (layout == 1 && C_out == C), "layout 1 (NCHW fp32) needs C_out == C");
  B200_REQUIRE(C_out >= C, "C_out < C");
  const size_t total = (size_t)N * H * W;
  b200_launch(augment_kernel, dim3(b200_grid_for(total, 256, 8)), dim3(256), 0, (cudaStream_t)stream, img_u8, offsets, out, N, H, W, C, C_out, pad, mean, inv_std, layout);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_trojan_stamp(float* x, int N, int C, int H, int W, const float* pattern, int pattern_channels,
                                 int p_h, int p_w, int loc_h, int loc_w, float alpha, const uint8_t* mask,
                                 void* stream) {
  B200_REQUIRE(x && pattern, "null argument");
  B200_REQUIRE(pattern_channels == 1 || pattern_channels == C, "pattern channels must be 1 or C");
  const size_t total = (size_t)N * pattern_channels * p_h * p_w;
  if (total == 0) return 0;
  b200_launch(trojan_stamp_kernel, dim3(b200_grid_for(total, 256, 8)), dim3(256), 0, (cudaStream_t)stream, x, N, C, H, W, pattern, pattern_channels, p_h, p_w, loc_h, loc_w, alpha, mask);
  B200_CHECK_LAUNCH();
  return 0;
}

template <typename T>
static int b200_colsum_accum_impl(const void* dy, size_t rows, int C, int c_valid, float* out, void* stream) {
  B200_REQUIRE(dy && out && C % 8 == 0 && c_valid <= C, "bad argument");
  if (rows == 0) return 0;
  size_t rb = rows / 64;
  if (rb < 1) rb = 1;
  if (rb > 64) rb = 64;
  dim3 grid((unsigned)rb, (unsigned)((C + 63) / 64), 1);
  b200_launch(colsum_accum_kernel<T>, dim3(grid), dim3(256), 0, (cudaStream_t)stream, (const T*)dy, rows, C, c_valid, out);
  B200_CHECK_LAUNCH();
  return 0;
}
B200_POOL_BOTH(b200_colsum_accum, (const void* dy, size_t rows, int C, int c_valid, float* out, void* stream),
               (dy, rows, C, c_valid, out, stream))
