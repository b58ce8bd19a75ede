// Fused optimizers over flat fp32 arenas.
//
// Replaces ATen's foreach SGD-with-momentum (notebooks/code/cifar10-distributed-native-cpu.py:144,156;
// 62 tensors / 11.69 M elements for resnet18) and Adam (notebooks/code/utils_basic.py:96).  Because
// b200ddp keeps parameters, gradients and optimizer state in flat arenas with identical layouts, the
// "multi-tensor" step is ONE streaming kernel: read p,g,m (+v) once, write p,m (+v), the bf16 shadow
// weights the tensor-core kernels consume, and (optionally) zero the gradient arena in the same pass.
#include "common.cuh"

namespace {

template <bool kShadow, bool kZero, bool kNesterov>
__global__ void __launch_bounds__(256)
sgd_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ m, uint2* __restrict__ shadow,
           size_t n4, float lr, float momentum, float one_minus_damp, float wd, float gscale) {
  pdl_trigger();
  pdl_wait();
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pv = p[i], gv = g[i];
    float pa[4] = {pv.x, pv.y, pv.z, pv.w};
    float ga[4] = {gv.x, gv.y, gv.z, gv.w};
    float ma[4] = {0.f, 0.f, 0.f, 0.f};
    if (momentum != 0.f) {
      float4 mv = m[i];
      ma[0] = mv.x; ma[1] = mv.y; ma[2] = mv.z; ma[3] = mv.w;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float d = fmaf(ga[k], gscale, wd * pa[k]);
      if (momentum != 0.f) {
        ma[k] = fmaf(momentum, ma[k], one_minus_damp * d);
        d = kNesterov ? fmaf(momentum, ma[k], d) : ma[k];
      }
      pa[k] = fmaf(-lr, d, pa[k]);
    }
    p[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    if (momentum != 0.f) m[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    if (kShadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pa[0], pa[1]), hi = __floats2bfloat162_rn(pa[2], pa[3]);
      uint2 s;
      s.x = *reinterpret_cast<uint32_t*>(&lo);
      s.y = *reinterpret_cast<uint32_t*>(&hi);
      shadow[i] = s;
    }
    if (kZero) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

template <bool kShadow, bool kZero>
__global__ void __launch_bounds__(256)
adam_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v,
            uint2* __restrict__ shadow, size_t n4, float lr, float b1, float b2, float eps, float wd,
            const int* __restrict__ step_dev, float gscale) {
  pdl_trigger();
  pdl_wait();
  const float t = (float)(*step_dev);
  const float bc1 = 1.f - powf(b1, t), bc2 = 1.f - powf(b2, t);
  const float step_size = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 pv = p[i], gv = g[i], mv = m[i], vv = v[i];
    float pa[4] = {pv.x, pv.y, pv.z, pv.w}, ga[4] = {gv.x, gv.y, gv.z, gv.w};
    float ma[4] = {mv.x, mv.y, mv.z, mv.w}, va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float d = fmaf(ga[k], gscale, wd * pa[k]);
      ma[k] = fmaf(b1, ma[k], (1.f - b1) * d);
      va[k] = fmaf(b2, va[k], (1.f - b2) * d * d);
      float denom = sqrtf(va[k]) * inv_sqrt_bc2 + eps;
      pa[k] -= step_size * ma[k] / denom;
    }
    p[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    m[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    v[i] = make_float4(va[0], va[1], va[2], va[3]);
    if (kShadow) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pa[0], pa[1]), hi = __floats2bfloat162_rn(pa[2], pa[3]);
      uint2 s;
      s.x = *reinterpret_cast<uint32_t*>(&lo);
      s.y = *reinterpret_cast<uint32_t*>(&hi);
      shadow[i] = s;
    }
    if (kZero) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void step_inc_kernel(int* s) {
  pdl_trigger();
  pdl_wait(); *s += 1; }

__global__ void __launch_bounds__(256) cast_kernel(const float4* __restrict__ src, uint2* __restrict__ dst, size_t n4) {
  pdl_trigger();
  pdl_wait();
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 a = src[i];
    __nv_bfloat162 lo = __floats2bfloat162_rn(a.x, a.y), hi = __floats2bfloat162_rn(a.z, a.w);
    uint2 s;
    s.x = *reinterpret_cast<uint32_t*>(&lo);
    s.y = *reinterpret_cast<uint32_t*>(&hi);
    dst[i] = s;
  }
}

}  // namespace

extern "C" int b200_sgd_momentum(float* p, float* g, float* m, void* shadow_bf16, size_t n, float lr, float momentum,
                                 float dampening, float weight_decay, int nesterov, float grad_scale, int zero_grad,
                                 void* stream) {
  B200_REQUIRE(p && g, "null p/g");
  B200_REQUIRE(momentum == 0.f || m != nullptr, "momentum buffer required");
  B200_REQUIRE(n % 4 == 0, "flat arenas are padded to a multiple of 4 elements");
  B200_REQUIRE(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m) % 16 == 0 && ((uintptr_t)shadow_bf16) % 8 == 0, "alignment");
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  const unsigned grid = b200_grid_for(n4, 256, 8);
  cudaStream_t s = (cudaStream_t)stream;
  float4 *p4 = (float4*)p, *g4 = (float4*)g, *m4 = (float4*)m;
  uint2* sh = (uint2*)shadow_bf16;
  const float omd = 1.f - dampening;
#define B200_SGD(S, Z, N) b200_launch(sgd_kernel<S, Z, N>, dim3(grid), dim3(256), 0, s, p4, g4, m4, sh, n4, lr, momentum, omd, weight_decay, grad_scale)
  const bool S = sh != nullptr, Z = zero_grad != 0, N = nesterov != 0;
  if (S && Z && N) B200_SGD(true, true, true);
  else if (S && Z) B200_SGD(true, true, false);
  else if (S && N) B200_SGD(true, false, true);
  else if (S) B200_SGD(true, false, false);
  else if (Z && N) B200_SGD(false, true, true);
  else if (Z) B200_SGD(false, true, false);
  else if (N) B200_SGD(false, false, true);
  else B200_SGD(false, false, false);
#undef B200_SGD
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_adam(float* p, float* g, float* m, float* v, void* shadow_bf16, size_t n, float lr, float beta1,
                         float beta2, float eps, float weight_decay, const int* step_dev, float grad_scale,
                         int zero_grad, void* stream) {
  B200_REQUIRE(p && g && m && v && step_dev, "null argument");
  B200_REQUIRE(n % 4 == 0, "flat arenas are padded to a multiple of 4 elements");
  if (n == 0) return 0;
  const size_t n4 = n / 4;
  const unsigned grid = b200_grid_for(n4, 256, 8);
  cudaStream_t s = (cudaStream_t)stream;
  uint2* sh = (uint2*)shadow_bf16;
#define B200_ADAM(S, Z) \
  b200_launch(adam_kernel<S, Z>, dim3(grid), dim3(256), 0, s, (float4*)p, (float4*)g, (float4*)m, (float4*)v, sh, n4, lr, beta1, beta2, eps, weight_decay, step_dev, grad_scale)
  const bool S = sh != nullptr, Z = zero_grad != 0;
  if (S && Z) B200_ADAM(true, true);
  else if (S) B200_ADAM(true, false);
  else if (Z) B200_ADAM(false, true);
  else B200_ADAM(false, false);
#undef B200_ADAM
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_step_inc(int* step_dev, void* stream) {
  B200_REQUIRE(step_dev, "null step");
  b200_launch(step_inc_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream, step_dev);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_cast_f32_to_bf16(const float* src, void* dst_bf16, size_t n, void* stream) {
  B200_REQUIRE(src && dst_bf16, "null argument");
  B200_REQUIRE(n % 4 == 0, "n must be a multiple of 4");
  if (n == 0) return 0;
  b200_launch(cast_kernel, dim3(b200_grid_for(n / 4, 256, 8)), dim3(256), 0, (cudaStream_t)stream, (const float4*)src, (uint2*)dst_bf16, n / 4);
  B200_CHECK_LAUNCH();
  return 0;
}
