// Eight consecutive channels of one NHWC activation row, for either activation dtype.
//
// The elementwise / normalisation / pooling kernels of this library all work on 8-channel units (one 16-byte vector
// of bf16).  With fp32 activations (the precision of the reference's own training run,
// notebooks/code/cifar10-distributed-smddp-gpu.py:145 — no autocast) the same unit is two 16-byte vectors.  The kernels
// are written once over the element type T in terms of the helpers below and instantiated for __nv_bfloat16 and float;
// the exported C entry points come in pairs, `b200_xxx` (bf16) and `b200_xxx_f32`.
#pragma once
#include "common.cuh"

struct alignas(16) f32x8 {
  float4 lo, hi;
};

template <typename T> struct Act8Sel;
template <> struct Act8Sel<__nv_bfloat16> { using type = bf16x8; };
template <> struct Act8Sel<float> { using type = f32x8; };
template <typename T> using Act8 = typename Act8Sel<T>::type;

__device__ __forceinline__ bf16x8 ld_act8(const __nv_bfloat16* p) { return *reinterpret_cast<const bf16x8*>(p); }
__device__ __forceinline__ f32x8 ld_act8(const float* p) {
  f32x8 v;
  v.lo = *reinterpret_cast<const float4*>(p);
  v.hi = *reinterpret_cast<const float4*>(p + 4);
  return v;
}
__device__ __forceinline__ void act8_to_float(const bf16x8& in, float* out) { bf16x8_to_float(in, out); }
__device__ __forceinline__ void act8_to_float(const f32x8& in, float* out) {
  out[0] = in.lo.x; out[1] = in.lo.y; out[2] = in.lo.z; out[3] = in.lo.w;
  out[4] = in.hi.x; out[5] = in.hi.y; out[6] = in.hi.z; out[7] = in.hi.w;
}
__device__ __forceinline__ void st_act8(__nv_bfloat16* p, const float* f) { *reinterpret_cast<bf16x8*>(p) = float_to_bf16x8(f); }
__device__ __forceinline__ void st_act8(float* p, const float* f) {
  *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(f[4], f[5], f[6], f[7]);
}
