// Fused losses and evaluation metrics.
//
// softmax_xent: log-softmax + NLL + mean + gradient in one pass over the logits; replaces ATen's
//   log_softmax / nll_loss forward+backward chain behind nn.CrossEntropyLoss
//   (notebooks/code/cifar10-distributed-native-cpu.py:143,152; model_lib/*_model.py `loss`).
// bce_logits:  binary_cross_entropy_with_logits fwd+bwd (model_lib/rtNLP_cnn_model.py:64,
//   meta_classifier.py:30).
// eval_count:  argmax + eq + sum, plus the reference's `F.nll_loss(logits, target, size_average=False)`
//   sum on raw logits for log-line parity (…native-cpu.py:185-187).
#include "common.cuh"

namespace {

template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f(float v);
template <>
__device__ __forceinline__ float from_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

// one warp per row; 4 rows per CTA.  Rows are short (10..1000 logits) so the three passes hit L1.
// The last CTA to finish (ticket counter in scratch) sums row_loss in index order -> deterministic mean.
template <typename T>
__global__ void __launch_bounds__(128)
softmax_xent_kernel(const T* __restrict__ logits, const int64_t* __restrict__ targets, int rows, int cols,
                    float* __restrict__ row_loss, float* __restrict__ loss_mean, T* __restrict__ dlogits,
                    float grad_scale, unsigned int* __restrict__ counter) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int row = blockIdx.x * 4 + w;
  if (row < rows) {
    const T* x = logits + (size_t)row * cols;
    float mx = -INFINITY;
    for (int c = lane; c < cols; c += 32) mx = fmaxf(mx, to_f(x[c]));
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < cols; c += 32) se += __expf(to_f(x[c]) - mx);
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    const int t = (int)targets[row];
    if (lane == 0) row_loss[row] = (t >= 0 && t < cols) ? lse - to_f(x[t]) : 0.f;
    if (dlogits != nullptr) {
      T* dx = dlogits + (size_t)row * cols;
      const float inv = 1.f / se;
      for (int c = lane; c < cols; c += 32) {
        float p = __expf(to_f(x[c]) - mx) * inv;
        dx[c] = from_f<T>((p - (c == t ? 1.f : 0.f)) * grad_scale);
      }
    }
  }
  // ---- deterministic mean by the last CTA
  __shared__ bool is_last;
  __shared__ float red[32];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned ticket = atomicAdd(counter, 1u);
    is_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    float s = 0.f;
    for (int r = threadIdx.x; r < rows; r += blockDim.x) s += __ldcg(row_loss + r);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
      *loss_mean = s / (float)rows;
      *counter = 0u;  // self-resetting scratch: ready for the next launch / graph replay
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
bce_logits_kernel(const T* __restrict__ x, const float* __restrict__ t, int n, float* __restrict__ loss_mean,
                  T* __restrict__ dx, float grad_scale) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = to_f(x[i]), y = t[i];
    s += fmaxf(v, 0.f) - v * y + log1pf(__expf(-fabsf(v)));
    if (dx != nullptr) dx[i] = from_f<T>((1.f / (1.f + __expf(-v)) - y) * grad_scale);
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) *loss_mean = s / (float)n;
}

template <typename T>
__global__ void __launch_bounds__(128)
eval_count_kernel(const T* __restrict__ logits, const int64_t* __restrict__ targets, int rows, int cols,
                  int* __restrict__ correct, float* __restrict__ nll_sum) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int row = blockIdx.x * 4 + w;
  if (row >= rows) return;
  const T* x = logits + (size_t)row * cols;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int c = lane; c < cols; c += 32) {
    const float v = to_f(x[c]);
    if (v > best) { best = v; bi = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) {
    const int t = (int)targets[row];
    if (bi == t) atomicAdd(correct, 1);
    if (nll_sum != nullptr && t >= 0 && t < cols) atomicAdd(nll_sum, -to_f(x[t]));
  }
}

}  // namespace

extern "C" int b200_softmax_xent(const void* logits, int dtype, const int64_t* targets, int rows, int cols,
                                 float* row_loss, float* loss_mean, void* dlogits, float grad_scale,
                                 unsigned int* scratch_counter, void* stream) {
  B200_REQUIRE(logits && targets && row_loss && loss_mean && scratch_counter, "null argument");
  B200_REQUIRE(rows > 0 && cols > 0, "empty problem");
  const unsigned grid = (rows + 3) / 4;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B200_DTYPE_F32)
    b200_launch(softmax_xent_kernel<float>, dim3(grid), dim3(128), 0, s, (const float*)logits, targets, rows, cols, row_loss, loss_mean,
                                                   (float*)dlogits, grad_scale, scratch_counter);
  else if (dtype == B200_DTYPE_BF16)
    b200_launch(softmax_xent_kernel<__nv_bfloat16>, dim3(grid), dim3(128), 0, s, (const __nv_bfloat16*)logits, targets, rows, cols, row_loss,
                                                           loss_mean, (__nv_bfloat16*)dlogits, grad_scale,
                                                           scratch_counter);
  else {
    b200_set_error("softmax_xent: bad dtype %d", dtype);
    return B200_ERR_ARG;
  }
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_bce_logits(const void* logits, int dtype, const float* targets, int n, float* loss_mean,
                               void* dlogits, float grad_scale, void* stream) {
  B200_REQUIRE(logits && targets && loss_mean && n > 0, "bad argument");
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B200_DTYPE_F32)
    b200_launch(bce_logits_kernel<float>, dim3(1), dim3(256), 0, s, (const float*)logits, targets, n, loss_mean, (float*)dlogits, grad_scale);
  else
    b200_launch(bce_logits_kernel<__nv_bfloat16>, dim3(1), dim3(256), 0, s, (const __nv_bfloat16*)logits, targets, n, loss_mean,
                                                      (__nv_bfloat16*)dlogits, grad_scale);
  B200_CHECK_LAUNCH();
  return 0;
}

extern "C" int b200_eval_count(const void* logits, int dtype, const int64_t* targets, int rows, int cols,
                               int* correct_out, float* nll_sum_out, void* stream) {
  B200_REQUIRE(logits && targets && correct_out && rows > 0 && cols > 0, "bad argument");
  const unsigned grid = (rows + 3) / 4;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B200_DTYPE_F32)
    b200_launch(eval_count_kernel<float>, dim3(grid), dim3(128), 0, s, (const float*)logits, targets, rows, cols, correct_out, nll_sum_out);
  else
    b200_launch(eval_count_kernel<__nv_bfloat16>, dim3(grid), dim3(128), 0, s, (const __nv_bfloat16*)logits, targets, rows, cols,
                                                         correct_out, nll_sum_out);
  B200_CHECK_LAUNCH();
  return 0;
}
