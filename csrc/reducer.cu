// Native gradient reducer: per-bucket ready counting, in-order launch of the fused all-reduce on the comm
// stream, event plumbing between the producer streams and the comm stream, end-of-backward join.
//
// This is the job torch's C++ `Reducer` does for the reference (DistributedDataParallel at
// notebooks/code/cifar10-distributed-native-cpu.py:141 and cifar10-distributed-smddp-gpu.py:148; job log
// "Reducer buckets have been rebuilt", notebooks/2_pytorch_dist_smddp_gpu.ipynb:1843), re-designed around
// the flat symmetric gradient arena: a bucket is a byte range, "copy into the bucket" does not exist, and
// launching a bucket is one kernel (csrc/comm.cu) behind one or two cudaStreamWaitEvent.  Everything here
// is stream-ordered and host-synchronisation free, so a whole backward pass — including these launches —
// can be captured in a CUDA graph.
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

struct b200_reducer {
  b200_comm comm;
  int dtype = B200_DTYPE_F32;
  float scale = 1.f;
  int max_blocks = 64;
  size_t oneshot_max = 64 * 1024, nvls_min = 64 * 1024;
  std::vector<int> param_bucket;
  std::vector<size_t> offset, count;       // per bucket: byte offset in the arena, element count
  std::vector<int> bucket_params, pending;
  std::vector<char> ready;                 // per parameter, this backward
  std::vector<int> algo_used;              // per bucket, last launch
  int next_bucket = 0;
  bool dirty = false;                      // something was launched on the comm stream since the last join
  cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_done = nullptr;
  // fused bucket step (b200_reducer_set_sgd): each bucket's launch is reduce-scatter + SGD + weight all-gather
  bool sgd = false;
  size_t grad_base = 0, param_base = 0, shadow_base = 0;   // arena byte offsets of element 0 of the flat arenas
  float* momentum_base = nullptr;
  float lr = 0.f, mom = 0.f, damp = 0.f, wd = 0.f;
  int nesterov = 0;
};

static size_t reducer_elem_size(int dtype) { return dtype == B200_DTYPE_BF16 ? 2 : 4; }

extern "C" b200_reducer* b200_reducer_create(const b200_comm* comm, int n_params, const int* param_bucket, int n_buckets,
                                             const size_t* bucket_offset_bytes, const size_t* bucket_count, int dtype,
                                             float scale, int max_blocks, size_t oneshot_max_bytes, size_t nvls_min_bytes) {
  if (!comm || n_params <= 0 || n_buckets <= 0 || !param_bucket || !bucket_offset_bytes || !bucket_count) {
    b200_set_error("b200_reducer_create: bad argument");
    return nullptr;
  }
  b200_reducer* r = new b200_reducer();
  r->comm = *comm;
  r->dtype = dtype;
  r->scale = scale;
  r->max_blocks = max_blocks;
  r->oneshot_max = oneshot_max_bytes;
  r->nvls_min = nvls_min_bytes;
  r->param_bucket.assign(param_bucket, param_bucket + n_params);
  r->offset.assign(bucket_offset_bytes, bucket_offset_bytes + n_buckets);
  r->count.assign(bucket_count, bucket_count + n_buckets);
  r->bucket_params.assign(n_buckets, 0);
  for (int i = 0; i < n_params; ++i) {
    if (param_bucket[i] < 0 || param_bucket[i] >= n_buckets) {
      b200_set_error("b200_reducer_create: parameter %d maps to bucket %d of %d", i, param_bucket[i], n_buckets);
      delete r;
      return nullptr;
    }
    r->bucket_params[param_bucket[i]]++;
  }
  r->pending = r->bucket_params;
  r->ready.assign(n_params, 0);
  r->algo_used.assign(n_buckets, 0);
  if (cudaEventCreateWithFlags(&r->ev_a, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&r->ev_b, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&r->ev_done, cudaEventDisableTiming) != cudaSuccess) {
    b200_set_error("b200_reducer_create: cudaEventCreate failed");
    delete r;
    return nullptr;
  }
  return r;
}

extern "C" void b200_reducer_destroy(b200_reducer* r) {
  if (!r) return;
  if (r->ev_a) cudaEventDestroy(r->ev_a);
  if (r->ev_b) cudaEventDestroy(r->ev_b);
  if (r->ev_done) cudaEventDestroy(r->ev_done);
  delete r;
}

extern "C" int b200_reducer_begin(b200_reducer* r) {
  B200_REQUIRE(r != nullptr, "null reducer");
  r->pending = r->bucket_params;
  std::fill(r->ready.begin(), r->ready.end(), 0);
  r->next_bucket = 0;
  return 0;
}

extern "C" int b200_reducer_was_ready(const b200_reducer* r, int param) {
  return (r && param >= 0 && param < (int)r->ready.size()) ? (int)r->ready[param] : 0;
}

static int reducer_pick_algo(const b200_reducer* r, size_t bytes) {
  return b200_pick_algo(r->comm.world, r->comm.mc != nullptr, bytes, r->oneshot_max, r->nvls_min);
}

// comm stream waits for everything queued so far on the producer stream(s), then runs the bucket's all-reduce
static int reducer_launch(b200_reducer* r, int b, cudaStream_t producer, cudaStream_t other, cudaStream_t comm) {
  B200_CHECK_CUDA(cudaEventRecord(r->ev_a, producer));
  B200_CHECK_CUDA(cudaStreamWaitEvent(comm, r->ev_a, 0));
  if (other != nullptr && other != producer) {
    B200_CHECK_CUDA(cudaEventRecord(r->ev_b, other));
    B200_CHECK_CUDA(cudaStreamWaitEvent(comm, r->ev_b, 0));
  }
  const size_t bytes = r->count[b] * reducer_elem_size(r->dtype);
  r->dirty = true;
  if (r->sgd) {
    // the bucket's position inside the flat gradient arena gives its position inside the parameter / shadow / momentum arenas
    const size_t elem0 = (r->offset[b] - r->grad_base) / sizeof(float);
    const int algo = (r->comm.mc != nullptr && r->comm.world > 2) ? B200_ALGO_NVLS : B200_ALGO_TWOSHOT;
    r->algo_used[b] = algo;
    // CTA budget: the fused step also streams the optimizer's traffic (p, m read/write of the owned shard, zeroing of the
    // bucket's gradients), so it gets more CTAs than a bare all-reduce; the LAST bucket completes with the last wgrad
    // kernel of backward, nothing is left to overlap with and it is on the critical path: it takes most of the machine.
    static const int fused_ctas = []() { const char* e = getenv("B200DDP_FUSED_CTAS"); return e ? atoi(e) : 16; }();
    static const int tail_ctas = []() { const char* e = getenv("B200DDP_FUSED_TAIL_CTAS"); return e ? atoi(e) : 64; }();
    const bool last = b + 1 == (int)r->count.size();
    const int blocks = std::max(r->max_blocks, last ? tail_ctas : fused_ctas);
    return b200_allreduce_sgd(&r->comm, r->offset[b], r->param_base + elem0 * sizeof(float),
                              r->shadow_base ? r->shadow_base + elem0 * 2 : 0, r->momentum_base ? r->momentum_base + elem0 : nullptr,
                              r->count[b], r->scale, r->lr, r->mom, r->damp, r->wd, r->nesterov, algo, blocks, comm);
  }
  const int algo = reducer_pick_algo(r, bytes);
  r->algo_used[b] = algo;
  // the last bucket has nothing left to overlap with: B200DDP_TAIL_CTAS (default 32: resnet18 step on 2 GPUs 1.098 -> 1.049 ms; on 8 GPUs it makes no difference)
  static const int tail_ctas = []() { const char* e = getenv("B200DDP_TAIL_CTAS"); return e ? atoi(e) : 32; }();
  if (tail_ctas > 0 && b + 1 == (int)r->count.size())
    return b200_allreduce(&r->comm, r->offset[b], r->count[b], r->dtype, r->scale, algo, std::max(r->max_blocks, tail_ctas), comm);
  return b200_allreduce(&r->comm, r->offset[b], r->count[b], r->dtype, r->scale, algo, r->max_blocks, comm);
}

// `producer`: the stream the gradient of `param` was written on; `other`: a second stream that may hold
// unfinished gradients of the same bucket (the wgrad aux stream / the main stream), or NULL.
// Returns the number of buckets launched by this call (>= 0) or a negative error.
extern "C" int b200_reducer_mark_ready(b200_reducer* r, int param, void* producer, void* other, void* comm_stream) {
  B200_REQUIRE(r != nullptr, "null reducer");
  B200_REQUIRE(param >= 0 && param < (int)r->ready.size(), "parameter index out of range");
  if (r->ready[param]) return 0;
  r->ready[param] = 1;
  if (r->comm.world <= 1) return 0;
  r->pending[r->param_bucket[param]]--;
  int launched = 0;
  while (r->next_bucket < (int)r->pending.size() && r->pending[r->next_bucket] <= 0) {   // static order on every rank
    int rc = reducer_launch(r, r->next_bucket, (cudaStream_t)producer, (cudaStream_t)other, (cudaStream_t)comm_stream);
    if (rc != 0) return rc < 0 ? rc : -rc;
    r->next_bucket++;
    launched++;
  }
  return launched;
}

// End of backward: flush buckets that never filled (unused parameters keep their zeroed slots), then make
// `main_stream` wait for the comm stream.
extern "C" int b200_reducer_finalize(b200_reducer* r, void* main_stream, void* other, void* comm_stream, int sync_grads) {
  B200_REQUIRE(r != nullptr, "null reducer");
  if (r->comm.world > 1 && sync_grads) {
    while (r->next_bucket < (int)r->pending.size()) {
      int rc = reducer_launch(r, r->next_bucket, (cudaStream_t)main_stream, (cudaStream_t)other, (cudaStream_t)comm_stream);
      if (rc != 0) return rc < 0 ? rc : -rc;
      r->next_bucket++;
    }
  }
  if (r->dirty) {
    B200_CHECK_CUDA(cudaEventRecord(r->ev_done, (cudaStream_t)comm_stream));
    B200_CHECK_CUDA(cudaStreamWaitEvent((cudaStream_t)main_stream, r->ev_done, 0));
    r->dirty = false;
  }
  return 0;
}

extern "C" int b200_reducer_bucket_algo(const b200_reducer* r, int bucket) {
  return (r && bucket >= 0 && bucket < (int)r->algo_used.size()) ? r->algo_used[bucket] : -1;
}

// Switch the reducer's per-bucket launch between the plain all-reduce and the fused bucket step (b200_allreduce_sgd).
// grad_base / param_base / shadow_base: arena byte offsets of element 0 of the flat gradient / fp32 parameter / bf16 shadow
// arenas (shadow_base 0 = none); momentum_base: this rank's local momentum buffer laid out like the parameters (or NULL).
// Hyper-parameters are passed by value: call again when the learning rate changes (and re-capture a CUDA graph).
extern "C" int b200_reducer_set_sgd(b200_reducer* r, int enabled, size_t grad_base, size_t param_base, size_t shadow_base,
                                    float* momentum_base, float lr, float momentum, float dampening, float weight_decay,
                                    int nesterov) {
  B200_REQUIRE(r != nullptr, "null reducer");
  B200_REQUIRE(!enabled || r->dtype == B200_DTYPE_F32, "the fused bucket step takes fp32 gradients");
  B200_REQUIRE(!enabled || momentum == 0.f || momentum_base != nullptr, "momentum buffer required");
  r->sgd = enabled != 0;
  r->grad_base = grad_base; r->param_base = param_base; r->shadow_base = shadow_base; r->momentum_base = momentum_base;
  r->lr = lr; r->mom = momentum; r->damp = dampening; r->wd = weight_decay; r->nesterov = nesterov;
  return 0;
}
