// Gradient-norm pieces of the step tail: sum of squares over a flat buffer, the
// clip_grad_norm_ coefficient, and the in-place scale.
//
// Replaces torch.nn.utils.clip_grad_norm_ as called at ref: ray-jobs/pytorch_llm_ray.py:277
// (max_norm 1.0) and by HF Trainer with MAX_GRAD_NORM (ref: ray-jobs/fine_tune_config.json:19).
// torch/nn/utils/clip_grad.py: total = ||(||g_k||_2)_k||_2 ; coef = clamp(max_norm/(total+1e-6), max=1) ;
// g_k *= coef. Here the per-tensor norms collapse into one flat sum of squares (fp32 per thread,
// double across blocks), which differs from torch's norm-of-norms only by rounding.
//
// HBM-bound: algorithmic bytes = n * sizeof(dtype) read (sumsq), + the same written (scale).
#include "b2_common.cuh"
#include <mutex>

// scratch layout for b2_sumsq: word 0 = arrival ticket (self-resetting), words [32, 32+blocks) = per-block partials
#define B2_SUMSQ_MAX_BLOCKS (B2_NUM_SMS * 8)
#define B2_SUMSQ_SCRATCH_WORDS (32 + B2_SUMSQ_MAX_BLOCKS)

namespace {

template <int DT>
__device__ __forceinline__ float sumsq_thread(const void* __restrict__ g, size_t n) {
  using G = Vec8<DT>;
  const size_t n8 = n >> 3;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  float acc0 = 0.f, acc1 = 0.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float x[8];
    G::template load<false>(g, i, x);
#pragma unroll
    for (int k = 0; k < 8; k += 2) { acc0 = fmaf(x[k], x[k], acc0); acc1 = fmaf(x[k + 1], x[k + 1], acc1); }
  }
  const size_t tail0 = n8 << 3;
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - tail0)) {
    float x = G::load1(g, tail0 + threadIdx.x);
    acc0 = fmaf(x, x, acc0);
  }
  return acc0 + acc1;
}

// Each block writes one partial; the last block to arrive adds them in index order (deterministic)
// and resets the ticket so the scratch can be reused by the next call on the same stream.
template <int DT>
__global__ void __launch_bounds__(256) sumsq_kernel(const void* __restrict__ g, size_t n, float* __restrict__ scratch,
                                                    float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  __shared__ bool last;
  float v = block_sum(sumsq_thread<DT>(g, n), red);
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
  float* partials = scratch + 32;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = v;
    __threadfence();
    last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double acc = 0.0;
  for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) acc += (double)__ldcg(partials + i);
  // block reduce in double via shuffles
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ double dred[8];
  if ((threadIdx.x & 31) == 0) dred[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += dred[w];
    out[0] = accumulate ? out[0] + (float)t : (float)t;
    *ticket = 0u;
  }
}

__global__ void clip_coef_kernel(const float* __restrict__ parts, int nparts, float max_norm, float* __restrict__ norm_out,
                                 float* __restrict__ coef_out) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double t = 0.0;
  for (int i = 0; i < nparts; ++i) t += (double)parts[i];
  const float total = (float)sqrt(t);
  if (norm_out) norm_out[0] = total;
  // clip_grad.py:165-183: clip_coef = max_norm / (total_norm + 1e-6); clamp(max=1.0)
  float coef = max_norm / (total + 1e-6f);
  coef = coef > 1.f ? 1.f : coef;          // NaN norm propagates, as in torch
  if (coef_out) coef_out[0] = coef;
}

template <int DT>
__global__ void __launch_bounds__(256) scale_kernel(void* __restrict__ g, size_t n, const float* __restrict__ coef) {
  using G = Vec8<DT>;
  const float c = __ldg(coef);
  const size_t n8 = n >> 3;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float x[8];
    G::template load<true>(g, i, x);
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = __fmul_rn(x[k], c);
    G::store(g, i, x);
  }
  const size_t tail0 = n8 << 3;
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - tail0)) {
    const size_t i = tail0 + threadIdx.x;
    G::store1(g, i, __fmul_rn(G::load1(g, i), c));
  }
}

unsigned grid_for(size_t n, size_t cap) {
  size_t blocks = ((n >> 3) + 255) / 256;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  return (unsigned)blocks;
}

}  // namespace

extern "C" size_t b2_sumsq_scratch_bytes(void) { return sizeof(float) * B2_SUMSQ_SCRATCH_WORDS; }

static int b2_sumsq_launch(const void* g, size_t n, int dtype, float* scratch, float* out, int accumulate, cudaStream_t stream) {
  const unsigned grid = grid_for(n, B2_SUMSQ_MAX_BLOCKS);
  if (dtype == B2_F32) sumsq_kernel<B2_F32><<<grid, 256, 0, stream>>>(g, n, scratch, out, accumulate);
  else sumsq_kernel<B2_BF16><<<grid, 256, 0, stream>>>(g, n, scratch, out, accumulate);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_sumsq(const void* g, size_t n, int dtype, float* scratch, float* out, int accumulate, cudaStream_t stream) {
  B2_REQUIRE(out && scratch, "b2_sumsq: null out/scratch");
  B2_REQUIRE(n == 0 || g, "b2_sumsq: null input");
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_sumsq: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(b2_aligned(g, 16) && b2_aligned(scratch, 16), "b2_sumsq: buffers must be 16-byte aligned");
  return b2_sumsq_launch(g, n, dtype, scratch, out, accumulate, stream);
}

extern "C" int b2_clip_coef(const float* sumsq_parts, int nparts, float max_norm, float* norm_out, float* coef_out,
                            cudaStream_t stream) {
  B2_REQUIRE(sumsq_parts && nparts >= 1, "b2_clip_coef: need at least one partial");
  clip_coef_kernel<<<1, 32, 0, stream>>>(sumsq_parts, nparts, max_norm, norm_out, coef_out);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_scale(void* g, size_t n, int dtype, const float* coef_dev, cudaStream_t stream) {
  if (n == 0) return B2_OK;
  B2_REQUIRE(g && coef_dev, "b2_scale: null pointer");
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_scale: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(b2_aligned(g, 16), "b2_scale: buffer must be 16-byte aligned");
  const unsigned grid = grid_for(n, (size_t)B2_NUM_SMS * 16);
  if (dtype == B2_F32) scale_kernel<B2_F32><<<grid, 256, 0, stream>>>(g, n, coef_dev);
  else scale_kernel<B2_BF16><<<grid, 256, 0, stream>>>(g, n, coef_dev);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// b2_sumsq_static — the same sum of squares for callers that have no way to hand us scratch memory.
//
// Who that is: a host that only holds the pointers it was given (a C, Go or JVM binding with no device allocator of
// its own); a process whose framework allocator has already claimed the whole GPU, so a raw cudaMalloc from a side
// library fails or fragments the pool; and code running where an allocation's implicit device-wide synchronisation
// is not acceptable, such as a gradient hook called from the autograd engine's thread in the middle of backward.
// For them the per-block partials and the arrival ticket live in the fixed array below, which is part of the library
// image: every GPU that loads libb2train.so carries these few KB until the process exits.
//
// Limits and rules:
//   * at most B2_SUMSQ_STATIC_MAX_N elements per call (larger inputs are rejected with B2_ERR_LIMIT — split the buffer
//     and use accumulate=1 for the later pieces);
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress, that device current)
//     and remembers it; make that first call before starting any stream capture;
//   * all calls on a device share the one array, which still holds whatever the previous call left there. The only
//     word that has to start at zero, the arrival ticket, is cleared at the start of every call; the partials are
//     overwritten before they are read;
//   * because the array is shared, two calls that may run on the GPU at the same time (different streams, or different
//     host threads on one device) race on it. Issue them on one stream, or order them with events.
// Nothing here calls cudaMalloc, cudaFree or any routine that allocates.
#define B2_SUMSQ_STATIC_BLOCKS (B2_NUM_SMS * 4)
#define B2_SUMSQ_STATIC_WORDS (32 + B2_SUMSQ_STATIC_BLOCKS)      /* 624 floats = 2496 bytes */
#define B2_SUMSQ_STATIC_MAX_N ((size_t)1 << 31)
#define B2_MAX_DEVICES 64

static __device__ __align__(256) float g_sumsq_static_scratch[B2_SUMSQ_STATIC_WORDS];

// In host code the name g_sumsq_static_scratch is only a handle for the symbol, not an address the GPU can use;
// the real device address has to be asked for, once per device.
static float* sumsq_static_scratch_ptr(int* rc_out) {
  static float* ptrs[B2_MAX_DEVICES] = {nullptr};
  static std::mutex mu;
  *rc_out = B2_OK;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= B2_MAX_DEVICES) {
    b2_set_error("b2_sumsq_static: cannot identify the current device (%s)", e == cudaSuccess ? "index out of range" : cudaGetErrorString(e));
    *rc_out = B2_ERR_CUDA;
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(mu);
  if (ptrs[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_sumsq_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      b2_set_error("b2_sumsq_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      *rc_out = B2_ERR_CUDA;
      return nullptr;
    }
    ptrs[dev] = p;
  }
  return ptrs[dev];
}

extern "C" size_t b2_sumsq_static_max_elems(void) { return B2_SUMSQ_STATIC_MAX_N; }

extern "C" int b2_sumsq_static(const void* g, size_t n, int dtype, float* out, int accumulate, cudaStream_t stream) {
  B2_REQUIRE(out, "b2_sumsq_static: null out");
  B2_REQUIRE(n == 0 || g, "b2_sumsq_static: null input");
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_sumsq_static: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(b2_aligned(g, 16), "b2_sumsq_static: input must be 16-byte aligned");
  if (n > B2_SUMSQ_STATIC_MAX_N) {
    b2_set_error("b2_sumsq_static: %zu elements exceed the limit of %zu per call; split the buffer and pass accumulate=1 "
                 "for the later pieces, or use b2_sumsq with caller-provided scratch", n, (size_t)B2_SUMSQ_STATIC_MAX_N);
    return B2_ERR_LIMIT;
  }
  int rc;
  float* scratch = sumsq_static_scratch_ptr(&rc);
  if (rc != B2_OK) return rc;
  // the ticket must read zero when the first block arrives, whatever an earlier (possibly failed) launch left behind
  B2_CHECK_CUDA(cudaMemsetAsync(scratch, 0, sizeof(unsigned), stream));
  const unsigned grid = grid_for(n, B2_SUMSQ_STATIC_BLOCKS);
  if (dtype == B2_F32) sumsq_kernel<B2_F32><<<grid, 256, 0, stream>>>(g, n, scratch, out, accumulate);
  else sumsq_kernel<B2_BF16><<<grid, 256, 0, stream>>>(g, n, scratch, out, accumulate);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}
