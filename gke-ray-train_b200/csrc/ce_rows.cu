// Cross-entropy over materialised logits rows, and the loss finalisation shared with the
// fused LM-head kernel.
//
// b2_ce_rows_* replace nn.CrossEntropyLoss()(outputs.reshape(-1,V), targets.reshape(-1)) as the
// reference calls it at ref: ray-jobs/pytorch_llm_ray.py:237,275 (fp32 logits, char-vocab-sized V,
// labels pre-shifted by TextDataset, ref :116-119). They also accept the HF in-loss shift
// (transformers/loss/loss_utils.py:45-67: labels = pad(labels,(0,1),-100)[...,1:]) via `shift_T`.
//
// HBM-bound: forward reads N*V*s bytes once; backward reads N*V*s and writes N*V*s.
#include "b2_common.cuh"
#include <math.h>

namespace {

__device__ __forceinline__ long long label_of(const long long* __restrict__ labels, long long i, int shift_T,
                                              long long ignore_index) {
  if (shift_T <= 0) return labels[i];
  return ((i % shift_T) + 1 < shift_T) ? labels[i + 1] : ignore_index;
}

// online (max, sum-exp) merge
__device__ __forceinline__ void lse_merge(float& m, float& s, float m2, float s2) {
  const float nm = fmaxf(m, m2);
  if (nm == -INFINITY) { m = nm; s = 0.f; return; }
  s = s * __expf(m - nm) + s2 * __expf(m2 - nm);
  m = nm;
}

template <int DT>
__global__ void __launch_bounds__(256) ce_rows_fwd_kernel(const void* __restrict__ logits, long long N, long long V,
                                                          long long ld, const long long* __restrict__ labels,
                                                          int shift_T, long long ignore_index, float* __restrict__ lse,
                                                          float* __restrict__ zt) {
  using L = Vec8<DT>;
  __shared__ float sm[8], ss[8];
  for (long long row = blockIdx.x; row < N; row += gridDim.x) {
    const size_t base = (size_t)row * ld;
    float m = -INFINITY, s = 0.f;
    for (long long j = threadIdx.x; j < V; j += blockDim.x) {
      const float x = L::load1(logits, base + j);
      if (x > m) { s = s * __expf(m - x) + 1.f; m = x; }
      else if (m != -INFINITY) s += __expf(x - m);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
      lse_merge(m, s, m2, s2);
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = m; ss[threadIdx.x >> 5] = s; }
    __syncthreads();
    if (threadIdx.x == 0) {
      float M = sm[0], S = ss[0];
      for (int w = 1; w < (int)(blockDim.x >> 5); ++w) lse_merge(M, S, sm[w], ss[w]);
      lse[row] = M + logf(S);
      const long long y = label_of(labels, row, shift_T, ignore_index);
      zt[row] = (y >= 0 && y < V) ? L::load1(logits, base + y) : 0.f;
    }
  }
}

template <int DT>
// (logits / dlogits carry no __restrict__: the ABI lets dlogits alias logits — the gradient may overwrite the logits in
// place — and a restrict-qualified const input could be read through the non-coherent path while being written)
__global__ void __launch_bounds__(256) ce_rows_bwd_kernel(const void* logits, long long N, long long V,
                                                          long long ld, const long long* __restrict__ labels,
                                                          int shift_T, long long ignore_index,
                                                          const float* __restrict__ lse, const float* __restrict__ gscale,
                                                          void* dlogits) {
  using L = Vec8<DT>;
  const float gs = __ldg(gscale);
  for (long long row = blockIdx.x; row < N; row += gridDim.x) {
    const size_t base = (size_t)row * ld;
    const long long y = label_of(labels, row, shift_T, ignore_index);
    const bool valid = (y != ignore_index) && y >= 0 && y < V;
    const float l = lse[row];
    for (long long j = threadIdx.x; j < V; j += blockDim.x) {
      float d = 0.f;
      if (valid) {
        const float x = L::load1(logits, base + j);
        d = (__expf(x - l) - (j == y ? 1.f : 0.f)) * gs;
      }
      L::store1(dlogits, base + j, d);
    }
  }
}

// One block: sums (lse - z) over rows whose label is not ignore_index, counts them, flags
// out-of-range labels (torch raises IndexError / device-asserts on those; we set *err and
// poison the loss with NaN so the failure cannot pass silently).
__global__ void __launch_bounds__(1024) ce_finalize_kernel(const float* __restrict__ lse, const float* __restrict__ zt,
                                                           const long long* __restrict__ labels, long long N, long long V,
                                                           int shift_T, long long ignore_index,
                                                           float* __restrict__ loss_out, long long* __restrict__ n_valid_out,
                                                           int* __restrict__ err_out) {
  __shared__ double dsum[32];
  __shared__ long long dcnt[32];
  __shared__ int derr[32];
  double acc = 0.0; long long cnt = 0; int bad = 0;
  for (long long i = threadIdx.x; i < N; i += blockDim.x) {
    const long long y = label_of(labels, i, shift_T, ignore_index);
    if (y == ignore_index) continue;
    if (y < 0 || y >= V) { bad = 1; continue; }
    acc += (double)(lse[i] - zt[i]);
    cnt += 1;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    bad |= __shfl_xor_sync(0xffffffffu, bad, o);
  }
  if ((threadIdx.x & 31) == 0) { dsum[threadIdx.x >> 5] = acc; dcnt[threadIdx.x >> 5] = cnt; derr[threadIdx.x >> 5] = bad; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0; long long c = 0; int e = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { t += dsum[w]; c += dcnt[w]; e |= derr[w]; }
    float sum = (float)t;
    float mean = (float)(t / (double)c);            // 0/0 -> NaN, like torch's mean over zero valid rows
    if (e) { sum = NAN; mean = NAN; }
    loss_out[0] = sum; loss_out[1] = mean;
    n_valid_out[0] = c;
    if (err_out) err_out[0] = e;
  }
}

}  // namespace

int b2_ce_finalize_launch(const float* lse, const float* zt, const long long* labels, long long N, long long V,
                          int shift_T, long long ignore_index, float* loss_out, long long* n_valid_out, int* err_out,
                          cudaStream_t stream) {
  ce_finalize_kernel<<<1, 1024, 0, stream>>>(lse, zt, labels, N, V, shift_T, ignore_index, loss_out, n_valid_out, err_out);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_ce_rows_fwd(const void* logits, int dtype, long long N, long long V, long long ld,
                              const long long* labels, int shift_T, long long ignore_index, float* lse, float* zt,
                              float* loss_out, long long* n_valid_out, int* err_out, cudaStream_t stream) {
  B2_REQUIRE(N >= 0 && V >= 1 && ld >= V, "b2_ce_rows_fwd: bad shape N=%lld V=%lld ld=%lld", N, V, ld);
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_ce_rows_fwd: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(loss_out && n_valid_out, "b2_ce_rows_fwd: null output");
  B2_REQUIRE(N == 0 || (logits && labels && lse && zt), "b2_ce_rows_fwd: null buffer");
  B2_REQUIRE(shift_T == 0 || (shift_T > 0 && N % shift_T == 0), "b2_ce_rows_fwd: N must be a multiple of shift_T");
  if (N > 0) {
    const unsigned grid = (unsigned)(N < (long long)B2_NUM_SMS * 16 ? N : (long long)B2_NUM_SMS * 16);
    if (dtype == B2_F32)
      ce_rows_fwd_kernel<B2_F32><<<grid, 256, 0, stream>>>(logits, N, V, ld, labels, shift_T, ignore_index, lse, zt);
    else
      ce_rows_fwd_kernel<B2_BF16><<<grid, 256, 0, stream>>>(logits, N, V, ld, labels, shift_T, ignore_index, lse, zt);
    B2_CHECK_CUDA(cudaGetLastError());
  }
  return b2_ce_finalize_launch(lse, zt, labels, N, V, shift_T, ignore_index, loss_out, n_valid_out, err_out, stream);
}

extern "C" int b2_ce_rows_bwd(const void* logits, int dtype, long long N, long long V, long long ld,
                              const long long* labels, int shift_T, long long ignore_index, const float* lse,
                              const float* gscale_dev, void* dlogits, cudaStream_t stream) {
  if (N == 0) return B2_OK;
  B2_REQUIRE(N > 0 && V >= 1 && ld >= V, "b2_ce_rows_bwd: bad shape N=%lld V=%lld ld=%lld", N, V, ld);
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_ce_rows_bwd: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(logits && labels && lse && gscale_dev && dlogits, "b2_ce_rows_bwd: null buffer");
  B2_REQUIRE(shift_T == 0 || (shift_T > 0 && N % shift_T == 0), "b2_ce_rows_bwd: N must be a multiple of shift_T");
  const unsigned grid = (unsigned)(N < (long long)B2_NUM_SMS * 16 ? N : (long long)B2_NUM_SMS * 16);
  if (dtype == B2_F32)
    ce_rows_bwd_kernel<B2_F32><<<grid, 256, 0, stream>>>(logits, N, V, ld, labels, shift_T, ignore_index, lse, gscale_dev, dlogits);
  else
    ce_rows_bwd_kernel<B2_BF16><<<grid, 256, 0, stream>>>(logits, N, V, ld, labels, shift_T, ignore_index, lse, gscale_dev, dlogits);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}
