// F3 — fused AdamW over a flat buffer: one pass, 128-bit loads/stores, fp32 math.
//
// Replaces torch.optim.AdamW.step() as the reference calls it
// (ref: ray-jobs/pytorch_llm_ray.py:236,278; HF path ref: ray-jobs/fine_tune_config.json:15-17).
// Arithmetic follows torch/optim/adam.py `_single_tensor_adam` / `_multi_tensor_adam`
// (decoupled weight decay): per element
//     p  <- p * (1 - lr*wd)
//     m  <- m + (1-b1) * (g - m)                     (lerp, weight < 0.5 branch)
//     v  <- v*b2 + (1-b2) * (g*g)                    (mul_ then addcmul_)
//     d  <- sqrt(v) / sqrt(1-b2^t) + eps
//     p  <- p + (-(lr/(1-b1^t))) * (m / d)           (addcdiv_)
// The two bias-correction scalars are computed by the host in double, as torch does, and
// passed as floats. The optional device scalar `gscale` is the clip_grad_norm_ coefficient
// (ref: pytorch_llm_ray.py:277): g is multiplied by it and rounded to the gradient dtype first,
// which is what clip_grad_norm_'s in-place `g.mul_(coef)` would have stored.
//
// HBM-bound: algorithmic bytes/element = read(p,g,m,v) + write(p,m,v)
//   = 28 B (all fp32), 14 B (all bf16), 22 B (bf16 p/g, fp32 m/v).
#include "b2_common.cuh"

namespace {

struct AdamWArgs {
  float one_minus_lr_wd;   // (float)(1 - lr*wd)
  float one_minus_b1;      // (float)(1 - beta1)
  float beta2;
  float one_minus_b2;
  float bc2_sqrt;          // (float)sqrt(1 - beta2^t)
  float eps;
  float neg_step_size;     // (float)(-(lr / (1 - beta1^t)))
};

__device__ __forceinline__ void adamw_elem(float& p, float g, float& m, float& v, const AdamWArgs& a) {
  p = __fmul_rn(p, a.one_minus_lr_wd);
  m = __fmaf_rn(a.one_minus_b1, __fsub_rn(g, m), m);
  v = __fmaf_rn(a.one_minus_b2, __fmul_rn(g, g), __fmul_rn(v, a.beta2));
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), a.bc2_sqrt), a.eps);
  p = __fmaf_rn(a.neg_step_size, __fdiv_rn(m, denom), p);
}

template <int PDT, int GDT, int SDT>
__global__ void __launch_bounds__(256) adamw_kernel(void* __restrict__ p, const void* __restrict__ g,
                                                    void* __restrict__ m, void* __restrict__ v, size_t n,
                                                    AdamWArgs a, const float* __restrict__ gscale) {
  using P = Vec8<PDT>; using G = Vec8<GDT>; using S = Vec8<SDT>;
  const bool scaled = gscale != nullptr;
  const float gs = scaled ? __ldg(gscale) : 1.f;
  const size_t n8 = n >> 3;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    float pp[8], gg[8], mm[8], vv[8];
    P::template load<true>(p, i, pp);
    G::template load<false>(g, i, gg);
    S::template load<true>(m, i, mm);
    S::template load<true>(v, i, vv);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float gk = scaled ? G::round(__fmul_rn(gg[k], gs)) : gg[k];
      adamw_elem(pp[k], gk, mm[k], vv[k], a);
    }
    P::store(p, i, pp);
    S::store(m, i, mm);
    S::store(v, i, vv);
  }
  // ragged tail (< 8 elements)
  const size_t tail0 = n8 << 3;
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - tail0)) {
    const size_t i = tail0 + threadIdx.x;
    float pp = P::load1(p, i), gg = G::load1(g, i), mm = S::load1(m, i), vv = S::load1(v, i);
    if (scaled) gg = G::round(__fmul_rn(gg, gs));
    adamw_elem(pp, gg, mm, vv, a);
    P::store1(p, i, pp); S::store1(m, i, mm); S::store1(v, i, vv);
  }
}

template <int PDT, int GDT, int SDT>
int launch_adamw(void* p, const void* g, void* m, void* v, size_t n, const AdamWArgs& a, const float* gscale,
                 cudaStream_t stream) {
  const size_t n8 = n >> 3;
  size_t blocks = (n8 + 255) / 256;
  // two resident waves of 256-thread blocks per SM saturate HBM; beyond that grid-stride.
  const size_t cap = (size_t)B2_NUM_SMS * 16;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) blocks = 1;
  adamw_kernel<PDT, GDT, SDT><<<(unsigned)blocks, 256, 0, stream>>>(p, g, m, v, n, a, gscale);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}


#define B2_ADAMW_MAX_GROUPS 8
struct AdamWGroupArgs { AdamWArgs a[B2_ADAMW_MAX_GROUPS]; };

// Same update, but hyper-parameters are looked up per 64-element chunk: chunk_group[c] selects one of up to 8
// parameter groups (HF Trainer uses two: weight decay on / off, $SP/transformers/trainer.py:1157-1170);
// 255 marks a chunk that is skipped (frozen parameters). n is a multiple of 64.
template <int PDT, int GDT, int SDT>
__global__ void __launch_bounds__(256) adamw_grouped_kernel(void* __restrict__ p, const void* __restrict__ g,
                                                            void* __restrict__ m, void* __restrict__ v, size_t n,
                                                            const uint8_t* __restrict__ chunk_group,
                                                            const __grid_constant__ AdamWGroupArgs ga,
                                                            const float* __restrict__ gscale) {
  using P = Vec8<PDT>; using G = Vec8<GDT>; using S = Vec8<SDT>;
  const bool scaled = gscale != nullptr;
  const float gs = scaled ? __ldg(gscale) : 1.f;
  const size_t n8 = n >> 3;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const unsigned gid = __ldg(chunk_group + (i >> 3));
    if (gid >= B2_ADAMW_MAX_GROUPS) continue;
    const AdamWArgs a = ga.a[gid];
    float pp[8], gg[8], mm[8], vv[8];
    P::template load<true>(p, i, pp);
    G::template load<false>(g, i, gg);
    S::template load<true>(m, i, mm);
    S::template load<true>(v, i, vv);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float gk = scaled ? G::round(__fmul_rn(gg[k], gs)) : gg[k];
      adamw_elem(pp[k], gk, mm[k], vv[k], a);
    }
    P::store(p, i, pp);
    S::store(m, i, mm);
    S::store(v, i, vv);
  }
}

AdamWArgs make_args(double lr, double beta1, double beta2, double eps, double weight_decay, long long step) {
  // host scalars in double, exactly as torch/optim/adam.py:414-547 forms them
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  AdamWArgs a;
  a.one_minus_lr_wd = (float)(1.0 - lr * weight_decay);
  a.one_minus_b1 = (float)(1.0 - beta1);
  a.beta2 = (float)beta2;
  a.one_minus_b2 = (float)(1.0 - beta2);
  a.bc2_sqrt = (float)sqrt(bc2);
  a.eps = (float)eps;
  a.neg_step_size = (float)(-(lr / bc1));
  return a;
}

template <int PDT, int GDT, int SDT>
int launch_adamw_grouped(void* p, const void* g, void* m, void* v, size_t n, const uint8_t* chunk_group,
                         const AdamWGroupArgs& ga, const float* gscale, cudaStream_t stream) {
  size_t blocks = ((n >> 3) + 255) / 256;
  const size_t cap = (size_t)B2_NUM_SMS * 16;
  if (blocks > cap) blocks = cap;
  adamw_grouped_kernel<PDT, GDT, SDT><<<(unsigned)blocks, 256, 0, stream>>>(p, g, m, v, n, chunk_group, ga, gscale);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}

}  // namespace

extern "C" int b2_adamw(void* p, const void* g, void* m, void* v, size_t n, int p_dtype, int g_dtype, int state_dtype,
                        double lr, double beta1, double beta2, double eps, double weight_decay, long long step,
                        const float* gscale_dev, cudaStream_t stream) {
  if (n == 0) return B2_OK;
  B2_REQUIRE(p && g && m && v, "b2_adamw: null buffer");
  B2_REQUIRE(b2_aligned(p, 16) && b2_aligned(g, 16) && b2_aligned(m, 16) && b2_aligned(v, 16),
             "b2_adamw: buffers must be 16-byte aligned");
  B2_REQUIRE(step >= 1, "b2_adamw: step must be >= 1 (got %lld)", step);
  B2_REQUIRE((p_dtype | 1) == 1 && (g_dtype | 1) == 1 && (state_dtype | 1) == 1, "b2_adamw: dtype must be 0 (f32) or 1 (bf16)");
  const AdamWArgs a = make_args(lr, beta1, beta2, eps, weight_decay, step);
#define B2_ADAMW_CASE(P, G, S)                                   \
  if (p_dtype == P && g_dtype == G && state_dtype == S)          \
    return launch_adamw<P, G, S>(p, g, m, v, n, a, gscale_dev, stream);
  B2_ADAMW_CASE(B2_F32, B2_F32, B2_F32)
  B2_ADAMW_CASE(B2_BF16, B2_BF16, B2_BF16)
  B2_ADAMW_CASE(B2_BF16, B2_BF16, B2_F32)
  B2_ADAMW_CASE(B2_F32, B2_BF16, B2_F32)
  B2_ADAMW_CASE(B2_BF16, B2_F32, B2_F32)
  B2_ADAMW_CASE(B2_F32, B2_F32, B2_BF16)
  B2_ADAMW_CASE(B2_F32, B2_BF16, B2_BF16)
  B2_ADAMW_CASE(B2_BF16, B2_F32, B2_BF16)
#undef B2_ADAMW_CASE
  b2_set_error("b2_adamw: unsupported dtype combination");
  return B2_ERR_INVALID;
}

struct B2AdamWGroup { double lr, beta1, beta2, eps, weight_decay; long long step; };

extern "C" int b2_adamw_grouped(void* p, const void* g, void* m, void* v, size_t n, int p_dtype, int g_dtype,
                                int state_dtype, const uint8_t* chunk_group_dev, const B2AdamWGroup* groups_host,
                                int ngroups, const float* gscale_dev, cudaStream_t stream) {
  if (n == 0) return B2_OK;
  B2_REQUIRE(p && g && m && v && chunk_group_dev && groups_host, "b2_adamw_grouped: null pointer");
  B2_REQUIRE(n % 64 == 0, "b2_adamw_grouped: n (%zu) must be a multiple of the 64-element chunk", n);
  B2_REQUIRE(ngroups >= 1 && ngroups <= B2_ADAMW_MAX_GROUPS, "b2_adamw_grouped: 1..%d groups supported (got %d)",
             B2_ADAMW_MAX_GROUPS, ngroups);
  B2_REQUIRE(b2_aligned(p, 16) && b2_aligned(g, 16) && b2_aligned(m, 16) && b2_aligned(v, 16),
             "b2_adamw_grouped: buffers must be 16-byte aligned");
  AdamWGroupArgs ga;
  for (int i = 0; i < B2_ADAMW_MAX_GROUPS; ++i) {
    const B2AdamWGroup& h = groups_host[i < ngroups ? i : 0];
    B2_REQUIRE(h.step >= 1, "b2_adamw_grouped: step must be >= 1");
    ga.a[i] = make_args(h.lr, h.beta1, h.beta2, h.eps, h.weight_decay, h.step);
  }
#define B2_ADAMW_GCASE(P, G, S)                                  \
  if (p_dtype == P && g_dtype == G && state_dtype == S)          \
    return launch_adamw_grouped<P, G, S>(p, g, m, v, n, chunk_group_dev, ga, gscale_dev, stream);
  B2_ADAMW_GCASE(B2_F32, B2_F32, B2_F32)
  B2_ADAMW_GCASE(B2_BF16, B2_BF16, B2_BF16)
  B2_ADAMW_GCASE(B2_BF16, B2_BF16, B2_F32)
  B2_ADAMW_GCASE(B2_F32, B2_BF16, B2_F32)
#undef B2_ADAMW_GCASE
  b2_set_error("b2_adamw_grouped: unsupported dtype combination");
  return B2_ERR_INVALID;
}
