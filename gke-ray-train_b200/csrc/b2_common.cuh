// Shared device/host helpers for the b2train kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#define B2_OK 0
#define B2_ERR_INVALID (-1)   // bad argument (null pointer, misalignment, unsupported dtype/shape)
#define B2_ERR_CUDA (-2)      // a CUDA runtime/driver call failed
#define B2_ERR_LIMIT (-3)     // a size exceeds a stated maximum
#define B2_ERR_STATE (-4)     // called in the wrong state (e.g. handle already closed)

#define B2_F32 0
#define B2_BF16 1

#define B2_NUM_SMS 148

// last-error string, thread local; see api.cu
void b2_set_error(const char* fmt, ...);

#define B2_CHECK_CUDA(expr)                                                            \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      b2_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr,                  \
                   cudaGetErrorString(_e));                                            \
      return B2_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

#define B2_REQUIRE(cond, ...)                                                          \
  do {                                                                                 \
    if (!(cond)) {                                                                     \
      b2_set_error(__VA_ARGS__);                                                       \
      return B2_ERR_INVALID;                                                           \
    }                                                                                  \
  } while (0)

static inline bool b2_aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }
static inline size_t b2_dtype_size(int dt) { return dt == B2_F32 ? 4 : 2; }

#ifdef __CUDACC__

// ---- 128-bit streaming loads/stores (data touched once per step: keep it out of L1) ----
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_rw(const void* p) {   // data we will overwrite: no .nc
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
// Two values rounded to bf16 precision (round-to-nearest-even, same result as round_bf16 on each) through ONE packed
// conversion (F2FP.BF16.F32.PACK_AB) plus two integer ops. The scalar form is an F2F per value, and F2F issues on the
// quarter-rate XU pipe that also serves MUFU.EX2 — in the softmax epilogues that pipe is the limiter.
__device__ __forceinline__ void round_bf16_pair(float& a, float& b) {
  const uint32_t w = pack_bf16(a, b);
  a = bf16_lo(w);
  b = bf16_hi(w);
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Exponent offset of the fused soft-max gradient: exp2(x * log2e + off) == exp(x - lse) * |gs|. One definition, so
// that every kernel producing d(logits) rounds this the same way.
__device__ __forceinline__ float ce_bwd_exp_offset(float gs, float lse) {
  return fmaf(-lse, 1.4426950408889634f, log2f(fabsf(gs)));
}

// load / store 8 consecutive elements of dtype f32 or bf16 as floats
template <int DT> struct Vec8;
template <> struct Vec8<B2_F32> {
  static constexpr int kBytes = 32;
  template <bool RW> __device__ __forceinline__ static void load(const void* base, size_t i8, float (&x)[8]) {
    const char* p = static_cast<const char*>(base) + i8 * 32;
    uint4 a = RW ? ld_rw(p) : ld_stream(p);
    uint4 b = RW ? ld_rw(p + 16) : ld_stream(p + 16);
    x[0] = __uint_as_float(a.x); x[1] = __uint_as_float(a.y); x[2] = __uint_as_float(a.z); x[3] = __uint_as_float(a.w);
    x[4] = __uint_as_float(b.x); x[5] = __uint_as_float(b.y); x[6] = __uint_as_float(b.z); x[7] = __uint_as_float(b.w);
  }
  __device__ __forceinline__ static void store(void* base, size_t i8, const float (&x)[8]) {
    char* p = static_cast<char*>(base) + i8 * 32;
    st_stream(p, make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]), __float_as_uint(x[2]), __float_as_uint(x[3])));
    st_stream(p + 16, make_uint4(__float_as_uint(x[4]), __float_as_uint(x[5]), __float_as_uint(x[6]), __float_as_uint(x[7])));
  }
  __device__ __forceinline__ static float load1(const void* base, size_t i) { return static_cast<const float*>(base)[i]; }
  __device__ __forceinline__ static void store1(void* base, size_t i, float x) { static_cast<float*>(base)[i] = x; }
  __device__ __forceinline__ static float round(float x) { return x; }
};
template <> struct Vec8<B2_BF16> {
  static constexpr int kBytes = 16;
  template <bool RW> __device__ __forceinline__ static void load(const void* base, size_t i8, float (&x)[8]) {
    const char* p = static_cast<const char*>(base) + i8 * 16;
    uint4 a = RW ? ld_rw(p) : ld_stream(p);
    x[0] = bf16_lo(a.x); x[1] = bf16_hi(a.x); x[2] = bf16_lo(a.y); x[3] = bf16_hi(a.y);
    x[4] = bf16_lo(a.z); x[5] = bf16_hi(a.z); x[6] = bf16_lo(a.w); x[7] = bf16_hi(a.w);
  }
  __device__ __forceinline__ static void store(void* base, size_t i8, const float (&x)[8]) {
    char* p = static_cast<char*>(base) + i8 * 16;
    st_stream(p, make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]), pack_bf16(x[6], x[7])));
  }
  __device__ __forceinline__ static float load1(const void* base, size_t i) {
    return __bfloat162float(static_cast<const __nv_bfloat16*>(base)[i]);
  }
  __device__ __forceinline__ static void store1(void* base, size_t i, float x) {
    static_cast<__nv_bfloat16*>(base)[i] = __float2bfloat16_rn(x);
  }
  __device__ __forceinline__ static float round(float x) { return round_bf16(x); }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide sum; result valid in thread 0. `red` must hold >= 32 floats.
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    v = (lane < (blockDim.x + 31) / 32) ? red[lane] : 0.f;
    v = warp_sum(v);
  }
  return v;
}

#endif  // __CUDACC__
