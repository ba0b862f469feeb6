// F2 — gradient all-reduce over peer memory (NVLink 5 / NVSwitch), fused with the 1/world scale, the
// dtype cast and the gradient sum-of-squares needed by clip_grad_norm_.
//
// Replaces the per-bucket ncclAllReduce that DistributedDataParallel issues during loss.backward()
// (installed by ray.train.torch.prepare_model, ref: ray-jobs/pytorch_llm_ray.py:230,276; c10d Reducer,
// $SP/torch/nn/parallel/distributed.py:1185-1280; default hook $SP/torch/distributed/algorithms/ddp_comm_hooks/
// default_hooks.py:19-31) together with the "grad / world_size" copy-in (Reducer::mark_variable_ready_dense).
// NCCL/gloo are used by the host only to exchange the IPC handles below (bootstrap).
//
// Every rank maps every other rank's flat gradient buffer (CUDA IPC). Rank r owns slice r of the buffer:
//   start barrier -> for each 16-byte vector of slice r: load it from all W buffers (W-1 loads cross NVLink),
//   add in fp32 in rank order, multiply by `scale`, round to the buffer dtype, accumulate its square,
//   store it back into all W buffers (W-1 stores cross NVLink) -> end barrier.
// Only rank r ever touches slice r between the two barriers, so the exchange is in place and needs no staging.
// Per GPU and direction this moves (W-1)/W * n * s bytes in and the same out — the reduce-scatter + all-gather
// volume; every rank ends with bit-identical values (each element is reduced exactly once).
// Modes: ALLREDUCE (above), REDUCE_SCATTER (result of slice r stored locally only — the sharded-optimizer path),
// and b2_peer_allgather (push slice r to every peer, used for the updated parameters).
//
// NVLink-bound: algorithmic bytes per GPU per direction = 2 * (W-1)/W * n * s for the all-reduce (NCCL busbw).
#include "b2_common.cuh"

#define B2_PEER_MAX_WORLD 8
// flag block (per rank, in that rank's symmetric memory): uint32 words
//   [0..7]   start-barrier arrivals, one word per source rank     [8..15] end-barrier arrivals
//   [16]     local block ticket (self-resetting)                  [17]    error word (timeout)
//   [32..63] float sumsq slots: [32 + src_rank]
//   [64..383] per-block sum-of-squares partials of the running kernel (added in block order by its last block)
#define B2_PEER_FLAG_WORDS 384
#define B2_PEER_PARTIALS 64
#define B2_PEER_MAX_BLOCKS 320

namespace {

struct PeerPtrs {
  void* buf[B2_PEER_MAX_WORLD];
  uint32_t* flags[B2_PEER_MAX_WORLD];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer(const void* p) {   // data another GPU wrote: never the non-coherent path
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_peer(void* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint64_t gtime_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Spin until *flag (written by a peer) reaches `epoch`. Ranks may legitimately be far apart (rank 0 writing a checkpoint
// while the others already run the next backward), so the wait is long — but a peer that never arrives (crashed worker)
// must not hang the GPU forever: after ~10 min (the order of NCCL's watchdog) record the failure and trap.
__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t epoch, uint32_t* err) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    if ((++spins & 0xfffu) == 0) {
      const uint64_t now = gtime_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 600000000000ull) {
        *err = 1;
        printf("b2train: peer barrier timed out waiting for epoch %u (block %d)\n", epoch, (int)blockIdx.x);
        __trap();
      }
    }
  }
}

// All blocks of every rank pass this point only after every rank's kernel has started (i.e. after every rank's
// preceding stream work — its backward pass — has finished).
__device__ __forceinline__ void start_barrier(const PeerPtrs& pp, int world, int rank, uint32_t epoch) {
  uint32_t* mine = pp.flags[rank];
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(pp.flags[threadIdx.x] + rank, epoch);        // tell peer `threadIdx.x` that `rank` is here
  }
  if (threadIdx.x < world) wait_flag(mine + threadIdx.x, epoch, mine + 17);
  __syncthreads();
}

// The last block of this rank to finish signals every peer, then waits for every peer's signal: when the kernel
// retires, all peers' stores into this rank's buffer are complete and visible.
__device__ __forceinline__ void end_barrier(const PeerPtrs& pp, int world, int rank, uint32_t epoch) {
  __shared__ bool last;
  uint32_t* mine = pp.flags[rank];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(mine + 16, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  __threadfence_system();
  if (threadIdx.x < world) {
    st_release_sys(pp.flags[threadIdx.x] + 8 + rank, epoch);
    wait_flag(mine + 8 + threadIdx.x, epoch, mine + 17);
  }
  if (threadIdx.x == 0) mine[16] = 0u;
  __syncthreads();
}

// Tail of the reducing kernels: the sum of squares of this rank's slice, then the end barrier.
// Each block leaves its partial in the flag block; the block that arrives last adds them IN BLOCK ORDER (double), so the
// result does not depend on the order in which blocks finish — the norm clip_grad_norm_ returns is reproducible run to
// run. want_sumsq 1: the slot restarts from this launch's total; 2: the total is added to the slot (next bucket of the
// same buffer). The finished slot is pushed to every peer before the end-barrier signal, so when the kernel retires all
// W slots of this rank's flag block are complete; they are then copied to `sumsq_out` (local, W floats) if given.
__device__ __forceinline__ void finish_reduce(const PeerPtrs& pp, int world, int rank, uint32_t epoch, int want_sumsq,
                                              float ss, float* red, float* sumsq_out) {
  __shared__ bool last;
  uint32_t* mine = pp.flags[rank];
  if (want_sumsq) {
    const float b = block_sum(ss, red);
    if (threadIdx.x == 0) reinterpret_cast<volatile float*>(mine + B2_PEER_PARTIALS)[blockIdx.x] = b;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(mine + 16, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  __threadfence_system();
  if (want_sumsq) {
    if (threadIdx.x == 0) {
      double tot = 0.0;
      for (unsigned b = 0; b < gridDim.x; ++b) tot += (double)reinterpret_cast<volatile float*>(mine + B2_PEER_PARTIALS)[b];
      volatile float* slot = reinterpret_cast<volatile float*>(mine + 32 + rank);
      *slot = (want_sumsq == 2) ? (float)((double)*slot + tot) : (float)tot;
      __threadfence_system();
    }
    __syncthreads();
  }
  if ((int)threadIdx.x < world) {
    if (want_sumsq && (int)threadIdx.x != rank) {
      const uint32_t bits = ld_acquire_sys(mine + 32 + rank);
      asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(pp.flags[threadIdx.x] + 32 + rank), "r"(bits) : "memory");
      __threadfence_system();
    }
    st_release_sys(pp.flags[threadIdx.x] + 8 + rank, epoch);
    wait_flag(mine + 8 + threadIdx.x, epoch, mine + 17);
    // peer `threadIdx.x` pushed its slot before that signal
    if (want_sumsq && sumsq_out != nullptr) sumsq_out[threadIdx.x] = *reinterpret_cast<volatile float*>(mine + 32 + threadIdx.x);
  }
  if (threadIdx.x == 0) mine[16] = 0u;
}

template <int DT> struct PeerVec;
template <> struct PeerVec<B2_F32> {
  static constexpr int kElems = 4;
  __device__ __forceinline__ static void unpack(const uint4& v, float (&x)[8]) {
    x[0] = __uint_as_float(v.x); x[1] = __uint_as_float(v.y); x[2] = __uint_as_float(v.z); x[3] = __uint_as_float(v.w);
    x[4] = x[5] = x[6] = x[7] = 0.f;
  }
  __device__ __forceinline__ static uint4 pack(const float (&x)[8]) {
    return make_uint4(__float_as_uint(x[0]), __float_as_uint(x[1]), __float_as_uint(x[2]), __float_as_uint(x[3]));
  }
};
template <> struct PeerVec<B2_BF16> {
  static constexpr int kElems = 8;
  __device__ __forceinline__ static void unpack(const uint4& v, float (&x)[8]) {
    x[0] = bf16_lo(v.x); x[1] = bf16_hi(v.x); x[2] = bf16_lo(v.y); x[3] = bf16_hi(v.y);
    x[4] = bf16_lo(v.z); x[5] = bf16_hi(v.z); x[6] = bf16_lo(v.w); x[7] = bf16_hi(v.w);
  }
  __device__ __forceinline__ static uint4 pack(const float (&x)[8]) {
    return make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]), pack_bf16(x[6], x[7]));
  }
};

// vectors [v0, v1) of 16 bytes belong to this rank
template <int DT, int WORLD, bool SCATTER_ONLY>
__global__ void __launch_bounds__(512, 2) peer_allreduce_kernel(const __grid_constant__ PeerPtrs pp, int rank, size_t v0,
                                                             size_t v1, float scale, uint32_t epoch, int want_sumsq,
                                                             float* sumsq_out) {
  using V = PeerVec<DT>;
  __shared__ float red[32];
  start_barrier(pp, WORLD, rank, epoch);

  float ss = 0.f;
  // Loads in flight per thread: WORLD per vector. With few ranks that is too little to cover the NVLink round trip, so
  // small worlds take several vectors per iteration (8 loads in flight for W <= 4). Element order, and with it every
  // result bit and the order of the sum of squares, is unchanged.
  constexpr int U = WORLD <= 2 ? 4 : (WORLD <= 4 ? 2 : 1);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = v0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < v1; i0 += stride * U) {
    uint4 in[U][WORLD];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i < v1) {
#pragma unroll
        for (int k = 0; k < WORLD; ++k) in[u][k] = ld_peer(static_cast<const uint4*>(pp.buf[k]) + i);   // all loads in flight
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i >= v1) break;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int k = 0; k < WORLD; ++k) {                                                           // fixed rank order
        float x[8];
        V::unpack(in[u][k], x);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += x[e];
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] *= scale;
      const uint4 out = V::pack(acc);
      if (want_sumsq) {
        float y[8];
        V::unpack(out, y);                                 // the rounded values, as clip_grad_norm_ would read them
#pragma unroll
        for (int e = 0; e < 8; ++e) ss = fmaf(y[e], y[e], ss);
      }
      if (SCATTER_ONLY) {
        st_peer(static_cast<uint4*>(pp.buf[rank]) + i, out);
      } else {
#pragma unroll
        for (int k = 0; k < WORLD; ++k) st_peer(static_cast<uint4*>(pp.buf[k]) + i, out);
      }
    }
  }

  finish_reduce(pp, WORLD, rank, epoch, want_sumsq, ss, red, sumsq_out);
}

template <int WORLD>
__global__ void __launch_bounds__(512) peer_allgather_kernel(const __grid_constant__ PeerPtrs pp, int rank, size_t v0,
                                                             size_t v1, uint32_t epoch) {
  start_barrier(pp, WORLD, rank, epoch);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = v0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < v1; i += stride) {
    const uint4 v = ld_rw(static_cast<const uint4*>(pp.buf[rank]) + i);
#pragma unroll
    for (int k = 0; k < WORLD; ++k)
      if (k != rank) st_peer(static_cast<uint4*>(pp.buf[k]) + i, v);
  }
  end_barrier(pp, WORLD, rank, epoch);
}

int fill_ptrs(PeerPtrs& pp, void* const* bufs, uint32_t* const* flags, int world, size_t byte_offset, const char* what) {
  B2_REQUIRE(world >= 1 && world <= B2_PEER_MAX_WORLD, "%s: world size %d not in 1..%d", what, world, B2_PEER_MAX_WORLD);
  B2_REQUIRE(bufs && flags, "%s: null pointer table", what);
  for (int k = 0; k < B2_PEER_MAX_WORLD; ++k) {
    pp.buf[k] = k < world ? static_cast<char*>(bufs[k]) + byte_offset : nullptr;
    pp.flags[k] = k < world ? flags[k] : nullptr;
    if (k < world) {
      B2_REQUIRE(bufs[k] && flags[k], "%s: null peer pointer for rank %d", what, k);
      B2_REQUIRE(b2_aligned(pp.buf[k], 16), "%s: peer buffer %d (+offset) must be 16-byte aligned", what, k);
    }
  }
  return B2_OK;
}

// slice of rank r in 16-byte vectors: contiguous, equal-sized (last one ragged)
// Slices start on 64-element boundaries (4 or 16 vectors... `align_vec` = 64 elements in vectors), the granularity of the
// AdamW chunk table, so that a sharded optimizer can update exactly the slice it reduced — also for a bucket (a sub-range
// of the flat buffer) whose length is not a multiple of 64 * world.
void slice_of(size_t nvec, int world, int rank, size_t align_vec, size_t& v0, size_t& v1) {
  size_t per = (nvec + world - 1) / world;
  per = (per + align_vec - 1) / align_vec * align_vec;
  v0 = per * rank < nvec ? per * rank : nvec;
  v1 = v0 + per < nvec ? v0 + per : nvec;
}

unsigned peer_grid(size_t nvec_slice, int max_blocks) {
  size_t blocks = (nvec_slice + 511) / 512;
  size_t cap = (size_t)B2_NUM_SMS * 2;           // 2 x 512 threads per SM: enough loads in flight to cover NVLink latency
  if (max_blocks > 0 && (size_t)max_blocks < cap) cap = (size_t)max_blocks;   // overlapped with compute: leave SMs free
  if (blocks > cap) blocks = cap;
  if (blocks > B2_PEER_MAX_BLOCKS) blocks = B2_PEER_MAX_BLOCKS;     // one sum-of-squares partial per block in the flag block
  if (blocks == 0) blocks = 1;
  return (unsigned)blocks;
}

}  // namespace

extern "C" size_t b2_peer_flags_bytes(void) { return sizeof(uint32_t) * B2_PEER_FLAG_WORDS; }

// ---- symmetric (peer-mappable) memory: cudaMalloc + CUDA IPC ------------------------------------------------
extern "C" int b2_symm_alloc(size_t bytes, void** dptr, unsigned char handle[64]) {
  B2_REQUIRE(dptr && handle && bytes > 0, "b2_symm_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  B2_CHECK_CUDA(cudaMalloc(dptr, bytes));
  cudaError_t e = cudaMemset(*dptr, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, *dptr);
  if (e != cudaSuccess) {
    cudaFree(*dptr);
    *dptr = nullptr;
    b2_set_error("b2_symm_alloc: %s", cudaGetErrorString(e));
    return B2_ERR_CUDA;
  }
  memcpy(handle, &h, 64);
  return B2_OK;
}
extern "C" int b2_symm_open(const unsigned char handle[64], void** dptr) {
  B2_REQUIRE(dptr && handle, "b2_symm_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  B2_CHECK_CUDA(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B2_OK;
}
extern "C" int b2_symm_close(void* dptr) {
  if (dptr) B2_CHECK_CUDA(cudaIpcCloseMemHandle(dptr));
  return B2_OK;
}
extern "C" int b2_symm_free(void* dptr) {
  if (dptr) B2_CHECK_CUDA(cudaFree(dptr));
  return B2_OK;
}

// ---- collectives ------------------------------------------------------------------------------------------------
extern "C" int b2_peer_allreduce(void* const* peer_bufs, uint32_t* const* peer_flags, int world, int rank,
                                 size_t offset_elems, size_t n, int dtype, float scale, unsigned epoch, int want_sumsq,
                                 int scatter_only, int max_blocks, float* sumsq_out, cudaStream_t stream) {
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_peer_allreduce: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(rank >= 0 && rank < world, "b2_peer_allreduce: rank %d outside world %d", rank, world);
  const size_t es = b2_dtype_size(dtype), per_vec = 16 / es;
  B2_REQUIRE(n % per_vec == 0, "b2_peer_allreduce: n (%zu) must be a multiple of %zu elements (16 bytes)", n, per_vec);
  PeerPtrs pp;
  int rc = fill_ptrs(pp, peer_bufs, peer_flags, world, offset_elems * es, "b2_peer_allreduce");
  if (rc != B2_OK) return rc;
  size_t v0, v1;
  slice_of(n / per_vec, world, rank, 64 / per_vec, v0, v1);
  const unsigned grid = peer_grid(v1 - v0, max_blocks);
#define B2_PEER_CASE(DT, W)                                                                                        \
  if (dtype == DT && world == W) {                                                                                 \
    if (scatter_only) peer_allreduce_kernel<DT, W, true><<<grid, 512, 0, stream>>>(pp, rank, v0, v1, scale, epoch, want_sumsq, sumsq_out); \
    else peer_allreduce_kernel<DT, W, false><<<grid, 512, 0, stream>>>(pp, rank, v0, v1, scale, epoch, want_sumsq, sumsq_out);  \
    B2_CHECK_CUDA(cudaGetLastError());                                                                             \
    return B2_OK;                                                                                                  \
  }
  B2_PEER_CASE(B2_BF16, 1) B2_PEER_CASE(B2_BF16, 2) B2_PEER_CASE(B2_BF16, 4) B2_PEER_CASE(B2_BF16, 8)
  B2_PEER_CASE(B2_F32, 1) B2_PEER_CASE(B2_F32, 2) B2_PEER_CASE(B2_F32, 4) B2_PEER_CASE(B2_F32, 8)
  B2_PEER_CASE(B2_BF16, 3) B2_PEER_CASE(B2_BF16, 5) B2_PEER_CASE(B2_BF16, 6) B2_PEER_CASE(B2_BF16, 7)
  B2_PEER_CASE(B2_F32, 3) B2_PEER_CASE(B2_F32, 5) B2_PEER_CASE(B2_F32, 6) B2_PEER_CASE(B2_F32, 7)
#undef B2_PEER_CASE
  b2_set_error("b2_peer_allreduce: unsupported configuration");
  return B2_ERR_INVALID;
}

extern "C" int b2_peer_allgather(void* const* peer_bufs, uint32_t* const* peer_flags, int world, int rank,
                                 size_t offset_elems, size_t n, int dtype, unsigned epoch, int max_blocks,
                                 cudaStream_t stream) {
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_peer_allgather: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(rank >= 0 && rank < world, "b2_peer_allgather: rank %d outside world %d", rank, world);
  const size_t es = b2_dtype_size(dtype), per_vec = 16 / es;
  B2_REQUIRE(n % per_vec == 0, "b2_peer_allgather: n (%zu) must be a multiple of %zu elements (16 bytes)", n, per_vec);
  PeerPtrs pp;
  int rc = fill_ptrs(pp, peer_bufs, peer_flags, world, offset_elems * es, "b2_peer_allgather");
  if (rc != B2_OK) return rc;
  size_t v0, v1;
  slice_of(n / per_vec, world, rank, 64 / per_vec, v0, v1);
  const unsigned grid = peer_grid(v1 - v0, max_blocks);
#define B2_AG_CASE(W)                                                                       \
  if (world == W) {                                                                         \
    peer_allgather_kernel<W><<<grid, 512, 0, stream>>>(pp, rank, v0, v1, epoch);            \
    B2_CHECK_CUDA(cudaGetLastError());                                                      \
    return B2_OK;                                                                           \
  }
  B2_AG_CASE(1) B2_AG_CASE(2) B2_AG_CASE(3) B2_AG_CASE(4) B2_AG_CASE(5) B2_AG_CASE(6) B2_AG_CASE(7) B2_AG_CASE(8)
#undef B2_AG_CASE
  b2_set_error("b2_peer_allgather: unsupported world size");
  return B2_ERR_INVALID;
}

// element range [begin, end) of the flat buffer that rank `rank` owns (the slice the kernels above reduce / push)
extern "C" int b2_peer_slice(size_t n, int dtype, int world, int rank, size_t* begin, size_t* end) {
  B2_REQUIRE(begin && end && world >= 1 && rank >= 0 && rank < world, "b2_peer_slice: bad arguments");
  const size_t per_vec = 16 / b2_dtype_size(dtype);
  size_t v0, v1;
  slice_of(n / per_vec, world, rank, 64 / per_vec, v0, v1);
  *begin = v0 * per_vec;
  *end = v1 * per_vec;
  return B2_OK;
}

#ifdef B2_BUILD_PROBES
// ---------------------------------------------------------------------------------------------------------------------
// Measurement aid (tools/peer_probe.py, not part of the documented ABI): raw peer copy in one direction with a chosen
// load flavour, unroll and grid, to find out what the NVLink path sustains from a kernel.
//   mode 0: remote read  -> local write      mode 1: local read -> remote write
//   flavour 0: ld.relaxed.sys   1: plain ld.global   2: ld.volatile
namespace {
template <int U, int FLAVOUR>
__global__ void __launch_bounds__(512) peer_copy_probe_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t nvec) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = i + (size_t)u * stride;
      if (j < nvec) {
        if (FLAVOUR == 0) v[u] = ld_peer(src + j);
        else if (FLAVOUR == 1) v[u] = src[j];
        else asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + j));
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = i + (size_t)u * stride;
      if (j < nvec) dst[j] = v[u];
    }
  }
}
}  // namespace

extern "C" int b2_debug_peer_copy(const void* src, void* dst, size_t bytes, int unroll, int flavour, int blocks, cudaStream_t stream) {
  const size_t nvec = bytes / 16;
#define B2_PROBE(U, F) if (unroll == U && flavour == F) { peer_copy_probe_kernel<U, F><<<blocks, 512, 0, stream>>>(static_cast<const uint4*>(src), static_cast<uint4*>(dst), nvec); B2_CHECK_CUDA(cudaGetLastError()); return B2_OK; }
  B2_PROBE(1, 0) B2_PROBE(2, 0) B2_PROBE(4, 0) B2_PROBE(8, 0)
  B2_PROBE(1, 1) B2_PROBE(2, 1) B2_PROBE(4, 1) B2_PROBE(8, 1)
  B2_PROBE(1, 2) B2_PROBE(4, 2)
#undef B2_PROBE
  b2_set_error("b2_debug_peer_copy: unsupported unroll/flavour");
  return B2_ERR_INVALID;
}

#endif  // B2_BUILD_PROBES

// ---------------------------------------------------------------------------------------------------------------------
// NVLS variant: the same in-place all-reduce, but the sum over ranks is formed INSIDE the NVSwitch.
// `mc` is a multicast address that aliases the same offset of every rank's buffer: multimem.ld_reduce returns the
// switch-side sum of all W copies (fp32 accumulation), multimem.st writes one value into all W copies. Each GPU then
// sends and receives about n*s bytes in total instead of 2*(W-1)/W*n*s per direction, which is what lets the result
// exceed the per-link roofline at W = 4 and 8 (reported against the same 2*(W-1)/W*n*s busbw formula, and labelled).
namespace {

template <int DT> struct McOps;
template <> struct McOps<B2_BF16> {
  __device__ __forceinline__ static uint4 ld_reduce(const void* p) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
  }
  __device__ __forceinline__ static void st(void* p, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
};
template <> struct McOps<B2_F32> {
  __device__ __forceinline__ static uint4 ld_reduce(const void* p) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
  }
  __device__ __forceinline__ static void st(void* p, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
};

template <int DT, bool SCATTER_ONLY>
__global__ void __launch_bounds__(512) peer_allreduce_mc_kernel(const __grid_constant__ PeerPtrs pp, void* mc, int world,
                                                                int rank, size_t v0, size_t v1, float scale,
                                                                uint32_t epoch, int want_sumsq, float* sumsq_out) {
  using V = PeerVec<DT>;
  __shared__ float red[32];
  start_barrier(pp, world, rank, epoch);
  float ss = 0.f;
  constexpr int U = 4;                               // switch round trips are long: keep four reductions in flight per thread
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = v0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < v1; i0 += stride * U) {
    uint4 sum[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i < v1) sum[u] = McOps<DT>::ld_reduce(static_cast<const uint4*>(mc) + i);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i >= v1) break;
      float x[8];
      V::unpack(sum[u], x);
#pragma unroll
      for (int e = 0; e < 8; ++e) x[e] *= scale;
      const uint4 out = V::pack(x);
      if (want_sumsq) {
        float y[8];
        V::unpack(out, y);
#pragma unroll
        for (int e = 0; e < 8; ++e) ss = fmaf(y[e], y[e], ss);
      }
      if (SCATTER_ONLY) st_peer(static_cast<uint4*>(pp.buf[rank]) + i, out);
      else McOps<DT>::st(static_cast<uint4*>(mc) + i, out);
    }
  }
  finish_reduce(pp, world, rank, epoch, want_sumsq, ss, red, sumsq_out);
}

}  // namespace

extern "C" int b2_peer_allreduce_mc(void* mc_buf, void* const* peer_bufs, uint32_t* const* peer_flags, int world, int rank,
                                    size_t offset_elems, size_t n, int dtype, float scale, unsigned epoch, int want_sumsq,
                                    int scatter_only, int max_blocks, float* sumsq_out, cudaStream_t stream) {
  B2_REQUIRE(mc_buf, "b2_peer_allreduce_mc: null multicast pointer");
  B2_REQUIRE(dtype == B2_F32 || dtype == B2_BF16, "b2_peer_allreduce_mc: dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(rank >= 0 && rank < world, "b2_peer_allreduce_mc: rank %d outside world %d", rank, world);
  const size_t es = b2_dtype_size(dtype), per_vec = 16 / es;
  B2_REQUIRE(n % per_vec == 0, "b2_peer_allreduce_mc: n (%zu) must be a multiple of %zu elements (16 bytes)", n, per_vec);
  PeerPtrs pp;
  int rc = fill_ptrs(pp, peer_bufs, peer_flags, world, offset_elems * es, "b2_peer_allreduce_mc");
  if (rc != B2_OK) return rc;
  void* mc = static_cast<char*>(mc_buf) + offset_elems * es;
  B2_REQUIRE(b2_aligned(mc, 16), "b2_peer_allreduce_mc: multicast pointer (+offset) must be 16-byte aligned");
  size_t v0, v1;
  slice_of(n / per_vec, world, rank, 64 / per_vec, v0, v1);
  // the switch does the adding: a few dozen CTAs keep it busy, more only contend (measured at W = 8: 32 CTAs 842 GB/s,
  // 296 CTAs 800 GB/s on 2 GiB; profiles/README.md)
  const int mc_cap = 32;
  const unsigned grid = peer_grid(v1 - v0, (max_blocks > 0 && max_blocks < mc_cap) ? max_blocks : mc_cap);
  if (dtype == B2_BF16) {
    if (scatter_only) peer_allreduce_mc_kernel<B2_BF16, true><<<grid, 512, 0, stream>>>(pp, mc, world, rank, v0, v1, scale, epoch, want_sumsq, sumsq_out);
    else peer_allreduce_mc_kernel<B2_BF16, false><<<grid, 512, 0, stream>>>(pp, mc, world, rank, v0, v1, scale, epoch, want_sumsq, sumsq_out);
  } else {
    if (scatter_only) peer_allreduce_mc_kernel<B2_F32, true><<<grid, 512, 0, stream>>>(pp, mc, world, rank, v0, v1, scale, epoch, want_sumsq, sumsq_out);
    else peer_allreduce_mc_kernel<B2_F32, false><<<grid, 512, 0, stream>>>(pp, mc, world, rank, v0, v1, scale, epoch, want_sumsq, sumsq_out);
  }
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}
