// Gradient landing: many separately allocated gradient tensors copied into their slices of the flat gradient buffer by
// ONE launch.
//
// Replaces the per-parameter copies with which torch's DistributedDataParallel fills its buckets during backward
// (Reducer::mark_variable_ready_dense -> bucket_view.copy_(grad), installed by ray.train.torch.prepare_model, ref:
// ray-jobs/pytorch_llm_ray.py:230,276) — there one copy kernel per parameter (GPT-2: 148 launches moving 249 MB at
// under 1 TB/s, most of them a few KB), here one grid over all segments. HBM-bound: sum(bytes) read + written once.
#include "b2_common.cuh"

#define B2_LAND_MAX_SEGS 112
#define B2_LAND_CHUNK (32 * 1024)      // bytes handled by one block at a time

namespace {

struct LandSeg {
  const char* src;
  unsigned long long dst_off;          // bytes from the flat buffer's base
  unsigned long long bytes;
  unsigned int chunk0;                 // index of this segment's first chunk in the launch
  unsigned int pad;
};
struct LandTable {                     // passed by value (kernel parameter space, < 4 KiB)
  LandSeg seg[B2_LAND_MAX_SEGS];
  int count;
  unsigned int total_chunks;
};

__global__ void __launch_bounds__(256) land_copy_kernel(const __grid_constant__ LandTable t, char* __restrict__ dst_base) {
  for (unsigned int c = blockIdx.x; c < t.total_chunks; c += gridDim.x) {
    // the segment that holds chunk c: last one whose first chunk is <= c (uniform over the block)
    int lo = 0, hi = t.count - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (t.seg[mid].chunk0 <= c) lo = mid; else hi = mid - 1;
    }
    const LandSeg& s = t.seg[lo];
    const unsigned long long off = (unsigned long long)(c - s.chunk0) * B2_LAND_CHUNK;
    const unsigned long long left = s.bytes - off;
    const unsigned int n = left < B2_LAND_CHUNK ? (unsigned int)left : (unsigned int)B2_LAND_CHUNK;
    const char* src = s.src + off;
    char* dst = dst_base + s.dst_off + off;
    const unsigned int nvec = n >> 4;
    const uint4* sv = reinterpret_cast<const uint4*>(src);
    uint4* dv = reinterpret_cast<uint4*>(dst);
    // four independent 16-byte loads per thread before the stores
    for (unsigned int i = threadIdx.x; i < nvec; i += 4 * blockDim.x) {
      uint4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned int j = i + k * blockDim.x;
        if (j < nvec) v[k] = ld_stream(sv + j);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const unsigned int j = i + k * blockDim.x;
        if (j < nvec) dv[j] = v[k];
      }
    }
    const unsigned int tail = n & 15u;               // a segment whose size is not a multiple of 16 bytes ends in its last chunk
    if (threadIdx.x < tail) dst[(nvec << 4) + threadIdx.x] = src[(nvec << 4) + threadIdx.x];
  }
}

}  // namespace

extern "C" int b2_land_copy(const void* const* srcs, const size_t* dst_offsets_bytes, const size_t* nbytes, int count,
                            void* dst_base, cudaStream_t stream) {
  if (count == 0) return B2_OK;
  B2_REQUIRE(count > 0 && srcs && dst_offsets_bytes && nbytes && dst_base, "b2_land_copy: bad arguments");
  B2_REQUIRE(b2_aligned(dst_base, 16), "b2_land_copy: destination base must be 16-byte aligned");
  int done = 0;
  while (done < count) {
    LandTable t;
    t.count = 0;
    t.total_chunks = 0;
    while (done < count && t.count < B2_LAND_MAX_SEGS) {
      const size_t b = nbytes[done];
      if (b == 0) { ++done; continue; }
      B2_REQUIRE(srcs[done] != nullptr, "b2_land_copy: null source %d", done);
      B2_REQUIRE(b2_aligned(srcs[done], 16) && dst_offsets_bytes[done] % 16 == 0,
                 "b2_land_copy: segment %d: source and destination offset must be 16-byte aligned", done);
      const size_t chunks = (b + B2_LAND_CHUNK - 1) / B2_LAND_CHUNK;
      B2_REQUIRE(chunks < (1ull << 31) && (size_t)t.total_chunks + chunks < (1ull << 31), "b2_land_copy: segment %d too large", done);
      LandSeg& s = t.seg[t.count++];
      s.src = static_cast<const char*>(srcs[done]);
      s.dst_off = dst_offsets_bytes[done];
      s.bytes = b;
      s.chunk0 = t.total_chunks;
      s.pad = 0;
      t.total_chunks += (unsigned int)chunks;
      ++done;
    }
    if (t.count == 0) break;
    unsigned int grid = t.total_chunks;
    if (grid > (unsigned int)B2_NUM_SMS * 8) grid = B2_NUM_SMS * 8;
    land_copy_kernel<<<grid, 256, 0, stream>>>(t, static_cast<char*>(dst_base));
    B2_CHECK_CUDA(cudaGetLastError());
  }
  return B2_OK;
}
