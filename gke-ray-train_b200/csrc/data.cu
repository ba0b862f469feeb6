// Input side of the step: stride-1 sliding windows gathered on the device.
//
// Replaces TextDataset.__getitem__ + the DataLoader collate + the pinned H2D copy that prepare_data_loader adds
// (ref: ray-jobs/pytorch_llm_ray.py:107-119, 206-216): with the token stream resident in HBM, a batch is
//   inputs[b, t]  = tokens[start_b + t]          targets[b, t] = tokens[start_b + t + 1]      (labels pre-shifted, ref :116-119)
// Integer copy, bit-exact by construction. HBM-bound: 2 * B * T * 8 bytes written, about as many read.
#include "b2_common.cuh"

namespace {
__global__ void __launch_bounds__(256) gather_windows_kernel(const long long* __restrict__ tokens, long long n_tokens,
                                                             const long long* __restrict__ starts, int B, int T,
                                                             long long* __restrict__ inputs, long long* __restrict__ targets,
                                                             int* __restrict__ err) {
  const long long total = (long long)B * T;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / T), t = (int)(i - (long long)b * T);
    const long long s = starts[b];
    if (s < 0 || s + T + 1 > n_tokens) {          // window (plus its shifted target) must lie inside the stream
      if (err) *err = 1;
      inputs[i] = 0; targets[i] = 0;
      continue;
    }
    inputs[i] = tokens[s + t];
    targets[i] = tokens[s + t + 1];
  }
}
}  // namespace

extern "C" int b2_gather_windows(const long long* tokens, long long n_tokens, const long long* starts, int B, int T,
                                 long long* inputs, long long* targets, int* err_out, cudaStream_t stream) {
  if (B == 0 || T == 0) return B2_OK;
  B2_REQUIRE(B > 0 && T > 0 && n_tokens > 0, "b2_gather_windows: bad shape B=%d T=%d n_tokens=%lld", B, T, n_tokens);
  B2_REQUIRE(tokens && starts && inputs && targets, "b2_gather_windows: null pointer");
  const long long total = (long long)B * T;
  long long blocks = (total + 255) / 256;
  if (blocks > B2_NUM_SMS * 8) blocks = B2_NUM_SMS * 8;
  gather_windows_kernel<<<(unsigned)blocks, 256, 0, stream>>>(tokens, n_tokens, starts, B, T, inputs, targets, err_out);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}
