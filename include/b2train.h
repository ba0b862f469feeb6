/* b2train.h — C ABI of libb2train.so: the B200 (sm_100a) kernels for the data-parallel training-step
 * tail of ericehanley/gke-ray-train (LM head + cross-entropy, gradient norm/clip, AdamW, peer-memory
 * gradient all-reduce).
 *
 * Conventions (all entry points):
 *   - plain pointers and sizes only; every pointer is a CUDA device pointer unless stated otherwise;
 *   - the caller owns every buffer; nothing is allocated or freed on the caller's behalf unless the
 *     function's name says alloc/open/free/close;
 *   - work is enqueued asynchronously on `stream` (a cudaStream_t passed as void*); the call returns
 *     before the kernels run;
 *   - return value: 0 = ok, <0 = error (B2_ERR_*); b2_last_error() gives the message for the calling
 *     thread. Nothing throws;
 *   - dtype codes: B2_F32 = 0, B2_BF16 = 1. Buffers must be 16-byte aligned.
 *
 * "ref:" citations name the file:line of the reference (ericehanley/gke-ray-train) call that the entry
 * point replaces; "$SP" is the site-packages tree of the third-party library that call lands in.
 */
#ifndef B2TRAIN_H
#define B2TRAIN_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_OK 0
#define B2_ERR_INVALID (-1)
#define B2_ERR_CUDA (-2)
#define B2_ERR_LIMIT (-3)
#define B2_ERR_STATE (-4)
#define B2_F32 0
#define B2_BF16 1

typedef void* b2_stream_t; /* cudaStream_t */

/* ---- library ------------------------------------------------------------------------------------ */
int b2_version(void);                 /* 100 = 0.1.0 */
const char* b2_last_error(void);      /* message of the last failing call on this thread */
int b2_device_count(void);            /* CUDA devices visible to this process (0 if none / no driver) */

/* ---- F3: fused AdamW --------------------------------------------------------------------------------
 * Replaces optimizer.step() of torch.optim.AdamW — ref: ray-jobs/pytorch_llm_ray.py:236,278 (lr 3e-4,
 * weight_decay 0.01) and the HF Trainer optimizer configured by ref: ray-jobs/fine_tune_config.json:15-17.
 * Arithmetic: $SP/torch/optim/adam.py:347-552 (decoupled weight decay, amsgrad off, maximize off).
 * One pass over n elements of p (params), g (grads), m (exp_avg), v (exp_avg_sq); `step` is the 1-based
 * step count AFTER the increment torch performs; bias corrections are formed in double on the host.
 * gscale_dev (nullable): device float; g is multiplied by it (and rounded to g's dtype) before use —
 * this is where clip_grad_norm_'s coefficient is applied without a separate pass over the gradients. */
int b2_adamw(void* p, const void* g, void* m, void* v, size_t n, int p_dtype, int g_dtype, int state_dtype,
             double lr, double beta1, double beta2, double eps, double weight_decay, long long step,
             const float* gscale_dev, b2_stream_t stream);

/* Grouped variant for optimizers with several param_groups (HF Trainer builds two: weight decay on / off,
 * $SP/transformers/trainer.py:1157-1170) laid out in ONE flat buffer: chunk_group_dev[c] (uint8, device) names the
 * group (0..ngroups-1, at most 8) of elements [64c, 64c+64); 255 = leave the chunk untouched. n % 64 == 0.
 * groups_host is a HOST array of ngroups hyper-parameter records. */
typedef struct B2AdamWGroup { double lr, beta1, beta2, eps, weight_decay; long long step; } B2AdamWGroup;
int b2_adamw_grouped(void* p, const void* g, void* m, void* v, size_t n, int p_dtype, int g_dtype, int state_dtype,
                     const uint8_t* chunk_group_dev, const B2AdamWGroup* groups_host, int ngroups,
                     const float* gscale_dev, b2_stream_t stream);

/* ---- gradient norm and clipping ------------------------------------------------------------------
 * Replace torch.nn.utils.clip_grad_norm_ — ref: ray-jobs/pytorch_llm_ray.py:277 (max_norm 1.0); HF
 * MAX_GRAD_NORM ref: ray-jobs/fine_tune_config.json:19. Arithmetic: $SP/torch/nn/utils/clip_grad.py:50-183.
 * b2_sumsq: out[0] (=, or += when accumulate!=0) sum of squares of g[0..n). `scratch` must hold
 *   b2_sumsq_scratch_bytes() bytes, be zero-filled once before its first use, and not be shared by calls
 *   that may overlap on the GPU.
 * b2_clip_coef: total = sqrt(sum of parts[0..nparts)); norm_out[0] = total;
 *   coef_out[0] = min(1, max_norm / (total + 1e-6)). Either output may be NULL.
 * b2_scale: g[0..n) *= coef_dev[0] in place. */
size_t b2_sumsq_scratch_bytes(void);
int b2_sumsq(const void* g, size_t n, int dtype, float* scratch, float* out, int accumulate, b2_stream_t stream);
int b2_clip_coef(const float* sumsq_parts, int nparts, float max_norm, float* norm_out, float* coef_out,
                 b2_stream_t stream);
int b2_scale(void* g, size_t n, int dtype, const float* coef_dev, b2_stream_t stream);

/* b2_sumsq for callers that cannot supply scratch memory. It has no counterpart in the reference (the reference never
 * manages device memory itself); it exists for hosts that only hold the pointers they were handed, for processes whose
 * framework allocator already owns the whole GPU, and for call sites — a gradient hook on the autograd thread, say —
 * where an allocation's implicit device synchronisation is unwelcome. The partial sums live in a fixed 2.5 KB array
 * inside the library image (one copy per GPU that loads the library, held until exit); the call allocates nothing.
 *   - n <= b2_sumsq_static_max_elems() (2^31) per call, else B2_ERR_LIMIT; split larger buffers and accumulate;
 *   - the first call on a device resolves the array's address (cudaGetSymbolAddress): do it before any stream capture;
 *   - every call on a device shares the array (it keeps the previous call's contents; the one word that must be zero
 *     is cleared per call), so calls that can overlap on the GPU race — keep them on one stream or order them.
 * out / accumulate as for b2_sumsq. */
size_t b2_sumsq_static_max_elems(void);
int b2_sumsq_static(const void* g, size_t n, int dtype, float* out, int accumulate, b2_stream_t stream);

/* ---- cross-entropy over materialised logits -------------------------------------------------------
 * Replace nn.CrossEntropyLoss()(outputs.reshape(-1,V), targets.reshape(-1)) — ref:
 * ray-jobs/pytorch_llm_ray.py:237,275 (labels pre-shifted by TextDataset, ref :116-119).
 * logits: [N, ld] row-major, first V columns used. labels: int64 [N].
 * shift_T: 0 = labels already aligned with rows; T>0 = rows are [B, T] and row (b,t) is scored against
 *   labels[b, t+1], the last position of each sequence being ignored (HF ForCausalLMLoss,
 *   $SP/transformers/loss/loss_utils.py:45-67).
 * Outputs: lse[N] (log-sum-exp per row), zt[N] (label logit per row), loss_out[0] = sum over non-ignored
 *   rows of (lse - z), loss_out[1] = that sum / n_valid (NaN when n_valid == 0, as torch's mean);
 *   n_valid_out[0] = number of rows whose label != ignore_index (int64, exact);
 *   err_out[0] (nullable) = 1 if any non-ignored label is outside [0, V) — torch raises there; here the
 *   loss is additionally set to NaN.
 * b2_ce_rows_bwd: dlogits[i, j] = (exp(logits[i,j] - lse[i]) - [j == label_i]) * gscale_dev[0] for
 *   non-ignored rows, 0 otherwise; dlogits may alias logits. gscale = grad_output / n_valid for 'mean'. */
int b2_ce_rows_fwd(const void* logits, int dtype, long long N, long long V, long long ld, const long long* labels,
                   int shift_T, long long ignore_index, float* lse, float* zt, float* loss_out,
                   long long* n_valid_out, int* err_out, b2_stream_t stream);
int b2_ce_rows_bwd(const void* logits, int dtype, long long N, long long V, long long ld, const long long* labels,
                   int shift_T, long long ignore_index, const float* lse, const float* gscale_dev, void* dlogits,
                   b2_stream_t stream);

/* ---- F1: LM head fused with cross-entropy (tcgen05 / TMEM / TMA) ------------------------------------
 * Replaces lm_head(hidden) followed by the causal-LM loss — ref: ray-jobs/fine_tune_llama_ray.py:333
 * (trainer.train() -> $SP/transformers/models/llama/modeling_llama.py:486-491 ->
 * $SP/transformers/loss/loss_utils.py:45-67) and BasicLLM.fc_out + criterion, ref:
 * ray-jobs/pytorch_llm_ray.py:87,104,237,275. b2_lmhead_ce_fwd never writes the [N, V] logits to
 * memory (b2_lmhead_ce_fwd_keep optionally leaves a bf16 copy for the backward; the fp32 [N, V] tensors of the reference
 * never exist).
 * H: [N, ldh] bf16 hidden states (D used columns); W: [V, ldw] bf16 (nn.Linear weight layout);
 * bias: fp32 [V] or NULL. ldh, ldw multiples of 8.
 * round_logits_bf16 != 0 rounds each logit to bf16 before the soft-max, which is what the reference's
 *   bf16 lm_head followed by .float() computes.
 * workspace: b2_lmhead_ce_workspace_bytes(N) bytes of scratch, contents irrelevant.
 * Outputs as for b2_ce_rows_fwd. */
size_t b2_lmhead_ce_workspace_bytes(long long N);
int b2_lmhead_ce_fwd(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                     const long long* labels, long long N, int D, int V, int shift_T, long long ignore_index,
                     int round_logits_bf16, void* workspace, size_t workspace_bytes, float* lse, float* zt,
                     float* loss_out, long long* n_valid_out, int* err_out, b2_stream_t stream);
/* Forward that also KEEPS the logits: as b2_lmhead_ce_fwd with round_logits_bf16 = 1, and additionally stores the
 * bf16-rounded logits (the values the loss was computed from) to logits_out [N, ldl] (ldl >= V, ldl % 8 == 0, 16-byte
 * aligned; columns [V, ldl) are padding, unspecified afterwards). For callers that can afford N * ldl * 2 bytes between
 * forward and backward: the backward then needs no second pass over the tensor cores, see b2_ce_dlogits_inplace.
 * Same reference lines as b2_lmhead_ce_fwd — there the materialised tensor is `logits` itself
 * ($SP/transformers/models/llama/modeling_llama.py:486; fp32 copy at $SP/transformers/loss/loss_utils.py:57). */
int b2_lmhead_ce_fwd_keep(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                          const long long* labels, long long N, int D, int V, int shift_T, long long ignore_index,
                          void* workspace, size_t workspace_bytes, float* lse, float* zt, float* loss_out,
                          long long* n_valid_out, int* err_out, void* logits_out, long long ldl, b2_stream_t stream);
/* Backward from kept logits, in place: logits[r, j] (bf16 [rows, ldl], as left by b2_lmhead_ce_fwd_keep; the pointer is
 * row row0 of that matrix) becomes dlogits[r, j] = (exp(logit - lse[row0+r]) - [j == label]) * gscale_dev[0], 0 for
 * ignored rows, bit-identical to what b2_lmhead_ce_bwd_dlogits writes; padding columns of the last 16-byte vector of a
 * row are zeroed. HBM-bound (one read + one write of the matrix). labels and lse are the full arrays. Replaces the
 * autograd backward of F.cross_entropy on the materialised logits ($SP/transformers/loss/loss_utils.py:28-42). */
int b2_ce_dlogits_inplace(void* logits, long long ldl, const long long* labels, long long row0, long long rows, int V,
                          int shift_T, long long ignore_index, const float* lse, const float* gscale_dev,
                          b2_stream_t stream);
/* Backward, step 1: recomputes the logits of rows [row0, row0+rows) and writes
 *   dlogits[r, j] = (exp(logit - lse[row0+r]) - [j == label]) * gscale_dev[0]   (bf16, [rows, ldd], ldd % 8 == 0;
 *   columns [V, ldd) are padding, unspecified afterwards)
 * H points at row row0 of the hidden states; labels and lse are the full arrays (indexed by row0 + r).
 * dH and dW then follow from two b2_gemm_bf16 calls on this chunk. */
int b2_lmhead_ce_bwd_dlogits(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                             const long long* labels, long long row0, long long rows, int D, int V, int shift_T,
                             long long ignore_index, int round_logits_bf16, const float* lse,
                             const float* gscale_dev, void* dlogits, long long ldd, b2_stream_t stream);
/* The LM-head backward of a chunk of rows behind ONE call — replaces `loss.backward()` through ForCausalLMLoss and
 * lm_head (ref: ray-jobs/pytorch_llm_ray.py:276; $SP/transformers/loss/loss_utils.py:28-67 and the autograd of
 * nn.Linear at $SP/transformers/models/llama/modeling_llama.py:486; bias: BasicLLM.fc_out, ref :87). Enqueues, in order:
 *   d(logits) for rows [row0, row0+rows) into `dlogits` (bf16 [rows, ldd]): kept != 0 — `dlogits` holds the logits that
 *     b2_lmhead_ce_fwd_keep left for these rows and is rewritten in place (b2_ce_dlogits_inplace); kept == 0 — recomputed
 *     on the tensor cores (b2_lmhead_ce_bwd_dlogits);
 *   dH [rows, lddh] bf16 = dlogits @ W                     (skipped when dH is NULL);
 *   dW [V, lddw] (dw_dtype f32 / bf16) = or += dlogits^T @ H   (skipped when dW is NULL; dw_accumulate != 0 adds);
 *   db [V] fp32 = or += column sums of dlogits              (skipped when db is NULL; fixed summation order).
 * H points at row row0 of the hidden states; labels and lse are the full arrays. gemm_workspace (optional):
 * b2_gemm_workspace_bytes(rows, D) bytes that let the dH GEMM split K. */
int b2_lmhead_ce_bwd(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                     const long long* labels, long long row0, long long rows, int D, int V, int shift_T,
                     long long ignore_index, int round_logits_bf16, const float* lse, const float* gscale_dev,
                     void* dlogits, long long ldd, int kept, void* dH, long long lddh, void* dW, int dw_dtype,
                     long long lddw, int dw_accumulate, float* db, int db_accumulate, void* gemm_workspace,
                     size_t gemm_workspace_bytes, b2_stream_t stream);
/* Plain bf16 GEMM on the same tcgen05 mainloop: C[M,N] (+)= A[M,K] * B[N,K]^T, fp32 accumulate.
 * a_mn_major = 0: A stored [M, lda] with K contiguous; 1: A stored [K, lda] with M contiguous. Same for B
 * with N. C row-major [M, ldc], fp32 or bf16; accumulate != 0 adds to the existing C.
 * Used for dH = dlogits * W and dW += dlogits^T * H, replacing the autograd GEMMs of nn.Linear
 * (ref: the loss.backward() at ray-jobs/pytorch_llm_ray.py:276). */
int b2_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major, long long ldb, void* C,
                 int c_dtype, long long ldc, int M, int N, int K, int accumulate, void* workspace,
                 size_t workspace_bytes, b2_stream_t stream);
/* Optional split-K scratch for b2_gemm_bf16: when `workspace` holds at least b2_gemm_workspace_bytes(M, N) bytes and
 * the output has too few 128x256 tiles to occupy the 148 SMs (e.g. dH: M = tokens, N = D, K = vocabulary), the K
 * range is split up to 6 ways into fp32 partials that a second kernel adds in a fixed order. NULL = never split. */
size_t b2_gemm_workspace_bytes(int M, int N);

/* ---- gradient landing: many gradient tensors copied into the flat gradient buffer by one launch -------------------
 * Replaces the per-parameter bucket copies of torch's DistributedDataParallel during backward (Reducer::
 * mark_variable_ready_dense: bucket_view.copy_(grad); wrapper installed by ray.train.torch.prepare_model, ref:
 * ray-jobs/pytorch_llm_ray.py:230,276). srcs / dst_offsets_bytes / nbytes: HOST arrays of `count` entries — segment k is
 * nbytes[k] bytes at device address srcs[k], copied to dst_base + dst_offsets_bytes[k]. Sources and destination offsets
 * 16-byte aligned; sizes arbitrary; segments must not overlap. Any number of segments (112 per launch internally). */
int b2_land_copy(const void* const* srcs, const size_t* dst_offsets_bytes, const size_t* nbytes, int count, void* dst_base,
                 b2_stream_t stream);

/* ---- input side: sliding token windows gathered on the device -------------------------------------------------
 * Replaces TextDataset.__getitem__ + DataLoader collate + the host-to-device copy of prepare_data_loader — ref:
 * ray-jobs/pytorch_llm_ray.py:107-119, 206-216. tokens: int64 [n_tokens] resident on the device; starts: int64 [B];
 * inputs[b,t] = tokens[starts[b]+t], targets[b,t] = tokens[starts[b]+t+1] (int64 [B,T] each). A window that does not
 * fit (start < 0 or start + T + 1 > n_tokens) sets err_out[0] = 1 (nullable) and yields zeros. */
int b2_gather_windows(const long long* tokens, long long n_tokens, const long long* starts, int B, int T,
                      long long* inputs, long long* targets, int* err_out, b2_stream_t stream);

/* ---- F2: gradient all-reduce over peer memory (NVLink 5 / NVSwitch) ----------------------------------
 * Replaces the bucket all-reduce (+ the grad / world_size copy-in) that DistributedDataParallel performs during
 * loss.backward() once ray.train.torch.prepare_model has wrapped the model — ref: ray-jobs/pytorch_llm_ray.py:230,276;
 * $SP/torch/nn/parallel/distributed.py:1185-1280, $SP/torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:19-31.
 *
 * Symmetric memory: b2_symm_alloc cudaMallocs `bytes` (zero-filled) and returns a 64-byte CUDA IPC handle that the
 * other processes of the job turn into a mapping of that memory with b2_symm_open (exchange the handles over the
 * process group the launcher already set up). The owner frees with b2_symm_free after every peer has
 * b2_symm_close'd its mapping.
 * Flag block: each rank owns b2_peer_flags_bytes() bytes of symmetric memory, zero-filled once, used for the
 * in-kernel barriers and the sum-of-squares exchange.
 *
 * peer_bufs / peer_flags: HOST arrays of `world` device pointers — entry k addresses rank k's buffer / flag block
 * as mapped in THIS process (entry `rank` is the local allocation). world <= 8.
 * b2_peer_allreduce, in place on elements [offset_elems, offset_elems + n) of every rank's buffer:
 *   buf[i] <- round_dtype(scale * sum_k buf_k[i]) on all ranks (scale = 1/world for DDP's mean), bit-identical
 *   everywhere. want_sumsq != 0 additionally leaves, in each rank's flag block, float slot 32 + k = sum of squares
 *   of the reduced values of rank k's slice (k = 0..world-1) — feed those `world` floats to b2_clip_coef.
 *   want_sumsq = 2 keeps adding to the slots instead of restarting them (several bucket calls of one backward pass).
 *   The per-block partials are added in block order by the last block to finish, so the sums are reproducible run to run.
 *   sumsq_out (may be NULL): local device array of `world` floats that receives a copy of the `world` slots once they
 *   are all complete (the kernel's last action) — no separate launch to collect them.
 *   scatter_only != 0 stores the reduced slice of rank r on rank r only (reduce-scatter; see b2_peer_slice).
 *   max_blocks > 0 caps the grid (a bucket reduced while backward kernels still run should not occupy every SM).
 * b2_peer_allgather: pushes this rank's slice of elements [offset_elems, offset_elems + n) to every peer (used for
 *   updated parameters; max_blocks as above).
 * Slices: rank r's slice of an n-element range is contiguous, the same on every rank, and starts on a 64-element
 *   boundary relative to the range (b2_peer_slice returns it) — the granularity of b2_adamw_grouped's chunk table.
 * `epoch` must increase by one on every collective call, identically on all ranks (start at 1).
 * Every rank must enqueue the same sequence of calls; the kernels wait for one another inside the GPU, so each
 * rank needs its own device. A peer that never arrives makes the kernel trap after ~10 min instead of hanging. */
size_t b2_peer_flags_bytes(void);
int b2_symm_alloc(size_t bytes, void** dptr, unsigned char handle[64]);
int b2_symm_open(const unsigned char handle[64], void** dptr);
int b2_symm_close(void* dptr);
int b2_symm_free(void* dptr);
int b2_peer_allreduce(void* const* peer_bufs, uint32_t* const* peer_flags, int world, int rank, size_t offset_elems,
                      size_t n, int dtype, float scale, unsigned epoch, int want_sumsq, int scatter_only,
                      int max_blocks, float* sumsq_out, b2_stream_t stream);
/* NVLS variant of b2_peer_allreduce: `mc_buf` is a multicast address aliasing the same offset of every rank's buffer
 * (cuMulticast* / torch.distributed._symmetric_memory); the sum over ranks is formed inside the NVSwitch
 * (multimem.ld_reduce, fp32 accumulation) and the result is broadcast with multimem.st. Same semantics and arguments
 * otherwise; peer_bufs[rank] must still be the local mapping of the buffer. */
int b2_peer_allreduce_mc(void* mc_buf, void* const* peer_bufs, uint32_t* const* peer_flags, int world, int rank,
                         size_t offset_elems, size_t n, int dtype, float scale, unsigned epoch, int want_sumsq,
                         int scatter_only, int max_blocks, float* sumsq_out, b2_stream_t stream);
int b2_peer_allgather(void* const* peer_bufs, uint32_t* const* peer_flags, int world, int rank, size_t offset_elems,
                      size_t n, int dtype, unsigned epoch, int max_blocks, b2_stream_t stream);
int b2_peer_slice(size_t n, int dtype, int world, int rank, size_t* begin, size_t* end);

#ifdef B2_BUILD_PROBES
/* ---- measurement aid (tools/peer_probe.py; no counterpart in the reference; only in a library built with
 * -DB2_BUILD_PROBES, i.e. `B2_BUILD_PROBES=1 python gke-ray-train_b200/build.py`) ---------------------------------
 * Raw one-directional copy between a local and a peer-mapped buffer with a chosen load flavour (0 ld.relaxed.sys,
 * 1 plain ld.global, 2 ld.volatile), unroll (1/2/4/8) and grid: what the NVLink path sustains from a kernel. */
int b2_debug_peer_copy(const void* src, void* dst, size_t bytes, int unroll, int flavour, int blocks, b2_stream_t stream);
#endif

#ifdef __cplusplus
}
#endif
#endif /* B2TRAIN_H */
