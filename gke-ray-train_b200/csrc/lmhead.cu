// F1 — LM head fused with causal cross-entropy on the 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
// Replaces, for bf16 models, the pair the reference executes at the end of every forward:
//   logits = lm_head(hidden)                  (transformers/models/llama/modeling_llama.py:486-491,
//                                              reached from trainer.train(), ref: ray-jobs/fine_tune_llama_ray.py:333;
//                                              BasicLLM.fc_out, ref: ray-jobs/pytorch_llm_ray.py:87,104)
//   loss   = ForCausalLMLoss(logits, labels)  (transformers/loss/loss_utils.py:45-67: .float(), shift, CE with
//                                              ignore_index=-100; nn.CrossEntropyLoss at ref: pytorch_llm_ray.py:237,275)
// The fp32 [tokens x vocab] logits are never written to memory: each 128x256 logits tile lives in TMEM, and
// the epilogue warps fold it into a per-row running (max, sum-exp) and pick out the label's logit.
// Backward recomputes the tile and writes d(logits) for a bounded chunk of rows only (bf16), which two
// plain GEMMs (same mainloop, MODE_GEMM) turn into dH and dW.
//
// One kernel template, four epilogues:
//   MODE_GEMM         C[M,N] (+)= A[M,K] * B[N,K]^T          (A/B each K-major or MN-major in memory)
//   MODE_CE_FWD       per (row, vocab split): (max, sumexp) partials + label logit
//   MODE_CE_BWD       dlogits[row, col] = (exp(logit - lse[row]) - [col == label]) * gscale   (0 for ignored rows)
//   MODE_CE_FWD_KEEP  MODE_CE_FWD that also stores the bf16-rounded logits; ce_dlogits_inplace_kernel then turns them
//                     into dlogits in place (HBM-bound), so the backward needs no MODE_CE_BWD pass. Used when the caller
//                     can hold [rows, V] bf16 between forward and backward (b200_train/loss.py: one chunk of rows).
//
// Tile 128 (M) x 256 (N) x 64 (K) per CTA, tcgen05.mma kind::f16 (bf16 in, fp32 out), fp32 accumulators double-buffered
// in TMEM (2 x 256 columns) so the epilogue of tile i overlaps the MMAs of tile i+1. Two shapes of the same template:
//   single CTAs (cta_group::1, M=128 N=256 K=16; 4-stage 48 KiB operand ring)         -> the two cross-entropy kernels
//   CTA pairs   (cta_group::2, M=256: each CTA holds its 128 rows of A and half of B;  -> the plain GEMMs
//                6-stage 32 KiB ring; the leader issues the MMAs and multicasts the commits)
// Warp roles (384 threads, 1 CTA/SM, persistent over work units): warp 0 = TMA producer, warp 1 = MMA issuer (one thread),
// warp 2 = TMEM allocator, warps 4-11 = epilogue (warp w owns TMEM lane quarter w%4 and one 128-column half; one thread =
// one row). The storing epilogues leave through 128B-swizzled shared memory and TMA stores, one 32x64 block per warp.
// Tensor-bound: algorithmic flops = 2*M*N*K per pass (measurements in profiles/README.md).
#include "b2_common.cuh"
#include "umma.cuh"
#include <math.h>
#include <stdlib.h>
#include <map>
#include <mutex>
#include <tuple>

int b2_ce_finalize_launch(const float* lse, const float* zt, const long long* labels, long long N, long long V,
                          int shift_T, long long ignore_index, float* loss_out, long long* n_valid_out, int* err_out,
                          cudaStream_t stream);

namespace {

using namespace umma;

constexpr int BM = 128, BN = 256, BK = 64, UK = 16;
constexpr int NUM_THREADS = 384;                    // warps: 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4-11 epilogue
constexpr int MAX_STAGES = 6;
constexpr int A_STAGE_BYTES = BM * BK * 2;            // 16 KiB: this CTA's 128 rows of A
// B per CTA: all 256 columns (32 KiB) when a CTA works alone, its half of them (16 KiB) in a CTA pair
constexpr int b_stage_bytes(bool pair) { return (pair ? BN / 2 : BN) * BK * 2; }
constexpr int stage_bytes(bool pair) { return A_STAGE_BYTES + b_stage_bytes(pair); }
// Operand ring: 192 KiB (4 x 48 KiB alone, 6 x 32 KiB in a pair), plus 32 KiB of staging for the storing modes: one 4 KiB
// [32 rows x 128 B] box per epilogue warp. (Round 2 tried two boxes per warp paid for with one ring stage — 3 / 5 stages:
// no gain for pairs, a loss for single CTAs; profiles/README.md.)
constexpr bool mode_has_staging(int mode) { return mode == 2 || mode == 3; }
constexpr int stages_for(int mode, bool pair) { return pair ? 6 : 4; }
constexpr int STAGING_BOX_BYTES = 32 * 128;           // [32 rows x 64 bf16]; the narrow form uses two 2 KiB halves of it
constexpr int STAGING_BYTES = 8 * STAGING_BOX_BYTES;
constexpr int smem_bytes_for(int mode, bool pair) {  // + 1 KiB slack to reach 1024-byte alignment
  return stages_for(mode, pair) * stage_bytes(pair) + 1024 + (mode_has_staging(mode) ? STAGING_BYTES : 0);
}
constexpr int BOX_MN_BYTES = 64 * BK * 2;             // one MN-major TMA box: 64 K-rows x 128 B
constexpr uint32_t TMEM_COLS = 512;
constexpr int MAX_SPLITS = 64;
#define B2_GEMM_MAX_KSPLIT 6
constexpr float LOG2E = 1.4426950408889634f;

enum { MODE_GEMM = 0, MODE_CE_FWD = 1, MODE_CE_BWD = 2, MODE_CE_FWD_KEEP = 3 };
// MODE_CE_FWD_KEEP: forward that ALSO leaves the bf16-rounded logits in memory, for the backward that transforms them in
// place (b2_ce_dlogits_inplace) instead of recomputing the GEMM
constexpr bool mode_is_fwd(int m) { return m == MODE_CE_FWD || m == MODE_CE_FWD_KEEP; }
constexpr bool mode_stores(int m) { return m == MODE_CE_BWD || m == MODE_CE_FWD_KEEP; }   // tiles leave through TMA stores

struct Params {
  int M, N, K;                       // rows (tokens), columns (vocab or features), reduction extent
  int num_m_blocks, num_n_blocks, num_k_blocks, n_splits;
  int k_splits, n_fastest;           // MODE_GEMM only: split-K factor; unit order (1: units sharing an A tile are adjacent)
  float* ws; long long ldws;         // split-K partial sums [k_splits][M][ldws] fp32
  // MODE_GEMM
  void* C; long long ldc; int c_dtype; int accumulate;
  // cross-entropy modes
  const long long* labels; int shift_T; long long ignore_index; long long row0;
  const float* bias; int round_bf16;
  float2* part; long long part_stride; float* zt;
  const float* lse; const float* gscale; __nv_bfloat16* dlogits; long long ldd;
  int store_wide;                    // storing modes: 1 = one [32 x 64] TMA store per two chunks, 0 = one [32 x 32] per chunk
  int debug_store;                   // measurement aid (B2_DEBUG_STORE): 1 = stage the tile but issue no TMA store, 2 = drop it
  int store_evict_normal;            // storing modes: 0 = evict-first hint on the tile stores (default), 1 = normal policy
  int store_direct;                  // storing modes: 1 = each thread writes its 64-byte row segment straight to global memory
                                     // (no shared-memory staging, no TMA store); needs p.dlogits / p.ldd
};

// Measurement aids (B2_DEBUG_STORE / B2_DEBUG_NOEPI, tools/f1_probe.py) exist only in a library built with
// -DB2_BUILD_PROBES (`B2_BUILD_PROBES=1 python build.py`); the shipped kernels carry none of their branches.
__device__ __forceinline__ bool probe_skips_store(const Params& p) {
#ifdef B2_BUILD_PROBES
  return p.debug_store != 0;
#else
  return false;
#endif
}

__device__ __forceinline__ long long label_at(const Params& p, long long i) {
  if (p.shift_T <= 0) return p.labels[i];
  return ((i % p.shift_T) + 1 < p.shift_T) ? p.labels[i + 1] : p.ignore_index;
}

struct Unit { int mb, s, ks, nb0, nb1, kb0, kb1; };

// Work unit u -> (128-row block, vocabulary/column range, K range). Units that are adjacent in u run at the same
// time on neighbouring CTAs, so the fastest-varying coordinate decides which operand tile they share through L2.
__device__ __forceinline__ Unit decode_unit(const Params& p, int u) {
  // 32-bit unsigned arithmetic throughout (all counts are far below 2^31): this runs on the single TMA and MMA threads
  // at every unit boundary, where a chain of 64-bit divisions would leave the tensor pipe idle.
  Unit t;
  const uint32_t per_k = (uint32_t)(p.num_m_blocks * p.n_splits), uu = (uint32_t)u;
  const uint32_t ks = uu / per_k, r = uu - ks * per_k;
  uint32_t mb, s;
  if (p.n_fastest) { mb = r / (uint32_t)p.n_splits; s = r - mb * (uint32_t)p.n_splits; }
  else { s = r / (uint32_t)p.num_m_blocks; mb = r - s * (uint32_t)p.num_m_blocks; }
  t.ks = (int)ks; t.mb = (int)mb; t.s = (int)s;
  if (p.n_splits == p.num_n_blocks) { t.nb0 = (int)s; t.nb1 = (int)s + 1; }
  else {
    t.nb0 = (int)(s * (uint32_t)p.num_n_blocks / (uint32_t)p.n_splits);
    t.nb1 = (int)((s + 1) * (uint32_t)p.num_n_blocks / (uint32_t)p.n_splits);
  }
  if (p.k_splits == 1) { t.kb0 = 0; t.kb1 = p.num_k_blocks; }
  else {
    t.kb0 = (int)(ks * (uint32_t)p.num_k_blocks / (uint32_t)p.k_splits);
    t.kb1 = (int)((ks + 1) * (uint32_t)p.num_k_blocks / (uint32_t)p.k_splits);
  }
  return t;
}

// PAIR = true: two CTAs of a cluster share every tile (tcgen05 cta_group::2, M = 256): each loads its own 128 rows of A
// and half of B, halving the shared-memory traffic per MMA; p.num_m_blocks then counts 256-row blocks.
template <int MODE, int A_MN, int B_MN, bool PAIR>
__global__ void __launch_bounds__(NUM_THREADS, 1)
umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmD, const Params p) {
  constexpr int STAGES = stages_for(MODE, PAIR);
  constexpr int NC = PAIR ? 2 : 1;
  constexpr int STAGE_BYTES = stage_bytes(PAIR);
  constexpr int BN_CTA = BN / NC;                   // columns of B this CTA keeps in shared memory
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[MAX_STAGES], empty_bar[MAX_STAGES], tfull_bar[2], tempty_bar[2];
  __shared__ uint32_t tmem_holder;

  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); if (mode_stores(MODE)) tma_prefetch_desc(&tmD); }
  const int crank = PAIR ? (int)cluster_ctarank() : 0;      // 0 = leader of the pair (issues the MMAs)
  const int cid = (int)blockIdx.x / NC, ncl = (int)gridDim.x / NC;   // this CTA('s pair) and the number of them
  if (threadIdx.x == 32) {
    // full: the leader's arrive.expect_tx (both CTAs' TMA bytes are credited to the leader's barrier)
    // empty / tfull: one tcgen05.commit each (multicast to both CTAs in a pair)
    // tempty: one arrival per epilogue warp of every CTA sharing the accumulator
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 8 * NC); }
    fence_mbar_init();
  }
  if (warp == 2) { if (PAIR) tmem_alloc_pair<TMEM_COLS>(&tmem_holder); else tmem_alloc<TMEM_COLS>(&tmem_holder); }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_holder;

  const int num_units = p.num_m_blocks * p.n_splits * p.k_splits;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    {
      // the whole warp runs the loops; one elected lane issues the expect_tx and the TMA loads (see umma::elect_one)
      const bool issuer = elect_one();
      const uint32_t full0 = pin_smem(&full_bar[0]), empty0 = pin_smem(&empty_bar[0]), ring0 = pin_smem(smem);
      int stage = 0; uint32_t phase = 0;
      for (int u = cid; u < num_units; u += ncl) {
        const Unit t = decode_unit(p, u);
        const int m0 = (t.mb * NC + crank) * BM;      // first row of A this CTA loads
        for (int nb = t.nb0; nb < t.nb1; ++nb) {
          const int n0 = nb * BN + crank * BN_CTA;    // first column of B this CTA loads
          for (int kb = t.kb0; kb < t.kb1; ++kb) {
            mbar_wait(empty0 + (uint32_t)stage * 8u, phase ^ 1);
            if (!issuer) { if (++stage == STAGES) { stage = 0; phase ^= 1; } continue; }
            const uint32_t fullb = full0 + (uint32_t)stage * 8u;
            if (crank == 0) mbar_arrive_expect_tx(fullb, NC * STAGE_BYTES);
            const uint32_t sa = ring0 + (uint32_t)(stage * STAGE_BYTES);
            const uint32_t sb = sa + A_STAGE_BYTES;
            auto load = [&](uint32_t dst, const CUtensorMap* m, int c0, int c1) {
              if (PAIR) tma_load_2d_pair(dst, m, fullb, c0, c1); else tma_load_2d(dst, m, fullb, c0, c1);
            };
            if (!A_MN) load(sa, &tmA, kb * BK, m0);
            else {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j) load(sa + j * BOX_MN_BYTES, &tmA, m0 + j * 64, kb * BK);
            }
            if (!B_MN) load(sb, &tmB, kb * BK, n0);
            else {
#pragma unroll
              for (int j = 0; j < BN_CTA / 64; ++j) load(sb + j * BOX_MN_BYTES, &tmB, n0 + j * 64, kb * BK);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread)
    if (crank == 0) {
      // the whole warp runs the loops (warp-uniform control flow); one elected lane issues the MMAs and commits
      const bool issuer = elect_one();
      constexpr uint32_t idesc = make_idesc_bf16(BM * NC, BN, A_MN, B_MN);
      int stage = 0; uint32_t phase = 0;
      int as = 0; uint32_t aphase = 0;
      const uint32_t full0 = pin_smem(&full_bar[0]), empty0 = pin_smem(&empty_bar[0]);
      const uint32_t tfull0 = pin_smem(&tfull_bar[0]), tempty0 = pin_smem(&tempty_bar[0]);
      const uint32_t a_addr0 = smem_u32(smem), b_addr0 = a_addr0 + A_STAGE_BYTES;
      const uint64_t adesc0 = A_MN ? make_sdesc_sw128(a_addr0, BOX_MN_BYTES, 1024) : make_sdesc_sw128(a_addr0, 16, 1024);
      const uint64_t bdesc0 = B_MN ? make_sdesc_sw128(b_addr0, BOX_MN_BYTES, 1024) : make_sdesc_sw128(b_addr0, 16, 1024);
      for (int u = cid; u < num_units; u += ncl) {
        const Unit t = decode_unit(p, u);
        for (int nb = t.nb0; nb < t.nb1; ++nb) {
          mbar_wait(tempty0 + (uint32_t)as * 8u, aphase ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + (uint32_t)(as * BN);
          for (int kb = t.kb0; kb < t.kb1; ++kb) {
            mbar_wait(full0 + (uint32_t)stage * 8u, phase);
            tc_fence_after();
            // descriptors differ from the stage-0 / k-0 ones only in the 14-bit start-address field (16-byte units)
            const uint64_t adesc_s = adesc0 + (uint64_t)((uint32_t)(stage * STAGE_BYTES) >> 4);
            const uint64_t bdesc_s = bdesc0 + (uint64_t)((uint32_t)(stage * STAGE_BYTES) >> 4);
            if (issuer) {
#pragma unroll
              for (int k = 0; k < BK / UK; ++k) {
                const uint64_t adesc = adesc_s + (uint64_t)((A_MN ? k * (UK * 128) : k * (UK * 2)) >> 4);
                const uint64_t bdesc = bdesc_s + (uint64_t)((B_MN ? k * (UK * 128) : k * (UK * 2)) >> 4);
                if (PAIR) mma_bf16_ss_pair(tmem_d, adesc, bdesc, idesc, (uint32_t)(kb > t.kb0 || k > 0));
                else mma_bf16_ss(tmem_d, adesc, bdesc, idesc, (uint32_t)(kb > t.kb0 || k > 0));
              }
              // frees the smem slot (in both CTAs of a pair) once these MMAs retire
              if (PAIR) mma_commit_pair(empty0 + (uint32_t)stage * 8u); else mma_commit(empty0 + (uint32_t)stage * 8u);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          // accumulator tile complete (each CTA of a pair reads its own 128 rows from its own TMEM)
          if (issuer) { if (PAIR) mma_commit_pair(tfull0 + (uint32_t)as * 8u); else mma_commit(tfull0 + (uint32_t)as * 8u); }
          as ^= 1; if (as == 0) aphase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue: TMEM -> registers -> result
    // Eight warps: warp w reads TMEM lane quarter q = w % 4 (hardware rule) and the column half (w - 4) / 4 of the
    // 256-column accumulator, so each SM sub-partition always has two epilogue warps to switch between.
    const int q = warp & 3;
    const int half = (warp - 4) >> 2;               // 0: tile columns [0,128)   1: [128,256)
    const int trow = q * 32 + lane;                 // row inside the tile
    int as = 0; uint32_t aphase = 0;
    uint32_t store_seq = 0;                         // storing modes: staging box toggle
    const bool lane_elect = elect_one();            // the lane that issues this warp's TMA stores / barrier arrivals
    // written once, read later by other kernels: by default the stores carry an evict-first hint so they do not displace
    // the W tiles being re-read from L2 (p.store_evict_normal: plain policy, A/B aid B2_STORE_POLICY=normal)
    const uint64_t store_policy = p.store_evict_normal ? l2_policy_evict_normal() : l2_policy_evict_first();
    for (int u = cid; u < num_units; u += ncl) {
      const Unit t = decode_unit(p, u);
      const int s = t.s, nb0 = t.nb0, nb1 = t.nb1;
      const int m0 = (t.mb * NC + crank) * BM;
      const long long row = (long long)m0 + trow;
      const bool row_ok = row < p.M;

      // per-row state for the cross-entropy epilogues
      float run_m = -INFINITY, run_s = 0.f, zval = 0.f;
      bool z_found = false;
      long long y = -1;
      float gs = 0.f, bwd_off = -INFINITY;
      if (MODE != MODE_GEMM && row_ok) {
        y = label_at(p, p.row0 + row);
        if (MODE == MODE_CE_BWD) {
          const bool valid = (y != p.ignore_index) && y >= 0 && y < p.N;
          gs = valid ? __ldg(p.gscale) : 0.f;
          if (gs != 0.f) bwd_off = ce_bwd_exp_offset(gs, p.lse[p.row0 + row]);
          if (!valid) y = -1;
        }
      }

      for (int nb = nb0; nb < nb1; ++nb) {
        mbar_wait(&tfull_bar[as], aphase);
        tc_fence_after();
        // This warp's part of the tile: columns [colbase, colbase + 128) in 32-column chunks, of which the first nch hold
        // real columns. The TMEM reads are software-pipelined: chunk cc+1 is requested before chunk cc is processed, and
        // the accumulator stage goes back to the MMA thread as soon as the LAST chunk sits in registers — not after its
        // arithmetic and stores — so the tensor pipe restarts on this buffer up to one chunk-time earlier.
        const int colbase = nb * BN + half * (BN / 2);
        int nch = (p.N - colbase + 31) / 32;
        nch = nch < 0 ? 0 : (nch > BN / 64 ? BN / 64 : nch);
#ifdef B2_BUILD_PROBES
        if (MODE == MODE_GEMM && p.c_dtype == 98) nch = 0;     // measurement aid (B2_DEBUG_NOEPI=2): not even the TMEM read
#endif
        const uint32_t taddr0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + half * (BN / 2));
        bool released = false;
        auto release = [&]() {                      // one arrival per warp hands the accumulator stage back to the MMA thread
          tc_fence_before();
          __syncwarp();
          if (lane_elect) {
            if (PAIR) mbar_arrive_leader(&tempty_bar[as]); else mbar_arrive(&tempty_bar[as]);
          }
          released = true;
        };
        uint32_t ra[32], rb[32];
        auto process = [&](const int cc, uint32_t (&r)[32], uint32_t (&rnext)[32]) {
          const int col0 = colbase + cc * 32;
          if (mode_stores(MODE)) {
            // the tile leaves through shared memory: each warp stages its own [32 rows x 32 columns] bf16 block (2 KiB,
            // 64B-swizzled) and one lane hands it to the TMA engine (coalesced store, clipped at the matrix edge by the
            // tensor map). Two blocks per warp alternate; nothing here synchronises across warps.
            // (store_wide: one [32 x 64] block, 4 KiB, 128B-swizzled, filled by two consecutive chunks and stored once.)
            if (p.store_direct) {
              // nothing staged: nothing to wait for
            } else if (p.store_wide) {
              if ((cc & 1) == 0) {
                if (lane_elect) bulk_wait_read_all();  // this warp's previous store has drained the box
                __syncwarp();
              }
            } else {
              if (lane_elect) bulk_wait_read_1();    // the store issued two blocks ago has drained its buffer
              __syncwarp();
            }
          }
          tmem_ld_wait_regs(r);
          if (cc + 1 < nch) tmem_ld_32x32(taddr0 + (uint32_t)((cc + 1) * 32), rnext);
          else release();

          // storing modes: this thread's 32 values (one row of the warp's [32 x 32] block) -> 64B-swizzled staging box ->
          // one TMA store per warp
          auto store_block = [&](const float (&v)[32]) {
#ifdef B2_BUILD_PROBES
            if (p.debug_store == 2) return;          // measurement aid: results are wrong, only the timing means something
#endif
            uint32_t w[16];
#pragma unroll
            for (int j = 0; j < 32; j += 2) w[j >> 1] = pack_bf16(v[j], v[j + 1]);
            if (p.store_direct) {
              // Straight from registers: this thread owns 32 consecutive columns of one row = 64 contiguous bytes = two
              // whole 32-byte sectors (the row pitch is a multiple of 16 elements, so every segment is sector-aligned).
              // Vectors that start beyond the pitch are dropped; columns in [N, ldd) are padding nobody reads.
              if (row_ok) {
                uint8_t* dst = reinterpret_cast<uint8_t*>(p.dlogits) + ((long long)row * p.ldd + col0) * 2;
                if (col0 + 32 <= p.ldd) {
                  st_global_v8_hint(dst, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7], store_policy);
                  st_global_v8_hint(dst + 32, w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15], store_policy);
                } else {
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    if (col0 + 8 * k + 8 <= p.ldd)
                      *reinterpret_cast<uint4*>(dst + 16 * k) = make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
                }
              }
              return;
            }
            if (p.store_wide) {
              uint8_t* box = smem + STAGES * STAGE_BYTES + (warp - 4) * STAGING_BOX_BYTES;
              uint8_t* rowp = box + lane * 128;
              const int h = cc & 1;                  // which 64-byte half of the 128-byte row this chunk fills
#pragma unroll
              for (int k = 0; k < 4; ++k)             // 16-byte chunk 4h+k of the row, SWIZZLE_128B: index ^ (row & 7)
                st_shared_v4(rowp + (((4 * h + k) ^ (lane & 7)) << 4), w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
              if (h == 1 || cc + 1 == nch) {         // block complete (a lone first half: the rest lies beyond the matrix edge)
                fence_proxy_async_smem();
                __syncwarp();
                if (lane_elect && !probe_skips_store(p)) {
                  tma_store_2d_hint(&tmD, box, colbase + (cc >> 1) * 64, m0 + q * 32, store_policy);
                  bulk_commit();
                }
              }
              return;
            }
            uint8_t* box = smem + STAGES * STAGE_BYTES + (warp - 4) * STAGING_BOX_BYTES + (int)(store_seq & 1) * (32 * 64);
            uint8_t* rowp = box + lane * 64;
#pragma unroll
            for (int k = 0; k < 4; ++k)               // 16-byte chunk k of the 64-byte row, SWIZZLE_64B: k ^ ((row >> 1) & 3)
              st_shared_v4(rowp + ((k ^ ((lane >> 1) & 3)) << 4), w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
            fence_proxy_async_smem();                // generic-proxy writes -> visible to the TMA engine
            __syncwarp();
            if (lane_elect) {
              // written once, read later by other kernels: do not let it displace the W tiles being re-read from L2
              tma_store_2d_hint(&tmD, box, col0, m0 + q * 32, store_policy);
              bulk_commit();
            }
            ++store_seq;
          };

          if (MODE == MODE_GEMM) {
#ifdef B2_BUILD_PROBES
            if (p.c_dtype == 99) {
              // measurement aid (B2_DEBUG_NOEPI=1): accumulator read and discarded, nothing stored
            } else
#endif
            if (row_ok && p.k_splits > 1) {
              // split-K: plain fp32 partial tile; splitk_reduce_kernel adds the splits in a fixed order
              float* dst = p.ws + ((long long)t.ks * p.M + row) * p.ldws + col0;     // ldws % 4 == 0, col0 % 32 == 0
              const int ncol = min(32, p.N - col0);
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                if (j < ncol)
                  __stcs(reinterpret_cast<float4*>(dst + j), make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                                                                         __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])));
            } else if (row_ok) {
              const int ncol = min(32, p.N - col0);
              if (p.c_dtype == B2_F32) {
                float* dst = static_cast<float*>(p.C) + row * p.ldc + col0;
                if (ncol == 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                  for (int j = 0; j < 32; j += 4) {
                    float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                    if (p.accumulate) { const float4 o = *reinterpret_cast<const float4*>(dst + j); v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
                    // streaming store: the output is written once and must not push the operand tiles (re-read by every
                    // unit of the wave) out of L2
                    __stcs(reinterpret_cast<float4*>(dst + j), v);
                  }
                } else {
#pragma unroll
                  for (int j = 0; j < 32; ++j)
                    if (j < ncol) dst[j] = __uint_as_float(r[j]) + (p.accumulate ? dst[j] : 0.f);
                }
              } else {
                __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(p.C) + row * p.ldc + col0;
                if (ncol == 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
                  for (int j = 0; j < 32; j += 8) {
                    float v[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[j + e]);
                    if (p.accumulate) {
                      const uint4 o = *reinterpret_cast<const uint4*>(dst + j);
                      v[0] += bf16_lo(o.x); v[1] += bf16_hi(o.x); v[2] += bf16_lo(o.y); v[3] += bf16_hi(o.y);
                      v[4] += bf16_lo(o.z); v[5] += bf16_hi(o.z); v[6] += bf16_lo(o.w); v[7] += bf16_hi(o.w);
                    }
                    __stcs(reinterpret_cast<uint4*>(dst + j), make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7])));
                  }
                } else {
#pragma unroll
                  for (int j = 0; j < 32; ++j) {
                    if (j < ncol) {
                      const float o = p.accumulate ? __bfloat162float(dst[j]) : 0.f;
                      dst[j] = __float2bfloat16_rn(__uint_as_float(r[j]) + o);
                    }
                  }
                }
              }
            }
          } else {
            // logits for this row, columns [col0, col0+32)
            float x[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(r[j]);
            if (p.bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 32; ++j) x[j] += (col0 + j < p.N) ? __ldg(p.bias + col0 + j) : 0.f;
            }
            if (p.round_bf16) {
#pragma unroll
              for (int j = 0; j < 32; j += 2) round_bf16_pair(x[j], x[j + 1]);
            }

            if (mode_is_fwd(MODE)) {
              if (MODE == MODE_CE_FWD_KEEP) {
                // the rounded logits themselves, before the padding columns are masked (the store clips them anyway)
                store_block(x);
              }
              if (col0 + 32 > p.N) {                 // last chunk of the vocabulary: mask the padding columns
#pragma unroll
                for (int j = 0; j < 32; ++j) if (col0 + j >= p.N) x[j] = -INFINITY;
              }
              float cmax = x[0];
#pragma unroll
              for (int j = 1; j < 32; ++j) cmax = fmaxf(cmax, x[j]);
              const float nm = fmaxf(run_m, cmax);
              const float nm_l2 = nm * LOG2E;
              float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
              for (int j = 0; j < 32; j += 2) {
                acc0 += fast_exp2(fmaf(x[j], LOG2E, -nm_l2));
                acc1 += fast_exp2(fmaf(x[j + 1], LOG2E, -nm_l2));
              }
              run_s = run_s * fast_exp2((run_m - nm) * LOG2E) + (acc0 + acc1);
              run_m = nm;
              if (y >= col0 && y < col0 + 32) {
                const int jj = (int)(y - col0);
#pragma unroll
                for (int j = 0; j < 32; ++j) if (j == jj) zval = x[j];
                z_found = true;
              }
            } else {  // MODE_CE_BWD
              // softmax * gscale in one FFMA + EX2 per element: the scale rides in the exponent offset
              // (bwd_off = log2|gscale| - lse*log2e; -inf for ignored rows, which therefore come out as exact zeros)
              float d[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) d[j] = fast_exp2(fmaf(x[j], LOG2E, bwd_off));
              if (y >= col0 && y < col0 + 32) {      // the label's column: subtract the one-hot term (once per row)
                const int jj = (int)(y - col0);
#pragma unroll
                for (int j = 0; j < 32; ++j) if (j == jj) d[j] -= fabsf(gs);
              }
              if (gs < 0.f) {                         // uniform: a negative upstream gradient flips every sign
#pragma unroll
                for (int j = 0; j < 32; ++j) d[j] = -d[j];
              }
              store_block(d);
            }
          }
        };
        if (nch > 0) tmem_ld_32x32(taddr0, ra);
#pragma unroll 1
        for (int cc = 0; cc < nch; cc += 2) {
          process(cc, ra, rb);
          if (cc + 1 < nch) process(cc + 1, rb, ra);
        }
        if (!released) release();
        as ^= 1; if (as == 0) aphase ^= 1;
      }

      if (mode_is_fwd(MODE) && row_ok) {
        // the two column halves of a unit are merged later like any two vocabulary splits
        p.part[(long long)(2 * s + half) * p.part_stride + row] = make_float2(run_m, run_s);
        if (z_found) p.zt[row] = zval;
      }
    }
    if (mode_stores(MODE) && !p.store_direct && lane_elect) bulk_wait_all();   // this warp's last stores have landed (same lane that issued them)
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();    // neither CTA of a pair may leave while the other still needs it
  if (warp == 2) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair<TMEM_COLS>(tmem_base); else tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

// merge the per-split (max, sumexp) partials of each row into lse[row]
__global__ void __launch_bounds__(256) ce_combine_kernel(const float2* __restrict__ part, long long part_stride, int n_splits,
                                                         long long N, float* __restrict__ lse) {
  const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= N) return;
  float m = -INFINITY, s = 0.f;
  for (int k = 0; k < n_splits; ++k) {
    const float2 v = part[(long long)k * part_stride + row];
    const float nm = fmaxf(m, v.x);
    s = s * exp2f((m - nm) * LOG2E) + v.y * exp2f((v.x - nm) * LOG2E);
    m = nm;
  }
  lse[row] = m + logf(s);
}

// Backward from KEPT logits: x[row, col] (bf16, rounded logits left by MODE_CE_FWD_KEEP) becomes, in place,
//   (exp(x - lse[row]) - [col == label]) * gscale          — the same arithmetic, in the same order, as the MODE_CE_BWD
// epilogue, so the two routes give the same bits. HBM-bound: one read and one write of the [rows, V] bf16 matrix.
// grid.y strides over rows, grid.x * block over the 16-byte vectors of a row.
__global__ void __launch_bounds__(256) ce_dlogits_inplace_kernel(__nv_bfloat16* __restrict__ x, long long ld,
                                                                 const long long* __restrict__ labels, long long row0,
                                                                 long long rows, int V, int shift_T, long long ignore_index,
                                                                 const float* __restrict__ lse,
                                                                 const float* __restrict__ gscale, int reverse) {
  const int nvec = (V + 7) >> 3;
  const float g = __ldg(gscale);
  for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
    // LAST rows first (blocks are scheduled in increasing blockIdx.y): the forward that left these logits wrote its last
    // rows last, so they are the ones still in L2; and the rows written last HERE are the first ones the dH GEMM reads
    const long long row = reverse ? rows - 1 - r : r;
    const long long i = row0 + row;
    long long y = (shift_T <= 0) ? labels[i] : (((i % shift_T) + 1 < shift_T) ? labels[i + 1] : ignore_index);
    const bool valid = (y != ignore_index) && y >= 0 && y < V;
    const float gs = valid ? g : 0.f;
    const float off = (gs != 0.f) ? ce_bwd_exp_offset(gs, lse[i]) : -INFINITY;
    const int yv = valid ? (int)(y >> 3) : -1, ye = (int)(y & 7);
    uint4* rp = reinterpret_cast<uint4*>(x + row * ld);
    // four independent 16-byte loads per thread before any arithmetic: enough bytes in flight to cover HBM latency
    for (int v0 = blockIdx.x * (blockDim.x * 4) + threadIdx.x; v0 < nvec; v0 += gridDim.x * (blockDim.x * 4)) {
      uint4 uu[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int v = v0 + k * (int)blockDim.x;
        if (v < nvec) uu[k] = rp[v];
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int v = v0 + k * (int)blockDim.x;
        if (v >= nvec) break;
        const uint4 u = uu[k];
        float f[8] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y), bf16_lo(u.z), bf16_hi(u.z), bf16_lo(u.w), bf16_hi(u.w)};
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = fast_exp2(fmaf(f[e], LOG2E, off));
        if (v == yv) {
#pragma unroll
          for (int e = 0; e < 8; ++e) if (e == ye) f[e] -= fabsf(gs);
        }
        if (gs < 0.f) {
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = -f[e];
        }
        if (v * 8 + 8 > V) {                         // the row's last vector: columns >= V are padding, left as zeros
#pragma unroll
          for (int e = 0; e < 8; ++e) if (v * 8 + e >= V) f[e] = 0.f;
        }
        rp[v] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      }
    }
  }
}

// C[M,N] (+)= sum over splits of ws[k][M][ldws]; 4 columns per thread
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ ws, int k_splits, int M, int N,
                                                            long long ldws, void* __restrict__ C, int c_dtype,
                                                            long long ldc, int accumulate) {
  const long long nq = ldws >> 2;
  const long long total = (long long)M * nq;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / nq;
    const int col = (int)(i - row * nq) * 4;
    if (col >= N) continue;
    float4 acc = *reinterpret_cast<const float4*>(ws + row * ldws + col);
    for (int k = 1; k < k_splits; ++k) {
      const float4 v = *reinterpret_cast<const float4*>(ws + ((long long)k * M + row) * ldws + col);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    const float a[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (col + j >= N) break;
      if (c_dtype == B2_F32) {
        float* d = static_cast<float*>(C) + row * ldc + col + j;
        *d = a[j] + (accumulate ? *d : 0.f);
      } else {
        __nv_bfloat16* d = static_cast<__nv_bfloat16*>(C) + row * ldc + col + j;
        *d = __float2bfloat16_rn(a[j] + (accumulate ? __bfloat162float(*d) : 0.f));
      }
    }
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
    else
      cudaGetLastError();
  });
  return fn;
}

// bf16 2-D row-major tensor [outer, inner] with row pitch `pitch_elems`; box = box_inner x box_outer, 128B swizzle,
// out-of-bounds elements read as zero.
int make_tmap(CUtensorMap* m, const void* ptr, long long inner, long long outer, long long pitch_elems, int box_inner,
              int box_outer, const char* what, CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { b2_set_error("cuTensorMapEncodeTiled is not available from this driver"); return B2_ERR_CUDA; }
  B2_REQUIRE(b2_aligned(ptr, 16), "%s: base pointer must be 16-byte aligned", what);
  B2_REQUIRE((pitch_elems * 2) % 16 == 0, "%s: row pitch (%lld elements) must be a multiple of 8 bf16 elements", what, pitch_elems);
  B2_REQUIRE(inner >= 1 && outer >= 1 && pitch_elems >= inner, "%s: bad extents inner=%lld outer=%lld pitch=%lld", what, inner, outer, pitch_elems);
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)pitch_elems * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { b2_set_error("%s: cuTensorMapEncodeTiled failed with CUresult %d", what, (int)r); return B2_ERR_CUDA; }
  return B2_OK;
}

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
      cudaGetLastError();
      n = B2_NUM_SMS;
    }
  }
  return n;
}

// Shape of the TMA stores of the storing epilogues: [32 x 64] (one per two chunks; default) or [32 x 32] (B2_STORE_BOX=32)
int debug_store() {
#ifdef B2_BUILD_PROBES
  static const int v = [] { const char* e = getenv("B2_DEBUG_STORE"); return e == nullptr ? 0 : atoi(e); }();
  return v;
#else
  return 0;
#endif
}
// B2_STORE_DIRECT=1: the storing epilogues write global memory straight from registers (A/B against the staged TMA stores)
int store_direct_env() {
  static const int v = [] { const char* e = getenv("B2_STORE_DIRECT"); return e == nullptr ? -1 : atoi(e); }();
  return v;
}
// direct stores need a 32-byte aligned buffer whose row pitch is a multiple of 16 elements
bool store_direct_ok(const void* ptr, long long pitch) {
  return store_direct_env() == 1 && pitch % 16 == 0 && b2_aligned(ptr, 32);
}
int store_evict_normal() {
  static const int v = [] { const char* e = getenv("B2_STORE_POLICY"); return (e != nullptr && e[0] == 'n') ? 1 : 0; }();
  return v;
}
// B2_TRANSFORM_REVERSE=0: the in-place soft-max gradient walks the rows front to back (A/B aid; default: back to front)
int transform_reverse() {
  static const int v = [] { const char* e = getenv("B2_TRANSFORM_REVERSE"); return e == nullptr ? 1 : (e[0] != '0'); }();
  return v;
}
bool store_wide() {
  static const bool wide = [] { const char* e = getenv("B2_STORE_BOX"); return e == nullptr ? true : atoi(e) == 64; }();
  return wide;
}
int make_store_tmap(CUtensorMap* m, const void* ptr, long long cols, long long rows, long long pitch, const char* what) {
  return store_wide() ? make_tmap(m, ptr, cols, rows, pitch, 64, 32, what, CU_TENSOR_MAP_SWIZZLE_128B)
                      : make_tmap(m, ptr, cols, rows, pitch, 32, 32, what, CU_TENSOR_MAP_SWIZZLE_64B);
}

// CTA pairs (tcgen05 cta_group::2, M = 256 per MMA: each CTA loads half of the B tile, a third less L2 -> SM operand traffic)
// or independent CTAs. Measured with the warp-uniform issue loops (profiles/README.md, round 2): pairs win for the plain
// GEMMs, for the forward that stores its logits at every shape, and for both cross-entropy kernels once K >= 1024 (Llama
// heads: forward 1.38 -> 1.31 ms, d(logits) 1.51 -> 1.33 ms at D = 2048); at K = 768 (GPT-2) independent CTAs are a few
// percent ahead for the two kernels that do not store. B2_PAIR=0 / B2_PAIR=1 forces one shape everywhere (A/B runs).
bool use_pair(int mode, int K) {
  static const int forced = [] { const char* e = getenv("B2_PAIR"); return e == nullptr ? -1 : (e[0] != '0'); }();
  if (forced >= 0) return forced != 0;
  if (mode == MODE_GEMM || mode == MODE_CE_FWD_KEEP) return true;
  return K >= 1024;
}
// rows of A covered by one work unit
int rows_per_unit(bool pair) { return pair ? 2 * BM : BM; }
// CTAs (single mode) or CTA pairs (pair mode) that run concurrently
int num_workers(bool pair) { return pair ? num_sms() / 2 : num_sms(); }

template <int MODE, int A_MN, int B_MN, bool PAIR>
int launch_impl(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const Params& p, cudaStream_t stream) {
  auto kern = umma_kernel<MODE, A_MN, B_MN, PAIR>;
  constexpr int SMEM_BYTES = smem_bytes_for(MODE, PAIR);
  B2_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  const int units = p.num_m_blocks * p.n_splits * p.k_splits;
  const int workers = units < num_workers(PAIR) ? units : num_workers(PAIR);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(workers * (PAIR ? 2 : 1)));
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = PAIR ? 2 : 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  B2_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, td, p));
  return B2_OK;
}

template <int MODE, int A_MN, int B_MN>
int launch(bool pair, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, const Params& p, cudaStream_t stream) {
  return pair ? launch_impl<MODE, A_MN, B_MN, true>(ta, tb, td, p, stream)
              : launch_impl<MODE, A_MN, B_MN, false>(ta, tb, td, p, stream);
}

// operand [rows = M or N extent, K] -> tensor map; K-major: memory is [extent, ld] with K contiguous;
// MN-major: memory is [K, ld] with the M/N extent contiguous.
int make_operand_tmap(CUtensorMap* m, const void* ptr, int mn_major, long long extent, long long K, long long ld,
                      int tile_extent, const char* what) {
  if (!mn_major) return make_tmap(m, ptr, K, extent, ld, BK, tile_extent, what);
  return make_tmap(m, ptr, extent, K, ld, 64, BK, what);
}

// Pick how many contiguous vocabulary ranges each 128-row block is split into so that the persistent
// grid finishes with the least idle time (units are dealt round-robin to CTAs).
int choose_splits(int num_m_blocks, int num_n_blocks, int grid) {
  // the search below costs ~1 ms of host time at GPT-2 sizes: remember the answer per shape
  static std::mutex mu;
  static std::map<std::tuple<int, int, int>, int> cache;
  const auto key = std::make_tuple(num_m_blocks, num_n_blocks, grid);
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
  }
  int best_s = 1; double best = 1e30;
  const int smax = num_n_blocks < MAX_SPLITS ? num_n_blocks : MAX_SPLITS;
  for (int S = 1; S <= smax; ++S) {
    const long long units = (long long)num_m_blocks * S;
    // CTA b handles units b, b+grid, ...; unit u has split s = u / num_m_blocks
    double worst = 0;
    const int ctas = units < grid ? (int)units : grid;
    for (int b = 0; b < ctas; ++b) {
      double t = 0;
      for (long long u = b; u < units; u += grid) {
        const int s = (int)(u / num_m_blocks);
        const int tiles = (int)((long long)(s + 1) * num_n_blocks / S - (long long)s * num_n_blocks / S);
        t += tiles + 0.15;   // per-unit start-up cost, in tile units
      }
      if (t > worst) worst = t;
    }
    if (worst < best - 1e-9) { best = worst; best_s = S; }
  }
  std::lock_guard<std::mutex> lk(mu);
  cache[key] = best_s;
  return best_s;
}

}  // namespace

// split-K factor for a plain GEMM: used when the output has too few 128x256 tiles to fill the 148 SMs
static int auto_k_splits(int tiles, int num_k_blocks, int sms) {
  if (tiles >= 6 * sms || num_k_blocks < 32) return 1;
  // units are dealt round-robin to `sms` persistent CTAs: pick the smallest split whose last wave is nearly full
  int best = 1; double best_eff = 0.0;
  for (int ks = 1; ks <= B2_GEMM_MAX_KSPLIT; ++ks) {
    if (ks > 1 && num_k_blocks / ks < 16) break;
    const long long units = (long long)tiles * ks;
    const long long waves = (units + sms - 1) / sms;
    const double eff = (double)units / (double)(waves * sms) - 0.01 * (ks - 1);     // small charge for the extra partials
    if (eff > best_eff + 1e-9) { best_eff = eff; best = ks; }
  }
  return best;
}

extern "C" size_t b2_gemm_workspace_bytes(int M, int N) {
  const size_t ldws = ((size_t)N + 3) / 4 * 4;
  return (size_t)B2_GEMM_MAX_KSPLIT * (size_t)M * ldws * sizeof(float);
}

extern "C" int b2_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major, long long ldb,
                            void* C, int c_dtype, long long ldc, int M, int N, int K, int accumulate, void* workspace,
                            size_t workspace_bytes, cudaStream_t stream) {
  B2_REQUIRE(M >= 1 && N >= 1 && K >= 1, "b2_gemm_bf16: bad shape M=%d N=%d K=%d", M, N, K);
  B2_REQUIRE(A && B && C, "b2_gemm_bf16: null pointer");
  B2_REQUIRE(c_dtype == B2_F32 || c_dtype == B2_BF16, "b2_gemm_bf16: c_dtype must be 0 (f32) or 1 (bf16)");
  B2_REQUIRE(ldc >= N, "b2_gemm_bf16: ldc < N");
  CUtensorMap ta, tb;
  int rc;
  if ((rc = make_operand_tmap(&ta, A, a_mn_major, M, K, lda, BM, "b2_gemm_bf16 A")) != B2_OK) return rc;
  const bool pair = use_pair(MODE_GEMM, K);
  if ((rc = make_operand_tmap(&tb, B, b_mn_major, N, K, ldb, pair ? BN / 2 : BN, "b2_gemm_bf16 B")) != B2_OK) return rc;
  Params p{};
  p.M = M; p.N = N; p.K = K;
  p.num_m_blocks = (M + rows_per_unit(pair) - 1) / rows_per_unit(pair); p.num_n_blocks = (N + BN - 1) / BN; p.num_k_blocks = (K + BK - 1) / BK;
  p.n_splits = p.num_n_blocks;
  // Unit order: the CTAs that run at the same time should share operand tiles through L2. Few column blocks, or a B
  // operand that fits L2 as a whole (dW at D = 4096: H^T is 33 MB, while one wave of distinct A strips is 150 MB and
  // was being re-read from HBM once per column block): let the units that share an A row-block run side by side.
  p.n_fastest = p.num_n_blocks <= 8 || (long long)N * K * 2 <= (48ll << 20);
  p.k_splits = 1;
  if (workspace != nullptr && workspace_bytes >= b2_gemm_workspace_bytes(M, N) && b2_aligned(workspace, 16))
    p.k_splits = auto_k_splits(p.num_m_blocks * p.num_n_blocks, p.num_k_blocks, num_workers(pair));
  p.ws = static_cast<float*>(workspace); p.ldws = ((long long)N + 3) / 4 * 4;
  p.C = C; p.ldc = ldc; p.c_dtype = c_dtype; p.accumulate = accumulate;
  // profiling only: mainloop ceiling without the epilogue stores (=1) or without even reading the accumulator (=2)
#ifdef B2_BUILD_PROBES
  static const int noepi = [] { const char* e = getenv("B2_DEBUG_NOEPI"); return e == nullptr ? 0 : (e[0] == '2' ? 2 : 1); }();
  if (noepi) { p.c_dtype = noepi == 2 ? 98 : 99; p.k_splits = 1; }
#endif
  if (!a_mn_major && !b_mn_major) rc = launch<MODE_GEMM, 0, 0>(pair, ta, tb, ta, p, stream);
  else if (!a_mn_major && b_mn_major) rc = launch<MODE_GEMM, 0, 1>(pair, ta, tb, ta, p, stream);
  else if (a_mn_major && !b_mn_major) rc = launch<MODE_GEMM, 1, 0>(pair, ta, tb, ta, p, stream);
  else rc = launch<MODE_GEMM, 1, 1>(pair, ta, tb, ta, p, stream);
  if (rc != B2_OK || p.k_splits == 1) return rc;
  const long long total = (long long)M * (p.ldws / 4);
  long long blocks = (total + 255) / 256;
  if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
  splitk_reduce_kernel<<<(unsigned)blocks, 256, 0, stream>>>(p.ws, p.k_splits, M, N, p.ldws, C, c_dtype, ldc, accumulate);
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" size_t b2_lmhead_ce_workspace_bytes(long long N) {
  const long long npad = (N + BM - 1) / BM * BM;
  return (size_t)2 * MAX_SPLITS * (size_t)npad * sizeof(float2);   // (max, sumexp) per row, split and column half
}

// logits_out == nullptr: the plain forward (MODE_CE_FWD); otherwise the forward that also stores the rounded logits
static int lmhead_ce_fwd_impl(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                              const long long* labels, long long N, int D, int V, int shift_T, long long ignore_index,
                              int round_logits_bf16, void* workspace, size_t workspace_bytes, float* lse, float* zt,
                              float* loss_out, long long* n_valid_out, int* err_out, void* logits_out, long long ldl,
                              cudaStream_t stream) {
  B2_REQUIRE(N >= 0 && D >= 1 && V >= 1, "b2_lmhead_ce_fwd: bad shape N=%lld D=%d V=%d", N, D, V);
  B2_REQUIRE(N < (1ll << 31) - BM, "b2_lmhead_ce_fwd: N too large");
  B2_REQUIRE(loss_out && n_valid_out, "b2_lmhead_ce_fwd: null output");
  B2_REQUIRE(shift_T == 0 || (shift_T > 0 && N % shift_T == 0), "b2_lmhead_ce_fwd: N must be a multiple of shift_T");
  if (N > 0) {
    B2_REQUIRE(H && W && labels && lse && zt && workspace, "b2_lmhead_ce_fwd: null buffer");
    B2_REQUIRE(workspace_bytes >= b2_lmhead_ce_workspace_bytes(N), "b2_lmhead_ce_fwd: workspace too small (%zu < %zu)",
               workspace_bytes, b2_lmhead_ce_workspace_bytes(N));
    B2_REQUIRE(b2_aligned(workspace, 16), "b2_lmhead_ce_fwd: workspace must be 16-byte aligned");
    CUtensorMap ta, tb;
    int rc;
    if ((rc = make_operand_tmap(&ta, H, 0, N, D, ldh, BM, "b2_lmhead_ce_fwd H")) != B2_OK) return rc;
    const bool pair = use_pair(logits_out == nullptr ? MODE_CE_FWD : MODE_CE_FWD_KEEP, D);
    if ((rc = make_operand_tmap(&tb, W, 0, V, D, ldw, pair ? BN / 2 : BN, "b2_lmhead_ce_fwd W")) != B2_OK) return rc;
    Params p{};
    p.M = (int)N; p.N = V; p.K = D;
    p.num_m_blocks = (int)((N + rows_per_unit(pair) - 1) / rows_per_unit(pair)); p.num_n_blocks = (V + BN - 1) / BN; p.num_k_blocks = (D + BK - 1) / BK;
    p.n_splits = choose_splits(p.num_m_blocks, p.num_n_blocks, num_workers(pair));
    p.k_splits = 1; p.n_fastest = 0;
    p.labels = labels; p.shift_T = shift_T; p.ignore_index = ignore_index; p.row0 = 0;
    p.bias = bias; p.round_bf16 = round_logits_bf16;
    p.part = static_cast<float2*>(workspace); p.part_stride = (N + BM - 1) / BM * BM; p.zt = zt;
    if (logits_out == nullptr) {
      if ((rc = launch<MODE_CE_FWD, 0, 0>(pair, ta, tb, ta, p, stream)) != B2_OK) return rc;
    } else {
      CUtensorMap td;
      if ((rc = make_store_tmap(&td, logits_out, V, N, ldl, "b2_lmhead_ce_fwd_keep logits")) != B2_OK) return rc;
      p.store_wide = store_wide(); p.debug_store = debug_store(); p.store_evict_normal = store_evict_normal();
      p.dlogits = static_cast<__nv_bfloat16*>(logits_out); p.ldd = ldl; p.store_direct = store_direct_ok(logits_out, ldl);
      if ((rc = launch<MODE_CE_FWD_KEEP, 0, 0>(pair, ta, tb, td, p, stream)) != B2_OK) return rc;
    }
    ce_combine_kernel<<<(unsigned)((N + 255) / 256), 256, 0, stream>>>(p.part, p.part_stride, 2 * p.n_splits, N, lse);
    B2_CHECK_CUDA(cudaGetLastError());
  }
  return b2_ce_finalize_launch(lse, zt, labels, N, V, shift_T, ignore_index, loss_out, n_valid_out, err_out, stream);
}

extern "C" int b2_lmhead_ce_fwd(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                                const long long* labels, long long N, int D, int V, int shift_T, long long ignore_index,
                                int round_logits_bf16, void* workspace, size_t workspace_bytes, float* lse, float* zt,
                                float* loss_out, long long* n_valid_out, int* err_out, cudaStream_t stream) {
  return lmhead_ce_fwd_impl(H, ldh, W, ldw, bias, labels, N, D, V, shift_T, ignore_index, round_logits_bf16, workspace,
                            workspace_bytes, lse, zt, loss_out, n_valid_out, err_out, nullptr, 0, stream);
}

extern "C" int b2_lmhead_ce_fwd_keep(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                                     const long long* labels, long long N, int D, int V, int shift_T,
                                     long long ignore_index, void* workspace, size_t workspace_bytes, float* lse,
                                     float* zt, float* loss_out, long long* n_valid_out, int* err_out, void* logits_out,
                                     long long ldl, cudaStream_t stream) {
  B2_REQUIRE(logits_out != nullptr || N == 0, "b2_lmhead_ce_fwd_keep: null logits buffer");
  B2_REQUIRE(ldl >= V && ldl % 8 == 0, "b2_lmhead_ce_fwd_keep: ldl must be >= V and a multiple of 8");
  B2_REQUIRE(b2_aligned(logits_out, 16), "b2_lmhead_ce_fwd_keep: logits buffer must be 16-byte aligned");
  // the kept values are the bf16-rounded logits, so the loss must be computed from exactly those: rounding is implied
  return lmhead_ce_fwd_impl(H, ldh, W, ldw, bias, labels, N, D, V, shift_T, ignore_index, 1, workspace, workspace_bytes,
                            lse, zt, loss_out, n_valid_out, err_out, N == 0 ? nullptr : logits_out, ldl, stream);
}

extern "C" int b2_ce_dlogits_inplace(void* logits, long long ldl, const long long* labels, long long row0, long long rows,
                                     int V, int shift_T, long long ignore_index, const float* lse,
                                     const float* gscale_dev, cudaStream_t stream) {
  if (rows == 0) return B2_OK;
  B2_REQUIRE(rows > 0 && row0 >= 0 && V >= 1, "b2_ce_dlogits_inplace: bad shape rows=%lld V=%d", rows, V);
  B2_REQUIRE(logits && labels && lse && gscale_dev, "b2_ce_dlogits_inplace: null buffer");
  B2_REQUIRE(ldl >= V && ldl % 8 == 0, "b2_ce_dlogits_inplace: ldl must be >= V and a multiple of 8");
  B2_REQUIRE(b2_aligned(logits, 16), "b2_ce_dlogits_inplace: logits buffer must be 16-byte aligned");
  const int nvec = (V + 7) / 8;
  // 4 vectors (64 B) per thread and row; rows are dealt to grid.y (at most 65535 of them, strided beyond that)
  const unsigned gx = (unsigned)((nvec + 256 * 4 - 1) / (256 * 4));
  const unsigned gy = (unsigned)(rows < 65535 ? rows : 65535);
  ce_dlogits_inplace_kernel<<<dim3(gx, gy), 256, 0, stream>>>(static_cast<__nv_bfloat16*>(logits), ldl, labels, row0, rows,
                                                             V, shift_T, ignore_index, lse, gscale_dev, transform_reverse());
  B2_CHECK_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_lmhead_ce_bwd_dlogits(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                                        const long long* labels, long long row0, long long rows, int D, int V, int shift_T,
                                        long long ignore_index, int round_logits_bf16, const float* lse,
                                        const float* gscale_dev, void* dlogits, long long ldd, cudaStream_t stream) {
  if (rows == 0) return B2_OK;
  B2_REQUIRE(rows > 0 && row0 >= 0 && D >= 1 && V >= 1, "b2_lmhead_ce_bwd_dlogits: bad shape rows=%lld D=%d V=%d", rows, D, V);
  B2_REQUIRE(rows < (1ll << 31) - BM, "b2_lmhead_ce_bwd_dlogits: too many rows");
  B2_REQUIRE(H && W && labels && lse && gscale_dev && dlogits, "b2_lmhead_ce_bwd_dlogits: null buffer");
  B2_REQUIRE(ldd >= V && ldd % 8 == 0, "b2_lmhead_ce_bwd_dlogits: ldd must be >= V and a multiple of 8");
  B2_REQUIRE(b2_aligned(dlogits, 16), "b2_lmhead_ce_bwd_dlogits: dlogits must be 16-byte aligned");
  CUtensorMap ta, tb;
  int rc;
  if ((rc = make_operand_tmap(&ta, H, 0, rows, D, ldh, BM, "b2_lmhead_ce_bwd_dlogits H")) != B2_OK) return rc;
  const bool pair = use_pair(MODE_CE_BWD, D);
  if ((rc = make_operand_tmap(&tb, W, 0, V, D, ldw, pair ? BN / 2 : BN, "b2_lmhead_ce_bwd_dlogits W")) != B2_OK) return rc;
  Params p{};
  p.M = (int)rows; p.N = V; p.K = D;
  p.num_m_blocks = (int)((rows + rows_per_unit(pair) - 1) / rows_per_unit(pair)); p.num_n_blocks = (V + BN - 1) / BN; p.num_k_blocks = (D + BK - 1) / BK;
  p.n_splits = p.num_n_blocks; p.k_splits = 1; p.n_fastest = 0;
  p.labels = labels; p.shift_T = shift_T; p.ignore_index = ignore_index; p.row0 = row0;
  p.bias = bias; p.round_bf16 = round_logits_bf16;
  p.lse = lse; p.gscale = gscale_dev; p.dlogits = static_cast<__nv_bfloat16*>(dlogits); p.ldd = ldd;
  CUtensorMap td;
  if ((rc = make_store_tmap(&td, dlogits, V, rows, ldd, "b2_lmhead_ce_bwd_dlogits dlogits")) != B2_OK) return rc;
  p.store_wide = store_wide(); p.debug_store = debug_store(); p.store_direct = store_direct_ok(dlogits, ldd);
  p.store_evict_normal = store_evict_normal();
  return launch<MODE_CE_BWD, 0, 0>(pair, ta, tb, td, p, stream);
}

// ---------------------------------------------------------------------------------------------------------------------
// db[j] (+)= sum over rows of dlogits[r, j]: the bias gradient of an LM head that has one (BasicLLM.fc_out,
// ref: ray-jobs/pytorch_llm_ray.py:87). One block per 64 columns: 8 threads cover one row's 128 bytes (whole lines), 32
// rows per pass; the 32 row-groups are then added in a fixed order — no atomics, reproducible.
namespace {
__global__ void __launch_bounds__(256) colsum_kernel(const __nv_bfloat16* __restrict__ x, long long ld, long long rows, int V,
                                                     float* __restrict__ db, int accumulate) {
  __shared__ float part[32][65];
  const int cg = threadIdx.x & 7, rg = threadIdx.x >> 3;       // 16-byte column group within the 64 columns; row group
  const int col0 = blockIdx.x * 64 + cg * 8;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col0 < V) {                                               // (ld is a multiple of 8: the vector is inside the row)
    for (long long r = rg; r < rows; r += 32) {
      const uint4 u = *reinterpret_cast<const uint4*>(x + r * ld + col0);
      acc[0] += bf16_lo(u.x); acc[1] += bf16_hi(u.x); acc[2] += bf16_lo(u.y); acc[3] += bf16_hi(u.y);
      acc[4] += bf16_lo(u.z); acc[5] += bf16_hi(u.z); acc[6] += bf16_lo(u.w); acc[7] += bf16_hi(u.w);
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) part[rg][cg * 8 + e] = acc[e];
  __syncthreads();
  if (threadIdx.x < 64) {
    const int col = blockIdx.x * 64 + threadIdx.x;
    if (col < V) {
      float s = 0.f;
      for (int g = 0; g < 32; ++g) s += part[g][threadIdx.x];
      db[col] = s + (accumulate ? db[col] : 0.f);
    }
  }
}
}  // namespace

// The whole LM-head backward for a chunk of rows behind one call (the entry SURVEY.md §8(b) proposed): d(logits) into
// `dlogits` (transformed in place from kept logits when kept != 0, recomputed on the tensor cores otherwise), then
// dH = dlogits @ W, dW (+)= dlogits^T @ H and, for a head with a bias, db (+)= column sums — all enqueued on `stream`.
extern "C" int b2_lmhead_ce_bwd(const void* H, long long ldh, const void* W, long long ldw, const float* bias,
                                const long long* labels, long long row0, long long rows, int D, int V, int shift_T,
                                long long ignore_index, int round_logits_bf16, const float* lse, const float* gscale_dev,
                                void* dlogits, long long ldd, int kept, void* dH, long long lddh, void* dW, int dw_dtype,
                                long long lddw, int dw_accumulate, float* db, int db_accumulate, void* gemm_workspace,
                                size_t gemm_workspace_bytes, cudaStream_t stream) {
  if (rows == 0) return B2_OK;
  B2_REQUIRE(dlogits != nullptr, "b2_lmhead_ce_bwd: null dlogits buffer");
  int rc;
  if (kept) rc = b2_ce_dlogits_inplace(dlogits, ldd, labels, row0, rows, V, shift_T, ignore_index, lse, gscale_dev, stream);
  else rc = b2_lmhead_ce_bwd_dlogits(H, ldh, W, ldw, bias, labels, row0, rows, D, V, shift_T, ignore_index,
                                     round_logits_bf16, lse, gscale_dev, dlogits, ldd, stream);
  if (rc != B2_OK) return rc;
  B2_REQUIRE(rows < (1ll << 31), "b2_lmhead_ce_bwd: too many rows");
  if (dH != nullptr) {        // dH[rows, D] = dL[rows, V] @ W[V, D]          (B = W read MN-major)
    const bool ws_ok = gemm_workspace != nullptr && gemm_workspace_bytes >= b2_gemm_workspace_bytes((int)rows, D);
    rc = b2_gemm_bf16(dlogits, 0, ldd, W, 1, ldw, dH, B2_BF16, lddh, (int)rows, D, V, 0, ws_ok ? gemm_workspace : nullptr,
                      ws_ok ? gemm_workspace_bytes : 0, stream);
    if (rc != B2_OK) return rc;
  }
  if (dW != nullptr) {        // dW[V, D] (+)= dL^T[V, rows] @ H[rows, D]      (A = dL and B = H both read MN-major)
    rc = b2_gemm_bf16(dlogits, 1, ldd, H, 1, ldh, dW, dw_dtype, lddw, V, D, (int)rows, dw_accumulate, nullptr, 0, stream);
    if (rc != B2_OK) return rc;
  }
  if (db != nullptr) {
    colsum_kernel<<<(unsigned)((V + 63) / 64), 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(dlogits), ldd, rows, V, db,
                                                               db_accumulate);
    B2_CHECK_CUDA(cudaGetLastError());
  }
  return B2_OK;
}
