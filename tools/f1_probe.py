"""Small driver for profiling: runs the fused LM-head+CE forward/backward (and optionally AdamW / sumsq) at the
BASELINE configs[1] head shape a few times and prints per-call CUDA-event timings. Used under ncu (see profiles/)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def timeit(fn, iters, flush=None):
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return ts


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--N", type=int, default=8192)
    ap.add_argument("--D", type=int, default=768)
    ap.add_argument("--V", type=int, default=50257)
    ap.add_argument("--T", type=int, default=1024)
    ap.add_argument("--chunk", type=int, default=8192)
    ap.add_argument("--adamw", action="store_true")
    ap.add_argument("--plain", action="store_true")
    args = ap.parse_args()
    from b200_train import ops
    dev = torch.device("cuda", 0)
    N, D, V, T = args.N, args.D, args.V, args.T
    g = torch.Generator(device="cpu").manual_seed(7)
    H = (torch.randn(N, D, generator=g) * 0.5).to(dev, torch.bfloat16)
    W = (torch.randn(V, D, generator=g) * 0.05).to(dev, torch.bfloat16)
    y = torch.randint(0, V, (N,), generator=g).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ldd = (V + 63) // 64 * 64          # row pitch: a multiple of 128 bytes
    chunk = min(args.chunk, N)
    dl = torch.empty(chunk, ldd, dtype=torch.bfloat16, device=dev)
    dH = torch.empty(N, D, dtype=torch.bfloat16, device=dev)
    dW = torch.empty(V, D, dtype=torch.float32, device=dev)
    gs = torch.tensor([1.0 / N], device=dev)
    res = {}
    r = ops.lmhead_ce_fwd(H, W, y, shift_T=T)
    res["fwd"] = timeit(lambda: ops.lmhead_ce_fwd(H, W, y, shift_T=T), args.iters, flush)
    res["bwd_dlogits(chunk)"] = timeit(lambda: ops.lmhead_ce_bwd_dlogits(H, W, y, r.lse, gs, 0, chunk, dl, shift_T=T), args.iters, flush)
    if chunk == N:
        res["fwd_keep_logits"] = timeit(lambda: ops.lmhead_ce_fwd(H, W, y, shift_T=T, keep_logits=dl), args.iters, flush)
        res["dlogits_inplace"] = timeit(lambda: ops.ce_dlogits_inplace_(dl, y, r.lse, gs, 0, N, V, shift_T=T), args.iters, flush)
        ops.lmhead_ce_bwd_dlogits(H, W, y, r.lse, gs, 0, chunk, dl, shift_T=T)          # real gradients for the GEMMs below
    res["gemm_dH(chunk)"] = timeit(lambda: ops.gemm_bf16(dl, W, dH[:chunk], chunk, D, V, b_mn_major=True), args.iters, flush)
    res["gemm_dW(chunk)"] = timeit(lambda: ops.gemm_bf16(dl, H[:chunk], dW, V, D, chunk, a_mn_major=True, b_mn_major=True), args.iters, flush)
    flops = 2.0 * chunk * V * D
    for k, ts in res.items():
        n = N if k.startswith("fwd") else chunk
        ms = sorted(ts)[len(ts) // 2]
        if k == "dlogits_inplace":
            print(f"{k:22s} median {ms:8.3f} ms   {4.0 * n * V / ms / 1e6:8.1f} GB/s      all={['%.3f' % t for t in ts]}")
            continue
        print(f"{k:22s} median {ms:8.3f} ms   {2.0 * n * V * D / ms / 1e9:8.1f} TFLOP/s   all={['%.3f' % t for t in ts]}")
    if args.plain:
        # fwd-shaped plain GEMM: logits[N, V] = H @ W^T (bf16 out) -- mainloop + the simplest epilogue
        C = torch.empty(N, ldd, dtype=torch.bfloat16, device=dev)
        ts = timeit(lambda: ops.gemm_bf16(H, W, C, N, V, D), args.iters, flush)
        ms = sorted(ts)[len(ts) // 2]
        print(f"plain gemm [N,V]=H@W^T  median {ms:8.3f} ms   {2.0 * N * V * D / ms / 1e9:8.1f} TFLOP/s   (B2_DEBUG_NOEPI={os.environ.get('B2_DEBUG_NOEPI')})")
        ts = timeit(lambda: torch.matmul(H, W.t(), out=C[:, :V]) if False else torch.matmul(H, W.t()), args.iters, flush)
        ms = sorted(ts)[len(ts) // 2]
        print(f"cuBLAS  H@W^T           median {ms:8.3f} ms   {2.0 * N * V * D / ms / 1e9:8.1f} TFLOP/s")
    if args.adamw:
        n = 1 << 27
        p = torch.randn(n, device=dev).to(torch.bfloat16); gr = (torch.randn(n, device=dev) * 0.01).to(torch.bfloat16)
        m = torch.zeros_like(p); v = torch.zeros_like(p)
        ts = timeit(lambda: ops.adamw_(p, gr, m, v, lr=2e-4, weight_decay=1e-3, step=1), args.iters)
        ms = sorted(ts)[len(ts) // 2]
        print(f"adamw bf16 n=2^27      median {ms:8.3f} ms   {14.0 * n / ms / 1e6:8.1f} GB/s")
        ts = timeit(lambda: ops.sumsq(gr), args.iters)
        ms = sorted(ts)[len(ts) // 2]
        print(f"sumsq bf16 n=2^27      median {ms:8.3f} ms   {2.0 * n / ms / 1e6:8.1f} GB/s")


if __name__ == "__main__":
    main()
