"""F2 on ONE GPU, for ncu: the W ranks' buffers all live on this GPU and each rank's kernel runs in turn with the barrier
flags pre-satisfied (the arrangement of tests/test_e2e_gpu.py::test_peer_allreduce_single_gpu_emulation). ncu profiles one
process on one GPU, so this is the only way to get `--set full` counters (DRAM / L2 bytes, issue stalls, occupancy) for
peer_allreduce_kernel; what it cannot show is NVLink — every "peer" access here is a local HBM access. The multi-GPU
numbers (busbw vs NCCL) come from bench.py at N > 1 and tools/f2_bench.py, timed with CUDA events.

    python tools/f2_emul_probe.py [--world 8] [--log2n 27]
"""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--log2n", type=int, default=27)
    ap.add_argument("--iters", type=int, default=3)
    a = ap.parse_args()
    from b200_train import _lib
    dev = torch.device("cuda", 0)
    n, W = 1 << a.log2n, a.world
    bufs = [(torch.randn(n, device=dev) * 0.01).to(torch.bfloat16) for _ in range(W)]
    flags = [torch.zeros(_lib.lib.b2_peer_flags_bytes() // 4, dtype=torch.int32, device=dev) for _ in range(W)]
    out = torch.zeros(W, dtype=torch.float32, device=dev)
    bt = (ctypes.c_void_p * W)(*[b.data_ptr() for b in bufs])
    ft = (ctypes.c_void_p * W)(*[f.data_ptr() for f in flags])
    stream = torch.cuda.current_stream().cuda_stream
    for it in range(a.iters):
        epoch = it + 1
        for f in flags:
            f[:16] = epoch
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.call("b2_peer_allreduce", bt, ft, W, 0, 0, n, _lib.BF16, 1.0 / W, epoch, 1, 0, 0, out.data_ptr(), stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        slice_bytes = n * 2 / W
        # rank 0's kernel: reads its slice from W buffers, writes it to W buffers
        print(f"rank-0 kernel, W={W}, n=2^{a.log2n} bf16: {ms:.3f} ms, {2 * W * slice_bytes / ms / 1e6:.0f} GB/s of local HBM traffic")


if __name__ == "__main__":
    main()
