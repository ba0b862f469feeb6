"""BASELINE configs[4] microbench (run under torch.distributed.run, one rank per GPU):
flat bf16 gradient buffer, peer-memory all-reduce (F2) vs NCCL all_reduce, and the fused tail
F2 + clip + AdamW vs NCCL all_reduce + clip_grad_norm_ + torch.optim.AdamW(fused / foreach).
Times are CUDA-event, max over ranks. busbw = 2 (W-1)/W * n * s / t (NCCL convention)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def timed(fn, iters, dev):
    ts = []
    for _ in range(iters):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t.item()))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2n", type=int, nargs="+", default=[20, 24, 27, 30])
    ap.add_argument("--iters", type=int, default=7)
    ap.add_argument("--tail", action="store_true")
    ap.add_argument("--sweep-blocks", type=int, nargs="*", default=[], help="also time the all-reduce on capped grids")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from b200_train import ops
    from b200_train.symm import PeerComm
    comm = PeerComm(dev)
    out = []
    for l2 in args.log2n:
        n = 1 << l2
        t = comm.alloc(n, torch.bfloat16, dev)
        t.copy_((torch.randn(n, device=dev) * 0.01).to(torch.bfloat16))
        ref = t.clone()
        for _ in range(2):
            comm.allreduce_(t, want_sumsq=True)
            dist.all_reduce(ref)
        ms_peer = timed(lambda: comm.allreduce_(t, want_sumsq=True), args.iters, dev)
        ms_nccl = timed(lambda: dist.all_reduce(ref), args.iters, dev)
        bus = 2.0 * (world - 1) / world * n * 2
        row = {"n": n, "bytes": 2 * n, "world": world, "peer_ms": ms_peer, "nccl_ms": ms_nccl,
               "peer_busbw_GBs": bus / ms_peer / 1e6, "nccl_busbw_GBs": bus / ms_nccl / 1e6, "nvls": bool(comm.use_nvls)}
        for nb in args.sweep_blocks:
            ms = timed(lambda: comm.allreduce_(t, want_sumsq=True, max_blocks=nb), args.iters, dev)
            row[f"peer_busbw_GBs_blocks{nb}"] = bus / ms / 1e6
        if args.tail and l2 >= 24:
            # fused tail: F2 (with sum of squares) -> clip coefficient -> fused AdamW
            p = (torch.randn(n, device=dev)).to(torch.bfloat16)
            m = torch.zeros_like(p); v = torch.zeros_like(p)

            def ours():
                parts = comm.allreduce_(t, want_sumsq=True)
                _, coef = ops.clip_coef(parts, 0.3)
                ops.adamw_(p, t, m, v, lr=2e-4, weight_decay=1e-3, step=1, gscale=coef)

            par = torch.nn.Parameter(p.clone())
            par.grad = ref
            opt_fused = torch.optim.AdamW([par], lr=2e-4, weight_decay=1e-3, fused=True)
            opt_foreach = torch.optim.AdamW([par], lr=2e-4, weight_decay=1e-3, foreach=True)

            def stock(opt):
                par.grad.div_(world)
                dist.all_reduce(par.grad)
                torch.nn.utils.clip_grad_norm_([par], 0.3)
                opt.step()

            ours(); stock(opt_fused); stock(opt_foreach)
            row["tail_ours_ms"] = timed(ours, args.iters, dev)
            row["tail_nccl_fused_adamw_ms"] = timed(lambda: stock(opt_fused), args.iters, dev)
            row["tail_nccl_foreach_adamw_ms"] = timed(lambda: stock(opt_foreach), args.iters, dev)
            del par, opt_fused, opt_foreach, p, m, v
        out.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        del t, ref
        torch.cuda.empty_cache()
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
