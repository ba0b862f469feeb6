"""Worker for tests/test_multigpu.py::test_comm_hook_on_stock_ddp — one process per GPU.

Stock torch DistributedDataParallel twice over the same tiny GPT-2: once with its default NCCL all-reduce, once with
b200_train.peer_allreduce_hook registered through DDP's own register_comm_hook (the plug-in point the
reference's wrapper already has). After each of three backward passes the gradients must agree: bit for bit at
world 2 (two addends: one rounding either way), to a few bf16 ulps of the largest gradient beyond that (NCCL's ring
rounds at every hop, our kernel sums in fp32 and rounds once). Prints 'HOOK_OK' from rank 0.
"""
import os
import sys

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as TorchDDP

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    import b200_train as b200
    import workloads as WL

    for dtype in (torch.bfloat16, torch.float32):
        kw = dict(dtype=dtype, seed=0, vocab_size=2003, n_embd=128, n_layer=2, n_head=4, n_positions=64,
                  resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
        # small bucket cap: several buckets per backward, so the staging buffer is reused within one pass
        stock = TorchDDP(WL.make_gpt2(**kw).to(dev).train(), device_ids=[local], bucket_cap_mb=0.25)
        ours = TorchDDP(WL.make_gpt2(**kw).to(dev).train(), device_ids=[local], bucket_cap_mb=0.25)
        state = b200.PeerAllreduceState()
        ours.register_comm_hook(state, b200.peer_allreduce_hook)
        try:
            ours.register_comm_hook(state, b200.peer_allreduce_hook)       # torch's own rule
        except RuntimeError:
            pass
        else:
            raise AssertionError("registering a second communication hook must fail")
        for it in range(3):
            ids, lab = WL.synthetic_hf_batch(4, 64, 2003, seed=70 + 10 * it + rank, device=dev)
            for m in (stock, ours):
                m.zero_grad(set_to_none=True)
                m(input_ids=ids, labels=lab).loss.backward()
            torch.cuda.synchronize()
            worst = 0.0
            for (name, a), b in zip(stock.module.named_parameters(), ours.module.parameters()):
                assert b.grad is not None and b.grad.dtype == a.grad.dtype, name
                if world == 2:
                    # same bits, except that torch halves BEFORE adding: a subnormal addend may lose its last bit there
                    err = float((a.grad.float() - b.grad.float()).abs().max())
                    assert err <= 1e-30, (str(dtype), it, name, err)
                else:
                    gmax = float(a.grad.float().abs().max())
                    # a bf16 ring rounds the running sum at every hop: up to world - 1 roundings of half an ulp (2^-9 relative) each
                    tol = gmax * (world * 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -20) + 1e-12
                    err = float((a.grad.float() - b.grad.float()).abs().max())
                    worst = max(worst, err / max(gmax, 1e-30))
                    assert err <= tol, (str(dtype), it, name, err, tol)
            # every rank holds the same bits
            flat = torch.cat([p.grad.reshape(-1) for p in ours.module.parameters()]).view(torch.uint8)
            f0 = flat.clone()
            dist.broadcast(f0, 0)
            assert torch.equal(flat, f0), "ranks hold different gradients after the hook"
        assert state.buckets_reduced >= 3 * 2, state.buckets_reduced       # several buckets per pass went through us
        dist.barrier()
        state.close()
        del stock, ours
    if rank == 0:
        print("HOOK_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
