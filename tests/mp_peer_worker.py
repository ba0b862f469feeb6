"""Worker for tests/test_multigpu.py — one process per GPU (launched by torch.distributed.run).

Checks, on real peers over NVLink:
  1. b2_peer_allreduce on a symmetric flat buffer == mean of the ranks' inputs (bit-identical on all ranks; equal to
     NCCL all_reduce(SUM)/W for W = 2), with the fused sum of squares;
  2. reduce-scatter + all-gather halves;
  3. prepare_model + AdamW + clip on a tiny GPT-2: every rank ends with identical parameters, equal (to bf16 tolerance)
     to a single-process run over the concatenated batch with the stock path.
Prints 'MP_OK' from rank 0 when everything passed.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    import b200_train as b200
    from b200_train.symm import PeerComm
    from oracle import step_tail as O
    from oracle import torch_ref as TR

    comm = PeerComm(dev)
    for dtype in (torch.bfloat16, torch.float32):
        n = 64 * 4099
        t = comm.alloc(n, dtype, dev)
        g = torch.Generator(device="cpu").manual_seed(100 + rank)
        src = (torch.randn(n, generator=g) * 0.1).to(dev, dtype)
        for it in range(3):                                  # repeated calls exercise the epoch flags
            t.copy_(src)
            ref = src.float().clone()
            dist.all_reduce(ref)                             # NCCL, checker only
            want = (ref * np.float32(1.0 / world)).to(dtype)
            parts = comm.allreduce_(t, want_sumsq=True)
            torch.cuda.synchronize()
            tol = 0 if (world == 2 or dtype == torch.float32) else float(want.float().abs().max()) * 2 ** -7
            if dtype == torch.float32 and world > 2:
                tol = 1e-6                                    # different fp32 summation order than NCCL's ring
            if comm.use_nvls and getattr(t._b2_symm[0], "mc_ptr", None):
                # the NVSwitch forms the sum; its rounding of the bf16 result is within one ulp of ours
                tol = max(tol, float(want.float().abs().max()) * 2 ** -7 if dtype == torch.bfloat16 else 1e-6)
            assert (t.float() - want.float()).abs().max() <= tol, (dtype, it, (t.float() - want.float()).abs().max())
            bits = t.view(torch.uint8).clone()
            b0 = bits.clone()
            dist.broadcast(b0, 0)
            assert torch.equal(bits, b0), "ranks hold different bits after the peer all-reduce"
            np.testing.assert_allclose(float(parts.sum().item()), t.float().double().pow(2).sum().item(), rtol=1e-5)
        # reduce-scatter, then all-gather
        t.copy_(src)
        comm.allreduce_(t, scatter_only=True)
        comm.allgather_(t)
        torch.cuda.synchronize()
        dist.barrier()
        assert (t.float() - want.float()).abs().max() <= tol

    # --- prepare_model path on a tiny GPT-2: replicated optimizer (DDP semantics) and sharded optimizer (ZeRO-1 style)
    kw = dict(dtype=torch.bfloat16, seed=0, vocab_size=2003, n_embd=128, n_layer=2, n_head=4, n_positions=64,
              resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
    batches = [TR.synthetic_hf_batch(4, 64, 2003, seed=50 + r, device=dev) for r in range(world)]

    def run_ddp(shard):
        torch.manual_seed(1000 + rank)                       # ranks start different; prepare_model syncs rank 0's weights
        model = TR.make_gpt2(**{**kw, "seed": rank}).train()
        b200.patch_causal_lm(model)
        model = b200.prepare_model(model, parallel_strategy_kwargs=shard if isinstance(shard, dict) else {"shard_optimizer": shard})
        assert hasattr(model, "module")
        opt = b200.AdamW(TR.hf_param_groups(model.module, 1e-3), lr=2e-4)
        for step in range(2):
            opt.zero_grad(set_to_none=True)
            ids, lab = batches[rank]
            loss = model(input_ids=ids, labels=lab).loss
            loss.backward()
            b200.clip_grad_norm_(model.parameters(), 0.3)
            opt.step()
        model.wait_parameters()          # (overlap_param_sync: the last all-gather may still be in flight)
        flat = torch.cat([p.detach().float().reshape(-1) for p in model.parameters()])
        f0 = flat.clone()
        dist.broadcast(f0, 0)
        assert torch.equal(flat, f0), "parameters diverged across ranks"
        sd = opt.state_dict()                                # collective when sharded
        return model, opt, flat, sd

    model, opt, flat, sd = run_ddp(False)
    model_s, opt_s, flat_s, sd_s = run_ddp(True)
    assert torch.equal(flat, flat_s), "sharded optimizer must give bit-identical parameters"
    # several buckets per buffer (per-bucket reduce-scatter / AdamW / all-gather), with the parameter all-gather running
    # under the next forward pass
    os.environ["B2_DDP_BUCKET_MB"] = "0.25"
    model_o, opt_o, flat_o, sd_o = run_ddp({"shard_optimizer": True, "overlap_param_sync": True})
    model_r, opt_r, flat_r, sd_r = run_ddp({"bucket_cap_mb": 0.25})
    del os.environ["B2_DDP_BUCKET_MB"]
    assert len(model_o._buckets) > 3 and model_o.overlap_param_sync
    # same buckets, sharded vs replicated: bit-identical (same reduced gradients, same per-rank Σg² partials, same AdamW)
    assert torch.equal(flat_r, flat_o), "bucketed sharded optimizer must give the bucketed replicated path's bits"
    for k in sd_r["state"]:
        assert torch.equal(sd_r["state"][k]["exp_avg"], sd_o["state"][k]["exp_avg"]), k
        assert torch.equal(sd_r["state"][k]["exp_avg_sq"], sd_o["state"][k]["exp_avg_sq"]), k
    # other bucket boundaries: the reduced gradients are the same bits, but Σg² is added up in another order, so the
    # clip coefficient may differ in its last bit and a parameter may land on the neighbouring bf16 value
    d = (flat - flat_r).abs()
    assert float(d.max()) <= 2.0 ** -7 * float(flat.abs().max()) and float((d > 0).float().mean()) < 1e-2, \
        ("bucket boundaries changed the result", float(d.max()), float((d > 0).float().mean()))
    model_o.close(); model_r.close()
    assert sd["state"].keys() == sd_s["state"].keys()
    for k in sd["state"]:
        assert float(sd["state"][k]["step"]) == float(sd_s["state"][k]["step"]) == 2.0
        assert torch.equal(sd["state"][k]["exp_avg"], sd_s["state"][k]["exp_avg"]), k
        assert torch.equal(sd["state"][k]["exp_avg_sq"], sd_s["state"][k]["exp_avg_sq"]), k
    # optimizer memory is really sharded
    full = sum(b.numel for b in opt.flat_buffers()); part = sum(b.exp_avg.numel() for b in opt_s.flat_buffers())
    # (slices are 64-element aligned per bucket: the first W-1 ranks round their share up, the last one takes the rest,
    # so a rank's share is within 64 * W elements per bucket of full / W)
    tot = torch.tensor([part], device=dev, dtype=torch.int64); dist.all_reduce(tot)
    assert int(tot) == full, (int(tot), full)
    assert abs(part * world - full) <= 64 * world * world * len(model_s._buckets), (part, full)
    model_s.close()
    # --- gradient accumulation: two micro-batches, the first under no_sync() (ref: fine_tune_config.json:14)
    torch.manual_seed(7)
    acc_model = b200.prepare_model(b200.patch_causal_lm(TR.make_gpt2(**kw).train()))
    acc_opt = b200.AdamW(TR.hf_param_groups(acc_model.module, 1e-3), lr=2e-4)
    mb = [TR.synthetic_hf_batch(2, 64, 2003, seed=900 + 10 * rank + k, device=dev) for k in range(2)]
    n_items = sum(int((lab[:, 1:] != -100).sum()) for _, lab in mb)
    acc_opt.zero_grad(set_to_none=True)
    with acc_model.no_sync():
        acc_model(input_ids=mb[0][0], labels=mb[0][1], num_items_in_batch=n_items).loss.backward()
    g_before = acc_model.flat_buffers[0].grad.clone()
    acc_model(input_ids=mb[1][0], labels=mb[1][1], num_items_in_batch=n_items).loss.backward()
    # after the synchronised backward every rank holds the same averaged gradient; it differs from the local partial one
    g = acc_model.flat_buffers[0].grad.clone()
    g0 = g.clone(); dist.broadcast(g0, 0)
    assert torch.equal(g, g0) and not torch.equal(g, g_before)
    # and equals the mean over ranks of the locally accumulated gradients, computed the slow way
    chk = b200.patch_causal_lm(TR.make_gpt2(**kw).to(dev).train())
    for ids, lab in mb:
        chk(input_ids=ids, labels=lab, num_items_in_batch=n_items).loss.backward()
    local = torch.cat([p.grad.reshape(-1).float() for p in chk.parameters()])
    dist.all_reduce(local); local /= world
    got = torch.cat([p.grad.reshape(-1).float() for p in acc_model.module.parameters()])
    assert (got - local).abs().max() <= 2.0 ** -7 * local.abs().max() + 1e-6
    acc_model.close()

    # --- gradients dropped BETWEEN forward and backward (out = model(x); optimizer.zero_grad(); loss.backward()) with a
    # foreign optimizer: autograd allocates fresh .grad tensors; they must reach the flat buffer before their bucket is
    # reduced, or ranks would average stale data and diverge silently
    torch.manual_seed(11)
    zg_model = b200.prepare_model(b200.patch_causal_lm(TR.make_gpt2(**kw).train()))
    zg_opt = torch.optim.AdamW(zg_model.parameters(), lr=1e-3)
    chk2 = b200.patch_causal_lm(TR.make_gpt2(**kw).to(dev).train())
    for step in range(2):
        ids, lab = batches[rank]
        out = zg_model(input_ids=ids, labels=lab)
        zg_opt.zero_grad(set_to_none=True)
        zg_model.flat_buffers[0].grad.fill_(5.0)               # whatever the buffer held before must not be averaged in
        out.loss.backward()
        if step == 0:
            chk2(input_ids=ids, labels=lab).loss.backward()
            local = torch.cat([p.grad.reshape(-1).float() for p in chk2.parameters()])
            dist.all_reduce(local); local /= world
            got = torch.cat([p.grad.reshape(-1).float() for p in zg_model.module.parameters()])
            assert (got - local).abs().max() <= 2.0 ** -7 * local.abs().max() + 1e-6, "stale gradients were reduced"
        zg_opt.step()
    flat_zg = torch.cat([p.detach().float().reshape(-1) for p in zg_model.parameters()])
    f0 = flat_zg.clone(); dist.broadcast(f0, 0)
    assert torch.equal(flat_zg, f0), "ranks diverged after forward -> zero_grad -> backward"
    zg_model.close()

    # stock single-process equivalent: DDP averages per-rank mean losses -> emulate by averaging gradients
    ref = TR.make_gpt2(**kw).to(dev).train()
    ropt = torch.optim.AdamW(TR.hf_param_groups(ref, 1e-3), lr=2e-4)
    before = [p.detach().clone() for p in ref.parameters()]
    for step in range(2):
        ropt.zero_grad(set_to_none=True)
        for ids, lab in batches:
            (ref(input_ids=ids, labels=lab).loss / world).backward()
        torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.3)
        ropt.step()
    for (n, a), b, b4 in zip(model.module.named_parameters(), ref.parameters(), before):
        got, want = a.detach().float().cpu().numpy(), b.detach().float().cpu().numpy()
        ulp = O.bf16_ulp(np.maximum(np.abs(want), np.abs(b4.float().cpu().numpy())))
        err = np.abs(got - want)
        assert (err <= 2 * ulp + 2.2 * 2e-4 * 2).all(), (n, err.max())      # 2 ulp + one Adam sign-flip per step
    dist.barrier()
    model.close()
    comm.close()
    if rank == 0:
        print("MP_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
